"""Host-side mirror of the reference's entry points over the C ABI
(include/pmcgpu.h).  Method names follow the Fortran they stand for:

    aero_state_sort      partmc/src/aero_state.F90:3218
    mc_coag              partmc/src/coagulation.F90:43
    aero_state_rebalance partmc/src/aero_state.F90:1765
    condense_particles   partmc/src/condense.F90:147
    trans_aero           interface/wrf_pmc_trans_aero.F90:40
    partmc_to_wrf        interface/wrf_pmc_driver.F90:1663

Only ctypes and numpy are used; no oracle code is imported here and nothing
runs on the CPU: a missing library or device raises.
"""
import ctypes as C
import os

import numpy as np

KERNEL_SEDI, KERNEL_ADDITIVE, KERNEL_CONSTANT, KERNEL_BROWN, KERNEL_ZERO = 1, 2, 3, 4, 5
KERNEL_BROWN_FREE, KERNEL_BROWN_CONT = 6, 7  # coag_kernel.F90:42-45
WEIGHT_NONE, WEIGHT_POWER = 1, 2

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


class PmcGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pmcgpu error %d: %s" % (code, msg))
        self.code = code


def comm_unique_id():
    """128-byte NCCL unique id (made on one rank, broadcast by the host: MPI_Bcast in WRF)"""
    buf = C.create_string_buffer(128)
    rc = load_library().pmcgpu_comm_unique_id(buf, C.c_int(128))
    if rc != 0:
        raise PmcGpuError(rc, "pmcgpu_comm_unique_id failed")
    return buf.raw


def library_path():
    return os.path.join(_PKG, "libpmcgpu.so")


def load_library():
    """Load libpmcgpu.so.  Fails loudly when it has not been built."""
    global _LIB
    if _LIB is None:
        path = library_path()
        if not os.path.exists(path):
            raise RuntimeError("%s is missing: build it with __graft_entry__.build() "
                               "(there is no CPU fallback for the aerosol step)" % path)
        _LIB = C.CDLL(path)
        _LIB.pmcgpu_last_error.restype = C.c_char_p
        _LIB.pmcgpu_last_error.argtypes = [C.c_void_p]
        _LIB.pmcgpu_launch_count.restype = C.c_longlong
        _LIB.pmcgpu_launch_count.argtypes = [C.c_void_p]
        _LIB.pmcgpu_stream.restype = C.c_void_p
        _LIB.pmcgpu_stream.argtypes = [C.c_void_p]
    return _LIB


class _Config(C.Structure):
    _fields_ = [("n_spec", C.c_int), ("density", C.POINTER(C.c_double)), ("kappa", C.POINTER(C.c_double)),
                ("molec_weight", C.POINTER(C.c_double)), ("num_ions", C.POINTER(C.c_int)), ("i_water", C.c_int),
                ("n_group", C.c_int), ("n_class", C.c_int), ("nx", C.c_int), ("nz", C.c_int), ("ny", C.c_int),
                ("global_nx", C.c_int), ("global_ny", C.c_int), ("i_offset", C.c_int), ("j_offset", C.c_int),
                ("periodic_x", C.c_int), ("periodic_y", C.c_int), ("n_part_ideal", C.c_double),
                ("cell_capacity", C.c_int), ("max_bins", C.c_int), ("seed", C.c_uint), ("device", C.c_int)]


class _TimestepArgs(C.Structure):
    _fields_ = [("temp", C.c_void_p), ("pressure", C.c_void_p), ("rel_humid", C.c_void_p),
                ("inverse_density", C.c_void_p), ("cell_volume", C.c_void_p), ("elapsed_time", C.c_double),
                ("dt", C.c_double), ("chem_dt", C.c_double), ("do_coagulation", C.c_int),
                ("coag_kernel_type", C.c_int), ("do_condensation", C.c_int), ("allow_doubling", C.c_int),
                ("allow_halving", C.c_int), ("step", C.c_uint), ("num_conc_out", C.c_void_p),
                ("rel_humid_out", C.c_void_p), ("no_wait", C.c_int)]


class _WrfFields(C.Structure):
    _fields_ = [("p", C.c_void_p), ("pb", C.c_void_p), ("t_2", C.c_void_p), ("qv", C.c_void_p), ("alt", C.c_void_p),
                ("z", C.c_void_p), ("z_at_w", C.c_void_p), ("msftx", C.c_void_p), ("msfty", C.c_void_p),
                ("rdx", C.c_float), ("rdy", C.c_float)]


class _BoundaryArgs(C.Structure):
    _fields_ = [("n_mode", C.c_int), ("mode_type", C.c_void_p), ("weight_class", C.c_void_p),
                ("char_radius", C.c_void_p), ("log10_std_dev", C.c_void_p), ("vol_frac", C.c_void_p),
                ("vol_frac_std", C.c_void_p), ("num_conc", C.c_void_p), ("outflow", C.c_void_p),
                ("throw_away", C.c_void_p)]


class _DepArgs(C.Structure):
    _fields_ = [("rmol", C.c_void_p), ("znt", C.c_void_p), ("ust", C.c_void_p), ("ivgtyp", C.c_void_p),
                ("height", C.c_void_p), ("dt", C.c_double), ("step", C.c_uint), ("no_wait", C.c_int)]


class _EmitArgs(C.Structure):
    _fields_ = [("n_mode", C.c_int), ("mode_type", C.c_void_p), ("weight_class", C.c_void_p),
                ("char_radius", C.c_void_p), ("log10_std_dev", C.c_void_p), ("vol_frac", C.c_void_p),
                ("vol_frac_std", C.c_void_p), ("num_conc", C.c_void_p), ("sample_prop", C.c_void_p),
                ("characteristic_factor", C.c_double), ("create_time", C.c_double), ("allow_doubling", C.c_int),
                ("allow_halving", C.c_int), ("step", C.c_uint), ("stream", C.c_uint), ("n_sample_bins", C.c_int),
                ("sample_radius", C.c_void_p), ("sample_num_conc", C.c_void_p)]


class _SeasaltArgs(C.Structure):
    _fields_ = [("u10", C.c_void_p), ("v10", C.c_void_p), ("xland", C.c_void_p), ("height", C.c_void_p),
                ("sea_salt_opt", C.c_int), ("i_na", C.c_int), ("i_cl", C.c_int), ("weight_class", C.c_void_p),
                ("dt", C.c_double), ("elapsed_time", C.c_double), ("allow_doubling", C.c_int),
                ("allow_halving", C.c_int), ("step", C.c_uint)]


class _TransArgs(C.Structure):
    _fields_ = [("u_flux", C.c_void_p), ("v_flux", C.c_void_p), ("w_flux", C.c_void_p), ("xkhh", C.c_void_p),
                ("alt", C.c_void_p), ("z", C.c_void_p), ("z_at_w", C.c_void_p), ("exch_h", C.c_void_p),
                ("dx", C.c_float), ("dy", C.c_float), ("dt", C.c_float), ("vertmix_onoff", C.c_int),
                ("allow_doubling", C.c_int), ("allow_halving", C.c_int), ("step", C.c_uint), ("no_wait", C.c_int)]


def _f64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


def _addr(a):
    """address of a numpy array, a torch tensor (host or device) or a raw integer pointer; None -> NULL"""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    return a.ctypes.data


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def _i64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int64)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class AeroData:
    """aero_data_t (aero_data.F90:55-93): the species table the path reads."""

    def __init__(self, density, kappa=None, molec_weight=None, num_ions=None, i_water=0):
        self.density = _f64(density)
        n = len(self.density)
        self.n_spec = n
        self.kappa = _f64(kappa if kappa is not None else np.zeros(n))
        self.molec_weight = _f64(molec_weight if molec_weight is not None else np.full(n, 18e-3))
        self.num_ions = _i32(num_ions if num_ions is not None else np.zeros(n))
        self.i_water = int(i_water)


class PartMCGPU:
    """All aero_state_t / env_state_t of one WRF patch, resident on one B200."""

    def __init__(self, aero_data, nx=1, nz=1, ny=1, n_group=1, n_class=1, n_part_ideal=1000.0, cell_capacity=None,
                 max_bins=64, seed=1, device=0, global_nx=None, global_ny=None, i_offset=0, j_offset=0,
                 periodic_x=True, periodic_y=True):
        self.L = load_library()
        self.data = aero_data
        self.nx, self.nz, self.ny = nx, nz, ny
        self.n_cells = nx * nz * ny
        self.n_group, self.n_class = n_group, n_class
        self.max_bins = max_bins
        if cell_capacity is None:
            cell_capacity = int(3 * n_part_ideal) + 64
        self.cell_capacity = cell_capacity
        cfg = _Config()
        cfg.n_spec = aero_data.n_spec
        cfg.density = aero_data.density.ctypes.data_as(C.POINTER(C.c_double))
        cfg.kappa = aero_data.kappa.ctypes.data_as(C.POINTER(C.c_double))
        cfg.molec_weight = aero_data.molec_weight.ctypes.data_as(C.POINTER(C.c_double))
        cfg.num_ions = aero_data.num_ions.ctypes.data_as(C.POINTER(C.c_int))
        cfg.i_water = aero_data.i_water
        cfg.n_group, cfg.n_class = n_group, n_class
        cfg.nx, cfg.nz, cfg.ny = nx, nz, ny
        cfg.global_nx = global_nx if global_nx is not None else nx
        cfg.global_ny = global_ny if global_ny is not None else ny
        cfg.i_offset, cfg.j_offset = i_offset, j_offset
        cfg.periodic_x, cfg.periodic_y = int(periodic_x), int(periodic_y)
        cfg.n_part_ideal = float(n_part_ideal)
        cfg.cell_capacity = cell_capacity
        cfg.max_bins = max_bins
        cfg.seed = seed
        cfg.device = device
        self.h = C.c_void_p()
        rc = self.L.pmcgpu_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            msg = self.L.pmcgpu_last_error(self.h).decode() if self.h else "pmcgpu_create failed (no CUDA device?)"
            raise PmcGpuError(rc, msg)

    def close(self):
        if getattr(self, "h", None):
            self.L.pmcgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise PmcGpuError(rc, self.L.pmcgpu_last_error(self.h).decode())

    # ---- state transfer
    def set_particles(self, cell_count, vol, wgroup=None, wclass=None, ids=None, n_primary=None, hyst=None, lct=None,
                      gct=None):
        cc = _i32(cell_count)
        assert len(cc) == self.n_cells
        vol = _f64(vol)
        a = [_i32(wgroup), _i32(wclass), _i64(ids), _i32(n_primary), _i32(hyst), _f64(lct), _f64(gct)]
        self._chk(self.L.pmcgpu_set_particles(self.h, _p(cc), _p(vol), *[_p(x) for x in a]))

    def counts(self):
        out = np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_get_counts(self.h, _p(out)))
        return out

    def get_particles(self):
        cnt = self.counts()
        N = int(cnt.sum())
        S = self.data.n_spec
        out = dict(count=cnt, vol=np.zeros((S, N)), wgroup=np.zeros(N, np.int32), wclass=np.zeros(N, np.int32),
                   id=np.zeros(N, np.int64), n_primary=np.zeros(N, np.int32), hyst=np.zeros(N, np.int32),
                   lct=np.zeros(N), gct=np.zeros(N))
        self._chk(self.L.pmcgpu_get_particles(self.h, C.c_longlong(N), _p(out["vol"]), _p(out["wgroup"]),
                                              _p(out["wclass"]), _p(out["id"]), _p(out["n_primary"]),
                                              _p(out["hyst"]), _p(out["lct"]), _p(out["gct"])))
        return out

    def set_next_id(self, next_id):
        a = _i64(next_id)
        self._chk(self.L.pmcgpu_set_next_id(self.h, _p(a)))

    def get_next_id(self):
        out = np.zeros(self.n_cells, np.int64)
        self._chk(self.L.pmcgpu_get_next_id(self.h, _p(out)))
        return out

    def set_weights(self, wtype, magnitude, exponent):
        n = self.n_cells * self.n_class * self.n_group
        t, m, e = _i32(wtype), _f64(magnitude), _f64(exponent)
        assert t.size == n and m.size == n and e.size == n
        self._chk(self.L.pmcgpu_set_weights(self.h, _p(t), _p(m), _p(e)))

    def get_weights(self):
        shp = (self.n_cells, self.n_class, self.n_group)
        t, m, e = np.zeros(shp, np.int32), np.zeros(shp), np.zeros(shp)
        self._chk(self.L.pmcgpu_get_weights(self.h, _p(t), _p(m), _p(e)))
        return t, m, e

    def set_env(self, temp=None, pressure=None, rel_humid=None, inverse_density=None, cell_volume=None):
        a = [_f64(x) for x in (temp, pressure, rel_humid, inverse_density, cell_volume)]
        for x in a:
            assert x is None or x.size == self.n_cells
        self._chk(self.L.pmcgpu_set_env(self.h, *[_p(x) for x in a]))

    def rel_humid(self):
        out = np.zeros(self.n_cells)
        self._chk(self.L.pmcgpu_get_rel_humid(self.h, _p(out)))
        return out

    # ---- the path
    def aero_state_sort(self):
        self._chk(self.L.pmcgpu_aero_state_sort(self.h))

    def kernel_bounds(self, kernel_type):
        self._chk(self.L.pmcgpu_kernel_bounds(self.h, C.c_int(kernel_type)))

    def mc_coag(self, kernel_type, del_t, step=0):
        ns, nc = np.zeros(self.n_cells, np.int32), np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_mc_coag(self.h, C.c_int(kernel_type), C.c_double(del_t), C.c_uint(step), _p(ns),
                                        _p(nc)))
        return ns, nc

    def condense_particles(self, temp_final, pressure_final, del_t, reltol=0.0, step=0):
        """condense_particles (condense.F90:147) for every cell; returns (rel_humid, stats[n][3], water_rel_error).
        `step` addresses the draws of aero_state_reweight (only drawn when weights depend on size)."""
        self._chk(self.L.pmcgpu_set_rng_step(self.h, C.c_uint(step)))
        T1, p1 = _f64(temp_final), _f64(pressure_final)
        assert T1.size == self.n_cells and p1.size == self.n_cells
        rh, st, wre = np.zeros(self.n_cells), np.zeros((self.n_cells, 3), np.int32), np.zeros(self.n_cells)
        self._chk(self.L.pmcgpu_condense_particles(self.h, _p(T1), _p(p1), C.c_double(del_t), C.c_double(reltol),
                                                   _p(rh), _p(st), _p(wre)))
        return rh, st, wre

    def set_condense_method(self, method):
        """5: CVODE-style BDF (default), 4: RODAS4 (cross-validation)"""
        self._chk(self.L.pmcgpu_set_condense_method(self.h, C.c_int(method)))

    def condense_equilib_particles(self, step=0):
        self._chk(self.L.pmcgpu_set_rng_step(self.h, C.c_uint(step)))
        self._chk(self.L.pmcgpu_condense_equilib_particles(self.h))

    def eval_condense_rates(self, inputs):
        a = _f64(inputs).reshape(-1, 9)
        out = np.zeros((a.shape[0], 6))
        self._chk(self.L.pmcgpu_eval_condense_rates(self.h, C.c_int(a.shape[0]), _p(a), _p(out)))
        return out

    def aero_state_rebalance(self, allow_doubling=True, allow_halving=True, step=0):
        self._chk(self.L.pmcgpu_aero_state_rebalance(self.h, C.c_int(int(allow_doubling)),
                                                     C.c_int(int(allow_halving)), C.c_uint(step)))

    def partmc_timestep(self, temp=None, pressure=None, rel_humid=None, inverse_density=None, cell_volume=None,
                        elapsed_time=0.0, dt=1.0, chem_dt=60.0, do_coagulation=True, coag_kernel_type=KERNEL_BROWN,
                        do_condensation=False, allow_doubling=True, allow_halving=True, step=0, num_conc_out=None,
                        rel_humid_out=None, no_wait=False):
        """partmc_timestep (interface/wrf_pmc_driver.F90:46) for every cell of the patch.  The met arrays are
        float64 [n_cells] buffers (numpy, pinned torch tensors, or device tensors); they are handed to the
        library as plain pointers and must stay alive until the call has completed."""
        a = _TimestepArgs()
        a.temp, a.pressure, a.rel_humid = _addr(temp), _addr(pressure), _addr(rel_humid)
        a.inverse_density, a.cell_volume = _addr(inverse_density), _addr(cell_volume)
        a.elapsed_time, a.dt, a.chem_dt = float(elapsed_time), float(dt), float(chem_dt)
        a.do_coagulation, a.coag_kernel_type = int(do_coagulation), int(coag_kernel_type)
        a.do_condensation = int(do_condensation)
        a.allow_doubling, a.allow_halving = int(allow_doubling), int(allow_halving)
        a.step = int(step)
        a.num_conc_out, a.rel_humid_out = _addr(num_conc_out), _addr(rel_humid_out)
        a.no_wait = int(no_wait)
        self._chk(self.L.pmcgpu_partmc_timestep(self.h, C.byref(a)))

    def wrf_to_partmc(self, p, pb, t_2, qv, alt, z_at_w, rdx, rdy, z=None, msftx=None, msfty=None):
        """wrf_to_partmc (interface/wrf_pmc_driver.F90:576) on the device from WRF's `real` fields of the patch;
        the result becomes the env_state of the next partmc_timestep called without met arrays.  Returns the
        staged fields as a dict of [n_cells] arrays."""
        a = _WrfFields()
        keep = []
        for k, x in (("p", p), ("pb", pb), ("t_2", t_2), ("qv", qv), ("alt", alt), ("z", z), ("z_at_w", z_at_w),
                     ("msftx", msftx), ("msfty", msfty)):
            if x is not None and isinstance(x, np.ndarray):
                x = np.ascontiguousarray(x, dtype=np.float32)
                keep.append(x)
            setattr(a, k, _addr(x))
        a.rdx, a.rdy = float(rdx), float(rdy)
        self._chk(self.L.pmcgpu_wrf_to_partmc(self.h, C.byref(a)))
        out = [np.zeros(self.n_cells) for _ in range(6)]
        self._chk(self.L.pmcgpu_get_env_staging(self.h, *[_p(x) for x in out]))
        return dict(zip(("temp", "pressure", "rel_humid", "inverse_density", "cell_volume", "height"), out))

    def wrf_pmc_trans_driver(self, fields, step=0, vertmix_onoff=1, allow_doubling=True, allow_halving=True,
                             no_wait=False):
        """wrf_pmc_trans_driver (interface/wrf_pmc_trans.F90:37): move probabilities, then trans_aero
        (interface/wrf_pmc_trans_aero.F90:40).  `fields` maps u_flux, v_flux, w_flux, xkhh, alt, z, z_at_w,
        exch_h to float32 slabs of this patch (numpy, pinned or device tensors; a missing key keeps the values
        of the previous call) and dx, dy, dt to scalars (synthetic.patch_fields builds such a dict)."""
        a = _TransArgs()
        keep = []
        for k in ("u_flux", "v_flux", "w_flux", "xkhh", "alt", "z", "z_at_w", "exch_h"):
            x = fields.get(k)
            if x is not None and isinstance(x, np.ndarray):
                x = np.ascontiguousarray(x, dtype=np.float32)
                keep.append(x)
            setattr(a, k, _addr(x))
        a.dx, a.dy, a.dt = float(fields["dx"]), float(fields["dy"]), float(fields["dt"])
        a.vertmix_onoff = int(vertmix_onoff)
        a.allow_doubling, a.allow_halving = int(allow_doubling), int(allow_halving)
        a.step = int(step)
        a.no_wait = int(no_wait)
        self._chk(self.L.pmcgpu_wrf_pmc_trans_driver(self.h, C.byref(a)))
        if no_wait and keep:
            self.synchronize()  # host copies made here must outlive the asynchronous upload

    def aero_state_add_aero_dist_sample(self, mode_type, weight_class, char_radius, log10_std_dev, vol_frac, num_conc,
                                        sample_prop, characteristic_factor, create_time=0.0, vol_frac_std=None,
                                        allow_doubling=True, allow_halving=True, step=0, stream=0,
                                        sample_radius=None, sample_num_conc=None):
        """aero_state_add_aero_dist_sample (aero_state.F90:758) for every cell: modes described by arrays
        [n_mode] (vol_frac [n_mode][n_spec]), num_conc [n_cells][n_mode], sample_prop [n_cells].
        Returns the number of particles added per cell."""
        a = _EmitArgs()
        ty, wc = _i32(mode_type), _i32(weight_class)
        cr, ls, vf = _f64(char_radius), _f64(log10_std_dev), _f64(vol_frac)
        nc, sp = _f64(num_conc), _f64(sample_prop)
        vs = _f64(vol_frac_std) if vol_frac_std is not None else None
        nm = ty.size
        assert wc.size == nm and cr.size == nm and ls.size == nm and vf.size == nm * self.data.n_spec
        assert nc.size == self.n_cells * nm and sp.size == self.n_cells
        a.n_mode = nm
        a.mode_type, a.weight_class, a.char_radius, a.log10_std_dev = _addr(ty), _addr(wc), _addr(cr), _addr(ls)
        a.vol_frac, a.vol_frac_std, a.num_conc, a.sample_prop = _addr(vf), _addr(vs), _addr(nc), _addr(sp)
        a.characteristic_factor, a.create_time = float(characteristic_factor), float(create_time)
        a.allow_doubling, a.allow_halving, a.step = int(allow_doubling), int(allow_halving), int(step)
        a.stream = int(stream)
        sr, sn = _f64(sample_radius), _f64(sample_num_conc)  # sampled modes (type 4): [n_mode][nb + 1], [n_mode][nb]
        if sr is not None:
            a.n_sample_bins = sn.size // nm
            assert sr.size == nm * (a.n_sample_bins + 1)
            a.sample_radius, a.sample_num_conc = _addr(sr), _addr(sn)
        n_add = np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_aero_state_add_aero_dist_sample(self.h, C.byref(a), _p(n_add)))
        return n_add

    def seasalt_emissions(self, u10, v10, xland, height, sea_salt_opt, i_na, i_cl, weight_class, dt, elapsed_time,
                          allow_doubling=True, allow_halving=True, step=0):
        """seasalt_emissions (interface/wrf_pmc_driver.F90:1767) as partmc_timestep calls it (:212-223): 2-D fields
        [ny][nx] u10, v10, xland (WRF reals) and the height of the lowest layer; returns particles added per cell."""
        a = _SeasaltArgs()
        keep = [np.ascontiguousarray(x, dtype=np.float32) for x in (u10, v10, xland)]
        keep.append(np.ascontiguousarray(height, dtype=np.float64))
        keep.append(_i32(weight_class))
        a.u10, a.v10, a.xland, a.height, a.weight_class = [_addr(x) for x in keep]
        a.sea_salt_opt, a.i_na, a.i_cl = int(sea_salt_opt), int(i_na), int(i_cl)
        a.dt, a.elapsed_time = float(dt), float(elapsed_time)
        a.allow_doubling, a.allow_halving, a.step = int(allow_doubling), int(allow_halving), int(step)
        n_add = np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_seasalt_emissions(self.h, C.byref(a), _p(n_add)))
        return n_add

    def aero_state_scale_weight(self, i_group, i_class, weight_ratio, allow_doubling=True, allow_halving=True, step=0):
        """aero_state_scale_weight (aero_state.F90:1832) for (i_group, i_class) of every cell; weight_ratio[n_cells]"""
        r = _f64(weight_ratio)
        assert r.size == self.n_cells
        self._chk(self.L.pmcgpu_aero_state_scale_weight(self.h, C.c_int(i_group), C.c_int(i_class), _p(r),
                                                        C.c_int(int(allow_doubling)), C.c_int(int(allow_halving)),
                                                        C.c_uint(step)))

    def wrf_pmc_dep_dry_driver(self, rmol, znt, ust, ivgtyp, height, dt, step=0):
        """wrf_pmc_dep_dry_driver (interface/wrf_pmc_dep.F90:31): dry deposition in the lowest layer.
        2-D fields [ny][nx]: rmol, znt, ust (WRF reals), ivgtyp (USGS land use), height of the lowest layer."""
        a = _DepArgs()
        keep = [np.ascontiguousarray(x, dtype=np.float32) for x in (rmol, znt, ust)]
        keep.append(np.ascontiguousarray(ivgtyp, dtype=np.int32))
        keep.append(np.ascontiguousarray(height, dtype=np.float64))
        for x in keep:
            assert x.size == self.nx * self.ny
        a.rmol, a.znt, a.ust, a.ivgtyp, a.height = [_addr(x) for x in keep]
        a.dt, a.step, a.no_wait = float(dt), int(step), 0
        self._chk(self.L.pmcgpu_wrf_pmc_dep_dry_driver(self.h, C.byref(a)))

    def eval_dep(self, inputs):
        a = _f64(inputs).reshape(-1, 10)
        out = np.zeros((a.shape[0], 3))
        self._chk(self.L.pmcgpu_eval_dep(self.h, C.c_int(a.shape[0]), _p(a), _p(out)))
        return out

    def trans_probs(self):
        """(adv[n_cells][C][6], diff[n_cells][C][4], vert[n_cells][C][nz]) of the last transport call"""
        n, Cc = self.n_cells, self.n_class
        adv, diff, vert = np.zeros((n, Cc, 6)), np.zeros((n, Cc, 4)), np.zeros((n, Cc, self.nz))
        self._chk(self.L.pmcgpu_get_trans_probs(self.h, _p(adv), _p(diff), _p(vert)))
        return adv, diff, vert

    def trans_moved(self):
        out = np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_get_trans_moved(self.h, _p(out)))
        return out

    def comm_init(self, unique_id, n_ranks, rank, px, py):
        """patch decomposition px x py of the periodic domain, one rank per patch (pmcgpu_comm_init)"""
        self._chk(self.L.pmcgpu_comm_init(self.h, unique_id, C.c_int(n_ranks), C.c_int(rank), C.c_int(px),
                                          C.c_int(py)))

    def set_boundary(self, mode_type, weight_class, char_radius, log10_std_dev, vol_frac, num_conc, outflow, throw_away,
                     vol_frac_std=None):
        """boundary conditions of an open domain (pmcgpu_set_boundary): background modes ([n_mode] arrays, vol_frac
        [n_mode][n_spec]), num_conc [n_cells][n_mode], outflow / throw_away [n_cells] of THIS patch's cells"""
        a = _BoundaryArgs()
        ty, wc = _i32(mode_type), _i32(weight_class)
        cr, ls, vf, vs = _f64(char_radius), _f64(log10_std_dev), _f64(vol_frac), _f64(vol_frac_std)
        nc, of, ta = _f64(num_conc), _i32(outflow), _i32(throw_away)
        assert nc.size == self.n_cells * len(ty) and of.size == self.n_cells and ta.size == self.n_cells
        a.n_mode = len(ty)
        a.mode_type, a.weight_class = ty.ctypes.data, wc.ctypes.data
        a.char_radius, a.log10_std_dev, a.vol_frac = cr.ctypes.data, ls.ctypes.data, vf.ctypes.data
        a.vol_frac_std = None if vs is None else vs.ctypes.data
        a.num_conc, a.outflow, a.throw_away = nc.ctypes.data, of.ctypes.data, ta.ctypes.data
        self._chk(self.L.pmcgpu_set_boundary(self.h, C.byref(a)))

    def track_components(self, n_opaque=0):
        """carry aero_particle_t%component(:) and n_opaque opaque doubles (optics) with every particle record"""
        self._chk(self.L.pmcgpu_track_components(self.h, C.c_int(n_opaque)))
        self.n_opaque = int(n_opaque)

    def set_components(self, n_comp, source, create_time, opaque=None):
        """n_comp [N], source / create_time [N][20], opaque [N][n_opaque], particles in get_particles order"""
        nc, src, tm = _i32(n_comp), _i32(source), _f64(create_time)
        oq = _f64(opaque) if opaque is not None else None
        assert src.size == nc.size * 20 and tm.size == nc.size * 20
        self._chk(self.L.pmcgpu_set_components(self.h, C.c_longlong(nc.size), _p(nc), _p(src), _p(tm),
                                               _p(oq) if oq is not None else None))

    def get_components(self):
        n = int(self.counts().sum())
        no = getattr(self, "n_opaque", 0)
        nc, src, tm, oq = np.zeros(n, np.int32), np.zeros((n, 20), np.int32), np.zeros((n, 20)), np.zeros((n, no))
        self._chk(self.L.pmcgpu_get_components(self.h, C.c_longlong(n), _p(nc), _p(src), _p(tm),
                                               _p(oq) if no else None))
        return nc, src, tm, oq

    def num_concs_by_source(self, n_source):
        """aero_state_num_concs_by_source (aero_state.F90:1310) per cell: [n_cells][n_source]"""
        out = np.zeros((self.n_cells, n_source))
        self._chk(self.L.pmcgpu_num_concs_by_source(self.h, C.c_int(n_source), _p(out)))
        return out

    def partmc_process(self, n_bin=100, d_min=1e-9, d_max=1e-3, static=False):
        """per-cell diagnostics of partmc_process (interface/wrf_pmc_driver.F90:826-944), reduced on the device;
        static=True: pmcgpu_partmc_process_static (no allocation, result table in a static device array)"""
        n, S, Cc = self.n_cells, self.data.n_spec, self.n_class
        out = dict(tot_num_conc=np.zeros(n), tot_wet_num_conc=np.zeros(n), tot_mass_conc=np.zeros(n),
                   num_dist=np.zeros((n, n_bin)), mass_dist=np.zeros((n, n_bin)), num_conc_by_class=np.zeros((n, Cc)),
                   species_mass_conc=np.zeros((n, S)))
        fn = self.L.pmcgpu_partmc_process_static if static else self.L.pmcgpu_partmc_process
        fn.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double] + [C.c_void_p] * 7
        self._chk(fn(self.h, n_bin, d_min, d_max, *[out[k].ctypes.data for k in (
            "tot_num_conc", "tot_wet_num_conc", "tot_mass_conc", "num_dist", "mass_dist", "num_conc_by_class",
            "species_mass_conc")]))
        return out

    def checkpoint_write(self, path):
        self._chk(self.L.pmcgpu_checkpoint_write(self.h, C.c_char_p(path.encode())))

    def checkpoint_read(self, path):
        self._chk(self.L.pmcgpu_checkpoint_read(self.h, C.c_char_p(path.encode())))

    def comm_destroy(self):
        self._chk(self.L.pmcgpu_comm_destroy(self.h))

    def set_n_part_ideal(self, n_part_ideal):
        """n_part_ideal(group, class) (aero_state.F90:81): [n_class][n_group]"""
        a = _f64(n_part_ideal)
        assert a.size == self.n_class * self.n_group
        self._chk(self.L.pmcgpu_set_n_part_ideal(self.h, _p(a)))

    def set_fractal(self, frac_dim=3.0, prime_radius=1e-8, vol_fill_factor=1.0):
        """aero_data%fractal (fractal.F90:55-82); no arguments = spherical.  Before the first sort of a run."""
        self._chk(self.L.pmcgpu_set_fractal(self.h, C.c_double(frac_dim), C.c_double(prime_radius),
                                            C.c_double(vol_fill_factor)))

    def check_status(self):
        self._chk(self.L.pmcgpu_check_status(self.h))

    def profile_enable(self, on=True):
        self._chk(self.L.pmcgpu_profile_enable(self.h, C.c_int(int(on))))

    def profile_read(self, reset=True):
        """{kernel name: (total ms, launches)} since the last reset"""
        buf = C.create_string_buffer(1 << 16)
        self._chk(self.L.pmcgpu_profile_read(self.h, buf, C.c_int(len(buf)), C.c_int(int(reset))))
        out = {}
        for line in buf.value.decode().splitlines():
            name, ms, n = line.split()
            out[name] = (float(ms), int(n))
        return out

    def measure_fp64_peak(self):
        v = C.c_double(0)
        self._chk(self.L.pmcgpu_measure_fp64_peak(self.h, C.byref(v)))
        return v.value

    def condense_stats(self):
        """[n_cells][3] internal steps, rejected steps, rate evaluations per particle of the last condensation"""
        out = np.zeros((self.n_cells, 3), np.int32)
        self._chk(self.L.pmcgpu_get_condense_stats(self.h, _p(out)))
        return out

    def coag_stats(self):
        """(tot_n_samp[n_cells], tot_n_coag[n_cells]) of the last mc_coag"""
        a, b = np.zeros(self.n_cells, np.int32), np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_get_coag_stats(self.h, _p(a), _p(b)))
        return a, b

    def work(self, reset=False):
        """device work counters since the last reset: dict(cond_rate_evals = rate evaluations x particles of
        condense_particles, cond_particles, particle_timesteps)"""
        w = np.zeros(4, np.uint64)
        self._chk(self.L.pmcgpu_get_work(self.h, _p(w), C.c_int(1 if reset else 0)))
        return dict(cond_rate_evals=int(w[0]), cond_particles=int(w[1]), particle_timesteps=int(w[2]))

    def update_density(self, new_inverse_density):
        a = _f64(new_inverse_density)
        self._chk(self.L.pmcgpu_update_density(self.h, _p(a)))

    def partmc_to_wrf(self):
        out = np.zeros((self.n_cells, self.n_class), np.float32)
        self._chk(self.L.pmcgpu_num_conc_by_class(self.h, _p(out)))
        return out

    # ---- inspection
    def grid(self):
        nb, gmin, gmax = np.zeros(self.n_cells, np.int32), np.zeros(self.n_cells), np.zeros(self.n_cells)
        self._chk(self.L.pmcgpu_get_grid(self.h, _p(nb), _p(gmin), _p(gmax)))
        return nb, gmin, gmax

    def bins(self):
        N = int(self.counts().sum())
        out = np.zeros(N, np.int32)
        self._chk(self.L.pmcgpu_get_bins(self.h, C.c_longlong(N), _p(out)))
        return out

    def bin_counts(self):
        out = np.zeros((self.n_cells, self.n_class, self.max_bins), np.int32)
        self._chk(self.L.pmcgpu_get_bin_counts(self.h, _p(out)))
        return out

    def kernel_max(self, cell):
        nb = int(self.grid()[0][cell])
        out = np.zeros((nb, nb))
        self._chk(self.L.pmcgpu_get_kernel_max(self.h, C.c_int(cell), _p(out)))
        return out

    def cell_status(self):
        out = np.zeros(self.n_cells, np.int32)
        self._chk(self.L.pmcgpu_get_cell_status(self.h, _p(out)))
        return out

    def launch_count(self):
        return int(self.L.pmcgpu_launch_count(self.h))

    def stream(self):
        return self.L.pmcgpu_stream(self.h)

    def synchronize(self):
        self._chk(self.L.pmcgpu_synchronize(self.h))

    def eval_kernel_brown(self, v1, d1, v2, d2, temp, pressure):
        a = [_f64(x) for x in (v1, d1, v2, d2, temp, pressure)]
        n = a[0].size
        out = np.zeros(n)
        self._chk(self.L.pmcgpu_eval_kernel_brown(self.h, C.c_int(n), *[_p(x) for x in a], _p(out)))
        return out

    def eval_philox(self, cell, step, phase, a, b, j):
        arrs = [np.ascontiguousarray(x, dtype=np.uint32) for x in (cell, step, phase, a, b, j)]
        n = arrs[0].size
        out = np.zeros(n)
        self._chk(self.L.pmcgpu_eval_philox(self.h, C.c_int(n), *[_p(x) for x in arrs], _p(out)))
        return out
