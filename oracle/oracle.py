"""ctypes binding of the CPU oracle (oracle/libpmc_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, by bench.py's cpu_baseline /
``--impl reference`` leg and by ``__graft_entry__.smoke()`` as the checker.
The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_lp = C.POINTER(C.c_longlong)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libpmc_oracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        L = _LIB
        L.pmco_create.restype = C.c_void_p
        L.pmco_last_error.restype = C.c_char_p
        L.pmco_get_next_id.restype = C.c_longlong
        L.pmco_get_rel_humid.restype = C.c_double
        L.pmco_kernel_brown.restype = C.c_double
        L.pmco_kernel_brown.argtypes = [C.c_double] * 6
        L.pmco_philox_uniform.restype = C.c_double
        L.pmco_poisson_pmf.restype = C.c_double
        L.pmco_poisson_pmf.argtypes = [C.c_double, C.c_int]
        L.pmco_binomial_pmf.restype = C.c_double
        L.pmco_binomial_pmf.argtypes = [C.c_int, C.c_double, C.c_int]
        L.pmco_mt_u32.restype = C.c_uint
        L.pmco_sat_vapor_pressure.restype = C.c_double
        L.pmco_sat_vapor_pressure.argtypes = [C.c_double]
    return _LIB


def _p(a, ct):
    return None if a is None else a.ctypes.data_as(C.POINTER(ct))


def _f64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def _i64(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int64)


KERNEL_SEDI, KERNEL_ADDITIVE, KERNEL_CONSTANT, KERNEL_BROWN, KERNEL_ZERO = 1, 2, 3, 4, 5
KERNEL_BROWN_FREE, KERNEL_BROWN_CONT = 6, 7  # coag_kernel.F90:42-45
WEIGHT_NONE, WEIGHT_POWER = 1, 2


class AeroData:
    """aero_data_t (aero_data.F90:55-93): species table read by the path."""

    def __init__(self, density, kappa=None, molec_weight=None, num_ions=None, i_water=0, names=None):
        self.density = _f64(density)
        n = len(self.density)
        self.n_spec = n
        self.kappa = _f64(kappa if kappa is not None else np.zeros(n))
        self.molec_weight = _f64(molec_weight if molec_weight is not None else np.full(n, 18e-3))
        self.num_ions = _i32(num_ions if num_ions is not None else np.zeros(n))
        self.i_water = int(i_water)
        self.names = names


class Oracle:
    """A domain of independent grid cells, each an aero_state_t + env_state_t."""

    def __init__(self, n_cells, data, n_group=1, n_class=1, n_part_ideal=1000.0, rng="philox", seed=1,
                 n_cells_global=None, global_cell=None):
        L = lib()
        self.L = L
        self.n_cells = n_cells
        self.data = data
        self.n_group, self.n_class = n_group, n_class
        gc = _i64(global_cell) if global_cell is not None else None
        self.h = C.c_void_p(L.pmco_create(
            C.c_int(n_cells), C.c_int(data.n_spec), _p(data.density, C.c_double), _p(data.kappa, C.c_double),
            _p(data.molec_weight, C.c_double), _p(data.num_ions, C.c_int), C.c_int(data.i_water),
            C.c_int(n_group), C.c_int(n_class), C.c_double(n_part_ideal), C.c_int(0 if rng == "philox" else 1),
            C.c_uint(seed), C.c_longlong(n_cells_global if n_cells_global is not None else n_cells),
            _p(gc, C.c_longlong)))

    def __del__(self):
        try:
            self.L.pmco_destroy(self.h)
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError("oracle error %d: %s" % (rc, self.L.pmco_last_error(self.h).decode()))

    def set_cell(self, cell, vol, wgroup=None, wclass=None, ids=None, n_primary=None, hyst=None, lct=None, gct=None):
        vol = _f64(vol)  # [S][n]
        assert vol.ndim == 2 and vol.shape[0] == self.data.n_spec
        n = vol.shape[1]
        a = [_i32(wgroup), _i32(wclass), _i64(ids), _i32(n_primary), _i32(hyst), _f64(lct), _f64(gct)]
        self._chk(self.L.pmco_set_cell(self.h, C.c_int(cell), C.c_int(n), _p(vol, C.c_double), _p(a[0], C.c_int),
                                       _p(a[1], C.c_int), _p(a[2], C.c_longlong), _p(a[3], C.c_int),
                                       _p(a[4], C.c_int), _p(a[5], C.c_double), _p(a[6], C.c_double)))

    def count(self, cell):
        return int(self.L.pmco_cell_count(self.h, C.c_int(cell)))

    def get_cell(self, cell):
        n = self.count(cell)
        S = self.data.n_spec
        out = dict(vol=np.zeros((S, n)), wgroup=np.zeros(n, np.int32), wclass=np.zeros(n, np.int32),
                   id=np.zeros(n, np.int64), n_primary=np.zeros(n, np.int32), hyst=np.zeros(n, np.int32),
                   lct=np.zeros(n), gct=np.zeros(n))
        self.L.pmco_get_cell(self.h, C.c_int(cell), _p(out["vol"], C.c_double), _p(out["wgroup"], C.c_int),
                             _p(out["wclass"], C.c_int), _p(out["id"], C.c_longlong), _p(out["n_primary"], C.c_int),
                             _p(out["hyst"], C.c_int), _p(out["lct"], C.c_double), _p(out["gct"], C.c_double))
        return out

    def set_components(self, cell, n_comp, source, create_time, opaque=None):
        """component(:) lists of a cell's particles: n_comp [n], source / create_time [n][20]; opaque [n][n_opq]"""
        nc, src, tm = _i32(n_comp), _i32(source), _f64(create_time)
        oq = _f64(opaque) if opaque is not None else None
        n_opq = 0 if oq is None or nc.size == 0 else oq.size // nc.size
        self._chk(self.L.pmco_set_components(self.h, C.c_int(cell), _p(nc, C.c_int), _p(src, C.c_int),
                                             _p(tm, C.c_double), C.c_int(n_opq),
                                             _p(oq, C.c_double) if oq is not None else None))

    def get_components(self, cell, n_opq=0):
        n = self.count(cell)
        nc, src, tm = np.zeros(n, np.int32), np.zeros((n, 20), np.int32), np.zeros((n, 20))
        oq = np.zeros((n, n_opq))
        self._chk(self.L.pmco_get_components(self.h, C.c_int(cell), _p(nc, C.c_int), _p(src, C.c_int),
                                             _p(tm, C.c_double), C.c_int(n_opq), _p(oq, C.c_double)))
        return nc, src, tm, oq

    def set_weights(self, cell, wtype, magnitude, exponent):
        t, m, e = _i32(wtype).ravel(), _f64(magnitude).ravel(), _f64(exponent).ravel()
        self.L.pmco_set_weights(self.h, C.c_int(cell), _p(t, C.c_int), _p(m, C.c_double), _p(e, C.c_double))

    def get_weights(self, cell):
        k = self.n_group * self.n_class
        t, m, e = np.zeros(k, np.int32), np.zeros(k), np.zeros(k)
        self.L.pmco_get_weights(self.h, C.c_int(cell), _p(t, C.c_int), _p(m, C.c_double), _p(e, C.c_double))
        return t, m, e

    def set_env(self, cell, T=288.0, p=1e5, RH=0.5, rrho=1.0, cell_volume=1.0, elapsed=0.0):
        self.L.pmco_set_env(self.h, C.c_int(cell), C.c_double(T), C.c_double(p), C.c_double(RH), C.c_double(rrho),
                            C.c_double(cell_volume), C.c_double(elapsed))

    def set_next_id(self, cell, v):
        self.L.pmco_set_next_id(self.h, C.c_int(cell), C.c_longlong(v))

    def sort(self, cell, force_resort=True, physical=True, edge_recheck=False):
        self._chk(self.L.pmco_sort(self.h, C.c_int(cell), C.c_int(int(force_resort)), C.c_int(int(physical)),
                                   C.c_int(int(edge_recheck))))

    def grid(self, cell, cap=512):
        nb, gmin, gmax = C.c_int(0), C.c_double(0), C.c_double(0)
        edges = np.zeros(cap)
        self.L.pmco_get_grid(self.h, C.c_int(cell), C.byref(nb), C.byref(gmin), C.byref(gmax), _p(edges, C.c_double),
                             C.c_int(cap))
        return nb.value, gmin.value, gmax.value, edges[: nb.value + 1].copy()

    def bins(self, cell):
        out = np.zeros(self.count(cell), np.int32)
        rc = self.L.pmco_get_bins(self.h, C.c_int(cell), _p(out, C.c_int))
        assert rc == 0
        return out

    def bin_counts(self, cell):
        nb = self.grid(cell)[0]
        out = np.zeros((self.n_class, nb), np.int32)
        self.L.pmco_get_bin_counts(self.h, C.c_int(cell), _p(out, C.c_int))
        return out

    def kernel_bounds(self, cell, kernel_type):
        nb = self.grid(cell)[0]
        kmin, kmax = np.zeros((nb, nb)), np.zeros((nb, nb))  # [b2][b1]
        self._chk(self.L.pmco_kernel_bounds(self.h, C.c_int(cell), C.c_int(kernel_type), _p(kmin, C.c_double),
                                            _p(kmax, C.c_double)))
        return kmin, kmax

    def mc_coag(self, cell, kernel_type, del_t, step=0):
        st = np.zeros(3, np.int32)
        self._chk(self.L.pmco_mc_coag(self.h, C.c_int(cell), C.c_int(kernel_type), C.c_double(del_t), C.c_uint(step),
                                      _p(st, C.c_int)))
        return int(st[0]), int(st[1]), int(st[2])

    def mc_coag_all(self, kernel_type, del_t, step=0, n_threads=1):
        tot = np.zeros(2, np.int64)
        rc = self.L.pmco_mc_coag_all(self.h, C.c_int(kernel_type), C.c_double(del_t), C.c_uint(step),
                                     C.c_int(n_threads), _p(tot, C.c_longlong))
        assert rc == 0, rc
        return int(tot[0]), int(tot[1])

    def rebalance(self, cell, allow_doubling=True, allow_halving=True, step=0):
        self._chk(self.L.pmco_rebalance(self.h, C.c_int(cell), C.c_int(int(allow_doubling)),
                                        C.c_int(int(allow_halving)), C.c_uint(step)))

    def env_rel_humid(self, cell):
        return float(self.L.pmco_get_rel_humid(self.h, C.c_int(cell)))

    def condense_equilib(self, cell, step=0):
        self._chk(self.L.pmco_condense_equilib(self.h, C.c_int(cell), C.c_uint(step)))

    def set_condense_method(self, method):
        """5: CVODE-style BDF (default), 4: RODAS4, 3: RODAS3"""
        self._chk(self.L.pmco_set_condense_method(self.h, C.c_int(method)))

    def condense_particles(self, cell, T1, p1, del_t, step=0, reltol=0.0, method=0):
        st = np.zeros(3, np.int32)
        wre = C.c_double(0)
        self._chk(self.L.pmco_condense_particles(self.h, C.c_int(cell), C.c_double(T1), C.c_double(p1),
                                                 C.c_double(del_t), C.c_uint(step), C.c_double(reltol),
                                                 C.c_int(method), _p(st, C.c_int), C.byref(wre)))
        return st, wre.value

    def partmc_timestep(self, cell_lo, cell_hi, temp=None, pressure=None, rel_humid=None, inverse_density=None,
                        elapsed_time=0.0, chem_dt=60.0, do_coagulation=True, coag_kernel_type=KERNEL_BROWN,
                        do_condensation=False, allow_doubling=True, allow_halving=True, step=0, n_threads=1):
        """partmc_timestep (wrf_pmc_driver.F90:46-290) for cells [cell_lo, cell_hi); met arrays are full length"""
        a = [_f64(x) for x in (temp, pressure, rel_humid, inverse_density)]
        tot = np.zeros(4, np.int64)
        self._chk(self.L.pmco_partmc_timestep(
            self.h, C.c_int(cell_lo), C.c_int(cell_hi), *[_p(x, C.c_double) for x in a], C.c_double(elapsed_time),
            C.c_double(chem_dt), C.c_int(int(do_coagulation)), C.c_int(coag_kernel_type),
            C.c_int(int(do_condensation)), C.c_int(int(allow_doubling)), C.c_int(int(allow_halving)),
            C.c_uint(step), C.c_int(n_threads), _p(tot, C.c_longlong)))
        return dict(n_samp=int(tot[0]), n_coag=int(tot[1]), cond_steps=int(tot[2]), cond_rates=int(tot[3]))

    # ---- transport (pmc_oracle_trans.cpp)
    def trans_setup(self, gnx, nz, gny, px=1, py=1, dx=100.0, dy=100.0, vertmix=1):
        self.trans_dims = (gnx, nz, gny)
        self._chk(self.L.pmco_trans_setup(self.h, C.c_int(gnx), C.c_int(nz), C.c_int(gny), C.c_int(px), C.c_int(py),
                                          C.c_float(dx), C.c_float(dy), C.c_int(vertmix)))

    def trans_set_boundary(self, open_domain, mode_type=None, weight_class=None, char_radius=None, log10_std_dev=None,
                           vol_frac=None, num_conc=None, outflow=None, throw_away=None, vol_frac_std=None):
        """open (non-periodic) domain and its boundary conditions: background modes (arrays [n_mode], vol_frac
        [n_mode][n_spec]), num_conc [n_cells][n_mode], outflow / throw_away [n_cells] (global cell order)"""
        if not open_domain:
            self._chk(self.L.pmco_trans_set_boundary(self.h, C.c_int(0), C.c_int(0), *([None] * 9)))
            return
        ty, wc = _i32(mode_type), _i32(weight_class)
        cr, ls, vf = _f64(char_radius), _f64(log10_std_dev), _f64(vol_frac)
        vs = _f64(vol_frac_std)
        nc, of, ta = _f64(num_conc), _i32(outflow), _i32(throw_away)
        assert nc.size == self.n_cells * len(ty) and of.size == self.n_cells and ta.size == self.n_cells
        self._chk(self.L.pmco_trans_set_boundary(self.h, C.c_int(1), C.c_int(len(ty)), _p(ty, C.c_int), _p(wc, C.c_int),
                                                 _p(cr, C.c_double), _p(ls, C.c_double), _p(vf, C.c_double),
                                                 _p(vs, C.c_double), _p(nc, C.c_double), _p(of, C.c_int),
                                                 _p(ta, C.c_int)))

    def trans_set_met(self, f):
        """f: the dict of synthetic.transport_fields (global float32 arrays)"""
        a = [np.ascontiguousarray(f[k], dtype=np.float32) for k in
             ("u_flux", "v_flux", "w_flux", "xkhh", "alt", "z", "z_at_w", "exch_h")]
        self._chk(self.L.pmco_trans_set_met(self.h, *[_p(x, C.c_float) for x in a]))

    def trans_probs(self, dt):
        """compute_advect_probs_wrf / compute_diffusion_probs / compute_vertical_probs / normalize_probs"""
        self._chk(self.L.pmco_trans_probs(self.h, C.c_float(dt)))

    def trans_get_probs(self, cell, cl):
        nz = self.trans_dims[1]
        adv, diff, vert = np.zeros(6), np.zeros(4), np.zeros(nz)
        self.L.pmco_trans_get_probs(self.h, C.c_int(cell), C.c_int(cl), _p(adv, C.c_double), _p(diff, C.c_double),
                                    _p(vert, C.c_double))
        return adv, diff, vert

    def trans_aero(self, step=0, allow_doubling=True, allow_halving=True, n_threads=1):
        """trans_aero (wrf_pmc_trans_aero.F90:40); returns the number of particles added to other cells"""
        m = C.c_longlong(0)
        self._chk(self.L.pmco_trans_aero(self.h, C.c_uint(step), C.c_int(int(allow_doubling)),
                                         C.c_int(int(allow_halving)), C.c_int(n_threads), C.byref(m)))
        return m.value

    def wrf_pmc_trans_driver(self, f, step=0, allow_doubling=True, allow_halving=True, n_threads=1):
        """wrf_pmc_trans_driver (wrf_pmc_trans.F90:37): probabilities, then trans_aero"""
        self.trans_set_met(f)
        self.trans_probs(float(f["dt"]))
        return self.trans_aero(step, allow_doubling, allow_halving, n_threads)

    def num_conc_by_class(self):
        """partmc_to_wrf (wrf_pmc_driver.F90:1704-1719): max(count * magnitude * alt, 1e-16) as float32"""
        out = np.zeros((self.n_cells, self.n_class), np.float32)
        self.L.pmco_num_conc_by_class(self.h, _p(out, C.c_float))
        return out

    def dep_dry_driver(self, nx, nz, ny, rmol, znt, ust, ivgtyp, height, dt, step=0, i_offset=0, j_offset=0,
                       global_nx=None, global_ny=None):
        """wrf_pmc_dep_dry_driver (wrf_pmc_dep.F90:31): dry deposition in the lowest layer; 2-D fields [ny][nx]"""
        a = [np.ascontiguousarray(x, np.float32).ravel() for x in (rmol, znt, ust)]
        iv = _i32(np.asarray(ivgtyp).ravel())
        hg = _f64(np.asarray(height).ravel())
        self._chk(self.L.pmco_dep_dry_driver(
            self.h, C.c_int(nx), C.c_int(nz), C.c_int(ny), C.c_int(i_offset), C.c_int(j_offset),
            C.c_int(global_nx if global_nx is not None else nx), C.c_int(global_ny if global_ny is not None else ny),
            _p(a[0], C.c_float), _p(a[1], C.c_float), _p(a[2], C.c_float), _p(iv, C.c_int), _p(hg, C.c_double),
            C.c_double(dt), C.c_uint(step)))

    def add_aero_dist_sample(self, cell, mode_type, weight_class, char_radius, log10_std_dev, vol_frac, num_conc,
                             sample_prop, characteristic_factor, create_time=0.0, vol_frac_std=None,
                             allow_doubling=True, allow_halving=True, step=0, sample_radius=None, sample_num_conc=None):
        """aero_state_add_aero_dist_sample (aero_state.F90:758) for one cell; num_conc [n_mode]; sampled modes
        (type 4): sample_radius [n_mode][nb + 1], sample_num_conc [n_mode][nb]"""
        ty, wc = _i32(mode_type), _i32(weight_class)
        cr, ls, vf, nc = _f64(char_radius), _f64(log10_std_dev), _f64(vol_frac), _f64(num_conc)
        vs = _f64(vol_frac_std) if vol_frac_std is not None else None
        n = C.c_int(0)
        self._chk(self.L.pmco_add_aero_dist_sample(
            self.h, C.c_int(cell), C.c_int(ty.size), _p(ty, C.c_int), _p(wc, C.c_int), _p(cr, C.c_double),
            _p(ls, C.c_double), _p(vf, C.c_double), _p(vs, C.c_double) if vs is not None else None,
            _p(nc, C.c_double), C.c_double(sample_prop), C.c_double(characteristic_factor), C.c_double(create_time),
            C.c_int(int(allow_doubling)), C.c_int(int(allow_halving)), C.c_uint(step), C.byref(n),
            C.c_int(0 if sample_num_conc is None else _f64(sample_num_conc).size // ty.size),
            _p(_f64(sample_radius), C.c_double) if sample_radius is not None else None,
            _p(_f64(sample_num_conc), C.c_double) if sample_num_conc is not None else None))
        return n.value

    def seasalt_emissions(self, nx, nz, ny, u10, v10, xland, height, sea_salt_opt, i_na, i_cl, weight_class, dt,
                          elapsed_time, allow_doubling=True, allow_halving=True, step=0):
        """the sea salt block of partmc_timestep (interface/wrf_pmc_driver.F90:212-223, :1767-1924)"""
        u, v, xl = [np.ascontiguousarray(x, np.float32) for x in (u10, v10, xland)]
        hg, wc = _f64(height), _i32(weight_class)
        n_add = np.zeros(nx * nz * ny, np.int32)
        self._chk(self.L.pmco_seasalt_emissions(
            self.h, C.c_int(nx), C.c_int(nz), C.c_int(ny), _p(u, C.c_float), _p(v, C.c_float), _p(xl, C.c_float),
            _p(hg, C.c_double), C.c_int(sea_salt_opt), C.c_int(i_na), C.c_int(i_cl), _p(wc, C.c_int), C.c_double(dt),
            C.c_double(elapsed_time), C.c_int(int(allow_doubling)), C.c_int(int(allow_halving)), C.c_uint(step),
            _p(n_add, C.c_int)))
        return n_add

    def scale_weight(self, cell, g, c, ratio, allow_doubling=True, allow_halving=True, step=0):
        self._chk(self.L.pmco_scale_weight(self.h, C.c_int(cell), C.c_int(g), C.c_int(c), C.c_double(ratio),
                                           C.c_int(int(allow_doubling)), C.c_int(int(allow_halving)), C.c_uint(step)))


def seasalt_modes(sea_salt_opt, n_spec, i_na, i_cl):
    """mode table of seasalt_emissions: char_radius, log10 sigma, number flux, vol_frac [n_modes][n_spec]"""
    L = lib()
    cr, ls, fl, vf = np.zeros(5), np.zeros(5), np.zeros(5), np.zeros(5 * n_spec)
    n = L.pmco_seasalt_modes(C.c_int(sea_salt_opt), C.c_int(n_spec), C.c_int(i_na), C.c_int(i_cl), _p(cr, C.c_double),
                             _p(ls, C.c_double), _p(fl, C.c_double), _p(vf, C.c_double))
    if n < 0:
        raise ValueError("invalid sea salt option")
    return cr[:n], ls[:n], fl[:n], vf[:n * n_spec].reshape(n, n_spec)


def dep_eval(inputs):
    """inputs [n][10] = diameter, density, temp, rrho, ustar, rmol, znt, ivgtyp, dt, dz ->
    outputs [n][3] = aer_res, vd, remove_prob (wrf_pmc_dep.F90:188, wrf_pmc_dep_aero.F90:204-275)"""
    a = _f64(inputs).reshape(-1, 10)
    out = np.zeros((a.shape[0], 3))
    lib().pmco_dep_eval(C.c_int(a.shape[0]), _p(a, C.c_double), _p(out, C.c_double))
    return out


def condense_rates(inputs):
    """inputs [n][9] = T, Tdot, H, p, pdot, V_comp, D, D_dry, kappa -> outputs [n][9] (condense.F90:74-84 order)"""
    a = _f64(inputs).reshape(-1, 9)
    out = np.zeros_like(a)
    lib().pmco_condense_rates(C.c_int(a.shape[0]), _p(a, C.c_double), _p(out, C.c_double))
    return out


def wrf_to_partmc(nx, nz, ny, p, pb, t_2, qv, alt, z_at_w, rdx, rdy, msftx=None, msfty=None):
    """wrf_to_partmc + rel_hum + get_grid_cell_volume (wrf_pmc_driver.F90:576-666, 2002-2020) for a patch of WRF
    `real` fields; returns dict(temp, pressure, rel_humid, inverse_density, cell_volume, height) of [n_cells]"""
    f32 = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float32)
    a = [f32(x) for x in (p, pb, t_2, qv, alt, z_at_w, msftx, msfty)]
    n = nx * nz * ny
    out = np.zeros((6, n))
    L = lib()
    L.pmco_wrf_to_partmc.argtypes = [C.c_int] * 3 + [C.c_void_p] * 8 + [C.c_float, C.c_float, C.c_void_p]
    L.pmco_wrf_to_partmc(nx, nz, ny, *[None if x is None else x.ctypes.data for x in a], float(rdx), float(rdy),
                         out.ctypes.data)
    return dict(zip(("temp", "pressure", "rel_humid", "inverse_density", "cell_volume", "height"), out))


def partmc_process_cell(vol, wclass, num_conc, density, i_water, n_class, n_bin=100, d_min=1e-9, d_max=1e-3):
    """the per-cell diagnostics of partmc_process (interface/wrf_pmc_driver.F90:826-944) for one cell, statement by
    statement in numpy: vol [S][n], wclass [n] (1-based), num_conc [n] (aero_state_num_concs), density [S]"""
    vol = np.asarray(vol, np.float64)
    density = np.asarray(density, np.float64)
    S, n = vol.shape
    masses = (vol * density[:, None]).sum(axis=0)                      # aero_state_masses
    dry = np.ones(S, bool)
    if i_water > 0:
        dry[i_water - 1] = False
    dry_diam = 2.0 * (3.0 * vol[dry].sum(axis=0) / (4.0 * np.pi)) ** (1.0 / 3.0)  # aero_state_dry_diameters
    out = dict(tot_num_conc=num_conc.sum(), tot_mass_conc=(masses * num_conc).sum(),
               tot_wet_num_conc=num_conc[vol[i_water - 1] > 0].sum() if i_water > 0 else 0.0)
    width = (np.log(d_max) - np.log(d_min)) / n_bin                    # bin_grid_make, log grid
    b = bin_grid_find(n_bin, d_min, d_max, dry_diam).astype(int)      # bin_grid_find (bin_grid.F90:143)
    ok = (b >= 1) & (b <= n_bin)
    out["num_dist"] = np.bincount(b[ok] - 1, weights=num_conc[ok] / width, minlength=n_bin)      # bin_grid_histogram_1d
    out["mass_dist"] = np.bincount(b[ok] - 1, weights=(num_conc * masses)[ok] / width, minlength=n_bin)
    out["num_conc_by_class"] = np.bincount(np.asarray(wclass) - 1, weights=num_conc, minlength=n_class)
    out["species_mass_conc"] = (vol * density[:, None] * num_conc[None, :]).sum(axis=1)
    return out


def run_sect(n_bin, d_min, d_max, density, total_num_conc, geom_mean_diam, log10_sigma, kernel_type, temp, pressure,
             t_max, del_t, t_output):
    """the reference's sectional coagulation solver (run_sect.F90:56-362) for one log-normal mode of a one-species
    aerosol; returns (num, mass) densities per ln(r), [n_out][n_bin], at t = 0, t_output, ..."""
    n_out = int(round(t_max / t_output)) + 1
    num, mass = np.zeros((n_out, n_bin)), np.zeros((n_out, n_bin))
    L = lib()
    L.pmco_run_sect.argtypes = [C.c_int] + [C.c_double] * 6 + [C.c_int] + [C.c_double] * 5 + [C.c_void_p, C.c_void_p, C.c_int]
    n = L.pmco_run_sect(n_bin, d_min, d_max, density, total_num_conc, geom_mean_diam, log10_sigma, kernel_type, temp,
                        pressure, t_max, del_t, t_output, num.ctypes.data, mass.ctypes.data, n_out)
    assert n == n_out, n
    return num, mass


def condense_rates_dt(inputs):
    """time derivatives of condense_rates at fixed (D, H): [n][3] = dDdot/dt, dHdot_i/dt, dHdot_env/dt"""
    a = _f64(inputs).reshape(-1, 9)
    out = np.zeros((a.shape[0], 3))
    lib().pmco_condense_rates_dt(C.c_int(a.shape[0]), _p(a, C.c_double), _p(out, C.c_double))
    return out


def condense_integrate(D, H, D_dry, kappa, num_conc, T0, p0, Tdot, pdot, del_t, reltol=1e-8, abstol_scale=1e-8,
                       abstol_H=1e-10, h_init=0.0, method=4):
    D = _f64(D).copy()
    Hc = C.c_double(H)
    st = np.zeros(3, np.int32)
    a = [_f64(D_dry), _f64(kappa), _f64(num_conc)]
    rc = lib().pmco_condense_integrate(C.c_int(len(D)), _p(D, C.c_double), C.byref(Hc), _p(a[0], C.c_double),
                                       _p(a[1], C.c_double), _p(a[2], C.c_double), C.c_double(T0), C.c_double(p0),
                                       C.c_double(Tdot), C.c_double(pdot), C.c_double(del_t), C.c_double(reltol),
                                       C.c_double(abstol_scale), C.c_double(abstol_H), C.c_double(h_init),
                                       C.c_int(method), _p(st, C.c_int))
    return rc, D, Hc.value, st


def kernel_sedi(v1, v2):
    """kernel_sedi_helper (coag_kernel_sedi.F90:79-103)"""
    L = lib()
    L.pmco_kernel_sedi.restype = C.c_double
    L.pmco_kernel_sedi.argtypes = [C.c_double, C.c_double]
    return L.pmco_kernel_sedi(float(v1), float(v2))


def set_fractal(frac_dim=3.0, prime_radius=1e-8, vol_fill_factor=1.0):
    """aero_data%fractal (fractal.F90:55-82) for the whole process; no arguments = spherical"""
    lib().pmco_set_fractal(C.c_double(frac_dim), C.c_double(prime_radius), C.c_double(vol_fill_factor))


def aero_vol2rad(v):
    L = lib()
    L.pmco_aero_vol2rad.restype = C.c_double
    return L.pmco_aero_vol2rad(C.c_double(v))


def aero_rad2vol(r):
    L = lib()
    L.pmco_aero_rad2vol.restype = C.c_double
    return L.pmco_aero_rad2vol(C.c_double(r))


def kernel_brown_free(v1, d1, v2, d2, T):
    """kernel_brown_free_helper (coag_kernel_brown_free.F90:118-147)"""
    L = lib()
    L.pmco_kernel_brown_free.restype = C.c_double
    return L.pmco_kernel_brown_free(*[C.c_double(x) for x in (v1, d1, v2, d2, T)])


def kernel_brown_cont(v1, d1, v2, d2, T):
    """kernel_brown_cont_helper (coag_kernel_brown_cont.F90:117-144)"""
    L = lib()
    L.pmco_kernel_brown_cont.restype = C.c_double
    return L.pmco_kernel_brown_cont(*[C.c_double(x) for x in (v1, d1, v2, d2, T)])


def kernel_brown(v1, d1, v2, d2, T, p):
    return lib().pmco_kernel_brown(v1, d1, v2, d2, T, p)


def bin_grid_make(n_bin, gmin, gmax):
    e, c = np.zeros(n_bin + 1), np.zeros(n_bin)
    lib().pmco_bin_grid_make(C.c_int(n_bin), C.c_double(gmin), C.c_double(gmax), _p(e, C.c_double), _p(c, C.c_double))
    return e, c


def bin_grid_find(n_bin, gmin, gmax, r):
    r = _f64(r)
    out = np.zeros(len(r), np.int32)
    lib().pmco_bin_grid_find(C.c_int(n_bin), C.c_double(gmin), C.c_double(gmax), C.c_int(len(r)), _p(r, C.c_double),
                             _p(out, C.c_int))
    return out


def philox_uniform(seed, cell, step, phase, a, b, j):
    return lib().pmco_philox_uniform(C.c_uint(seed), C.c_uint(cell), C.c_uint(step), C.c_uint(phase), C.c_uint(a),
                                     C.c_uint(b), C.c_uint(j))


def philox_raw(ctr, key):
    c = np.asarray(ctr, np.uint32)
    k = np.asarray(key, np.uint32)
    o = np.zeros(4, np.uint32)
    lib().pmco_philox_raw(_p(c, C.c_uint), _p(k, C.c_uint), _p(o, C.c_uint))
    return o


def sample_poisson(mean, n_samp, seed=1, rng="mt", hist_len=200):
    h = np.zeros(hist_len, np.int64)
    lib().pmco_sample_poisson(C.c_double(mean), C.c_int(n_samp), C.c_uint(seed), C.c_int(0 if rng == "philox" else 1),
                              _p(h, C.c_longlong), C.c_int(hist_len))
    return h


def sample_binomial(n, p, n_samp, seed=1, rng="mt", hist_len=200):
    h = np.zeros(hist_len, np.int64)
    lib().pmco_sample_binomial(C.c_int(n), C.c_double(p), C.c_int(n_samp), C.c_uint(seed),
                               C.c_int(0 if rng == "philox" else 1), _p(h, C.c_longlong), C.c_int(hist_len))
    return h
