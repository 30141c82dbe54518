/* pmcgpu.h -- C ABI of the B200-native PartMC per-grid-cell aerosol step.
 *
 * This is the drop-in boundary: a Fortran host (WRF-PartMC's interface/ layer)
 * binds these symbols through ISO_C_BINDING (wrf-partmc_b200/fortran/pmc_gpu.F90,
 * INTEGRATION.md) in place of its serial per-cell loops.  Conventions follow the
 * reference's own C boundary (partmc/src/condense.F90:183-195 <->
 * condense_solver.c:84): scalars by value, arrays as plain pointers, int return
 * code, 0 = success; a non-zero code is what the Fortran side hands to
 * die_msg(code, ...) (util.F90:134).
 *
 * All particles of one WRF patch live in device memory behind the opaque
 * handle.  Cell order everywhere is WRF memory order: i fastest, then k, then j
 * (aero_states(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je),
 * interface/wrf_pmc_driver.F90:46-87), cell = i + nx*(k + nz*j), zero based.
 * Weight arrays are (group, class) with group fastest, like
 * aero_weight_array_t%weight(n_group, n_class) (aero_weight_array.F90:35-39).
 * Bins, classes and groups are 1-based in every array that crosses this
 * boundary.  There is no CPU fallback: every entry point fails with
 * PMCGPU_ERR_CUDA when no CUDA device is usable.
 */
#ifndef PMCGPU_H
#define PMCGPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pmcgpu_ctx pmcgpu_ctx;

enum {
  PMCGPU_OK = 0,
  PMCGPU_ERR_CUDA = 1,            /* CUDA runtime error (see pmcgpu_last_error) */
  PMCGPU_ERR_ARG = 2,             /* bad argument */
  PMCGPU_ERR_CAPACITY = 3,        /* a cell outgrew cell_capacity */
  PMCGPU_ERR_BIN_RANGE = 134572570, /* coagulated particle outside bin grid (aero_sorted.F90:434 die_msg code) */
  PMCGPU_ERR_NBINS = 5,           /* adaptive grid needs more than max_bins bins */
  PMCGPU_ERR_ACCEL = 6,           /* reserved (try_per_particle_coag, coagulation.F90:150, is built: never returned) */
  PMCGPU_ERR_CONDENSE = 7,        /* condensation integrator failed (condense.F90:331-363 codes) */
  PMCGPU_ERR_PROB = 414897881,    /* transport probabilities invalid (wrf_pmc_trans.F90:702-771 assert codes) */
  PMCGPU_ERR_NCCL = 9,
  PMCGPU_ERR_REWEIGHT = 10        /* reserved (aero_state_reweight, aero_state.F90:597, is built: never returned) */
};

/* coag_kernel.F90:24-45 type codes */
enum { PMCGPU_KERNEL_SEDI = 1, PMCGPU_KERNEL_ADDITIVE = 2, PMCGPU_KERNEL_CONSTANT = 3, PMCGPU_KERNEL_BROWN = 4, PMCGPU_KERNEL_ZERO = 5,
       PMCGPU_KERNEL_BROWN_FREE = 6, PMCGPU_KERNEL_BROWN_CONT = 7 };
/* aero_weight.F90:25-30 */
enum { PMCGPU_WEIGHT_NONE = 1, PMCGPU_WEIGHT_POWER = 2, PMCGPU_WEIGHT_MFA = 3 };

typedef struct {
  /* aero_data_t (aero_data.F90:55-93) */
  int n_spec;
  const double* density;       /* [n_spec] kg/m^3 */
  const double* kappa;         /* [n_spec] */
  const double* molec_weight;  /* [n_spec] kg/mol */
  const int* num_ions;         /* [n_spec] */
  int i_water;                 /* 1-based, 0 = none */
  /* aero_weight_array_t dims; WRF uses n_group = 1 (wrf_pmc_init.F90:251-254) */
  int n_group, n_class;
  /* patch geometry (cells owned by this process) and the global domain */
  int nx, nz, ny;              /* local cells: i, k, j extents */
  int global_nx, global_ny;    /* global horizontal extents */
  int i_offset, j_offset;      /* global index of local (0,0) */
  int periodic_x, periodic_y;
  double n_part_ideal;         /* grid%num_particles, split evenly over (group, class) (aero_state.F90:241);
                                  per-(group, class) values: pmcgpu_set_n_part_ideal */
  int cell_capacity;           /* particle slots per cell (>= 2 * n_part_ideal + slack) */
  int max_bins;                /* capacity of the per-cell adaptive bin grid (<= 255) */
  unsigned int seed;           /* random_seed (wrf_pmc_init.F90:131) */
  int device;                  /* CUDA device ordinal */
} pmcgpu_config;

/* ---- lifetime -------------------------------------------------------- */
int pmcgpu_create(const pmcgpu_config* cfg, pmcgpu_ctx** out);
int pmcgpu_destroy(pmcgpu_ctx* ctx);
const char* pmcgpu_last_error(pmcgpu_ctx* ctx);
/* number of kernels this library has launched so far (bench.py's gpu_launches) */
long long pmcgpu_launch_count(pmcgpu_ctx* ctx);
/* the CUDA stream every kernel of this context is launched on (for event timing) */
void* pmcgpu_stream(pmcgpu_ctx* ctx);
int pmcgpu_synchronize(pmcgpu_ctx* ctx);

/* ---- aero_state upload / download (host SoA views) --------------------
 * cell_count[n_cells]; the per-particle arrays are the concatenation of all
 * cells in cell order, N = sum(cell_count).  vol is species-major: vol[s*N + p].
 * Any of wgroup..gct may be NULL on upload (defaults: group 1, class 1, id 0,
 * n_primary 1, hyst 0, times 0). */
int pmcgpu_set_particles(pmcgpu_ctx* ctx, const int* cell_count, const double* vol, const int* wgroup,
                         const int* wclass, const int64_t* id, const int* n_primary, const int* hyst,
                         const double* least_create_time, const double* greatest_create_time);
int pmcgpu_get_counts(pmcgpu_ctx* ctx, int* cell_count);
int pmcgpu_get_particles(pmcgpu_ctx* ctx, long long n_total, double* vol, int* wgroup, int* wclass, int64_t* id,
                         int* n_primary, int* hyst, double* least_create_time, double* greatest_create_time);
/* The cold part of the particle record (aero_particle.F90:38-54), opt-in because it more than doubles the record:
 * component(:) = up to PMCGPU_MAX_COMPONENTS {source_id, create_time} pairs (MAX_AERO_COMPONENT_SIZE, :22) and
 * n_opaque doubles per particle that the device moves but never interprets (the optical fields absorb_cross_sect,
 * scatter_cross_sect, asymmetry, refract_shell, refract_core, core_vol of do_optical runs).  After
 * pmcgpu_track_components every record move carries them: sort, rebalance, reweight, scale_weight, deposition,
 * transport (also across ranks: the halo record grows by 4 + 20 * 12 + 8 n_opaque bytes), checkpoints; new particles
 * get one component {weight class of their mode as source (one class per source in WRF), create time} and zero
 * optics (aero_particle_set_component, aero_state.F90:823); coagulation concatenates the two lists or keeps 20 of
 * them split in proportion to n_primary_parts (aero_particle_coagulate, aero_particle.F90:955-972, and the
 * accelerated path's resampling, coagulation.F90:472-508) and zeroes the optics (:920-929).
 * Call once per context, before pmcgpu_comm_init and before the first transport step; particles already uploaded get
 * the default one component {weight class, least_create_time}, like every later pmcgpu_set_particles. */
#define PMCGPU_MAX_COMPONENTS 20
int pmcgpu_track_components(pmcgpu_ctx* ctx, int n_opaque /* 0..64 */);
/* packed like pmcgpu_get_particles (cell by cell, particle order): n_comp [n_total], source and create_time
 * [n_total][PMCGPU_MAX_COMPONENTS] (entries beyond n_comp ignored / returned as 0), opaque [n_total][n_opaque]
 * (may be NULL).  n_total must equal the device population. */
int pmcgpu_set_components(pmcgpu_ctx* ctx, long long n_total, const int* n_comp, const int* source,
                          const double* create_time, const double* opaque);
int pmcgpu_get_components(pmcgpu_ctx* ctx, long long n_total, int* n_comp, int* source, double* create_time,
                          double* opaque);
/* aero_state_num_concs_by_source (aero_state.F90:1310-1347) for every cell: out [n_cells][n_source] (host or device),
 * sources 1..n_source; every component credits its source with num_conc * (n_primary_parts / n_components), the
 * quotient being the Fortran's integer division.  Needs pmcgpu_track_components. */
int pmcgpu_num_concs_by_source(pmcgpu_ctx* ctx, int n_source, double* out);
/* per-cell id counters (aero_particle.F90:64 next_id, one per cell on the device) */
int pmcgpu_set_next_id(pmcgpu_ctx* ctx, const int64_t* next_id);
int pmcgpu_get_next_id(pmcgpu_ctx* ctx, int64_t* next_id);
/* aero_weight_array per cell: [n_cells][n_class][n_group] */
int pmcgpu_set_weights(pmcgpu_ctx* ctx, const int* type, const double* magnitude, const double* exponent);
int pmcgpu_get_weights(pmcgpu_ctx* ctx, int* type, double* magnitude, double* exponent);
/* env_state per cell (env_state.F90:29-76; wrf_to_partmc, wrf_pmc_driver.F90:576-644). NULL = keep. */
int pmcgpu_set_env(pmcgpu_ctx* ctx, const double* temp, const double* pressure, const double* rel_humid,
                   const double* inverse_density, const double* cell_volume);
int pmcgpu_get_rel_humid(pmcgpu_ctx* ctx, double* rel_humid);
/* aero_state%n_part_ideal(n_group, n_class) (aero_state.F90:81, :241): one value per (group, class) for all
 * cells, [(class - 1) * n_group + (group - 1)]; the default is pmcgpu_config.n_part_ideal split evenly. */
int pmcgpu_set_n_part_ideal(pmcgpu_ctx* ctx, const double* n_part_ideal);
/* aero_data%fractal (fractal.F90:55-66; spec file do_fractal / frac_dim / prime_radius / vol_fill_factor,
 * spec_file_read_fractal fractal.F90:546-566).  Default: spherical particles (3, 1e-8, 1), which is what WRF-PartMC
 * runs with.  A non-spherical setting changes the geometric radius (fractal_vol2rad / fractal_rad2vol,
 * fractal.F90:106-158) that the sort, the majorant tables and mc_coag use, for the kernels brown_free, brown_cont,
 * additive, constant and zero, and the volume of the particles pmcgpu_aero_state_add_aero_dist_sample creates from
 * their sampled geometric radius (aero_state.F90:817).  NOT built for fractal particles: the mobility radius
 * (fractal.F90:239-440), i.e. the Brownian and sedimentation kernels (mc_coag returns PMCGPU_ERR_ARG) and
 * condensation, dry deposition, sea salt, transport and the diagnostics, which return PMCGPU_ERR_ARG while a
 * non-spherical setting is active.  Call before the first sort of a run; every cell gets a new bin grid at its
 * next sort. */
int pmcgpu_set_fractal(pmcgpu_ctx* ctx, double frac_dim, double prime_radius, double vol_fill_factor);

/* ---- the path ---------------------------------------------------------- */
/* aero_state_sort (aero_state.F90:3218) for every cell: adaptive bin grid
 * (aero_sorted.F90:214-373), bin of every particle (bin_grid.F90:143), and the
 * (class, bin)-ordered particle layout that replaces integer_rmap2. */
int pmcgpu_aero_state_sort(pmcgpu_ctx* ctx);
/* est_k_minmax_binned_unweighted (coag_kernel.F90:253) for cells whose bounds are stale. */
int pmcgpu_kernel_bounds(pmcgpu_ctx* ctx, int coag_kernel_type);
/* mc_coag (coagulation.F90:43) for every cell.  tot_n_samp / tot_n_coag are
 * per-cell outputs [n_cells], may be NULL.  step addresses the random draws. */
int pmcgpu_mc_coag(pmcgpu_ctx* ctx, int coag_kernel_type, double del_t, unsigned int step, int* tot_n_samp,
                   int* tot_n_coag);
/* aero_state_scale_weight (aero_state.F90:1832-1895) for (i_group, i_class) of every cell:
 * weight_ratio[n_cells] (host or device; <= 0 or exactly 1 leaves a cell alone).  ratio > 1: magnitude x ratio,
 * prob_round(n (1 - 1/ratio)) random particles of the class removed; ratio < 1: magnitude x ratio, every
 * particle of the class duplicated prob_round(1/ratio) - 1 times.  Used by the reference for preweighting in
 * transport (ratio > 1, also built into pmcgpu_wrf_pmc_trans_driver) and by aero_state_prepare_weight_for_add
 * before emissions (:703-757). */
int pmcgpu_aero_state_scale_weight(pmcgpu_ctx* ctx, int i_group, int i_class, const double* weight_ratio,
                                   int allow_doubling, int allow_halving, unsigned step);

/* aero_state_add_aero_dist_sample (aero_state.F90:758-831) for every cell: emission / inflow of new
 * particles sampled from a distribution of modes (the reference's callers: emissions
 * wrf_pmc_driver.F90:1767-1924, boundary inflow wrf_pmc_trans_aero.F90:689-766).  Per (group, mode):
 * aero_state_prepare_weight_for_add (:703-757) rescales the class weight when the new count would leave
 * [ideal/2, 2 ideal]; a Poisson number of particles is created with radii from the mode (aero_mode.F90:549)
 * and volume fractions from aero_mode_sample_vols (:633).  All pointers host or device. */
typedef struct {
  int n_mode;
  const int* mode_type;         /* [n_mode] 1 log-normal, 2 exponential, 3 mono-disperse, 4 sampled (aero_mode.F90:26-37) */
  const int* weight_class;      /* [n_mode] weight class of the mode's source, 1-based */
  const double* char_radius;    /* [n_mode] geometric mean / mean-volume / mono radius (m) */
  const double* log10_std_dev;  /* [n_mode] log10 of the geometric standard deviation (log-normal modes) */
  const double* vol_frac;       /* [n_mode][n_spec] species volume fractions */
  const double* vol_frac_std;   /* [n_mode][n_spec] their standard deviations, NULL = 0 */
  const double* num_conc;       /* [n_cells][n_mode] number concentration of each mode in each cell (m^-3) */
  const double* sample_prop;    /* [n_cells] fraction to sample, e.g. emission_rate_scale * dt / height */
  double characteristic_factor; /* sample_timescale / dt (:771) */
  double create_time;
  int allow_doubling, allow_halving;
  unsigned int step;            /* addresses the random draws */
  unsigned int stream;          /* added to the (group, mode) part of the draw address: 0 for scenario emissions; callers
                                   that sample in the same cells in the same step use different values (sea salt: 1024) */
  /* sampled (tabulated) modes, mode_type 4 (aero_mode.F90:456-491, 533-538, 581-608): every mode has a row of
   * n_sample_bins + 1 bin-edge radii (m) and n_sample_bins number concentrations (m^-3) (rows of other mode types are
   * ignored; all pointers may be NULL / 0 when no mode is sampled); num_conc[cell][mode] of a sampled mode is a
   * dimensionless factor on its table (1 = as tabulated) */
  int n_sample_bins;
  const double* sample_radius;    /* [n_mode][n_sample_bins + 1] */
  const double* sample_num_conc;  /* [n_mode][n_sample_bins] */
} pmcgpu_emit_args;
int pmcgpu_aero_state_add_aero_dist_sample(pmcgpu_ctx* ctx, const pmcgpu_emit_args* args, int* n_part_add /* [n_cells] or NULL */);

/* seasalt_emissions (interface/wrf_pmc_driver.F90:1767-1924) as partmc_timestep calls it (:212-223): for every cell
 * of the lowest level (k = 1) over water (grid%xland > 1.5), wind_speed = (u10^2 + v10^2)**.5 in single precision,
 * log-normal sea salt modes of Na / Cl (volume fractions mw_na / (mw_na + mw_cl) and the rest) sampled with
 * p = emission_rate_scale * dt / env_state%height and characteristic_factor = max(1, min(3600, elapsed_time)) / dt
 * through aero_state_add_aero_dist_sample.  sea_salt_opt (grid%partmc_seasalt_param): 1 = Gong (2003) matched to the
 * two MAM3 modes (number fluxes 73.7 / 10.2 x wind_speed**3.41), 2 = Ovadnevaite et al. (2014), five modes with the
 * Reynolds number the reference fixes at 3.1e6.  weight_class [2 or 5]: classes of the sources "sea_salt_01" ...
 * Fields are 2-D [ny][nx], i fastest, host or device pointers.  Call after pmcgpu_partmc_timestep's emissions slot;
 * the draws use stream 1024 of the emission phase. */
typedef struct {
  const float* u10;        /* grid%u10 */
  const float* v10;        /* grid%v10 */
  const float* xland;      /* grid%xland: 1 land, 2 water */
  const double* height;    /* env_states(i,1,j)%height */
  int sea_salt_opt;
  int i_na, i_cl;          /* species numbers of "Na" and "Cl", 1-based */
  const int* weight_class; /* [n_modes] 1-based */
  double dt;               /* grid%dt */
  double elapsed_time;     /* env_state%elapsed_time: creation time of the new particles */
  int allow_doubling, allow_halving;
  unsigned int step;       /* addresses the random draws */
} pmcgpu_seasalt_args;
int pmcgpu_seasalt_emissions(pmcgpu_ctx* ctx, const pmcgpu_seasalt_args* args, int* n_part_add /* [n_cells] or NULL */);

/* Step number that addresses the random draws of the stand-alone condensation calls below (the
 * prob_round draws of aero_state_reweight, aero_state.F90:597-653, when weights depend on size);
 * pmcgpu_partmc_timestep uses its own `step` argument instead.  The reference has one global
 * stream (rand.F90), which a parallel device cannot reproduce: DESIGN.md section 2. */
int pmcgpu_set_rng_step(pmcgpu_ctx* ctx, unsigned step);

/* condense_particles (condense.F90:147) for every cell: env_state_initial is the cell's current
 * env_state (pmcgpu_set_env), env_state_final = (temp_final, pressure_final) [n_cells]; on return the
 * device env is the final one with rel_humid from the integrator (condense.F90:263).  reltol <= 0
 * selects the reference's 1e-8.  Optional outputs [n_cells]: rel_humid_out; stats_out [n_cells][3] =
 * internal steps, rejected steps, rate evaluations per particle; water_rel_error_out (condense.F90:315). */
int pmcgpu_condense_particles(pmcgpu_ctx* ctx, const double* temp_final, const double* pressure_final,
                              double del_t, double reltol, double* rel_humid_out, int* stats_out,
                              double* water_rel_error_out);
/* time stepper of condense_particles: 4 (default) = RODAS4, a stiffly accurate L-stable Rosenbrock method;
 * 5 = restatement of CVODE's variable-order BDF method, the solver the reference calls
 * (condense_solver.c:131-166).  Both use the reference's tolerances and its exact arrow-head solve; they agree
 * to the integration tolerance.  RODAS4 is the default because its few accept / reject decisions make the
 * device reproduce the CPU restatement to 1e-10, whereas the many threshold decisions of a variable-order
 * BDF code let two correct implementations drift apart to the tolerance level. */
int pmcgpu_set_condense_method(pmcgpu_ctx* ctx, int method);
/* condense_equilib_particles (condense.F90:744) with each cell's current env_state. */
int pmcgpu_condense_equilib_particles(pmcgpu_ctx* ctx);
/* evaluation of condense_rates (condense.F90:371) on the device, for known-answer tests:
 * in [n][9] = T,Tdot,H,p,pdot,V_comp,D,D_dry,kappa ; out [n][6] = Ddot,Hdot_i,dDdot_dD,dDdot_dH,dHdoti_dD,dHdoti_dH */
int pmcgpu_eval_condense_rates(pmcgpu_ctx* ctx, int n, const double* in, double* out);
/* aero_state_rebalance (aero_state.F90:1765) for every cell. */
int pmcgpu_aero_state_rebalance(pmcgpu_ctx* ctx, int allow_doubling, int allow_halving, unsigned int step);
/* scenario_update_aero_state, WRF branch (scenario.F90:331-334): awa *= old_rrho / rrho */
int pmcgpu_update_density(pmcgpu_ctx* ctx, const double* new_inverse_density);
/* partmc_to_wrf (wrf_pmc_driver.F90:1704-1719): out[cell*n_class + c] = max(count*magnitude*alt, 1e-16) */
int pmcgpu_num_conc_by_class(pmcgpu_ctx* ctx, float* out);

/* wrf_to_partmc (interface/wrf_pmc_driver.F90:576-644) + rel_hum (:649-666) + get_grid_cell_volume (:2002-2020)
 * for every cell of the patch, on the device: WRF's single-precision fields in, the env_state of every cell out.
 * Arithmetic types follow the Fortran statement by statement: p + pb and t_2 + t0 are single-precision sums,
 * temperature = theta (p / p0)**rcp in double with WRF's single-precision constants (t0 = 300., p0 = 100000.,
 * rcp = r_d / cp = 287. / (7. * 287. / 2.), share/module_model_constants.F:23-43), rh =
 * qv / (3.8 exp(17.27 (T - 273) / (T - 36)) / (.01 p)) with the literals 17.27 and .01 in single precision,
 * clamped to [0.1, 0.95]; height = z_at_w(k+1) - z_at_w(k) (single); cell_volume = dx dy dz with
 * dx = (1 / rdx) / msftx.  All arrays are WRF `real`, patch-sized, WRF memory order, host or device pointers:
 *   p, pb, t_2, qv, alt, z [nx nz ny];  z_at_w [nx (nz+1) ny];  msftx, msfty [nx ny] (NULL = 1).
 * The new fields are kept in a staging area and become the env_state of the NEXT pmcgpu_partmc_timestep whose
 * temp / pressure / rel_humid / inverse_density / cell_volume pointers are NULL (the old state is then kept as
 * old_env_states, :615, exactly as when host arrays are passed).  pmcgpu_get_env_staging copies them out
 * ([n_cells] each, any pointer may be NULL) for inspection. */
typedef struct {
  const float* p; const float* pb; const float* t_2; const float* qv; const float* alt; const float* z;
  const float* z_at_w; const float* msftx; const float* msfty;
  float rdx, rdy;                 /* grid%rdx, grid%rdy */
} pmcgpu_wrf_fields;
int pmcgpu_wrf_to_partmc(pmcgpu_ctx* ctx, const pmcgpu_wrf_fields* f);
int pmcgpu_get_env_staging(pmcgpu_ctx* ctx, double* temp, double* pressure, double* rel_humid,
                           double* inverse_density, double* cell_volume, double* height);

/* partmc_timestep (interface/wrf_pmc_driver.F90:46-290) for every cell of the patch, as one
 * asynchronous chain on the context's stream: wrf_to_partmc (new env_state from the host arrays below,
 * the previous one becomes old_env_states), mc_coag when mod(elapsed_time, chem_dt) == 0 and
 * do_coagulation (:189-197), scenario_update_aero_state's WRF branch (awa *= old_rrho / rrho),
 * water uptake by condense_particles(old_env, env, chem_dt) in the slot of mosaic_timestep (:235) when
 * do_condensation, aero_state_rebalance (:248) and partmc_to_wrf (:1704-1719).
 * Arrays are [n_cells] in cell order, in host memory (what the Fortran side passes) or in device memory
 * (unified addressing tells them apart); NULL keeps the device value / skips the output. */
typedef struct {
  const double* temp;             /* env_state%temp            (wrf_to_partmc, :576-644) */
  const double* pressure;         /* env_state%pressure */
  const double* rel_humid;        /* env_state%rel_humid; NULL keeps the integrator's value.  Installed only in steps
                                     without condensation: condense_particles starts from old_env_states' value
                                     (condense.F90:237) and overwrites the new one (:263) */
  const double* inverse_density;  /* grid%alt */
  const double* cell_volume;
  double elapsed_time;            /* seconds since simulation start (:126) */
  double dt;                      /* grid%dt */
  double chem_dt;                 /* grid%partmc_chem_dt */
  int do_coagulation;             /* grid%do_coagulation */
  int coag_kernel_type;           /* WRF: PMCGPU_KERNEL_BROWN (:192) */
  int do_condensation;
  int allow_doubling, allow_halving;
  unsigned int step;              /* addresses the random draws of this call */
  float* num_conc_out;            /* [n_cells][n_class] partmc_to_wrf, optional */
  double* rel_humid_out;          /* [n_cells], optional */
  int no_wait;                    /* 1: return after enqueueing; call pmcgpu_check_status later */
} pmcgpu_timestep_args;
int pmcgpu_partmc_timestep(pmcgpu_ctx* ctx, const pmcgpu_timestep_args* args);
/* waits for the stream and returns the first non-zero per-cell status (0 = ok) */
int pmcgpu_check_status(pmcgpu_ctx* ctx);

/* wrf_pmc_trans_driver (interface/wrf_pmc_trans.F90:37-200) for the patch: the move probabilities
 * (compute_advect_probs_wrf :205, compute_diffusion_probs :358, compute_vertical_probs :414, normalize_probs
 * :692) from WRF's per-class advective fluxes and mixing coefficients, then trans_aero
 * (interface/wrf_pmc_trans_aero.F90:40-971: preweighting, multinomial sampling of the moves, ghost exchange,
 * add + rebalance + sort).  Doubly periodic domains, or open domains (see pmcgpu_set_boundary).
 * Field slabs are WRF `real` arrays of the patch in WRF memory order (i fastest, then k, then j; the class
 * index slowest), in host or device memory; NULL keeps the previous call's values:
 *   u_flux [(nx+1) nz ny n_class]   grid%u_flux(i, k, j, 1 + class), staggered in x (faces i .. i+1)
 *   v_flux [nx nz (ny+1) n_class]   grid%v_flux, staggered in y
 *   w_flux [nx (nz+1) ny n_class]   grid%w_flux, staggered in z
 *   xkhh, alt [(nx+2) nz (ny+2)]    grid%xkhh, grid%alt with WRF's one-cell halo in x and y
 *   z [nx nz ny], z_at_w [nx (nz+1) ny], exch_h [nx (nz+1) ny]   grid%z, grid%z_at_w, grid%exch_h
 * cell volumes and inverse densities are the env_state values (pmcgpu_set_env / pmcgpu_partmc_timestep). */
typedef struct {
  const float* u_flux; const float* v_flux; const float* w_flux;
  const float* xkhh; const float* alt;
  const float* z; const float* z_at_w; const float* exch_h;
  float dx, dy, dt;               /* grid%dx, grid%dy, grid%dt */
  int vertmix_onoff;              /* config_flags%vertmix_onoff */
  int allow_doubling, allow_halving;
  unsigned int step;              /* addresses the random draws of this call (use a value no other call of
                                     the same WRF step uses) */
  int no_wait;
} pmcgpu_trans_args;
int pmcgpu_wrf_pmc_trans_driver(pmcgpu_ctx* ctx, const pmcgpu_trans_args* args);

/* Open (non-periodic) domains: pmcgpu_config.periodic_x = periodic_y = 0.  The outermost ring of cells of the GLOBAL
 * domain then holds the boundary condition, like in the reference: pmcgpu_partmc_timestep runs the aerosol step
 * on interior cells only (wrf_pmc_driver.F90:157-167), pmcgpu_wrf_pmc_trans_driver samples the multinomial moves of
 * interior cells only (wrf_pmc_trans.F90:116-136, wrf_pmc_trans_aero.F90:497-506), feeds the first interior cells
 * from the boundary cells (boundary inflow, :689-766), and finally refreshes every boundary-edge cell (:811-955):
 *   outflow[cell] != 0 (the wind blows out of the domain there: u_2 < 0 on the west face, ...):
 *       aero_state_dup_all_particles (aero_state.F90:3529) from the interior neighbour;
 *   else aero_state_resample (wrf_pmc_trans_aero.F90:976-1098) towards the background distribution, with
 *       throw_away[cell] != 0 (wind reversed since the last step, or the background changed) discarding the cell
 *       first.  density = 1 / inverse_density, allow_doubling = allow_halving = true as at the reference's call sites.
 * The background (scenario%aero_background interpolated to the current time, aero_dist_interp_1d) is a list of modes
 * like pmcgpu_emit_args; num_conc [n_cells][n_mode], outflow and throw_away [n_cells] are read for boundary-edge
 * cells only.  The fields xkhh / alt of pmcgpu_trans_args carry WRF's halo values beyond the domain faces.
 * Call before every pmcgpu_wrf_pmc_trans_driver whose boundary data changed (host or device pointers). */
typedef struct {
  int n_mode;
  const int* mode_type;         /* [n_mode] 1 log-normal, 2 exponential, 3 mono-disperse */
  const int* weight_class;      /* [n_mode] 1-based */
  const double* char_radius;    /* [n_mode] */
  const double* log10_std_dev;  /* [n_mode] */
  const double* vol_frac;       /* [n_mode][n_spec] */
  const double* vol_frac_std;   /* [n_mode][n_spec], NULL = 0 */
  const double* num_conc;       /* [n_cells][n_mode] */
  const int* outflow;           /* [n_cells] */
  const int* throw_away;        /* [n_cells] */
} pmcgpu_boundary_args;
int pmcgpu_set_boundary(pmcgpu_ctx* ctx, const pmcgpu_boundary_args* args);

/* wrf_pmc_dep_dry_driver (interface/wrf_pmc_dep.F90:31-97, called from wrf_pmc_trans_driver when
 * grid%do_deposition, wrf_pmc_trans.F90:183-197): dry deposition of particles in the lowest layer.
 * Aerodynamic resistance from depvel_partmc (:188-213), per-particle deposition velocity and removal
 * probability (dt / dz) v_d from Zhang et al. (2001) as coded in wrf_pmc_dep_aero.F90:204-470, land-use
 * mapping of dep_aero_dry_driver (:31-150), backward removal loop of aero_particles_remove_by_dep (:280).
 * All fields are 2-D, [ny][nx] with i fastest (host or device pointers); cells on the outermost ring of
 * the GLOBAL domain are skipped like in the reference (:113-114). */
typedef struct {
  const float* rmol;     /* grid%rmol: inverse Monin-Obukhov length */
  const float* znt;      /* grid%znt: roughness length (m) */
  const float* ust;      /* grid%ust: friction velocity (m/s) */
  const int* ivgtyp;     /* grid%ivgtyp: USGS land-use category */
  const double* height;  /* env_states(i,1,j)%height = z_at_w(i,2,j) - z_at_w(i,1,j) (wrf_pmc_driver.F90:627) */
  double dt;             /* grid%dt */
  unsigned int step;     /* addresses the random draws */
  int no_wait;           /* 1: return without waiting for the device */
} pmcgpu_dep_args;
int pmcgpu_wrf_pmc_dep_dry_driver(pmcgpu_ctx* ctx, const pmcgpu_dep_args* args);
/* unit evaluation of the deposition formulas (parity tests): in[n][10] = diameter, density, temp,
 * inverse_density, ustar, rmol, znt, ivgtyp, dt, dz; out[n][3] = aer_res, v_d, removal probability */
int pmcgpu_eval_dep(pmcgpu_ctx* ctx, int n, const double* in, double* out);
/* patch decomposition of the periodic domain: px x py patches, rank = rx + px * ry, one process and one
 * GPU per patch (WRF's 2-D decomposition; MPI_Cart_shift of wrf_pmc_trans_aero.F90:153-159).  unique_id is
 * the 128-byte NCCL id made by pmcgpu_comm_unique_id on one rank and distributed by the host (MPI_Bcast in
 * WRF).  Without this call the context is a single periodic patch. */
int pmcgpu_comm_unique_id(char* out, int len);
int pmcgpu_comm_init(pmcgpu_ctx* ctx, const char* unique_id, int n_ranks, int rank, int px, int py);
int pmcgpu_comm_destroy(pmcgpu_ctx* ctx);
/* inspection: probabilities of the last call ([n_cells][n_class][6 | 4 | nz]; advection -x +x -y +y -z +z,
 * diffusion -x +x -y +y, vertical row of B* A) and the number of arrivals per cell */
int pmcgpu_get_trans_probs(pmcgpu_ctx* ctx, double* adv, double* diff, double* vert);
int pmcgpu_get_trans_moved(pmcgpu_ctx* ctx, int* n_moved);

/* The per-cell diagnostics of partmc_process (interface/wrf_pmc_driver.F90:826-944) as device reductions, so that
 * a history alarm does not pull the particles to the host: total number concentration (:840), wet number
 * concentration (aero_state_total_num_conc_wet), total mass concentration (:846-848), number and mass size
 * distributions over the DRY diameter on the logarithmic grid bin_grid_make(LOG, n_bin, d_min, d_max) (the
 * reference uses 100, 1d-9, 1d-3; bin_grid_histogram_1d, bin_grid.F90:225-248: weights divided by the bin width in
 * ln d), number concentration per weight class (= source in WRF, :858-861) and mass concentration of every species
 * (sum of species mass x number concentration; :865-930 picks its named species from this table).
 * Outputs are [n_cells], [n_cells][n_bin], [n_cells][n_class], [n_cells][n_spec]; host or device; any may be NULL. */
int pmcgpu_partmc_process(pmcgpu_ctx* ctx, int n_bin, double d_min, double d_max, double* tot_num_conc,
                          double* tot_wet_num_conc, double* tot_mass_conc, double* num_dist, double* mass_dist,
                          double* num_conc_by_class, double* species_mass_conc);
/* pmcgpu_partmc_process without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, same results.  The per-cell result table the kernel fills before it is copied to the caller's
 * arrays lives in a file-scope device array of PMCGPU_PROCESS_STATIC_MAX_DOUBLES doubles (8 MiB, held by every GPU
 * that loads the library until exit): n_cells (3 + 2 n_bin + n_class + n_spec) beyond that returns
 * PMCGPU_ERR_CAPACITY with a message in pmcgpu_last_error.  All calls on one GPU share the array; it keeps the last
 * call's contents (every entry used by a call is rewritten by that call), and calls that overlap on the GPU -- two
 * contexts on one device, two host threads -- race on it: serialise them. */
#define PMCGPU_PROCESS_STATIC_MAX_DOUBLES (1 << 20)
int pmcgpu_partmc_process_static(pmcgpu_ctx* ctx, int n_bin, double d_min, double d_max, double* tot_num_conc,
                                 double* tot_wet_num_conc, double* tot_mass_conc, double* num_dist,
                                 double* mass_dist, double* num_conc_by_class, double* species_mass_conc);

/* Checkpoint / restart of the whole device state of a context (what WRF's restart alarm needs; the reference
 * writes netCDF, partmc/src/output.F90:827-967): one flat binary file per rank with the particle records in cell
 * order, weights, env_state, id counters, bin grids and majorant tables.  Draws are counter-based, so a context of
 * the same geometry that reads the file continues bit for bit. */
int pmcgpu_checkpoint_write(pmcgpu_ctx* ctx, const char* path);
int pmcgpu_checkpoint_read(pmcgpu_ctx* ctx, const char* path);

/* ---- measurement ---------------------------------------------------------
 * per-kernel CUDA-event timing on the context's stream; pmcgpu_profile_read writes one
 * "kernel_name total_ms launches" line per kernel into buf. */
int pmcgpu_profile_enable(pmcgpu_ctx* ctx, int on);
int pmcgpu_profile_read(pmcgpu_ctx* ctx, char* buf, int buf_len, int reset);
/* FP64 FMA throughput of the device (TFLOP/s), the roofline denominator of the FP64-bound kernels */
int pmcgpu_measure_fp64_peak(pmcgpu_ctx* ctx, double* tflops);

/* ---- inspection (parity tests) ----------------------------------------- */
int pmcgpu_get_grid(pmcgpu_ctx* ctx, int* n_bin, double* grid_min, double* grid_max); /* [n_cells] each */
int pmcgpu_get_bins(pmcgpu_ctx* ctx, long long n_total, int* bin_of_particle);        /* packed like particles */
int pmcgpu_get_bin_counts(pmcgpu_ctx* ctx, int* counts); /* [n_cells][n_class][max_bins] */
int pmcgpu_get_kernel_max(pmcgpu_ctx* ctx, int cell, double* kmax); /* [n_bin][n_bin], (b1,b2) at [(b2-1)*n_bin+b1-1] */
int pmcgpu_get_cell_status(pmcgpu_ctx* ctx, int* status); /* [n_cells] first error per cell */
/* counters of the last condensation ([n_cells][3]: steps, rejected steps, rate evaluations per particle;
 * condense.F90:135-138 condense_count_vf / _solve) and of the last mc_coag (tot_n_samp, tot_n_coag) */
int pmcgpu_get_condense_stats(pmcgpu_ctx* ctx, int* stats);
int pmcgpu_get_coag_stats(pmcgpu_ctx* ctx, int* tot_n_samp, int* tot_n_coag);
/* work counters accumulated on the device since the last reset (bench.py's roofline and metric):
 * work[0] particle rate evaluations of condense_particles (rate evaluations x particles, over cells and calls),
 * work[1] particles condense_particles integrated, work[2] particles at the start of every
 * pmcgpu_partmc_timestep (particle-timesteps), work[3] unused */
int pmcgpu_get_work(pmcgpu_ctx* ctx, unsigned long long* work, int reset);
/* stand-alone evaluations of device functions, for known-answer tests */
int pmcgpu_eval_kernel_brown(pmcgpu_ctx* ctx, int n, const double* v1, const double* d1, const double* v2,
                             const double* d2, const double* temp, const double* pressure, double* out);
int pmcgpu_eval_philox(pmcgpu_ctx* ctx, int n, const unsigned int* cell, const unsigned int* step,
                       const unsigned int* phase, const unsigned int* a, const unsigned int* b,
                       const unsigned int* j, double* out);

#ifdef __cplusplus
}
#endif
#endif /* PMCGPU_H */
