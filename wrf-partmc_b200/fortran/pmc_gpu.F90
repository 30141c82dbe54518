! pmc_gpu.F90 -- ISO_C_BINDING layer between WRF-PartMC's interface/ code and
! libpmcgpu.so (include/pmcgpu.h).
!
! This is the only host code a WRF-PartMC build adds: it keeps the argument lists of
!   partmc_timestep       (interface/wrf_pmc_driver.F90:46-87)
!   wrf_pmc_trans_driver  (interface/wrf_pmc_trans.F90:37-76)
! and forwards them to the device library, which owns all particles of the patch.
! Conventions follow the reference's own C boundary (partmc/src/condense.F90:183-195):
! scalars by value, arrays through c_loc of a contiguous target, int return code with
! 0 = success, a non-zero code is handed to die_msg (util.F90:134).
!
! NOTE: no Fortran compiler exists in the build image of this repository (gfortran,
! flang, nvfortran, ifx all absent), so this file is written against the Fortran 2003
! standard and the header but has NOT been compiled; tests drive the same C symbols
! through wrf-partmc_b200/host/pmcgpu.py.  INTEGRATION.md lists the build line.
module pmc_gpu

  use, intrinsic :: iso_c_binding
  implicit none
  private

  public :: pmcgpu_config_t, pmcgpu_timestep_args_t, pmcgpu_trans_args_t
  public :: pmc_gpu_init, pmc_gpu_finalize, pmc_gpu_upload_aero_states, &
       pmc_gpu_download_aero_states, pmc_gpu_partmc_timestep, &
       pmc_gpu_wrf_pmc_trans_driver, pmc_gpu_check
  public :: pmc_gpu_handle

  integer, parameter :: dp = c_double

  !> coag_kernel.F90:24-40
  integer(c_int), parameter, public :: PMCGPU_KERNEL_ADDITIVE = 2, &
       PMCGPU_KERNEL_CONSTANT = 3, PMCGPU_KERNEL_BROWN = 4, PMCGPU_KERNEL_ZERO = 5

  !> pmcgpu_config (include/pmcgpu.h); member order must match the C struct.
  type, bind(c) :: pmcgpu_config_t
     integer(c_int) :: n_spec
     type(c_ptr) :: density, kappa, molec_weight, num_ions
     integer(c_int) :: i_water
     integer(c_int) :: n_group, n_class
     integer(c_int) :: nx, nz, ny
     integer(c_int) :: global_nx, global_ny
     integer(c_int) :: i_offset, j_offset
     integer(c_int) :: periodic_x, periodic_y
     real(c_double) :: n_part_ideal
     integer(c_int) :: cell_capacity
     integer(c_int) :: max_bins
     integer(c_int) :: seed          ! unsigned int on the C side (same size)
     integer(c_int) :: device
  end type pmcgpu_config_t

  !> pmcgpu_timestep_args (include/pmcgpu.h)
  type, bind(c) :: pmcgpu_timestep_args_t
     type(c_ptr) :: temp, pressure, rel_humid, inverse_density, cell_volume
     real(c_double) :: elapsed_time, dt, chem_dt
     integer(c_int) :: do_coagulation, coag_kernel_type, do_condensation
     integer(c_int) :: allow_doubling, allow_halving
     integer(c_int) :: step
     type(c_ptr) :: num_conc_out, rel_humid_out
     integer(c_int) :: no_wait
  end type pmcgpu_timestep_args_t

  !> pmcgpu_trans_args (include/pmcgpu.h)
  type, bind(c) :: pmcgpu_trans_args_t
     type(c_ptr) :: u_flux, v_flux, w_flux, xkhh, alt, z, z_at_w, exch_h
     real(c_float) :: dx, dy, dt
     integer(c_int) :: vertmix_onoff, allow_doubling, allow_halving
     integer(c_int) :: step
     integer(c_int) :: no_wait
  end type pmcgpu_trans_args_t

  !> pmcgpu_dep_args (include/pmcgpu.h)
  type, bind(c) :: pmcgpu_dep_args_t
     type(c_ptr) :: rmol, znt, ust, ivgtyp, height
     real(c_double) :: dt
     integer(c_int) :: step
     integer(c_int) :: no_wait
  end type pmcgpu_dep_args_t

  !> Opaque device context of this MPI rank (one GPU per rank).
  type(c_ptr), save :: pmc_gpu_handle = c_null_ptr
  !> Counter that addresses the random draws: two values per WRF step.
  integer(c_int), save :: pmc_gpu_step = 0

  interface
     integer(c_int) function pmcgpu_create(cfg, ctx) bind(c, name='pmcgpu_create')
       import :: c_int, c_ptr, pmcgpu_config_t
       type(pmcgpu_config_t), intent(in) :: cfg
       type(c_ptr), intent(out) :: ctx
     end function pmcgpu_create
     integer(c_int) function pmcgpu_destroy(ctx) bind(c, name='pmcgpu_destroy')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx
     end function pmcgpu_destroy
     type(c_ptr) function pmcgpu_last_error(ctx) bind(c, name='pmcgpu_last_error')
       import :: c_ptr
       type(c_ptr), value :: ctx
     end function pmcgpu_last_error
     integer(c_int) function pmcgpu_set_particles(ctx, cell_count, vol, wgroup, wclass, id, &
          n_primary, hyst, lct, gct) bind(c, name='pmcgpu_set_particles')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx, cell_count, vol, wgroup, wclass, id, n_primary, hyst, lct, gct
     end function pmcgpu_set_particles
     integer(c_int) function pmcgpu_get_counts(ctx, cell_count) bind(c, name='pmcgpu_get_counts')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx, cell_count
     end function pmcgpu_get_counts
     integer(c_int) function pmcgpu_get_particles(ctx, n_total, vol, wgroup, wclass, id, &
          n_primary, hyst, lct, gct) bind(c, name='pmcgpu_get_particles')
       import :: c_int, c_ptr, c_long_long
       type(c_ptr), value :: ctx
       integer(c_long_long), value :: n_total
       type(c_ptr), value :: vol, wgroup, wclass, id, n_primary, hyst, lct, gct
     end function pmcgpu_get_particles
     integer(c_int) function pmcgpu_set_weights(ctx, wtype, magnitude, exponent) &
          bind(c, name='pmcgpu_set_weights')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx, wtype, magnitude, exponent
     end function pmcgpu_set_weights
     integer(c_int) function pmcgpu_get_weights(ctx, wtype, magnitude, exponent) &
          bind(c, name='pmcgpu_get_weights')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx, wtype, magnitude, exponent
     end function pmcgpu_get_weights
     integer(c_int) function pmcgpu_set_env(ctx, temp, pressure, rel_humid, inverse_density, &
          cell_volume) bind(c, name='pmcgpu_set_env')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx, temp, pressure, rel_humid, inverse_density, cell_volume
     end function pmcgpu_set_env
     integer(c_int) function pmcgpu_partmc_timestep(ctx, args) &
          bind(c, name='pmcgpu_partmc_timestep')
       import :: c_int, c_ptr, pmcgpu_timestep_args_t
       type(c_ptr), value :: ctx
       type(pmcgpu_timestep_args_t), intent(in) :: args
     end function pmcgpu_partmc_timestep
     integer(c_int) function pmcgpu_wrf_pmc_trans_driver(ctx, args) &
          bind(c, name='pmcgpu_wrf_pmc_trans_driver')
       import :: c_int, c_ptr, pmcgpu_trans_args_t
       type(c_ptr), value :: ctx
       type(pmcgpu_trans_args_t), intent(in) :: args
     end function pmcgpu_wrf_pmc_trans_driver
     integer(c_int) function pmcgpu_wrf_pmc_dep_dry_driver(ctx, args) &
          bind(c, name='pmcgpu_wrf_pmc_dep_dry_driver')
       import :: c_int, c_ptr, pmcgpu_dep_args_t
       type(c_ptr), value :: ctx
       type(pmcgpu_dep_args_t), intent(in) :: args
     end function pmcgpu_wrf_pmc_dep_dry_driver
     integer(c_int) function pmcgpu_check_status(ctx) bind(c, name='pmcgpu_check_status')
       import :: c_int, c_ptr
       type(c_ptr), value :: ctx
     end function pmcgpu_check_status
     integer(c_int) function pmcgpu_comm_unique_id(out, len) bind(c, name='pmcgpu_comm_unique_id')
       import :: c_int, c_char
       character(kind=c_char), intent(out) :: out(*)
       integer(c_int), value :: len
     end function pmcgpu_comm_unique_id
     integer(c_int) function pmcgpu_comm_init(ctx, unique_id, n_ranks, rank, px, py) &
          bind(c, name='pmcgpu_comm_init')
       import :: c_int, c_ptr, c_char
       type(c_ptr), value :: ctx
       character(kind=c_char), intent(in) :: unique_id(*)
       integer(c_int), value :: n_ranks, rank, px, py
     end function pmcgpu_comm_init
  end interface

contains

  !> Map a library status to the reference's fatal-error path (util.F90:134 die_msg).
  subroutine pmc_gpu_check(code, where)
    use pmc_util, only: die_msg
    integer(c_int), intent(in) :: code
    character(len=*), intent(in) :: where
    if (code /= 0) then
       call die_msg(int(code), 'pmcgpu: ' // trim(where) // ' failed')
    end if
  end subroutine pmc_gpu_check

  !> Create the device context of this rank's patch (called once from
  !! partmc_init, interface/wrf_pmc_init.F90, after aero_data is read).
  subroutine pmc_gpu_init(aero_data, n_class, n_part_ideal, pmc_is, pmc_ie, pmc_js, pmc_je, &
       pmc_ks, pmc_ke, global_nx, global_ny, periodic_x, periodic_y, seed, local_rank)
    use pmc_aero_data
    type(aero_data_t), intent(in), target :: aero_data
    integer, intent(in) :: n_class
    real(kind=dp), intent(in) :: n_part_ideal
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    integer, intent(in) :: global_nx, global_ny
    logical, intent(in) :: periodic_x, periodic_y
    integer, intent(in) :: seed, local_rank

    type(pmcgpu_config_t) :: cfg
    real(c_double), allocatable, target :: density(:), kappa(:), molec_weight(:)
    integer(c_int), allocatable, target :: num_ions(:)

    density = aero_data%density
    kappa = aero_data%kappa
    molec_weight = aero_data%molec_weight
    num_ions = aero_data%num_ions
    cfg%n_spec = size(density)
    cfg%density = c_loc(density)
    cfg%kappa = c_loc(kappa)
    cfg%molec_weight = c_loc(molec_weight)
    cfg%num_ions = c_loc(num_ions)
    cfg%i_water = aero_data%i_water
    cfg%n_group = 1                      ! wrf_pmc_init.F90:251-254
    cfg%n_class = n_class
    cfg%nx = pmc_ie - pmc_is + 1
    cfg%nz = pmc_ke - pmc_ks + 1
    cfg%ny = pmc_je - pmc_js + 1
    cfg%global_nx = global_nx
    cfg%global_ny = global_ny
    cfg%i_offset = pmc_is - 1
    cfg%j_offset = pmc_js - 1
    cfg%periodic_x = merge(1, 0, periodic_x)
    cfg%periodic_y = merge(1, 0, periodic_y)
    cfg%n_part_ideal = n_part_ideal
    cfg%cell_capacity = int(2.5d0 * n_part_ideal) + 64
    cfg%max_bins = 64
    cfg%seed = seed
    cfg%device = local_rank
    call pmc_gpu_check(pmcgpu_create(cfg, pmc_gpu_handle), 'pmcgpu_create')
  end subroutine pmc_gpu_init

  subroutine pmc_gpu_finalize()
    if (c_associated(pmc_gpu_handle)) then
       call pmc_gpu_check(pmcgpu_destroy(pmc_gpu_handle), 'pmcgpu_destroy')
       pmc_gpu_handle = c_null_ptr
    end if
  end subroutine pmc_gpu_finalize

  !> aero_states(i,k,j) -> device SoA (after partmc_init and after a restart read).
  subroutine pmc_gpu_upload_aero_states(aero_states, env_states, pmc_is, pmc_ie, pmc_js, &
       pmc_je, pmc_ks, pmc_ke)
    use pmc_aero_state
    use pmc_env_state
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), &
         intent(in) :: aero_states
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), &
         intent(in) :: env_states

    integer :: i, j, k, c, p, n_cells, n_spec, n_class, n_tot, q, i_class
    integer(c_int), allocatable, target :: counts(:), wgroup(:), wclass(:), nprim(:), hyst(:), wtype(:)
    integer(c_int64_t), allocatable, target :: ids(:)
    real(c_double), allocatable, target :: vol(:,:), lct(:), gct(:), wmag(:), wexp(:)
    real(c_double), allocatable, target :: temp(:), pres(:), rh(:), alt(:), cvol(:)

    n_cells = size(aero_states)
    n_spec = 0
    allocate(counts(n_cells))
    c = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       counts(c) = aero_state_n_part(aero_states(i,k,j))
       if (counts(c) > 0 .and. n_spec == 0) &
            n_spec = size(aero_states(i,k,j)%apa%particle(1)%vol)
    end do
    end do
    end do
    n_tot = sum(counts)
    n_class = size(aero_states(pmc_is,pmc_ks,pmc_js)%awa%weight, 2)
    allocate(vol(n_tot, n_spec), wgroup(n_tot), wclass(n_tot), ids(n_tot), nprim(n_tot), &
         hyst(n_tot), lct(n_tot), gct(n_tot))
    allocate(wtype(n_cells * n_class), wmag(n_cells * n_class), wexp(n_cells * n_class))
    allocate(temp(n_cells), pres(n_cells), rh(n_cells), alt(n_cells), cvol(n_cells))
    c = 0
    q = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       do p = 1, counts(c)
          q = q + 1
          vol(q, :) = aero_states(i,k,j)%apa%particle(p)%vol     ! species-major: vol(q, s)
          wgroup(q) = aero_states(i,k,j)%apa%particle(p)%weight_group
          wclass(q) = aero_states(i,k,j)%apa%particle(p)%weight_class
          ids(q) = aero_states(i,k,j)%apa%particle(p)%id
          nprim(q) = aero_states(i,k,j)%apa%particle(p)%n_primary_parts
          hyst(q) = aero_states(i,k,j)%apa%particle(p)%water_hyst_leg
          lct(q) = aero_states(i,k,j)%apa%particle(p)%least_create_time
          gct(q) = aero_states(i,k,j)%apa%particle(p)%greatest_create_time
       end do
       do i_class = 1, n_class
          wtype((c - 1) * n_class + i_class) = aero_states(i,k,j)%awa%weight(1, i_class)%type
          wmag((c - 1) * n_class + i_class) = aero_states(i,k,j)%awa%weight(1, i_class)%magnitude
          wexp((c - 1) * n_class + i_class) = aero_states(i,k,j)%awa%weight(1, i_class)%exponent
       end do
       temp(c) = env_states(i,k,j)%temp
       pres(c) = env_states(i,k,j)%pressure
       rh(c) = env_states(i,k,j)%rel_humid
       alt(c) = env_states(i,k,j)%inverse_density
       cvol(c) = env_states(i,k,j)%cell_volume
    end do
    end do
    end do
    call pmc_gpu_check(pmcgpu_set_particles(pmc_gpu_handle, c_loc(counts), c_loc(vol), &
         c_loc(wgroup), c_loc(wclass), c_loc(ids), c_loc(nprim), c_loc(hyst), c_loc(lct), &
         c_loc(gct)), 'pmcgpu_set_particles')
    call pmc_gpu_check(pmcgpu_set_weights(pmc_gpu_handle, c_loc(wtype), c_loc(wmag), &
         c_loc(wexp)), 'pmcgpu_set_weights')
    call pmc_gpu_check(pmcgpu_set_env(pmc_gpu_handle, c_loc(temp), c_loc(pres), c_loc(rh), &
         c_loc(alt), c_loc(cvol)), 'pmcgpu_set_env')
  end subroutine pmc_gpu_upload_aero_states

  !> Device SoA -> aero_states(i,k,j): refresh of the Fortran mirror when a WRF history or
  !! restart alarm needs it (partmc_before_solve_io, wrf_pmc_driver.F90:369-384).
  subroutine pmc_gpu_download_aero_states(aero_states, aero_data, pmc_is, pmc_ie, pmc_js, &
       pmc_je, pmc_ks, pmc_ke)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_aero_particle
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), &
         intent(inout) :: aero_states
    type(aero_data_t), intent(in) :: aero_data

    integer :: i, j, k, c, p, q, n_cells, n_spec, n_class, n_tot, i_class
    integer(c_int), allocatable, target :: counts(:), wgroup(:), wclass(:), nprim(:), hyst(:), wtype(:)
    integer(c_int64_t), allocatable, target :: ids(:)
    real(c_double), allocatable, target :: vol(:,:), lct(:), gct(:), wmag(:), wexp(:)
    type(aero_particle_t) :: particle

    n_cells = size(aero_states)
    n_spec = aero_data_n_spec(aero_data)
    n_class = size(aero_states(pmc_is,pmc_ks,pmc_js)%awa%weight, 2)
    allocate(counts(n_cells))
    call pmc_gpu_check(pmcgpu_get_counts(pmc_gpu_handle, c_loc(counts)), 'pmcgpu_get_counts')
    n_tot = sum(counts)
    allocate(vol(n_tot, n_spec), wgroup(n_tot), wclass(n_tot), ids(n_tot), nprim(n_tot), &
         hyst(n_tot), lct(n_tot), gct(n_tot))
    allocate(wtype(n_cells * n_class), wmag(n_cells * n_class), wexp(n_cells * n_class))
    call pmc_gpu_check(pmcgpu_get_particles(pmc_gpu_handle, int(n_tot, c_long_long), c_loc(vol), &
         c_loc(wgroup), c_loc(wclass), c_loc(ids), c_loc(nprim), c_loc(hyst), c_loc(lct), &
         c_loc(gct)), 'pmcgpu_get_particles')
    call pmc_gpu_check(pmcgpu_get_weights(pmc_gpu_handle, c_loc(wtype), c_loc(wmag), &
         c_loc(wexp)), 'pmcgpu_get_weights')
    c = 0
    q = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       call aero_state_zero(aero_states(i,k,j))        ! keeps awa and n_part_ideal
       do i_class = 1, n_class
          aero_states(i,k,j)%awa%weight(1, i_class)%magnitude = wmag((c - 1) * n_class + i_class)
       end do
       do p = 1, counts(c)
          q = q + 1
          call aero_particle_zero(particle, aero_data)
          particle%vol = vol(q, :)
          particle%weight_group = wgroup(q)
          particle%weight_class = wclass(q)
          particle%id = ids(q)
          particle%n_primary_parts = nprim(q)
          particle%water_hyst_leg = hyst(q)
          particle%least_create_time = lct(q)
          particle%greatest_create_time = gct(q)
          call aero_state_add_particle(aero_states(i,k,j), particle, aero_data, &
               allow_resort=.false.)
       end do
    end do
    end do
    end do
  end subroutine pmc_gpu_download_aero_states

  !> Drop-in for partmc_timestep (interface/wrf_pmc_driver.F90:46-290): same argument list
  !! (gas_data / gas_states / scenario are passed through untouched: gas chemistry, emissions
  !! and MOSAIC are outside this library, SURVEY.md section 8f).
  subroutine pmc_gpu_partmc_timestep(grid, config_flags, env_states, pmc_is, pmc_ie, pmc_js, &
       pmc_je, pmc_ks, pmc_ke, global_nx, global_ny, global_nz)
    use module_domain, only: domain, domain_get_time_since_sim_start
    use module_configure, only: grid_config_rec_type, p_qv
    use module_model_constants, only: t0, p0 => p1000mb, rcp
    use pmc_env_state
    type(domain), intent(inout) :: grid
    type(grid_config_rec_type), intent(in) :: config_flags
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), &
         intent(inout) :: env_states
    integer, intent(in) :: global_nx, global_ny, global_nz

    type(pmcgpu_timestep_args_t) :: args
    integer :: i, j, k, c, n_cells, n_class
    real(c_double), allocatable, target :: temp(:), pres(:), rh(:), alt(:), cvol(:)
    real(c_float), allocatable, target :: num_conc(:)
    real(kind=dp) :: pressure, qsat
    real :: real_time

    n_cells = (pmc_ie - pmc_is + 1) * (pmc_ke - pmc_ks + 1) * (pmc_je - pmc_js + 1)
    n_class = size(grid%u_flux, 4) - 1
    allocate(temp(n_cells), pres(n_cells), rh(n_cells), alt(n_cells), cvol(n_cells))
    allocate(num_conc(n_cells * n_class))
    ! wrf_to_partmc (wrf_pmc_driver.F90:576-644): the host part is the conversion of WRF's
    ! perturbation fields; old_env_states is kept by the library
    c = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       pressure = grid%p(i,k,j) + grid%pb(i,k,j)
       pres(c) = pressure
       temp(c) = (grid%t_2(i,k,j) + t0) * (pressure / p0)**rcp
       qsat = 3.8d0 * exp(17.27 * (temp(c) - 273.0d0) / (temp(c) - 36.0d0)) / (.01 * pressure)
       rh(c) = max(.1d0, min(real(grid%moist(i,k,j,p_qv), dp) / qsat, .95d0))   ! rel_hum, :649-666
       alt(c) = grid%alt(i,k,j)
       cvol(c) = env_states(i,k,j)%cell_volume
    end do
    end do
    end do
    args%temp = c_loc(temp)
    args%pressure = c_loc(pres)
    args%rel_humid = c_null_ptr          ! water uptake carries RH from step to step
    args%inverse_density = c_loc(alt)
    args%cell_volume = c_loc(cvol)
    args%elapsed_time = real(real_time(domain_get_time_since_sim_start(grid)), kind=dp)
    args%dt = real(grid%dt, dp)
    args%chem_dt = real(grid%partmc_chem_dt, dp)
    args%do_coagulation = merge(1, 0, grid%do_coagulation)
    args%coag_kernel_type = PMCGPU_KERNEL_BROWN   ! wrf_pmc_driver.F90:192
    args%do_condensation = merge(1, 0, grid%do_mosaic)
    args%allow_doubling = merge(1, 0, grid%allow_doubling)
    args%allow_halving = merge(1, 0, grid%allow_halving)
    args%step = pmc_gpu_step
    args%num_conc_out = c_loc(num_conc)
    args%rel_humid_out = c_null_ptr
    args%no_wait = 0
    pmc_gpu_step = pmc_gpu_step + 1
    call pmc_gpu_check(pmcgpu_partmc_timestep(pmc_gpu_handle, args), 'pmcgpu_partmc_timestep')
    ! partmc_to_wrf (wrf_pmc_driver.F90:1704-1719)
    c = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       grid%chem(i,k,j,p_NUM_CONC_a01:p_NUM_CONC_a01 + n_class - 1) = &
            num_conc(c * n_class + 1:(c + 1) * n_class)
       c = c + 1
    end do
    end do
    end do
  end subroutine pmc_gpu_partmc_timestep

  !> Drop-in for wrf_pmc_trans_driver (interface/wrf_pmc_trans.F90:37-200): WRF's per-class
  !! advective fluxes and mixing coefficients of the patch (memory order i, k, j, class) go to
  !! the device; probabilities, sampling, ghost exchange and add + rebalance + sort run there.
  subroutine pmc_gpu_wrf_pmc_trans_driver(grid, config_flags, pmc_is, pmc_ie, pmc_js, pmc_je, &
       pmc_ks, pmc_ke, global_nx, global_ny, global_nz)
    use module_domain, only: domain
    use module_configure, only: grid_config_rec_type
    type(domain), intent(inout) :: grid
    type(grid_config_rec_type), intent(in) :: config_flags
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    integer, intent(in) :: global_nx, global_ny, global_nz

    type(pmcgpu_trans_args_t) :: args
    integer :: n_class
    real(c_float), allocatable, target :: u(:,:,:,:), v(:,:,:,:), w(:,:,:,:)
    real(c_float), allocatable, target :: xkhh(:,:,:), alt(:,:,:), z(:,:,:), zw(:,:,:), kv(:,:,:)

    n_class = size(grid%u_flux, 4) - 1
    ! contiguous copies of the patch sections (the flux arrays carry the class index last; index 1
    ! of that dimension is WRF's dummy scalar, wrf_pmc_trans.F90:229)
    u = grid%u_flux(pmc_is:pmc_ie + 1, pmc_ks:pmc_ke, pmc_js:pmc_je, 2:n_class + 1)
    v = grid%v_flux(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je + 1, 2:n_class + 1)
    w = grid%w_flux(pmc_is:pmc_ie, pmc_ks:pmc_ke + 1, pmc_js:pmc_je, 2:n_class + 1)
    xkhh = grid%xkhh(pmc_is - 1:pmc_ie + 1, pmc_ks:pmc_ke, pmc_js - 1:pmc_je + 1)
    alt = grid%alt(pmc_is - 1:pmc_ie + 1, pmc_ks:pmc_ke, pmc_js - 1:pmc_je + 1)
    z = grid%z(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)
    zw = grid%z_at_w(pmc_is:pmc_ie, pmc_ks:pmc_ke + 1, pmc_js:pmc_je)
    kv = grid%exch_h(pmc_is:pmc_ie, pmc_ks:pmc_ke + 1, pmc_js:pmc_je)
    args%u_flux = c_loc(u)
    args%v_flux = c_loc(v)
    args%w_flux = c_loc(w)
    args%xkhh = c_loc(xkhh)
    args%alt = c_loc(alt)
    args%z = c_loc(z)
    args%z_at_w = c_loc(zw)
    args%exch_h = c_loc(kv)
    args%dx = grid%dx
    args%dy = grid%dy
    args%dt = grid%dt
    args%vertmix_onoff = config_flags%vertmix_onoff
    args%allow_doubling = merge(1, 0, grid%allow_doubling)
    args%allow_halving = merge(1, 0, grid%allow_halving)
    args%step = pmc_gpu_step
    args%no_wait = 0
    pmc_gpu_step = pmc_gpu_step + 1
    call pmc_gpu_check(pmcgpu_wrf_pmc_trans_driver(pmc_gpu_handle, args), &
         'pmcgpu_wrf_pmc_trans_driver')
    ! Do deposition on lowest layer (wrf_pmc_trans.F90:183-197)
    if (pmc_ks == 1 .and. grid%do_deposition) then
       call pmc_gpu_wrf_pmc_dep_dry_driver(grid, pmc_is, pmc_ie, pmc_js, pmc_je)
    end if
  end subroutine pmc_gpu_wrf_pmc_trans_driver

  !> Drop-in for wrf_pmc_dep_dry_driver (interface/wrf_pmc_dep.F90:31): only the 2-D surface fields
  !> cross the boundary; the particles of the lowest layer are already on the device.
  subroutine pmc_gpu_wrf_pmc_dep_dry_driver(grid, pmc_is, pmc_ie, pmc_js, pmc_je)
    use module_domain, only: domain
    type(domain), intent(inout) :: grid
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je

    type(pmcgpu_dep_args_t) :: args
    real(c_float), allocatable, target :: rmol(:,:), znt(:,:), ust(:,:)
    integer(c_int), allocatable, target :: ivgtyp(:,:)
    real(c_double), allocatable, target :: height(:,:)

    rmol = grid%rmol(pmc_is:pmc_ie, pmc_js:pmc_je)
    znt = grid%znt(pmc_is:pmc_ie, pmc_js:pmc_je)
    ust = grid%ust(pmc_is:pmc_ie, pmc_js:pmc_je)
    ivgtyp = grid%ivgtyp(pmc_is:pmc_ie, pmc_js:pmc_je)
    ! env_states(i,1,j)%height (wrf_pmc_driver.F90:627)
    height = real(grid%z_at_w(pmc_is:pmc_ie, 2, pmc_js:pmc_je) &
         - grid%z_at_w(pmc_is:pmc_ie, 1, pmc_js:pmc_je), c_double)
    args%rmol = c_loc(rmol)
    args%znt = c_loc(znt)
    args%ust = c_loc(ust)
    args%ivgtyp = c_loc(ivgtyp)
    args%height = c_loc(height)
    args%dt = real(grid%dt, c_double)
    args%step = pmc_gpu_step
    args%no_wait = 0
    pmc_gpu_step = pmc_gpu_step + 1
    call pmc_gpu_check(pmcgpu_wrf_pmc_dep_dry_driver(pmc_gpu_handle, args), &
         'pmcgpu_wrf_pmc_dep_dry_driver')
  end subroutine pmc_gpu_wrf_pmc_dep_dry_driver

end module pmc_gpu
