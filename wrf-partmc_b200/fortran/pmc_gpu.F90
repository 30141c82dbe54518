! pmc_gpu.F90 -- the host side of the drop-in: Fortran routines with the reference's OWN names and argument
! lists that forward to libpmcgpu.so (include/pmcgpu.h) through the generated ISO_C_BINDING mirror
! pmc_gpu_c.F90.  A WRF-PartMC build compiles pmc_gpu_c.F90 and this file and replaces the `use` of the
! reference modules for these routines by `use pmc_gpu`:
!
!   WRF level (all cells of the patch, the library owns the particles)
!     partmc_timestep       interface/wrf_pmc_driver.F90:46-87       -> pmcgpu_wrf_to_partmc + pmcgpu_partmc_timestep
!     wrf_pmc_trans_driver  interface/wrf_pmc_trans.F90:37-76        -> pmcgpu_wrf_pmc_trans_driver (+ dry deposition)
!     trans_aero            interface/wrf_pmc_trans_aero.F90:40-76   -> (reached through wrf_pmc_trans_driver)
!     wrf_pmc_dep_dry_driver interface/wrf_pmc_dep.F90:31            -> pmcgpu_wrf_pmc_dep_dry_driver
!   PartMC level (ONE aero_state, e.g. run_part.F90:583,599): a one-cell context per call
!     mc_coag               partmc/src/coagulation.F90:43-59         -> pmcgpu_mc_coag
!     condense_particles    partmc/src/condense.F90:147-161          -> pmcgpu_condense_particles
!     aero_state_sort       partmc/src/aero_state.F90:3218           -> pmcgpu_aero_state_sort
!     aero_state_rebalance  partmc/src/aero_state.F90:1765           -> pmcgpu_aero_state_rebalance
!
! Conventions follow the reference's own C boundary (partmc/src/condense.F90:183-195): scalars by value, arrays
! through c_loc of a contiguous target, int return code with 0 = success, a non-zero code goes to die_msg
! (util.F90:134).  The cold part of the record (DESIGN.md section 2) -- aero_particle_t%component(:) source lists and
! the optical fields (absorb/scatter_cross_sect, asymmetry, refract_*, core_vol) -- crosses the boundary only with
! pmc_gpu_track_cold = .true.; otherwise the downloaded mirror has one component per particle and zero optics and the
! upload reports how many particles that concerns.  Gas chemistry, emissions, MOSAIC and optics keep
! running on the host (SURVEY.md section 8f).
!
! NOTE: no Fortran compiler exists in the build image of this repository (gfortran, flang, nvfortran, ifx all
! absent), so this file is written against the Fortran 2003 standard and the generated interfaces but has NOT
! been compiled; tests/test_abi.py checks the struct layouts and symbol coverage of pmc_gpu_c.F90, and the
! tests drive the same C symbols through wrf-partmc_b200/host/pmcgpu.py.  INTEGRATION.md lists the build line.
module pmc_gpu

  use, intrinsic :: iso_c_binding
  use pmc_gpu_c
  implicit none
  private

  public :: pmc_gpu_init, pmc_gpu_finalize, pmc_gpu_upload_aero_states, &
       pmc_gpu_download_aero_states, pmc_gpu_check, pmc_gpu_handle
  ! the reference's entry points
  public :: partmc_timestep, wrf_pmc_trans_driver, trans_aero, wrf_pmc_dep_dry_driver
  public :: mc_coag, condense_particles, aero_state_sort, aero_state_rebalance

  integer, parameter :: dp = c_double

  !> coag_kernel.F90:24-40
  integer(c_int), parameter, public :: PMCGPU_KERNEL_ADDITIVE = 2, &
       PMCGPU_KERNEL_CONSTANT = 3, PMCGPU_KERNEL_BROWN = 4, PMCGPU_KERNEL_ZERO = 5

  !> Opaque device context of this MPI rank (one GPU per rank).
  type(c_ptr), save :: pmc_gpu_handle = c_null_ptr
  !> Counter that addresses the random draws: two values per WRF step.
  integer(c_int), save :: pmc_gpu_step = 0
  !> The Fortran aero_states array is a mirror: .true. once the device holds the particles.
  logical, save :: pmc_gpu_resident = .false.
  !> Carry aero_particle_t%component(:) and the optical fields on the device (pmcgpu_track_components): set it
  !! before pmc_gpu_upload_aero_states.  n_swbands doubles each for absorb / scatter_cross_sect and asymmetry,
  !! 2 n_swbands each for refract_shell / refract_core, one for core_vol.
  logical, save, public :: pmc_gpu_track_cold = .false.
  integer, save :: pmc_gpu_n_swbands = 0

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
    ! aero_data%fractal (fractal.F90:55-66): spherical in WRF-PartMC; a do_fractal run keeps coagulation with the
    ! brown_free / brown_cont kernels and is refused by the calls that need the mobility radius (include/pmcgpu.h)
    call pmc_gpu_check(pmcgpu_set_fractal(pmc_gpu_handle, aero_data%fractal%frac_dim, &
         aero_data%fractal%prime_radius, aero_data%fractal%vol_fill_factor), 'pmcgpu_set_fractal')
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
    use pmc_aero_particle, only: aero_particle_n_components
    use pmc_env_state
    use pmc_util, only: warn_assert_msg, integer_to_string
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), &
         intent(in) :: aero_states
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), &
         intent(in) :: env_states

    integer :: i, j, k, c, p, n_cells, n_spec, n_class, n_tot, q, i_class
    integer :: n_lost_comp, n_lost_optics, n_opaque, n_comp_p, b
    integer(c_int), allocatable, target :: n_comp(:), comp_src(:,:)
    real(c_double), allocatable, target :: comp_time(:,:), opaque(:,:)
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
    if (pmc_gpu_track_cold) then
       pmc_gpu_n_swbands = 0
       if (n_tot > 0) then
          do j = pmc_js, pmc_je
          do k = pmc_ks, pmc_ke
          do i = pmc_is, pmc_ie
             if (aero_state_n_part(aero_states(i,k,j)) > 0 .and. pmc_gpu_n_swbands == 0) then
                if (allocated(aero_states(i,k,j)%apa%particle(1)%absorb_cross_sect)) &
                     pmc_gpu_n_swbands = size(aero_states(i,k,j)%apa%particle(1)%absorb_cross_sect)
             end if
          end do
          end do
          end do
       end if
       n_opaque = 7 * pmc_gpu_n_swbands + 1
       allocate(n_comp(n_tot), comp_src(20, n_tot), comp_time(20, n_tot), opaque(n_opaque, n_tot))
       comp_src = 0
       comp_time = 0d0
       opaque = 0d0
       call pmc_gpu_check(pmcgpu_track_components(pmc_gpu_handle, int(n_opaque, c_int)), 'pmcgpu_track_components')
    end if
    c = 0
    q = 0
    n_lost_comp = 0
    n_lost_optics = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       do p = 1, counts(c)
          q = q + 1
          ! what the device record does not carry (DESIGN.md section 2): count it, report it below
          if (aero_particle_n_components(aero_states(i,k,j)%apa%particle(p)) > 1) &
               n_lost_comp = n_lost_comp + 1
          if (allocated(aero_states(i,k,j)%apa%particle(p)%absorb_cross_sect)) then
             if (any(aero_states(i,k,j)%apa%particle(p)%absorb_cross_sect /= 0d0) &
                  .or. any(aero_states(i,k,j)%apa%particle(p)%scatter_cross_sect /= 0d0)) &
                  n_lost_optics = n_lost_optics + 1
          end if
          if (pmc_gpu_track_cold) then
             n_comp_p = max(0, aero_particle_n_components(aero_states(i,k,j)%apa%particle(p)))
             n_comp(q) = n_comp_p
             do b = 1, n_comp_p
                comp_src(b, q) = aero_states(i,k,j)%apa%particle(p)%component(b)%source_id
                comp_time(b, q) = aero_states(i,k,j)%apa%particle(p)%component(b)%create_time
             end do
             call pmc_gpu_pack_optics(aero_states(i,k,j)%apa%particle(p), pmc_gpu_n_swbands, opaque(:, q))
          end if
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
    ! non-fatal, the reference's convention for recoverable conditions (util.F90 warn_assert_msg): the device
    ! state keeps least / greatest_create_time and n_primary_parts, not the per-component source lists and optics
    if (pmc_gpu_track_cold) then
       n_lost_comp = 0
       n_lost_optics = 0
    end if
    call warn_assert_msg(731204655, n_lost_comp == 0, &
         'pmc_gpu_upload_aero_states: ' // trim(integer_to_string(n_lost_comp)) &
         // ' particles carry more than one aero_component; the device record keeps one (DESIGN.md section 2)')
    call warn_assert_msg(731204656, n_lost_optics == 0, &
         'pmc_gpu_upload_aero_states: ' // trim(integer_to_string(n_lost_optics)) &
         // ' particles carry optical properties; they are recomputed on the host after a download')
    call pmc_gpu_check(pmcgpu_set_particles(pmc_gpu_handle, c_loc(counts), c_loc(vol), &
         c_loc(wgroup), c_loc(wclass), c_loc(ids), c_loc(nprim), c_loc(hyst), c_loc(lct), &
         c_loc(gct)), 'pmcgpu_set_particles')
    ! the C side is [n_total][20] row-major = (20, n_total) column-major
    if (pmc_gpu_track_cold) call pmc_gpu_check(pmcgpu_set_components(pmc_gpu_handle, int(n_tot, c_long_long), &
         c_loc(n_comp), c_loc(comp_src), c_loc(comp_time), c_loc(opaque)), 'pmcgpu_set_components')
    call pmc_gpu_check(pmcgpu_set_weights(pmc_gpu_handle, c_loc(wtype), c_loc(wmag), &
         c_loc(wexp)), 'pmcgpu_set_weights')
    call pmc_gpu_check(pmcgpu_set_env(pmc_gpu_handle, c_loc(temp), c_loc(pres), c_loc(rh), &
         c_loc(alt), c_loc(cvol)), 'pmcgpu_set_env')
  end subroutine pmc_gpu_upload_aero_states

  !> The optical fields of a particle as the opaque blob of the device record and back (aero_particle.F90:38-49).
  subroutine pmc_gpu_pack_optics(particle, n_sw, blob)
    use pmc_aero_particle
    type(aero_particle_t), intent(in) :: particle
    integer, intent(in) :: n_sw
    real(c_double), intent(out) :: blob(:)
    blob = 0d0
    if (n_sw > 0 .and. allocated(particle%absorb_cross_sect)) then
       blob(1:n_sw) = particle%absorb_cross_sect
       blob(n_sw+1:2*n_sw) = particle%scatter_cross_sect
       blob(2*n_sw+1:3*n_sw) = particle%asymmetry
       blob(3*n_sw+1:4*n_sw) = real(particle%refract_shell, c_double)
       blob(4*n_sw+1:5*n_sw) = aimag(particle%refract_shell)
       blob(5*n_sw+1:6*n_sw) = real(particle%refract_core, c_double)
       blob(6*n_sw+1:7*n_sw) = aimag(particle%refract_core)
    end if
    blob(7*n_sw+1) = particle%core_vol
  end subroutine pmc_gpu_pack_optics

  subroutine pmc_gpu_unpack_optics(blob, n_sw, particle)
    use pmc_aero_particle
    real(c_double), intent(in) :: blob(:)
    integer, intent(in) :: n_sw
    type(aero_particle_t), intent(inout) :: particle
    if (n_sw > 0) then
       particle%absorb_cross_sect = blob(1:n_sw)
       particle%scatter_cross_sect = blob(n_sw+1:2*n_sw)
       particle%asymmetry = blob(2*n_sw+1:3*n_sw)
       particle%refract_shell = cmplx(blob(3*n_sw+1:4*n_sw), blob(4*n_sw+1:5*n_sw), kind=c_double)
       particle%refract_core = cmplx(blob(5*n_sw+1:6*n_sw), blob(6*n_sw+1:7*n_sw), kind=c_double)
    end if
    particle%core_vol = blob(7*n_sw+1)
  end subroutine pmc_gpu_unpack_optics

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
    integer :: n_opaque, b
    integer(c_int), allocatable, target :: n_comp(:), comp_src(:,:)
    real(c_double), allocatable, target :: comp_time(:,:), opaque(:,:)

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
    if (pmc_gpu_track_cold) then
       n_opaque = 7 * pmc_gpu_n_swbands + 1
       allocate(n_comp(n_tot), comp_src(20, n_tot), comp_time(20, n_tot), opaque(n_opaque, n_tot))
       call pmc_gpu_check(pmcgpu_get_components(pmc_gpu_handle, int(n_tot, c_long_long), c_loc(n_comp), &
            c_loc(comp_src), c_loc(comp_time), c_loc(opaque)), 'pmcgpu_get_components')
    end if
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
          if (pmc_gpu_track_cold) then
             if (allocated(particle%component)) deallocate(particle%component)
             allocate(particle%component(n_comp(q)))
             do b = 1, n_comp(q)
                particle%component(b)%source_id = comp_src(b, q)
                particle%component(b)%create_time = comp_time(b, q)
             end do
             call pmc_gpu_unpack_optics(opaque(:, q), pmc_gpu_n_swbands, particle)
          else
             ! one component: the weight class as source (one class per source in WRF), least_create_time
             call aero_particle_set_component(particle, wclass(q), lct(q))
          end if
          call aero_state_add_particle(aero_states(i,k,j), particle, aero_data, &
               allow_resort=.false.)
       end do
    end do
    end do
    end do
  end subroutine pmc_gpu_download_aero_states

  !> partmc_timestep (interface/wrf_pmc_driver.F90:46-87), same name and argument list.  scenario, gas_data
  !! and gas_states are passed through untouched: gas chemistry, emissions and MOSAIC are outside this
  !! library (SURVEY.md section 8f) and keep their host code.
  subroutine partmc_timestep(grid, config_flags, scenario, env_states, aero_data, aero_states, gas_data, &
       gas_states, pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke, global_nx, global_ny, global_nz)
    use module_domain, only: domain, domain_get_time_since_sim_start
    use module_configure, only: grid_config_rec_type, p_qv
    use pmc_env_state
    use pmc_aero_data
    use pmc_aero_state
    use pmc_gas_data
    use pmc_gas_state
    use pmc_scenario
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(domain), intent(inout) :: grid
    type(grid_config_rec_type), intent(in) :: config_flags
    type(scenario_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: scenario
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: env_states
    type(aero_data_t), intent(in) :: aero_data
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: aero_states
    type(gas_data_t), intent(in) :: gas_data
    type(gas_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: gas_states
    integer, intent(in) :: global_nx, global_ny, global_nz

    type(pmcgpu_timestep_args_t) :: args
    type(pmcgpu_wrf_fields_t) :: wf
    integer :: i, j, k, c, n_cells, n_class
    real(c_float), allocatable, target :: p(:,:,:), pb(:,:,:), t_2(:,:,:), qv(:,:,:), alt(:,:,:), z(:,:,:)
    real(c_float), allocatable, target :: zw(:,:,:), msftx(:,:), msfty(:,:)
    real(c_double), allocatable, target :: temp(:), pres(:), rh(:), rrho(:), cvol(:), height(:)
    real(c_float), allocatable, target :: num_conc(:)
    real :: real_time

    n_cells = (pmc_ie - pmc_is + 1) * (pmc_ke - pmc_ks + 1) * (pmc_je - pmc_js + 1)
    n_class = size(aero_states(pmc_is,pmc_ks,pmc_js)%awa%weight, 2)
    if (.not. pmc_gpu_resident) then   ! first call: the particles move to the device and stay there
       call pmc_gpu_upload_aero_states(aero_states, env_states, pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke)
       pmc_gpu_resident = .true.
    end if
    ! wrf_to_partmc (wrf_pmc_driver.F90:576-644) on the device: contiguous copies of WRF's patch sections in
    p = grid%p(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)
    pb = grid%pb(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)
    t_2 = grid%t_2(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)
    qv = grid%moist(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je, p_qv)
    alt = grid%alt(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)
    z = grid%z(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)
    zw = grid%z_at_w(pmc_is:pmc_ie, pmc_ks:pmc_ke + 1, pmc_js:pmc_je)
    msftx = grid%msftx(pmc_is:pmc_ie, pmc_js:pmc_je)
    msfty = grid%msfty(pmc_is:pmc_ie, pmc_js:pmc_je)
    wf%p = c_loc(p)
    wf%pb = c_loc(pb)
    wf%t_2 = c_loc(t_2)
    wf%qv = c_loc(qv)
    wf%alt = c_loc(alt)
    wf%z = c_loc(z)
    wf%z_at_w = c_loc(zw)
    wf%msftx = c_loc(msftx)
    wf%msfty = c_loc(msfty)
    wf%rdx = grid%rdx
    wf%rdy = grid%rdy
    call pmc_gpu_check(pmcgpu_wrf_to_partmc(pmc_gpu_handle, wf), 'pmcgpu_wrf_to_partmc')
    ! the host mirror of env_states (gas chemistry and emissions on the host read it)
    allocate(temp(n_cells), pres(n_cells), rh(n_cells), rrho(n_cells), cvol(n_cells), height(n_cells))
    call pmc_gpu_check(pmcgpu_get_env_staging(pmc_gpu_handle, c_loc(temp), c_loc(pres), c_loc(rh), &
         c_loc(rrho), c_loc(cvol), c_loc(height)), 'pmcgpu_get_env_staging')
    c = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       env_states(i,k,j)%temp = temp(c)
       env_states(i,k,j)%pressure = pres(c)
       env_states(i,k,j)%rel_humid = rh(c)
       env_states(i,k,j)%inverse_density = rrho(c)
       env_states(i,k,j)%cell_volume = cvol(c)
       env_states(i,k,j)%height = height(c)
       env_states(i,k,j)%altitude = grid%z(i,k,j)
       env_states(i,k,j)%z_min = grid%z_at_w(i,k,j)
       env_states(i,k,j)%z_max = grid%z_at_w(i,k+1,j)
       env_states(i,k,j)%elapsed_time = real(real_time(domain_get_time_since_sim_start(grid)), kind=dp)
    end do
    end do
    end do
    allocate(num_conc(n_cells * n_class))
    ! NULL met pointers: the env_state staged by pmcgpu_wrf_to_partmc is installed, the previous one is kept as
    ! old_env_states (wrf_pmc_driver.F90:615)
    args%temp = c_null_ptr
    args%pressure = c_null_ptr
    args%rel_humid = c_null_ptr
    args%inverse_density = c_null_ptr
    args%cell_volume = c_null_ptr
    args%elapsed_time = real(real_time(domain_get_time_since_sim_start(grid)), kind=dp)
    args%dt = real(grid%dt, dp)
    args%chem_dt = real(grid%partmc_chem_dt, dp)
    args%do_coagulation = merge(1, 0, grid%do_coagulation)
    args%coag_kernel_type = PMCGPU_KERNEL_BROWN   ! wrf_pmc_driver.F90:192
    args%do_condensation = merge(1, 0, grid%do_mosaic)
    args%allow_doubling = 1                       ! aero_state_rebalance(..., .true., .true., ...), :248
    args%allow_halving = 1
    args%step = pmc_gpu_step
    args%num_conc_out = c_loc(num_conc)
    args%rel_humid_out = c_loc(rh)
    args%no_wait = 0
    pmc_gpu_step = pmc_gpu_step + 1
    call pmc_gpu_check(pmcgpu_partmc_timestep(pmc_gpu_handle, args), 'pmcgpu_partmc_timestep')
    ! partmc_to_wrf (wrf_pmc_driver.F90:1704-1719), and the relative humidity the water uptake leaves behind
    c = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       grid%chem(i,k,j,p_NUM_CONC_a01:p_NUM_CONC_a01 + n_class - 1) = &
            num_conc(c * n_class + 1:(c + 1) * n_class)
       c = c + 1
       env_states(i,k,j)%rel_humid = rh(c)
    end do
    end do
    end do
  end subroutine partmc_timestep

  !> wrf_pmc_trans_driver (interface/wrf_pmc_trans.F90:37-76), same name and argument list: WRF's per-class
  !! advective fluxes and mixing coefficients of the patch (memory order i, k, j, class) go to
  !! the device; probabilities, sampling, ghost exchange and add + rebalance + sort run there.
  subroutine wrf_pmc_trans_driver(grid, config_flags, env_states, scenario, aero_data, aero_states, &
       gas_states, pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke, global_nx, global_ny, global_nz)
    use module_domain, only: domain
    use module_configure, only: grid_config_rec_type
    use pmc_env_state
    use pmc_aero_data
    use pmc_aero_state
    use pmc_gas_state
    use pmc_scenario
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(domain), intent(inout) :: grid
    type(grid_config_rec_type), intent(in) :: config_flags
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: env_states
    type(scenario_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: scenario
    type(aero_data_t), intent(in) :: aero_data
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: aero_states
    type(gas_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: gas_states
    integer, intent(in) :: global_nx, global_ny, global_nz

    type(pmcgpu_trans_args_t) :: args
    integer :: n_class
    real(c_float), allocatable, target :: u(:,:,:,:), v(:,:,:,:), w(:,:,:,:)
    real(c_float), allocatable, target :: xkhh(:,:,:), alt(:,:,:), z(:,:,:), zw(:,:,:), kv(:,:,:)

    n_class = size(grid%u_flux, 4) - 1
    if (.not. (config_flags%periodic_x .and. config_flags%periodic_y)) then
       call pmc_gpu_boundary_from_wrf(grid, scenario, env_states, pmc_is, pmc_ie, pmc_js, pmc_je, &
            pmc_ks, pmc_ke, global_nx, global_ny)
    end if
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
       call pmc_gpu_dep_dry(grid, pmc_is, pmc_ie, pmc_js, pmc_je)
    end if
  end subroutine wrf_pmc_trans_driver

  !> Drop-in for wrf_pmc_dep_dry_driver (interface/wrf_pmc_dep.F90:31): only the 2-D surface fields
  !> cross the boundary; the particles of the lowest layer are already on the device.
  subroutine pmc_gpu_dep_dry(grid, pmc_is, pmc_ie, pmc_js, pmc_je)
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
  end subroutine pmc_gpu_dep_dry

  !> The sea salt block of partmc_timestep (interface/wrf_pmc_driver.F90:212-223 -> seasalt_emissions, :1767):
  !> the surface fields cross the boundary once per step; wind speed, sampling proportion and the sampling of the
  !> new Na / Cl particles run on the device for all water cells of the lowest level.
  subroutine pmc_gpu_seasalt_emissions(grid, aero_data, env_states, pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke)
    use module_domain, only: domain
    use pmc_env_state
    use pmc_aero_data
    use pmc_util, only: string_array_find
    type(domain), intent(inout) :: grid
    type(aero_data_t), intent(in) :: aero_data
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(env_state_t), intent(in) :: env_states(pmc_is:pmc_ie, pmc_ks:pmc_ke, pmc_js:pmc_je)

    type(pmcgpu_seasalt_args_t) :: args
    real(c_float), allocatable, target :: u10(:,:), v10(:,:), xland(:,:)
    real(c_double), allocatable, target :: height(:,:)
    integer(c_int), target :: weight_class(5)
    character(len=AERO_SOURCE_NAME_LEN) :: mode_name
    integer :: i_mode, n_modes

    if (grid%partmc_seasalt_param <= 0) return
    n_modes = merge(2, 5, grid%partmc_seasalt_param == 1)
    do i_mode = 1,n_modes
       write(mode_name,'(a,i2.2)') 'sea_salt_', i_mode
       weight_class(i_mode) = string_array_find(aero_data%weight_class_name, mode_name)
    end do
    u10 = grid%u10(pmc_is:pmc_ie, pmc_js:pmc_je)
    v10 = grid%v10(pmc_is:pmc_ie, pmc_js:pmc_je)
    xland = grid%xland(pmc_is:pmc_ie, pmc_js:pmc_je)
    height = env_states(pmc_is:pmc_ie, pmc_ks, pmc_js:pmc_je)%height
    args%u10 = c_loc(u10)
    args%v10 = c_loc(v10)
    args%xland = c_loc(xland)
    args%height = c_loc(height)
    args%sea_salt_opt = grid%partmc_seasalt_param
    args%i_na = aero_data_spec_by_name(aero_data, "Na")
    args%i_cl = aero_data_spec_by_name(aero_data, "Cl")
    args%weight_class = c_loc(weight_class)
    args%dt = real(grid%dt, c_double)
    args%elapsed_time = env_states(pmc_is, pmc_ks, pmc_js)%elapsed_time
    args%allow_doubling = merge(1, 0, grid%allow_doubling)
    args%allow_halving = merge(1, 0, grid%allow_halving)
    args%step = pmc_gpu_step
    pmc_gpu_step = pmc_gpu_step + 1
    call pmc_gpu_check(pmcgpu_seasalt_emissions(pmc_gpu_handle, args, c_null_ptr), 'pmcgpu_seasalt_emissions')
  end subroutine pmc_gpu_seasalt_emissions

  !> Open domain: the boundary conditions trans_aero derives per step (wrf_pmc_trans_aero.F90:811-955) --
  !! which boundary cells see outflow (u_2 < 0 on the west face, u_2(i+1) > 0 on the east face, v_2 likewise),
  !! which are thrown away (the wind changed sign since the last step, or the background changed), and the
  !! background aerosol interpolated to the current time (aero_dist_interp_1d) -- handed to pmcgpu_set_boundary.
  subroutine pmc_gpu_boundary_from_wrf(grid, scenario, env_states, pmc_is, pmc_ie, pmc_js, pmc_je, &
       pmc_ks, pmc_ke, global_nx, global_ny)
    use module_domain, only: domain
    use pmc_env_state
    use pmc_scenario
    use pmc_aero_dist
    use pmc_aero_mode
    use pmc_util, only: find_1d
    type(domain), intent(in) :: grid
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke, global_nx, global_ny
    type(scenario_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(in) :: scenario
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(in) :: env_states

    type(pmcgpu_boundary_args_t) :: b
    type(aero_dist_t) :: background
    real(kind=dp) :: dilution_rate, u2, u1
    integer :: i, j, k, c, m, n_cells, n_mode, n_spec, p_t, p_tm1
    logical :: edge, change_bc
    integer(c_int), allocatable, target :: mode_type(:), weight_class(:), outflow(:), throw_away(:)
    real(c_double), allocatable, target :: char_radius(:), log10_std(:), vol_frac(:,:), vol_frac_std(:,:), &
         num_conc(:,:)

    n_cells = (pmc_ie - pmc_is + 1) * (pmc_ke - pmc_ks + 1) * (pmc_je - pmc_js + 1)
    call aero_dist_interp_1d(scenario(pmc_is,pmc_ks,pmc_js)%aero_background, &
         scenario(pmc_is,pmc_ks,pmc_js)%aero_dilution_time, scenario(pmc_is,pmc_ks,pmc_js)%aero_dilution_rate, &
         env_states(pmc_is,pmc_ks,pmc_js)%elapsed_time, background, dilution_rate)
    n_mode = aero_dist_n_mode(background)
    n_spec = size(background%mode(1)%vol_frac)
    allocate(mode_type(n_mode), weight_class(n_mode), char_radius(n_mode), log10_std(n_mode), &
         vol_frac(n_spec, n_mode), vol_frac_std(n_spec, n_mode), num_conc(n_mode, n_cells), &
         outflow(n_cells), throw_away(n_cells))
    do m = 1, n_mode
       mode_type(m) = background%mode(m)%type          ! aero_mode.F90:26-34 (1 log-normal, 2 exp, 3 mono)
       weight_class(m) = background%mode(m)%weight_class
       char_radius(m) = background%mode(m)%char_radius
       log10_std(m) = background%mode(m)%log10_std_dev_radius
       vol_frac(:, m) = background%mode(m)%vol_frac
       vol_frac_std(:, m) = background%mode(m)%vol_frac_std
    end do
    outflow = 0
    throw_away = 0
    num_conc = 0d0
    c = 0
    do j = pmc_js, pmc_je
    do k = pmc_ks, pmc_ke
    do i = pmc_is, pmc_ie
       c = c + 1
       edge = .true.
       if (i == 1 .and. j > 1 .and. j < global_ny) then
          u2 = grid%u_2(i,k,j); u1 = grid%u_1(i,k,j)
          if (u2 < 0.0d0) outflow(c) = 1
       else if (i == global_nx .and. j > 1 .and. j < global_ny) then
          u2 = grid%u_2(i+1,k,j); u1 = grid%u_1(i+1,k,j)
          if (u2 > 0.0d0) outflow(c) = 1
       else if (j == 1 .and. i > 1 .and. i < global_nx) then
          u2 = grid%v_2(i,k,j); u1 = grid%v_1(i,k,j)
          if (u2 < 0.0d0) outflow(c) = 1
       else if (j == global_ny .and. i > 1 .and. i < global_nx) then
          u2 = grid%v_2(i,k,j+1); u1 = grid%v_1(i,k,j+1)
          if (u2 > 0.0d0) outflow(c) = 1
       else
          edge = .false.
       end if
       if (.not. edge) cycle
       ! change_boundary_conditions (:816-826) and the wind reversal test (:836)
       p_tm1 = find_1d(size(scenario(i,k,j)%aero_background), scenario(i,k,j)%aero_dilution_time, &
            env_states(i,k,j)%elapsed_time - real(grid%dt, dp))
       p_t = find_1d(size(scenario(i,k,j)%aero_background), scenario(i,k,j)%aero_dilution_time, &
            env_states(i,k,j)%elapsed_time)
       change_bc = (p_t > p_tm1 .and. p_t > 1)
       if (u2 * u1 < 0.0d0 .or. change_bc) throw_away(c) = 1
       call aero_dist_interp_1d(scenario(i,k,j)%aero_background, scenario(i,k,j)%aero_dilution_time, &
            scenario(i,k,j)%aero_dilution_rate, env_states(i,k,j)%elapsed_time, background, dilution_rate)
       do m = 1, n_mode
          num_conc(m, c) = aero_mode_total_num_conc(background%mode(m))
       end do
    end do
    end do
    end do
    b%n_mode = n_mode
    b%mode_type = c_loc(mode_type)
    b%weight_class = c_loc(weight_class)
    b%char_radius = c_loc(char_radius)
    b%log10_std_dev = c_loc(log10_std)
    b%vol_frac = c_loc(vol_frac)
    b%vol_frac_std = c_loc(vol_frac_std)
    b%num_conc = c_loc(num_conc)
    b%outflow = c_loc(outflow)
    b%throw_away = c_loc(throw_away)
    call pmc_gpu_check(pmcgpu_set_boundary(pmc_gpu_handle, b), 'pmcgpu_set_boundary')
  end subroutine pmc_gpu_boundary_from_wrf

  !> wrf_pmc_dep_dry_driver (interface/wrf_pmc_dep.F90:31-60), same name and argument list.
  subroutine wrf_pmc_dep_dry_driver(grid, config_flags, aero_data, aero_states, gas_states, env_states, &
       pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke)
    use module_domain, only: domain
    use module_configure, only: grid_config_rec_type
    use pmc_env_state
    use pmc_aero_data
    use pmc_aero_state
    use pmc_gas_state
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(domain), intent(inout) :: grid
    type(grid_config_rec_type), intent(in) :: config_flags
    type(aero_data_t), intent(in) :: aero_data
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: aero_states
    type(gas_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: gas_states
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: env_states
    call pmc_gpu_dep_dry(grid, pmc_is, pmc_ie, pmc_js, pmc_je)
  end subroutine wrf_pmc_dep_dry_driver

  !> trans_aero (interface/wrf_pmc_trans_aero.F90:40-76), same name and argument list (note k before j).  The
  !! move probabilities it reads from module state in the reference (wrf_pmc_trans.F90:140-176) are computed
  !! inside pmcgpu_wrf_pmc_trans_driver from the fields of `grid`, so this entry forwards to the driver.
  subroutine trans_aero(grid, config_flags, aero_data, scenario, env_states, aero_states, pmc_is, pmc_ie, &
       pmc_ks, pmc_ke, pmc_js, pmc_je, global_nx, global_ny, global_nz)
    use module_domain, only: domain
    use module_configure, only: grid_config_rec_type
    use pmc_env_state
    use pmc_aero_data
    use pmc_aero_state
    use pmc_gas_state
    use pmc_scenario
    integer, intent(in) :: pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke
    type(domain), intent(inout) :: grid
    type(grid_config_rec_type), intent(in) :: config_flags
    type(aero_data_t), intent(in) :: aero_data
    type(scenario_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: scenario
    type(env_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: env_states
    type(aero_state_t), dimension(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je), intent(inout) :: aero_states
    integer, intent(in) :: global_nx, global_ny, global_nz
    type(gas_state_t), allocatable :: no_gas(:,:,:)
    allocate(no_gas(pmc_is:pmc_ie,pmc_ks:pmc_ke,pmc_js:pmc_je))
    call wrf_pmc_trans_driver(grid, config_flags, env_states, scenario, aero_data, aero_states, no_gas, &
         pmc_is, pmc_ie, pmc_js, pmc_je, pmc_ks, pmc_ke, global_nx, global_ny, global_nz)
  end subroutine trans_aero

  ! ------------------------------------------------------------------------------------------------
  ! PartMC-level entry points: one aero_state_t in, the same aero_state_t out.  Each call builds a one-cell
  ! context (pmcgpu_create with nx = nz = ny = 1), uploads the state, runs the device routine for that cell
  ! and copies the result back: the form run_part.F90 (:583 mc_coag, :599 condense_particles) needs.  A WRF
  ! build never comes through here (its particles stay resident behind pmc_gpu_handle).
  ! ------------------------------------------------------------------------------------------------

  !> aero_state_t + env_state_t -> a fresh one-cell device context
  subroutine box_open(aero_state, aero_data, env_state, box)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_env_state
    type(aero_state_t), intent(in) :: aero_state
    type(aero_data_t), intent(in), target :: aero_data
    type(env_state_t), intent(in) :: env_state
    type(c_ptr), intent(out) :: box

    type(pmcgpu_config_t) :: cfg
    integer :: p, n, n_spec, n_group, n_class, g, c
    real(c_double), allocatable, target :: density(:), kappa(:), molec_weight(:), vol(:,:), lct(:), gct(:)
    real(c_double), allocatable, target :: wmag(:), wexp(:), ideal(:), env1(:)
    integer(c_int), allocatable, target :: num_ions(:), counts(:), wgroup(:), wclass(:), nprim(:), hyst(:), wtype(:)
    integer(c_int64_t), allocatable, target :: ids(:)

    n = aero_state_n_part(aero_state)
    n_spec = aero_data_n_spec(aero_data)
    n_group = size(aero_state%awa%weight, 1)
    n_class = size(aero_state%awa%weight, 2)
    density = aero_data%density
    kappa = aero_data%kappa
    molec_weight = aero_data%molec_weight
    num_ions = aero_data%num_ions
    cfg%n_spec = n_spec
    cfg%density = c_loc(density)
    cfg%kappa = c_loc(kappa)
    cfg%molec_weight = c_loc(molec_weight)
    cfg%num_ions = c_loc(num_ions)
    cfg%i_water = aero_data%i_water
    cfg%n_group = n_group
    cfg%n_class = n_class
    cfg%nx = 1
    cfg%nz = 1
    cfg%ny = 1
    cfg%global_nx = 1
    cfg%global_ny = 1
    cfg%i_offset = 0
    cfg%j_offset = 0
    cfg%periodic_x = 1
    cfg%periodic_y = 1
    cfg%n_part_ideal = sum(aero_state%n_part_ideal)
    cfg%cell_capacity = max(4 * n, int(2.5d0 * cfg%n_part_ideal)) + 64
    cfg%max_bins = 255
    cfg%seed = 1
    cfg%device = 0
    call pmc_gpu_check(pmcgpu_create(cfg, box), 'pmcgpu_create')
    call pmc_gpu_check(pmcgpu_set_fractal(box, aero_data%fractal%frac_dim, aero_data%fractal%prime_radius, &
         aero_data%fractal%vol_fill_factor), 'pmcgpu_set_fractal')
    allocate(counts(1), vol(max(n, 1), n_spec), wgroup(max(n, 1)), wclass(max(n, 1)), ids(max(n, 1)), &
         nprim(max(n, 1)), hyst(max(n, 1)), lct(max(n, 1)), gct(max(n, 1)))
    counts(1) = n
    do p = 1, n
       vol(p, :) = aero_state%apa%particle(p)%vol
       wgroup(p) = aero_state%apa%particle(p)%weight_group
       wclass(p) = aero_state%apa%particle(p)%weight_class
       ids(p) = aero_state%apa%particle(p)%id
       nprim(p) = aero_state%apa%particle(p)%n_primary_parts
       hyst(p) = aero_state%apa%particle(p)%water_hyst_leg
       lct(p) = aero_state%apa%particle(p)%least_create_time
       gct(p) = aero_state%apa%particle(p)%greatest_create_time
    end do
    call pmc_gpu_check(pmcgpu_set_particles(box, c_loc(counts), c_loc(vol), c_loc(wgroup), c_loc(wclass), &
         c_loc(ids), c_loc(nprim), c_loc(hyst), c_loc(lct), c_loc(gct)), 'pmcgpu_set_particles')
    allocate(wtype(n_group * n_class), wmag(n_group * n_class), wexp(n_group * n_class), ideal(n_group * n_class))
    do c = 1, n_class
       do g = 1, n_group      ! [class][group], group fastest (include/pmcgpu.h)
          wtype((c - 1) * n_group + g) = aero_state%awa%weight(g, c)%type
          wmag((c - 1) * n_group + g) = aero_state%awa%weight(g, c)%magnitude
          wexp((c - 1) * n_group + g) = aero_state%awa%weight(g, c)%exponent
          ideal((c - 1) * n_group + g) = aero_state%n_part_ideal(g, c)
       end do
    end do
    call pmc_gpu_check(pmcgpu_set_weights(box, c_loc(wtype), c_loc(wmag), c_loc(wexp)), 'pmcgpu_set_weights')
    call pmc_gpu_check(pmcgpu_set_n_part_ideal(box, c_loc(ideal)), 'pmcgpu_set_n_part_ideal')
    allocate(env1(5))
    env1 = (/ env_state%temp, env_state%pressure, env_state%rel_humid, env_state%inverse_density, &
         env_state%cell_volume /)
    call pmc_gpu_check(pmcgpu_set_env(box, c_loc(env1(1:1)), c_loc(env1(2:2)), c_loc(env1(3:3)), &
         c_loc(env1(4:4)), c_loc(env1(5:5))), 'pmcgpu_set_env')
  end subroutine box_open

  !> device cell -> aero_state_t (particles and weight magnitudes), then the context is destroyed
  subroutine box_close(box, aero_state, aero_data)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_aero_particle
    type(c_ptr), intent(inout) :: box
    type(aero_state_t), intent(inout) :: aero_state
    type(aero_data_t), intent(in) :: aero_data

    integer :: p, n, n_spec, n_group, n_class, g, c
    integer(c_int), allocatable, target :: counts(:), wgroup(:), wclass(:), nprim(:), hyst(:), wtype(:)
    integer(c_int64_t), allocatable, target :: ids(:)
    real(c_double), allocatable, target :: vol(:,:), lct(:), gct(:), wmag(:), wexp(:)
    type(aero_particle_t) :: particle

    n_spec = aero_data_n_spec(aero_data)
    n_group = size(aero_state%awa%weight, 1)
    n_class = size(aero_state%awa%weight, 2)
    allocate(counts(1))
    call pmc_gpu_check(pmcgpu_get_counts(box, c_loc(counts)), 'pmcgpu_get_counts')
    n = counts(1)
    allocate(vol(max(n, 1), n_spec), wgroup(max(n, 1)), wclass(max(n, 1)), ids(max(n, 1)), nprim(max(n, 1)), &
         hyst(max(n, 1)), lct(max(n, 1)), gct(max(n, 1)))
    allocate(wtype(n_group * n_class), wmag(n_group * n_class), wexp(n_group * n_class))
    call pmc_gpu_check(pmcgpu_get_particles(box, int(n, c_long_long), c_loc(vol), c_loc(wgroup), c_loc(wclass), &
         c_loc(ids), c_loc(nprim), c_loc(hyst), c_loc(lct), c_loc(gct)), 'pmcgpu_get_particles')
    call pmc_gpu_check(pmcgpu_get_weights(box, c_loc(wtype), c_loc(wmag), c_loc(wexp)), 'pmcgpu_get_weights')
    call aero_state_zero(aero_state)        ! keeps awa and n_part_ideal
    do c = 1, n_class
       do g = 1, n_group
          aero_state%awa%weight(g, c)%magnitude = wmag((c - 1) * n_group + g)
       end do
    end do
    do p = 1, n
       call aero_particle_zero(particle, aero_data)
       particle%vol = vol(p, :)
       particle%weight_group = wgroup(p)
       particle%weight_class = wclass(p)
       particle%id = ids(p)
       particle%n_primary_parts = nprim(p)
       particle%water_hyst_leg = hyst(p)
       particle%least_create_time = lct(p)
       particle%greatest_create_time = gct(p)
       call aero_state_add_particle(aero_state, particle, aero_data, allow_resort=.false.)
    end do
    call pmc_gpu_check(pmcgpu_destroy(box), 'pmcgpu_destroy')
    box = c_null_ptr
  end subroutine box_close

  !> mc_coag (partmc/src/coagulation.F90:43-59), same name and argument list.
  subroutine mc_coag(coag_kernel_type, env_state, aero_data, aero_state, del_t, tot_n_samp, tot_n_coag)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_env_state
    integer, intent(in) :: coag_kernel_type
    type(env_state_t), intent(in) :: env_state
    type(aero_data_t), intent(in) :: aero_data
    type(aero_state_t), intent(inout) :: aero_state
    real(kind=dp) :: del_t
    integer, intent(out) :: tot_n_samp
    integer, intent(out) :: tot_n_coag
    type(c_ptr) :: box
    integer(c_int), target :: ns(1), nc(1)
    call box_open(aero_state, aero_data, env_state, box)
    call pmc_gpu_check(pmcgpu_mc_coag(box, int(coag_kernel_type, c_int), del_t, pmc_gpu_step, c_loc(ns), &
         c_loc(nc)), 'pmcgpu_mc_coag')
    pmc_gpu_step = pmc_gpu_step + 1
    tot_n_samp = ns(1)
    tot_n_coag = nc(1)
    call box_close(box, aero_state, aero_data)
    aero_state%valid_sort = .false.
  end subroutine mc_coag

  !> condense_particles (partmc/src/condense.F90:147-161), same name and argument list;
  !! env_state_final%rel_humid is overwritten with the integrator's value (:263).
  subroutine condense_particles(aero_state, aero_data, env_state_initial, env_state_final, del_t)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_env_state
    type(aero_state_t), intent(inout) :: aero_state
    type(aero_data_t), intent(in) :: aero_data
    type(env_state_t), intent(in) :: env_state_initial
    type(env_state_t), intent(inout) :: env_state_final
    real(kind=dp), intent(in) :: del_t
    type(c_ptr) :: box
    real(c_double), target :: t1(1), p1(1), rh(1)
    call box_open(aero_state, aero_data, env_state_initial, box)
    t1(1) = env_state_final%temp
    p1(1) = env_state_final%pressure
    call pmc_gpu_check(pmcgpu_set_rng_step(box, pmc_gpu_step), 'pmcgpu_set_rng_step')
    pmc_gpu_step = pmc_gpu_step + 1
    call pmc_gpu_check(pmcgpu_condense_particles(box, c_loc(t1), c_loc(p1), del_t, 0d0, c_loc(rh), &
         c_null_ptr, c_null_ptr), 'pmcgpu_condense_particles')
    env_state_final%rel_humid = rh(1)
    call box_close(box, aero_state, aero_data)
    aero_state%valid_sort = .false.     ! condense.F90:759
  end subroutine condense_particles

  !> aero_state_sort (partmc/src/aero_state.F90:3218), same name and argument list.  The device keeps its own
  !! (class, bin)-ordered layout; on return the particle ARRAY is in that order and the host-side
  !! aero_sorted of the mirror is rebuilt by the reference's own routine on next use (valid_sort = .false.).
  !! A caller-supplied bin_grid / all_procs_same is not forwarded: the device always uses the adaptive
  !! per-cell grid of aero_sorted_remake_if_needed (:214-373).
  subroutine aero_state_sort(aero_state, aero_data, bin_grid, all_procs_same)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_bin_grid
    use pmc_env_state
    type(aero_state_t), intent(inout) :: aero_state
    type(aero_data_t), intent(in) :: aero_data
    type(bin_grid_t), optional, intent(in) :: bin_grid
    logical, optional, intent(in) :: all_procs_same
    type(c_ptr) :: box
    type(env_state_t) :: env_state
    env_state%temp = 288d0
    env_state%pressure = 1d5
    env_state%rel_humid = 0.5d0
    env_state%inverse_density = 1d0
    env_state%cell_volume = 1d0
    call box_open(aero_state, aero_data, env_state, box)
    call pmc_gpu_check(pmcgpu_aero_state_sort(box), 'pmcgpu_aero_state_sort')
    call box_close(box, aero_state, aero_data)
    aero_state%valid_sort = .false.
  end subroutine aero_state_sort

  !> aero_state_rebalance (partmc/src/aero_state.F90:1765), same name and argument list.
  subroutine aero_state_rebalance(aero_state, aero_data, allow_doubling, allow_halving, initial_state_warning)
    use pmc_aero_state
    use pmc_aero_data
    use pmc_env_state
    type(aero_state_t), intent(inout) :: aero_state
    type(aero_data_t), intent(in) :: aero_data
    logical, intent(in) :: allow_doubling
    logical, intent(in) :: allow_halving
    logical, intent(in) :: initial_state_warning
    type(c_ptr) :: box
    type(env_state_t) :: env_state
    env_state%temp = 288d0
    env_state%pressure = 1d5
    env_state%rel_humid = 0.5d0
    env_state%inverse_density = 1d0
    env_state%cell_volume = 1d0
    call box_open(aero_state, aero_data, env_state, box)
    call pmc_gpu_check(pmcgpu_aero_state_rebalance(box, merge(1, 0, allow_doubling), merge(1, 0, allow_halving), &
         pmc_gpu_step), 'pmcgpu_aero_state_rebalance')
    pmc_gpu_step = pmc_gpu_step + 1
    call box_close(box, aero_state, aero_data)
    aero_state%valid_sort = .false.
  end subroutine aero_state_rebalance

end module pmc_gpu
