! gmcore_gpu_mod — ISO_C_BINDING interface to libgmcore_b200.so (include/gmcore_gpu.h).
!
! Drop-in for the body of time_integrate (src/dynamics/gmcore_mod.F90:577-591), the
! pre-loop block of gmcore_run (:101-136) and diagnose (:213-341).  The Fortran host keeps
! namelist parsing, mesh/filter/vertical-coordinate set-up, initial conditions, the clock and
! all fiona/netCDF I/O; this module only hands pointers to its own arrays across the C ABI.
!
! NOT compiled in this repository (no Fortran compiler on the build image): shipped as the
! source a GMCORE maintainer adds to src/dynamics/ and lists in CMakeLists.txt, linking
! -lgmcore_b200 -lcudart.  See INTEGRATION.md for the call sites.
module gmcore_gpu_mod

  use, intrinsic :: iso_c_binding
  use const_mod
  use namelist_mod
  use block_mod
  use mesh_mod
  use state_mod
  use static_mod
  use filter_types_mod
  use vert_coord_mod
  use process_mod, only: proc

  implicit none

  private

  public gmcore_gpu_init
  public gmcore_gpu_pin_state
  public gmcore_gpu_upload
  public gmcore_gpu_prepare_run
  public gmcore_gpu_time_integrate
  public gmcore_gpu_diagnose
  public gmcore_gpu_download
  public gmcore_gpu_final

  ! struct gmc_config (include/gmcore_gpu.h); member order and types must match the header.
  type, bind(c) :: gmc_config
    integer(c_int) nlon, nlat, nlev
    integer(c_int) hx, hy
    integer(c_int) lat_ibeg, lat_iend
    integer(c_int) device
    integer(c_int) rank, nranks
    integer(c_int) baroclinic
    integer(c_int) time_scheme
    integer(c_int) ke_scheme
    integer(c_int) pv_scheme
    integer(c_int) upwind_order_pv
    integer(c_int) pv_pole_stokes
    integer(c_int) coriolis_scheme
    integer(c_int) use_div_damp, div_damp_order, div_damp_k0
    integer(c_int) use_smag_damp
    integer(c_int) test_forcing
    integer(c_int) ntracers
    integer(c_int) limiter_type
    integer(c_int) ffsl_flux_type
    integer(c_int) tvd_limiter_type
    real(c_double) ke_cell_wgt, upwind_wgt_pv
    real(c_double) div_damp_coef2, div_damp_coef4, div_damp_top, div_damp_pole, div_damp_lat0
    real(c_double) smag_damp_coef
    real(c_double) dt_dyn, dt_adv
    real(c_double) omega, radius, g, Rd, cpd, p0, Rd_o_cpd
    real(c_double) ptop, hybrid_p0
  end type gmc_config

  ! struct gmc_mesh_vectors: 34 const double* in header order.
  type, bind(c) :: gmc_mesh_vectors
    type(c_ptr) full_lat, half_lat, full_cos_lat, half_cos_lat, full_sin_lat, half_sin_lat
    type(c_ptr) area_cell, area_lon, area_lon_west, area_lon_east, area_lon_north, area_lon_south
    type(c_ptr) area_lat, area_lat_west, area_lat_east, area_lat_north, area_lat_south, area_vtx
    type(c_ptr) area_subcell
    type(c_ptr) de_lon, le_lon, de_lat, le_lat
    type(c_ptr) full_f, half_f
    type(c_ptr) full_tangent_wgt, half_tangent_wgt
    type(c_ptr) half_lon
    type(c_ptr) hyai, hybi, full_lev, half_lev, full_dlev, half_dlev
  end type gmc_mesh_vectors

  interface
    integer(c_int) function gmc_create(cfg, h) bind(c, name='gmc_create')
      import; type(gmc_config), intent(in) :: cfg; type(c_ptr), intent(out) :: h
    end function
    subroutine gmc_destroy(h) bind(c, name='gmc_destroy')
      import; type(c_ptr), value :: h
    end subroutine
    integer(c_int) function gmc_set_mesh(h, mv) bind(c, name='gmc_set_mesh')
      import; type(c_ptr), value :: h; type(gmc_mesh_vectors), intent(in) :: mv
    end function
    integer(c_int) function gmc_set_filters(h, which, ngrid_lon, wgt_lon, ld_lon, ngrid_lat, wgt_lat, ld_lat) &
        bind(c, name='gmc_set_filters')
      import; type(c_ptr), value :: h, ngrid_lon, wgt_lon, ngrid_lat, wgt_lat
      integer(c_int), value :: which, ld_lon, ld_lat
    end function
    integer(c_int) function gmc_set_static(h, gzs) bind(c, name='gmc_set_static')
      import; type(c_ptr), value :: h, gzs
    end function
    integer(c_int) function gmc_upload_state(h, itime, u_lon, v_lat, pt_or_gz, phs) bind(c, name='gmc_upload_state')
      import; type(c_ptr), value :: h, u_lon, v_lat, pt_or_gz, phs; integer(c_int), value :: itime
    end function
    integer(c_int) function gmc_download_state(h, itime, u_lon, v_lat, pt_or_gz, phs) bind(c, name='gmc_download_state')
      import; type(c_ptr), value :: h, u_lon, v_lat, pt_or_gz, phs; integer(c_int), value :: itime
    end function
    integer(c_int) function gmc_set_tracers(h, q) bind(c, name='gmc_set_tracers')
      import; type(c_ptr), value :: h, q
    end function
    integer(c_int) function gmc_get_tracers(h, q) bind(c, name='gmc_get_tracers')
      import; type(c_ptr), value :: h, q
    end function
    integer(c_int) function gmc_prepare_run(h, old_idx) bind(c, name='gmc_prepare_run')
      import; type(c_ptr), value :: h; integer(c_int), value :: old_idx
    end function
    integer(c_int) function gmc_time_integrate(h, old_idx, new_idx, dt, time_step) bind(c, name='gmc_time_integrate')
      import; type(c_ptr), value :: h; integer(c_int), value :: old_idx, new_idx, time_step; real(c_double), value :: dt
    end function
    integer(c_int) function gmc_diagnose(h, itime, out) bind(c, name='gmc_diagnose')
      import; type(c_ptr), value :: h; integer(c_int), value :: itime; real(c_double), intent(out) :: out(8)
    end function
    ! same sums as gmc_diagnose, scratch in the library's static device array: allocates nothing (single-band handles)
    integer(c_int) function gmc_diagnose_static(h, itime, out) bind(c, name='gmc_diagnose_static')
      import; type(c_ptr), value :: h; integer(c_int), value :: itime; real(c_double), intent(out) :: out(8)
    end function
    integer(c_int) function gmc_get_field(h, itime, name, buf) bind(c, name='gmc_get_field')
      import; type(c_ptr), value :: h, buf; integer(c_int), value :: itime; character(kind=c_char), intent(in) :: name(*)
    end function
    integer(c_int) function gmc_interp_pressure_level(h, itime, var, po, logp, fill, y2d) bind(c, name='gmc_interp_pressure_level')
      import; type(c_ptr), value :: h, y2d; integer(c_int), value :: itime, logp; real(c_double), value :: po, fill
      character(kind=c_char), intent(in) :: var(*)
    end function
    integer(c_int) function gmc_nccl_unique_id(id) bind(c, name='gmc_nccl_unique_id')
      import; character(kind=c_char), intent(out) :: id(128)
    end function
    integer(c_int) function gmc_nccl_init(h, id) bind(c, name='gmc_nccl_init')
      import; type(c_ptr), value :: h; character(kind=c_char), intent(in) :: id(128)
    end function
    integer(c_int) function gmc_pin_host(h, ptr, bytes) bind(c, name='gmc_pin_host')
      import; type(c_ptr), value :: h, ptr; integer(c_size_t), value :: bytes
    end function
    integer(c_int) function gmc_unpin_host(h, ptr) bind(c, name='gmc_unpin_host')
      import; type(c_ptr), value :: h, ptr
    end function
    integer(c_int) function gmc_last_error(h, buf, len) bind(c, name='gmc_last_error')
      import; type(c_ptr), value :: h; character(kind=c_char), intent(out) :: buf(*); integer(c_int), value :: len
    end function
  end interface

  type(c_ptr) :: handle = c_null_ptr

contains

  subroutine check(rc, where)
    integer(c_int), intent(in) :: rc
    character(*), intent(in) :: where
    character(kind=c_char) buf(512)
    integer i, n
    character(512) msg
    if (rc == 0) return
    n = gmc_last_error(handle, buf, 512_c_int)
    msg = ''
    do i = 1, 512
      if (buf(i) == c_null_char) exit
      msg(i:i) = buf(i)
    end do
    ! same behaviour as the reference's log_error guards (flogger aborts the run)
    write(*, *) '[Error]: ', trim(where), ': ', trim(msg)
    call process_stop_gpu()
  end subroutine check

  subroutine process_stop_gpu()
    use mpi
    integer ierr
    call MPI_ABORT(proc%comm, 1, ierr)
  end subroutine process_stop_gpu

  ! gmcore_init tail (gmcore_mod.F90:45-95): one handle per block/rank, after
  ! process_create_blocks so that mesh, filters and static exist.
  subroutine gmcore_gpu_init(block, device, ntracers)

    type(block_type), intent(in), target :: block
    integer, intent(in) :: device, ntracers

    type(gmc_config) cfg
    type(gmc_mesh_vectors) mv
    type(mesh_type), pointer :: mesh
    character(kind=c_char) id(128)
    integer ierr

    mesh => block%mesh
    cfg%nlon = global_mesh%num_full_lon; cfg%nlat = global_mesh%num_full_lat; cfg%nlev = global_mesh%num_full_lev
    cfg%hx = mesh%lon_halo_width; cfg%hy = mesh%lat_halo_width
    cfg%lat_ibeg = mesh%full_lat_ibeg; cfg%lat_iend = mesh%full_lat_iend
    cfg%device = device
    cfg%rank = proc%id; cfg%nranks = proc%np          ! 1-D latitude decomposition (process_mod.F90:122-125)
    cfg%baroclinic = merge(1, 0, baroclinic)
    select case (time_scheme)
    case ('euler');  cfg%time_scheme = 0
    case ('pc2');    cfg%time_scheme = 1
    case ('wrfrk3'); cfg%time_scheme = 2
    end select
    cfg%ke_scheme = ke_scheme
    select case (pv_scheme)
    case ('midpoint'); cfg%pv_scheme = 0
    case ('upwind');   cfg%pv_scheme = 1
    case ('tvd');      cfg%pv_scheme = 2
    case default;      cfg%pv_scheme = -1          ! 'ffsl': rejected by gmc_create
    end select
    cfg%upwind_order_pv = upwind_order_pv
    cfg%pv_pole_stokes = merge(1, 0, pv_pole_stokes)
    cfg%coriolis_scheme = coriolis_scheme
    cfg%use_div_damp = merge(1, 0, use_div_damp); cfg%div_damp_order = div_damp_order; cfg%div_damp_k0 = div_damp_k0
    cfg%use_smag_damp = merge(1, 0, use_smag_damp)
    cfg%test_forcing = merge(1, 0, test_case == 'held_suarez')
    cfg%ntracers = ntracers
    select case (limiter_type)
    case ('none'); cfg%limiter_type = 0
    case ('mono'); cfg%limiter_type = 1
    case ('pd');   cfg%limiter_type = 2
    end select
    cfg%ffsl_flux_type = merge(1, 0, ffsl_flux_type == 'van_leer')
    select case (tvd_limiter_type)
    case ('none');     cfg%tvd_limiter_type = 0
    case ('van_leer'); cfg%tvd_limiter_type = 1
    case ('mc');       cfg%tvd_limiter_type = 2
    end select
    cfg%ke_cell_wgt = ke_cell_wgt; cfg%upwind_wgt_pv = upwind_wgt_pv
    cfg%div_damp_coef2 = div_damp_coef2; cfg%div_damp_coef4 = div_damp_coef4
    cfg%div_damp_top = div_damp_top; cfg%div_damp_pole = div_damp_pole; cfg%div_damp_lat0 = div_damp_lat0
    cfg%smag_damp_coef = smag_damp_coef
    cfg%dt_dyn = dt_dyn; cfg%dt_adv = dt_adv
    cfg%omega = omega; cfg%radius = radius; cfg%g = g; cfg%Rd = Rd; cfg%cpd = cpd; cfg%p0 = p0; cfg%Rd_o_cpd = Rd_o_cpd
    cfg%ptop = ptop; cfg%hybrid_p0 = hybrid_coord_p0
    call check(gmc_create(cfg, handle), 'gmc_create')

    ! mesh_type vectors: GLOBAL vectors with the reference's bounds (the library slices its band)
    mv%full_lat = c_loc(global_mesh%full_lat); mv%half_lat = c_loc(global_mesh%half_lat)
    mv%full_cos_lat = c_loc(global_mesh%full_cos_lat); mv%half_cos_lat = c_loc(global_mesh%half_cos_lat)
    mv%full_sin_lat = c_loc(global_mesh%full_sin_lat); mv%half_sin_lat = c_loc(global_mesh%half_sin_lat)
    mv%area_cell = c_loc(global_mesh%area_cell); mv%area_lon = c_loc(global_mesh%area_lon)
    mv%area_lon_west = c_loc(global_mesh%area_lon_west); mv%area_lon_east = c_loc(global_mesh%area_lon_east)
    mv%area_lon_north = c_loc(global_mesh%area_lon_north); mv%area_lon_south = c_loc(global_mesh%area_lon_south)
    mv%area_lat = c_loc(global_mesh%area_lat); mv%area_lat_west = c_loc(global_mesh%area_lat_west)
    mv%area_lat_east = c_loc(global_mesh%area_lat_east); mv%area_lat_north = c_loc(global_mesh%area_lat_north)
    mv%area_lat_south = c_loc(global_mesh%area_lat_south); mv%area_vtx = c_loc(global_mesh%area_vtx)
    mv%area_subcell = c_loc(global_mesh%area_subcell)
    mv%de_lon = c_loc(global_mesh%de_lon); mv%le_lon = c_loc(global_mesh%le_lon)
    mv%de_lat = c_loc(global_mesh%de_lat); mv%le_lat = c_loc(global_mesh%le_lat)
    mv%full_f = c_loc(global_mesh%full_f); mv%half_f = c_loc(global_mesh%half_f)
    mv%full_tangent_wgt = c_loc(global_mesh%full_tangent_wgt); mv%half_tangent_wgt = c_loc(global_mesh%half_tangent_wgt)
    mv%half_lon = c_loc(global_mesh%half_lon)
    if (baroclinic) then
      mv%hyai = c_loc(hyai); mv%hybi = c_loc(hybi)
      mv%full_lev = c_loc(global_mesh%full_lev); mv%half_lev = c_loc(global_mesh%half_lev)
      mv%full_dlev = c_loc(global_mesh%full_dlev); mv%half_dlev = c_loc(global_mesh%half_dlev)
    else
      mv%hyai = c_null_ptr; mv%hybi = c_null_ptr; mv%full_lev = c_null_ptr; mv%half_lev = c_null_ptr
      mv%full_dlev = c_null_ptr; mv%half_dlev = c_null_ptr
    end if
    call check(gmc_set_mesh(handle, mv), 'gmc_set_mesh')

    ! The library takes GLOBAL per-latitude filter rows (a band recomputes derived fields on rows of its neighbours),
    ! while block%big_filter etc. (block_mod.F90:62-64) hold the block's own rows.  The filters are pure functions of
    ! the mesh and dt_dyn (filter_types_mod.F90:49-147), so every rank builds them once more on global_mesh -- which
    ! exists on every rank (mesh_mod.F90:130) and already carries the final lon_halo_width.
    block
      type(filter_type) gf
      call gf%init(global_mesh, 'big_filter'   ); call set_filter(0, gf)
      call gf%init(global_mesh, 'small_filter1'); call set_filter(1, gf)
      call gf%init(global_mesh, 'small_filter2'); call set_filter(2, gf)
      call gf%clear()
    end block
    call check(gmc_set_static(handle, c_loc(block%static%gzs)), 'gmc_set_static')

    if (proc%np > 1) then
      ! NCCL bootstrap: rank 0 makes the id, MPI broadcasts its 128 bytes
      if (proc%id == 0) call check(gmc_nccl_unique_id(id), 'gmc_nccl_unique_id')
      call mpi_bcast_id(id)
      call check(gmc_nccl_init(handle, id), 'gmc_nccl_init')
    end if

  contains

    subroutine set_filter(which, f)
      integer, intent(in) :: which
      type(filter_type), intent(in), target :: f
      call check(gmc_set_filters(handle, int(which, c_int), c_loc(f%ngrid_lon), c_loc(f%wgt_lon), &
                 int(size(f%wgt_lon, 1), c_int), c_loc(f%ngrid_lat), c_loc(f%wgt_lat), int(size(f%wgt_lat, 1), c_int)), &
                 'gmc_set_filters')
    end subroutine set_filter

    subroutine mpi_bcast_id(id)
      use mpi
      character(kind=c_char), intent(inout) :: id(128)
      call MPI_BCAST(id, 128, MPI_BYTE, 0, proc%comm, ierr)
    end subroutine mpi_bcast_id

  end subroutine gmcore_gpu_init

  ! once after gmcore_gpu_init: page-lock the prognostic arrays that travel in gmcore_gpu_upload / _download
  ! (5x the transfer rate of pageable memory; optional)
  subroutine gmcore_gpu_pin_state(block, itime)
    type(block_type), intent(inout), target :: block
    integer, intent(in) :: itime
    call pin(c_loc(block%state(itime)%u_lon), size(block%state(itime)%u_lon))
    call pin(c_loc(block%state(itime)%v_lat), size(block%state(itime)%v_lat))
    if (baroclinic) then
      call pin(c_loc(block%state(itime)%pt), size(block%state(itime)%pt))
      call pin(c_loc(block%state(itime)%phs), size(block%state(itime)%phs))
    else
      call pin(c_loc(block%state(itime)%gz), size(block%state(itime)%gz))
    end if
  contains
    subroutine pin(p, n)
      type(c_ptr), intent(in) :: p
      integer, intent(in) :: n
      call check(gmc_pin_host(handle, p, int(n, c_size_t) * 8_c_size_t), 'gmc_pin_host')
    end subroutine pin
  end subroutine gmcore_gpu_pin_state

  ! after set_ic / restart_read: state(itime) -> device
  subroutine gmcore_gpu_upload(block, itime)
    type(block_type), intent(in), target :: block
    integer, intent(in) :: itime
    if (baroclinic) then
      call check(gmc_upload_state(handle, int(itime, c_int), c_loc(block%state(itime)%u_lon), c_loc(block%state(itime)%v_lat), &
                 c_loc(block%state(itime)%pt), c_loc(block%state(itime)%phs)), 'gmc_upload_state')
      if (allocated(block%adv_batches)) then
        if (size(block%adv_batches) > 0) call check(gmc_set_tracers(handle, c_loc(block%adv_batches(1)%q)), 'gmc_set_tracers')
      end if
    else
      call check(gmc_upload_state(handle, int(itime, c_int), c_loc(block%state(itime)%u_lon), c_loc(block%state(itime)%v_lat), &
                 c_loc(block%state(itime)%gz), c_null_ptr), 'gmc_upload_state')
    end if
  end subroutine gmcore_gpu_upload

  ! replaces gmcore_mod.F90:101-136
  subroutine gmcore_gpu_prepare_run(old)
    integer, intent(in) :: old
    call check(gmc_prepare_run(handle, int(old, c_int)), 'gmc_prepare_run')
  end subroutine gmcore_gpu_prepare_run

  ! replaces the body of time_integrate (gmcore_mod.F90:577-591)
  subroutine gmcore_gpu_time_integrate(old, new, dt, time_step)
    integer, intent(in) :: old, new, time_step
    real(8), intent(in) :: dt
    call check(gmc_time_integrate(handle, int(old, c_int), int(new, c_int), real(dt, c_double), int(time_step, c_int)), &
               'gmc_time_integrate')
  end subroutine gmcore_gpu_time_integrate

  ! replaces the sums of diagnose (gmcore_mod.F90:213-341); already globally reduced
  subroutine gmcore_gpu_diagnose(itime, tm, te, tav, tpe, te_ke, te_ie, te_pe, tpt)
    integer, intent(in) :: itime
    real(8), intent(out) :: tm, te, tav, tpe, te_ke, te_ie, te_pe, tpt
    real(c_double) out(8)
    call check(gmc_diagnose(handle, int(itime, c_int), out), 'gmc_diagnose')
    tm = out(1); te = out(2); tav = out(3); tpe = out(4); te_ke = out(5); te_ie = out(6); te_pe = out(7); tpt = out(8)
  end subroutine gmcore_gpu_diagnose

  ! before history_write / restart_write: device -> state(itime)
  subroutine gmcore_gpu_download(block, itime)
    type(block_type), intent(inout), target :: block
    integer, intent(in) :: itime
    if (baroclinic) then
      call check(gmc_download_state(handle, int(itime, c_int), c_loc(block%state(itime)%u_lon), c_loc(block%state(itime)%v_lat), &
                 c_loc(block%state(itime)%pt), c_loc(block%state(itime)%phs)), 'gmc_download_state')
      ! h0 diagnostics (history_mod.F90:519-597)
      call check(gmc_get_field(handle, int(itime, c_int), 't' // c_null_char, c_loc(block%state(itime)%t)), 'gmc_get_field t')
      call check(gmc_get_field(handle, int(itime, c_int), 'div' // c_null_char, c_loc(block%state(itime)%div)), 'gmc_get_field div')
      ! z and ph of the h0 set are not stored on the device: the library re-forms them on demand (gz = g z)
      call check(gmc_get_field(handle, int(itime, c_int), 'gz' // c_null_char, c_loc(block%state(itime)%gz)), 'gmc_get_field gz')
      call check(gmc_get_field(handle, int(itime, c_int), 'ph' // c_null_char, c_loc(block%state(itime)%ph)), 'gmc_get_field ph')
      call check(gmc_get_field(handle, int(itime, c_int), 'gz_lev' // c_null_char, c_loc(block%state(itime)%gz_lev)), 'gmc_get_field gz_lev')
      ! vor: history writes state%vor, the copy left by the last forward pass (one stage old, SURVEY T7); the device
      ! keeps pv of that pass instead, from which the host can form vor = pv * m_vtx - half_f if it writes 'vor'
      call check(gmc_get_field(handle, int(itime, c_int), 'pv' // c_null_char, c_loc(block%state(itime)%pv)), 'gmc_get_field pv')
    else
      call check(gmc_download_state(handle, int(itime, c_int), c_loc(block%state(itime)%u_lon), c_loc(block%state(itime)%v_lat), &
                 c_loc(block%state(itime)%gz), c_null_ptr), 'gmc_download_state')
      call check(gmc_get_field(handle, int(itime, c_int), 'pv' // c_null_char, c_loc(block%state(itime)%pv)), 'gmc_get_field pv')
    end if
  end subroutine gmcore_gpu_download

  ! diag_state_run, pressure_levels branch (utils/diag_state_mod.F90:241-258): one variable ('u', 'v', 'pt', 't', 'z')
  ! on one pressure surface into the caller's 2-D buffer (diag_state%buf), interpolated on the device
  subroutine gmcore_gpu_interp_pressure_level(itime, var, po, buf, logp)
    integer, intent(in) :: itime
    character(*), intent(in) :: var
    real(r8), intent(in) :: po
    real(r8), intent(inout), target :: buf(:,:)
    logical, intent(in), optional :: logp
    integer(c_int) lp
    lp = 0; if (present(logp)) lp = merge(1, 0, logp)
    call check(gmc_interp_pressure_level(handle, int(itime, c_int), trim(var) // c_null_char, real(po, c_double), lp, &
               real(buf(lbound(buf,1),lbound(buf,2)), c_double), c_loc(buf)), 'gmc_interp_pressure_level')
  end subroutine gmcore_gpu_interp_pressure_level

  subroutine gmcore_gpu_final()
    if (c_associated(handle)) call gmc_destroy(handle)
    handle = c_null_ptr
  end subroutine gmcore_gpu_final

end module gmcore_gpu_mod
