! thetal_b200_f2c_mod -- ISO_C_BINDING interfaces of the entry points libthetal_b200.so adds to the
! HOMMEXX C interface (include/thetal_b200.h).  The entry points HOMMEXX already has
! (init_simulation_params_c, init_elements_2d_c, prim_run_subcycle_c, ... ) keep the interface blocks of
! components/homme/src/theta-l_kokkos/theta_f2c_mod.F90:10-236 and
! components/homme/src/share/cxx/prim_cxx_driver_base.F90:46-196 unchanged: same names, same arguments.
!
! Shipped as source only: neither the build container nor the GPU box has a Fortran compiler, so this
! file has never been compiled here (INTEGRATION.md section 1).  Time levels are 0-based in every
! thetal_b200_* / *_c routine below except where noted.
module thetal_b200_f2c_mod
  use iso_c_binding, only: c_int, c_double, c_char, c_ptr, c_long
  implicit none
  interface

    ! ---- session set-up that HOMMEXX gets from MPI and from compile-time constants ----------------
    subroutine thetal_b200_set_nlev(nlev) bind(c)
      import :: c_int
      integer(c_int), intent(in) :: nlev
    end subroutine thetal_b200_set_nlev

    subroutine thetal_b200_get_nccl_id(id) bind(c)             ! rank 0 only; broadcast the 128 bytes with MPI
      import :: c_char
      character(kind=c_char), intent(out) :: id(128)
    end subroutine thetal_b200_get_nccl_id

    subroutine thetal_b200_init_comm(rank, nranks, device, id) bind(c)
      import :: c_int, c_char
      integer(c_int), intent(in) :: rank, nranks, device
      character(kind=c_char), intent(in) :: id(128)
    end subroutine thetal_b200_init_comm

    subroutine thetal_b200_set_abort_on_error(flag) bind(c)    ! 1 (default): abort like abortmp; 0: latch
      import :: c_int
      integer(c_int), intent(in) :: flag
    end subroutine thetal_b200_set_abort_on_error

    function thetal_b200_last_error() bind(c) result(code)
      import :: c_int
      integer(c_int) :: code
    end function thetal_b200_last_error

    ! ---- the stage routines of prim_advance_exp, for drivers that keep prim_run_subcycle in Fortran --
    subroutine prim_advance_exp_c(dt, nm1, n0, np1, nstep, compute_diagnostics) bind(c)   ! prim_advance_mod.F90:81
      import :: c_int, c_double
      real(c_double), intent(in) :: dt
      integer(c_int), intent(in) :: nm1, n0, np1, nstep, compute_diagnostics
    end subroutine prim_advance_exp_c

    subroutine caar_c(np1, nm1, n0, dt2, eta_ave_w, scale1, scale2, scale3) bind(c)       ! prim_advance_mod.F90:1023
      import :: c_int, c_double
      integer(c_int), intent(in) :: np1, nm1, n0
      real(c_double), intent(in) :: dt2, eta_ave_w, scale1, scale2, scale3
    end subroutine caar_c

    subroutine dirk_c(nm1, alphadt_nm1, n0, alphadt_n0, np1, dt2, itercount, itererr) bind(c)  ! imex_mod.F90:60
      import :: c_int, c_double
      integer(c_int), intent(in) :: nm1, n0, np1
      real(c_double), intent(in) :: alphadt_nm1, alphadt_n0, dt2
      integer(c_int), intent(out) :: itercount
      real(c_double), intent(out) :: itererr
    end subroutine dirk_c

    subroutine advance_hypervis_c(nt, dt2, eta_ave_w) bind(c)                             ! prim_advance_mod.F90:546
      import :: c_int, c_double
      integer(c_int), intent(in) :: nt
      real(c_double), intent(in) :: dt2, eta_ave_w
    end subroutine advance_hypervis_c

    subroutine vertical_remap_c(np1) bind(c)                                              ! vertremap_mod.F90:23
      import :: c_int
      integer(c_int), intent(in) :: np1
    end subroutine vertical_remap_c

    subroutine thetal_b200_set_vertical_remap(on) bind(c)      ! 0: prim_run_subcycle_c leaves the remap to the caller
      import :: c_int
      integer(c_int), intent(in) :: on
    end subroutine thetal_b200_set_vertical_remap

    ! ---- diagnostics without a state copy (prim_state_mod.F90:822-939 + global_integral) -----------
    subroutine thetal_b200_global_diagnostics(tl, out) bind(c)
      import :: c_int, c_double
      integer(c_int), intent(in) :: tl
      real(c_double), intent(out) :: out(8)    ! min ps, max ps, <ps>, <KE>, <PE>, <IE>, <KE+PE+IE>, nelemd
    end subroutine thetal_b200_global_diagnostics
    ! no counterpart in the reference: same numbers, static device scratch, allocates nothing (<= 16384 elements, 1 rank)
    subroutine thetal_b200_global_diagnostics_noalloc(tl, out) bind(c)
      import :: c_int, c_double
      integer(c_int), intent(in) :: tl
      real(c_double), intent(out) :: out(8)    ! min ps, max ps, <ps>, <KE>, <PE>, <IE>, <KE+PE+IE>, nelemd
    end subroutine thetal_b200_global_diagnostics_noalloc

    subroutine thetal_b200_sync() bind(c)
    end subroutine thetal_b200_sync

    ! ---- plain theta-l build: state inside the array of element structs (element_mod.F90:32-144) ----
    ! call init_elements_states_strided_c(c_loc(elem(1)%state%v), c_loc(elem(1)%state%w_i), ..., &
    !                                     int(storage_size(elem(1))/8, c_long))
    subroutine init_elements_states_strided_c(v, w_i, vtheta_dp, phinh_i, dp3d, elem_stride_bytes) bind(c)
      import :: c_ptr, c_long
      type(c_ptr), intent(in) :: v, w_i, vtheta_dp, phinh_i, dp3d
      integer(c_long), intent(in) :: elem_stride_bytes
    end subroutine init_elements_states_strided_c

    subroutine apply_cam_forcing_dynamics_c(nt, dt) bind(c)                               ! prim_advance_mod.F90:517
      import :: c_int, c_double
      integer(c_int), intent(in) :: nt
      real(c_double), intent(in) :: dt
    end subroutine apply_cam_forcing_dynamics_c

  end interface

contains

  ! What prim_init1 (theta-l_kokkos/prim_driver_mod.F90:60) calls before initialize_hommexx_session().
  subroutine thetal_b200_start(par_rank, par_nprocs, par_comm, gpus_per_node, nlev)
    use mpi
    integer, intent(in) :: par_rank, par_nprocs, par_comm, gpus_per_node, nlev
    character(kind=c_char) :: id(128)
    integer :: ierr
    if (par_rank == 0) call thetal_b200_get_nccl_id(id)
    call MPI_Bcast(id, 128, MPI_CHARACTER, 0, par_comm, ierr)
    call thetal_b200_init_comm(int(par_rank, c_int), int(par_nprocs, c_int), &
                               int(mod(par_rank, gpus_per_node), c_int), id)
    call thetal_b200_set_nlev(int(nlev, c_int))
  end subroutine thetal_b200_start

end module thetal_b200_f2c_mod
