!> dartray_gpu.f90 -- ISO_C_BINDING interface of the B200 ray-tracing engine (include/dartray_engine.h).
!!
!! UNVERIFIED: the image this repository is built in has no Fortran compiler, MPI or HDF5, so this file and
!! rt_wrappers.f90 have never been compiled.  The same entry points, with the same argument shapes (column-major
!! arrays, an array of c_ptr for the ragged scaspe_arr(k)%a, dre_params by reference), are exercised from a plain C
!! host in tests/abi_host.c and from Python in tests/.
!!
!! Link: h5fc ... -L<repo>/dart-ray_b200 -ldartray_engine  (NCCL is dlopen'ed when nranks > 1).
module dartray_gpu
  use iso_c_binding
  implicit none

  ! rt_type ids are the reference's own (rt_routines.f90:193-231).
  ! dre_download / dre_upload / dre_zero array ids (include/dartray_engine.h)
  integer(c_int), parameter :: DRE_ARR_U_FEST = 0, DRE_ARR_U_FINAL = 1, DRE_ARR_SCASPE = 2, DRE_ARR_SCASPE_PREV = 3, &
       DRE_ARR_SCASPE_TOT = 4, DRE_ARR_LUM_LOST = 5, DRE_ARR_PSEL_AV = 6, DRE_ARR_LUMCELL = 7, DRE_ARR_I_OBS = 8, &
       DRE_ARR_I_OBS_IN = 9

  type, bind(C) :: dre_params
     integer(c_int32_t) :: bm_par, bm_par_max, pabs_max, print_psel_av
     real(c_double)     :: rad_lim, en_lim, cs, modelsize
     integer(c_int32_t) :: wall_on(3), reserved0
     real(c_double)     :: wall_coord(6)
  end type dre_params

  type, bind(C) :: dre_counters
     integer(c_int64_t) :: sources, rays, intersections, deposits, beam_tests, splits, merges, &
                           sca_elems, active_lambda, steps
     real(c_double)     :: kernel_ms, pass_ms
     integer(c_int64_t) :: isect_hash, active_lambda_dep, descent_loads, launches, trace_launches
  end type dre_counters

  type(c_ptr) :: eng = c_null_ptr      !< one engine per MPI rank = one GPU

  interface
     ! ---- lifetime ---------------------------------------------------------------------------------------------
     integer(c_int) function dre_get_unique_id(id) bind(C, name='dre_get_unique_id')
       import; character(kind=c_char) :: id(128)
     end function
     integer(c_int) function dre_create(e, device, rank, nranks, id) bind(C, name='dre_create')
       import; type(c_ptr) :: e; integer(c_int), value :: device, rank, nranks; type(c_ptr), value :: id
     end function
     subroutine dre_destroy(e) bind(C, name='dre_destroy')
       import; type(c_ptr), value :: e
     end subroutine
     function dre_last_error(e) bind(C, name='dre_last_error')
       import; type(c_ptr), value :: e; type(c_ptr) :: dre_last_error
     end function
     ! ---- static inputs ----------------------------------------------------------------------------------------
     integer(c_int) function dre_set_grid(e, n, max_lvl, base, cchild, lvl, cindex, ccoord, csize, &
                                          csize_arr, carea_arr, cvol_arr) bind(C, name='dre_set_grid')
       import; type(c_ptr), value :: e; integer(c_int32_t), value :: n, max_lvl
       integer(c_int32_t) :: base(2), cchild(*), lvl(*); integer(c_int64_t) :: cindex(*)
       real(c_double) :: ccoord(3,*), csize(*), csize_arr(*), carea_arr(*), cvol_arr(*)
     end function
     integer(c_int) function dre_set_opacity(e, lnum, dens_arr, ksca, gsca, kp_sca, ndir, dir_i_out) &
          bind(C, name='dre_set_opacity')
       import; type(c_ptr), value :: e; integer(c_int32_t), value :: lnum, ndir
       real(c_double) :: dens_arr(lnum,*), ksca(*), gsca(*), dir_i_out(ndir,2); integer(c_int32_t) :: kp_sca(*)
     end function
     integer(c_int) function dre_set_sources(e, lumcell, s, ccoord_p_src, lum_p_src, src_cell, cell_src) &
          bind(C, name='dre_set_sources')
       import; type(c_ptr), value :: e, lumcell, ccoord_p_src, lum_p_src, src_cell, cell_src
       integer(c_int32_t), value :: s
     end function
     integer(c_int) function dre_set_params(e, p) bind(C, name='dre_set_params')
       import; type(c_ptr), value :: e; type(dre_params) :: p
     end function
     integer(c_int) function dre_set_option(e, name, value) bind(C, name='dre_set_option')
       import; type(c_ptr), value :: e; character(kind=c_char) :: name(*); real(c_double), value :: value
     end function
     ! ---- device-resident mirrors of u_fest_arr, u_final_arr, scaspe_arr, ... ------------------------------------
     ! planes: c_loc(array) for the (lnum, tot_ncell) arrays, c_loc(array of c_ptr) for the scaspe arrays
     integer(c_int) function dre_upload(e, which, planes) bind(C, name='dre_upload')
       import; type(c_ptr), value :: e, planes; integer(c_int), value :: which
     end function
     integer(c_int) function dre_download(e, which, planes) bind(C, name='dre_download')
       import; type(c_ptr), value :: e, planes; integer(c_int), value :: which
     end function
     integer(c_int) function dre_zero(e, which) bind(C, name='dre_zero')
       import; type(c_ptr), value :: e; integer(c_int), value :: which
     end function
     ! ---- passes -----------------------------------------------------------------------------------------------
     integer(c_int) function dre_run_pass_resident(e, rt_type, iterations, cnt) bind(C, name='dre_run_pass_resident')
       import; type(c_ptr), value :: e; integer(c_int), value :: rt_type, iterations; type(dre_counters) :: cnt
     end function
     integer(c_int) function dre_run_pass_2d(e, rt_type, iterations, cnt) bind(C, name='dre_run_pass_2d')
       import; type(c_ptr), value :: e; integer(c_int), value :: rt_type, iterations; type(dre_counters) :: cnt
     end function
     integer(c_int) function dre_dust_emission(e, lnum_stars, lnum_dust, kabs_stars, kabs_dust, lambda_si_dust, dl_stars, dl_dust, &
          kext_ref, dens_ref, rf_stars, rf_dust, iterations_dustem, dens_stars, t_dust, to_lumcell) bind(C, name='dre_dust_emission')
       import; type(c_ptr), value :: e, rf_dust, t_dust; integer(c_int32_t), value :: lnum_stars, lnum_dust, iterations_dustem, to_lumcell
       real(c_double) :: kabs_stars(*), kabs_dust(*), lambda_si_dust(*), dl_stars(*), dl_dust(*), dens_ref(*), rf_stars(lnum_stars,*), &
                         dens_stars(lnum_dust,*)
       real(c_double), value :: kext_ref
     end function
     integer(c_int) function dre_run_pass(e, rt_type, iterations, u_fest, u_accum, scaspe, scaspe_prev, &
                                          lum_lost, psel_av, cnt) bind(C, name='dre_run_pass')
       import; type(c_ptr), value :: e, u_fest, u_accum, scaspe, scaspe_prev, lum_lost, psel_av
       integer(c_int), value :: rt_type, iterations; type(dre_counters) :: cnt
     end function
     integer(c_int) function dre_run_iobs(e, rt_type, scaspe_tot, ndir_in, ccoord_obs, i_obs, i_obs_in, cnt) &
          bind(C, name='dre_run_iobs')
       import; type(c_ptr), value :: e, scaspe_tot, ccoord_obs, i_obs, i_obs_in
       integer(c_int), value :: rt_type; integer(c_int32_t), value :: ndir_in; type(dre_counters) :: cnt
     end function
     integer(c_int) function dre_allreduce(e, rt_type) bind(C, name='dre_allreduce')
       import; type(c_ptr), value :: e; integer(c_int), value :: rt_type
     end function
     integer(c_int) function dre_prepare_scatt(e, iterations, conv_en_lim, tot_rad_en, cnflag) &
          bind(C, name='dre_prepare_scatt')
       import; type(c_ptr), value :: e; integer(c_int), value :: iterations; real(c_double), value :: conv_en_lim
       real(c_double) :: tot_rad_en(*); integer(c_int32_t) :: cnflag
     end function
     ! ---- products of i_obs ------------------------------------------------------------------------------------
     integer(c_int) function dre_calc_sed(e, dist_obs, units, lambda_si, sed) bind(C, name='dre_calc_sed')
       import; type(c_ptr), value :: e, lambda_si, sed; real(c_double), value :: dist_obs; integer(c_int), value :: units
     end function
     integer(c_int) function dre_make_maps(e, npixel_maps, map_size_factor, units, lambda_si, maps) &
          bind(C, name='dre_make_maps')
       import; type(c_ptr), value :: e, lambda_si, maps; integer(c_int), value :: npixel_maps, units
       real(c_double), value :: map_size_factor
     end function
     integer(c_int) function dre_make_maps_in(e, ndir_in, ccoord_obs, kp_maps, units, lambda_si, maps_in) &
          bind(C, name='dre_make_maps_in')
       import; type(c_ptr), value :: e, ccoord_obs, lambda_si, maps_in; integer(c_int32_t), value :: ndir_in
       integer(c_int), value :: kp_maps, units
     end function
  end interface

contains

  !> Stops the program with the engine's message, like the reference's own STOP conditions (stop_prc).
  subroutine dre_check(rc, where)
    integer(c_int), intent(in) :: rc
    character(len=*), intent(in) :: where
    character(kind=c_char), pointer :: msg(:)
    integer :: n
    if (rc == 0) return
    call c_f_pointer(dre_last_error(eng), msg, [512])
    n = 0
    do while (n < 512)
       if (msg(n+1) == c_null_char) exit
       n = n + 1
    end do
    print *, 'dartray_gpu: ', where, ': ', msg(1:n)
    call stop_prc
  end subroutine dre_check

end module dartray_gpu
