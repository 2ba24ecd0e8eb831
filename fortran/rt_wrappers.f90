!> rt_wrappers.f90 -- the procedures of the DART-Ray host that the engine replaces, as thin wrappers around the C ABI.
!!
!! UNVERIFIED (no Fortran compiler in the build image; see dartray_gpu.f90).  These bodies replace, inside module
!! rt_routines / io_routines / dartray_hub, the procedures named beside each one; they use the module variables of the
!! reference exactly as the original bodies do (smooth_grid_routines.f90:41-230, rt_routines.f90:60-330).
!!
!! Protocol ("resident"): the accumulators live in HBM between the calls; the Fortran arrays are refreshed where the host
!! reads them -- in front of make_output / calc_sed / make_maps / restart files -- by dre_sync_host_arrays below.  Only the
!! main MPI process writes output files when no_communications is set (io_routines.f90:1396), so only it downloads.
!! A slower drop-in alternative that keeps the host arrays authoritative after every pass is dre_run_pass with host
!! arrays (rt_loop_hostarrays below).

!> rt_loop  (rt_routines.f90:985-1058; call sites dartray_hub.f90:22, 24, 31, 33, 49)
subroutine rt_loop
  use dartray_gpu
  type(dre_counters) :: cnt
  type(dre_params)   :: par
  if (test_run) return
  call fill_dre_params(par)
  call dre_check(dre_set_params(eng, par), 'dre_set_params')
  ! lumcell changes between the direct light (dens_stars*csize^3) and every scattering iteration (rt_prepare_scatt_gpu
  ! leaves the new one on the device, so nothing is uploaded then)
  if (rt_type /= rtt_scatt) then
     call dre_check(dre_set_sources(eng, c_loc(lumcell), tot_p_src, c_loc(ccoord_p_src), c_loc(lum_p_src_arr), &
                                    c_loc(src_cell), c_loc(cell_src)), 'dre_set_sources')
  end if
  call dre_check(dre_run_pass_resident(eng, rt_type, iterations, cnt), 'dre_run_pass_resident')
end subroutine rt_loop

!> rt_loop_2D  (rt_routines.f90:1061-1140; call sites dartray_hub.f90:80, 89, 105): both source loops and fix_symmetry on the device
subroutine rt_loop_2D
  use dartray_gpu
  type(dre_counters) :: cnt
  type(dre_params)   :: par
  if (test_run) return
  call fill_dre_params(par)
  call dre_check(dre_set_params(eng, par), 'dre_set_params')
  if (rt_type /= rtt_scatt) then
     call dre_check(dre_set_sources(eng, c_loc(lumcell), tot_p_src, c_loc(ccoord_p_src), c_loc(lum_p_src_arr), &
                                    c_loc(src_cell), c_loc(cell_src)), 'dre_set_sources')
  end if
  call dre_check(dre_run_pass_2d(eng, rt_type, iterations, cnt), 'dre_run_pass_2d')
end subroutine rt_loop_2D

!> set_dust_emission, dust_heating_type 'eff'  (sed_routines.f90:249-276; call site dartray_hub.f90:225): the emissivity of the next
!! dust-heating iteration from the radiation fields; u_final of the last dust pass is read on the device, lumcell is left there.
subroutine set_dust_emission_gpu
  use dartray_gpu
  integer :: i0
  i0 = i_lambda_dust(0)
  call dre_check(dre_dust_emission(eng, lnum_stars, lnum_dust, kabs_arr(0:lnum_stars-1), kabs_arr(i0:lnum_tot-1), &
       lambda_arr_SI(i0:lnum_tot-1), delta_lambda_bin_stars, delta_lambda_bin_dust, kext_ref, dens_ref, u_final_uv_opt, &
       c_null_ptr, iterations_dustem, dens_stars_arr, c_null_ptr, 1), 'dre_dust_emission')
  call reduce_dens_stars_arr
  call check_dens_stars_arr
end subroutine set_dust_emission_gpu

!> rt_mpi  (io_routines.f90:3667-3705 with reduce_u_fest_arr / reduce_u_final_arr / reduce_scaspe_arr / reduce_lum_lost /
!! reduce_psel_av_arr :3708-3881; call sites dartray_hub.f90:25, 34, 50): NCCL all-reduce of the device arrays.
subroutine rt_mpi
  use dartray_gpu
  call dre_check(dre_allreduce(eng, rt_type), 'dre_allreduce')
end subroutine rt_mpi

!> rt_prepare, case (rtt_scatt)  (dartray_hub.f90:432-457 with calc_total_luminosity_sca rt_routines.f90:663-758)
subroutine rt_prepare_scatt_gpu
  use dartray_gpu
  integer(c_int32_t) :: flag
  iterations = iterations + 1
  if (iterations == 1) bm_par = bm_par_sca
  ! u_fest <- u_final, lum_lost_prev <- lum_lost, scaspe_prev <- scaspe, lumcell = sum of the HEALPix part of scaspe,
  ! tot_rad_en at iteration 1, scaspe_tot += scaspe (iterations > 1), scaspe = 0: all on the device
  call dre_check(dre_prepare_scatt(eng, iterations, conv_en_lim, tot_rad_en, flag), 'dre_prepare_scatt')
  cnflag = (flag /= 0)
  if (limit_scattering_iterations .and. iterations > max_sca_iterations) cnflag = .true.
  if (.not. cnflag) then
     if (main_prc) print *, 'start scattering iteration number', iterations
     call expand_psel_av_arr
  end if
end subroutine rt_prepare_scatt_gpu

!> rt_loop_iobs  (rt_routines.f90:1470-1603; call sites dartray_hub.f90:37, 39, 55, 142, 144, 162).
!! i_obs_arr / i_obs_in_arr are "+=" arrays in the reference; the engine keeps them on the device (calc_sed / make_maps read
!! them there) and adds into the host arrays it is given.
subroutine rt_loop_iobs
  use dartray_gpu
  type(dre_counters) :: cnt
  type(dre_params)   :: par
  type(c_ptr) :: p_in, p_obs
  if (tot_ndir == 0 .and. tot_ndir_in == 0) return
  if (lnum_node == 0) return
  if (no_communications .and. .not. main_prc) return
  if (test_run) return
  call fill_dre_params(par)
  call dre_check(dre_set_params(eng, par), 'dre_set_params')
  p_in = c_null_ptr;  p_obs = c_null_ptr
  if (tot_ndir_in > 0) then
     p_in = c_loc(i_obs_in_arr);  p_obs = c_loc(ccoord_obs)
  end if
  ! scaspe_tot is already on the device (dre_prepare_scatt): no host planes are passed
  call dre_check(dre_run_iobs(eng, rt_type, c_null_ptr, tot_ndir_in, p_obs, c_loc(i_obs_arr), p_in, cnt), 'dre_run_iobs')
end subroutine rt_loop_iobs

!> Refreshes the Fortran arrays from the device in front of make_output / restart files (io_routines.f90:1112-1178,
!! 1244-1646).  Called by the main process only.
subroutine dre_sync_host_arrays
  use dartray_gpu
  type(c_ptr), allocatable, target :: sp(:)
  real(c_double), allocatable, target :: psel_slice(:)
  integer :: k, i0
  if (no_communications .and. .not. main_prc) return
  select case (rt_type)
  case (rtt_output_part1)
     call dre_check(dre_download(eng, DRE_ARR_U_FEST, c_loc(u_fest_arr)), 'download u_fest')
  case default
     call dre_check(dre_download(eng, DRE_ARR_U_FINAL, c_loc(u_final_arr)), 'download u_final')
     call dre_check(dre_download(eng, DRE_ARR_LUM_LOST, c_loc(lum_lost)), 'download lum_lost')
     allocate(sp(0:lnum-1))
     if (allocated(scaspe_tot_arr)) then
        do k = 0, lnum-1
           sp(k) = c_loc(scaspe_tot_arr(k)%a)
        end do
        call dre_check(dre_download(eng, DRE_ARR_SCASPE_TOT, c_loc(sp)), 'download scaspe_tot')
     else
        do k = 0, lnum-1
           sp(k) = c_loc(scaspe_arr(k)%a)
        end do
        call dre_check(dre_download(eng, DRE_ARR_SCASPE, c_loc(sp)), 'download scaspe')
     end if
     if (print_psel_av) then
        ! psel_av_arr(i0, iterations, :) is a STRIDED slice (the cell index is the slowest dimension,
        ! io_routines.f90:3874): download into a contiguous temporary and copy
        i0 = iterations_dustem - 1
        if (iterations_dustem == 0) i0 = 0
        allocate(psel_slice(0:tot_ncell+tot_p_src-1))
        call dre_check(dre_download(eng, DRE_ARR_PSEL_AV, c_loc(psel_slice)), 'download psel_av')
        psel_av_arr(i0, iterations, :) = psel_slice
        deallocate(psel_slice)
     end if
     deallocate(sp)
  end select
end subroutine dre_sync_host_arrays

!> Drop-in alternative to rt_loop that keeps the host arrays authoritative: every pass uploads its inputs and downloads
!! its outputs (dre_run_pass).  rt_mpi / rt_prepare / output stay the reference's own code in this mode.
subroutine rt_loop_hostarrays
  use dartray_gpu
  type(dre_counters) :: cnt
  type(dre_params)   :: par
  type(c_ptr), allocatable, target :: sp(:), spp(:)
  real(c_double), allocatable, target :: psel_slice(:)
  integer :: k, i0
  if (test_run) return
  call fill_dre_params(par)
  call dre_check(dre_set_params(eng, par), 'dre_set_params')
  call dre_check(dre_set_sources(eng, c_loc(lumcell), tot_p_src, c_loc(ccoord_p_src), c_loc(lum_p_src_arr), &
                                 c_loc(src_cell), c_loc(cell_src)), 'dre_set_sources')
  select case (rt_type)
  case (rtt_precalc_cell, rtt_precalc_src)
     call dre_check(dre_run_pass(eng, rt_type, iterations, c_null_ptr, c_loc(u_fest_arr), c_null_ptr, c_null_ptr, &
                                 c_null_ptr, c_null_ptr, cnt), 'dre_run_pass')
  case default
     allocate(sp(0:lnum-1), spp(0:lnum-1), psel_slice(0:tot_ncell+tot_p_src-1))   ! ragged scaspe_arr(k)%a -> array of c_ptr
     do k = 0, lnum-1
        sp(k) = c_loc(scaspe_arr(k)%a)
        spp(k) = c_null_ptr
        if (allocated(scaspe_prev_arr)) spp(k) = c_loc(scaspe_prev_arr(k)%a)
     end do
     i0 = iterations_dustem - 1
     if (iterations_dustem == 0) i0 = 0
     psel_slice = psel_av_arr(i0, iterations, :)          ! strided slice -> contiguous temporary (io_routines.f90:3874)
     call dre_check(dre_run_pass(eng, rt_type, iterations, c_loc(u_fest_arr), c_loc(u_final_arr), c_loc(sp), c_loc(spp), &
                                 c_loc(lum_lost), c_loc(psel_slice), cnt), 'dre_run_pass')
     psel_av_arr(i0, iterations, :) = psel_slice          ! only the entries of this pass's sources were rewritten
     deallocate(sp, spp, psel_slice)
  end select
end subroutine rt_loop_hostarrays

!> Scalars of the path -> dre_params (module variables of rt_routines.f90:16-26, 121, 268, 499-515; io_routines.f90:678)
subroutine fill_dre_params(par)
  use dartray_gpu
  type(dre_params), intent(out) :: par
  par%bm_par = bm_par                 ! already bm_par_sca during the scattering iterations (dartray_hub.f90:442)
  par%bm_par_max = bm_par_max
  par%pabs_max = pabs_max
  par%print_psel_av = merge(1, 0, print_psel_av)
  par%rad_lim = rad_lim;  par%en_lim = en_lim;  par%cs = cs;  par%modelsize = modelsize
  par%wall_on = [merge(1, 0, x_wall_on), merge(1, 0, y_wall_on), merge(1, 0, z_wall_on)]
  par%reserved0 = 0
  par%wall_coord = [x_wall_coord(1), x_wall_coord(2), y_wall_coord(1), y_wall_coord(2), z_wall_coord(1), z_wall_coord(2)]
end subroutine fill_dre_params

!> Once per run, from initialize_mpi (io_routines.f90:3610) and grid_initialize (dartray_hub.f90:472-549):
!!   if (id_mpi == 0) rc = dre_get_unique_id(nccl_id);  call MPI_Bcast(nccl_id, 128, MPI_CHARACTER, 0, MPI_COMM_WORLD, ierr)
!!   rc = dre_create(eng, local_rank, id_mpi, np_mpi, c_loc(nccl_id))
!!   rc = dre_set_grid(eng, tot_ncell, max_lvl, base, cchild, lvl, cindex, ccoord, csize, csize_arr, carea_arr, cvol_arr)
!!   rc = dre_set_opacity(eng, lnum, dens_arr, ksca_arr_norm(i0:), gsca_arr(i0:), kp_sca_arr(i0:), tot_ndir_scaspe, dir_i_out)
!! and again dre_set_opacity from dartray_dust when lnum and the opacity slices change (set_i_opacity_arrays :5074).
