! xtant_b200_shim.f90 -- ISO_C_BINDING stub that puts libxtb_b200.so (include/xtb_b200.h) behind the
! reference's own procedure names and argument lists for the tight-binding hot path.
!
! How it is used (INTEGRATION.md has the Makefile lines): add this file to OBJS after Objects.f90 and
! Universal_constants.f90, and in TB.f90 replace
!     use TB_NRL,      only : Construct_Vij_NRL, construct_TB_H_NRL, get_dHij_drij_NRL, Attract_TB_Forces_Press_NRL
!     use TB_Pettifor, only : construct_TB_H_Pettifor, dHij_s, Attract_TB_Forces_Press
!     use Atomic_tools,only : get_near_neighbours
!     use TB_DFTB,     only : Construct_Vij_DFTB, construct_TB_H_DFTB, get_dHij_drij_DFTB, Attract_TB_Forces_Press_DFTB
!     use Nonadiabatic,only : Electron_ion_coupling_Mij
! by   use XTANT_B200_shim   for those names (the rest of each module stays).  The `select type` dispatch in
! get_Hamilonian_and_E / create_and_diagonalize_H / Get_pressure (TB.f90:247-317,459-487,5551-5556) is unchanged:
! each `type is (TB_H_NRL)` / `type is (TB_H_Pettifor)` arm calls the same name with the same arguments as before.
!
! Status: no Fortran compiler exists in the build image nor on the GPU box (probed: gfortran, ifx, ifort,
! nvfortran, flang absent), so this file is checked by parsing (tests/test_abi_cpu.py: interfaces vs the header, struct
! layouts, the reference's argument lists) and its call sequence is executed by a plain-C program (tests/abi_driver.c).
!
! Data contract (SURVEY.md 8b): derived types cannot cross bind(C), so each wrapper flattens
!   Scell(NSC)%MDAtoms(:)%S, %KOA, Scell(NSC)%supce and the TB_H_*(KOA1,KOA2) scalars into contiguous buffers,
! and results are written straight into the reference's own allocatables (Near_neighbor_*, H_non, Sij, Aij,
! MDatoms(k)%forces%att, SCforce%att), all column-major as they already are.
module XTANT_B200_shim
use, intrinsic :: iso_c_binding
use Objects, only : Super_cell, Numerics_param, solid, Error_handling, Save_error_details, &
                    TB_H_NRL, TB_H_Pettifor, TB_Rep_Pettifor, TB_H_DFTB, TB_Rep_DFTB, TB_Exp_wall_simple, &
                    TB_vdW, TB_vdW_Girifalco, TB_vdW_LJ_cut, TB_vdW_ILJ_cut, TB_Coulomb, TB_Coulomb_cut, TB_Short_Rep
implicit none
private
public :: get_near_neighbours, Construct_Vij_NRL, construct_TB_H_NRL, Construct_Aij, Construct_Aij_x_En, &
          get_dHij_drij_NRL, Attract_TB_Forces_Press_NRL, construct_TB_H_Pettifor, dHij_s, Attract_TB_Forces_Press, &
          get_Erep_s, dErdr_s, dErdr_Pressure_s, xtb_b200_init, xtb_b200_finalize, &
          Construct_Vij_DFTB, construct_TB_H_DFTB, get_dHij_drij_DFTB, Attract_TB_Forces_Press_DFTB, &
          Electron_ion_coupling_Mij, xtb_b200_set_partition, xtb_b200_assemble_blocks, xtb_b200_expand_dense, &
          xtb_b200_density_mirror, xtb_b200_eigensolve, &
          get_Exp_wall_s, d_Exp_wall_pot_s, d_Exp_wall_Pressure_s, get_vdW_s, vdW_forces, get_mirror_cell_num, get_Coulomb_s, Coulomb_forces, &
          get_short_range_rep_s, d_Short_range_pot_s, d_Short_range_Pressure_s, &
          xtb_b200_comm_init, xtb_b200_exchange_blocks, &
          xtb_b200_exchange_forces, xtb_b200_allreduce_scforce, xtb_b200_step, xtb_b200_fermi, xtb_b200_md_half_step, &
          xtb_b200_md_set_velocities, xtb_b200_md_get_state, xtb_b200_check_pending, xtb_b200_dense_density

! ---- mirror of struct xtb_nrl_pair / xtb_pettifor_pair (include/xtb_b200.h) ----
type, bind(C) :: xtb_nrl_pair
   integer(c_int32_t) :: ind_split, ind_overlap
   real(c_double) :: lambd, Rc, lden
   real(c_double) :: al(4), bl(4), cl(4), dl(4)
   real(c_double) :: ellm(10), fllm(10), gllm(10), hllm(10)
   real(c_double) :: pllm(10), qllm(10), rllm(10), sllm(10)
end type xtb_nrl_pair

type, bind(C) :: xtb_pettifor_pair
   real(c_double) :: Es, Ep
   real(c_double) :: nc(4), rc(4), c0(4), c1(4), c2(4), c3(4), V0(4)
   real(c_double) :: r0, n, r1, rm
end type xtb_pettifor_pair

type, bind(C) :: xtb_pettifor_rep_pair      ! type(TB_Rep_Pettifor), Objects.f90:95-100
   real(c_double) :: E0_TB
   real(c_double) :: a0(5)
   real(c_double) :: phi0, d0, m, mc, dc, d1, dm
   real(c_double) :: c0(4)
end type xtb_pettifor_rep_pair
type, bind(C) :: xtb_dftb_pair              ! type(TB_H_DFTB), Objects.f90:202-222 (tables by address: they stay where they are)
   real(c_double) :: rcut, d, Ed, Ep, Es
   integer(c_int) :: n_r, pad_
   type(c_ptr) :: Rr, Vr, Sr
end type

type, bind(C) :: xtb_wall_pair              ! type(TB_Exp_wall_simple), Objects.f90:458-466
   real(c_double) :: C, r0, d0, dd
end type xtb_wall_pair

type, bind(C) :: xtb_short_pair             ! type(TB_Short_Rep), Objects.f90:437-454 (tables by address)
   real(c_double) :: d0, dd
   integer(c_int32_t) :: use_exp, use_inv_exp, use_zbl, use_nlh
   real(c_double) :: exp_Phi, exp_r0, exp_a
   real(c_double) :: inv_C, inv_r0
   integer(c_int32_t) :: n_pow, n_tab
   real(c_double) :: pow_Phi(8), pow_r0(8), pow_m(8)
   real(c_double) :: nlh_a(3), nlh_b(3)
   integer(c_int32_t) :: tab_spline, pad_
   type(c_ptr) :: tab_R, tab_E, tab_a, tab_b, tab_c, tab_d
end type xtb_short_pair

type, bind(C) :: xtb_vdw_pair               ! class(TB_vdW) Objects.f90:43-46 + Girifalco :302-328 / LJ_cut :336-343 / ILJ_cut :346-354
   real(c_double) :: d0_cut, dd_cut
   real(c_double) :: C12, C6, dm, d_cut, a, b, c, d
   real(c_double) :: dsm, ds_cut, as, bs, cs, ds, es, fs
   real(c_double) :: eps, r0, n, m
   real(c_double) :: d0_short, dd_short
end type xtb_vdw_pair

type, bind(C) :: xtb_coulomb_pair           ! type(TB_Coulomb_cut), Objects.f90:359-366
   real(c_double) :: k, dm, dd
end type xtb_coulomb_pair

type, bind(C) :: xtb_step_io                ! struct xtb_step_io (include/xtb_b200.h); c_null_ptr = not given / keep on the device
   type(c_ptr) :: S, supce, fe
   real(c_double) :: n_electrons, Te_eV
   type(c_ptr) :: Ei, fe_out, mu, E_band, F_att, SCforce_att
   type(c_ptr) :: E_rep, F_rep, SCforce_rep
   type(c_ptr) :: E_wall, F_wall, SCforce_wall
   type(c_ptr) :: E_vdw, F_vdw, SCforce_vdw
   type(c_ptr) :: E_coul, F_coul, SCforce_coul
end type xtb_step_io

type(c_ptr), save :: ctx = c_null_ptr      ! one context per process (single caller thread, TB.f90:227-272)
logical, save :: short_pushed = .false.     ! the General short-range class was pushed (needs matter%Atoms%Z: done by its own wrappers)
logical, save :: vdw_active = .false.       ! a usable class(TB_vdW) was pushed (push_params_vdw)
logical, save :: eigen_current = .false.   ! Ha, Ei of this step already on the device (uploaded once, by construct_TB_H_*)
logical, save :: density_current = .false. ! rho and rho_E of this step already contracted (ONE pass serves Construct_Aij and _x_En)
type(c_ptr), save :: ha_loc = c_null_ptr   ! address of the Ha that was uploaded
real(8), allocatable, save :: fe_used(:)   ! occupations the current density pass was made with
logical, save :: xtb_b200_dense_density = .true.   ! .false.: every consumer of Aij / Aij_x_Ei is grafted -> skip the two (N,N) copies to the host
character(512), save :: pending_error = ''  ! raised by a wrapper whose reference signature has no Err argument

interface
   integer(c_int) function xtb_create(out, device) bind(C, name="xtb_create")
      import; type(c_ptr), intent(out) :: out; integer(c_int), value :: device
   end function
   integer(c_int) function xtb_destroy(c) bind(C, name="xtb_destroy")
      import; type(c_ptr), value :: c
   end function
   type(c_ptr) function xtb_last_error(c) bind(C, name="xtb_last_error")
      import; type(c_ptr), value :: c
   end function
   integer(c_int) function xtb_set_params_nrl(c, n_kinds, p) bind(C, name="xtb_set_params_nrl")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds; type(xtb_nrl_pair), intent(in) :: p(*)
   end function
   integer(c_int) function xtb_set_params_pettifor(c, n_kinds, p) bind(C, name="xtb_set_params_pettifor")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds; type(xtb_pettifor_pair), intent(in) :: p(*)
   end function
   integer(c_int) function xtb_set_params_pettifor_rep(c, n_kinds, p) bind(C, name="xtb_set_params_pettifor_rep")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds; type(xtb_pettifor_rep_pair), intent(in) :: p(*)
   end function
   integer(c_int) function xtb_repulsive_energy(c, E_rep, Epot) bind(C, name="xtb_repulsive_energy")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: E_rep; real(c_double), intent(out) :: Epot(*)
   end function
   integer(c_int) function xtb_repulsive_forces(c, F_rep) bind(C, name="xtb_repulsive_forces")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: F_rep(3,*)
   end function
   integer(c_int) function xtb_repulsive_pressure(c, rep) bind(C, name="xtb_repulsive_pressure")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: rep(3,3)
   end function
   integer(c_int) function xtb_set_system(c, na, koa, S, supce) bind(C, name="xtb_set_system")
      import; type(c_ptr), value :: c; integer(c_int), value :: na
      integer(c_int32_t), intent(in) :: koa(*); real(c_double), intent(in) :: S(3,*), supce(3,3)
   end function
   integer(c_int) function xtb_set_partition(c, first, last) bind(C, name="xtb_set_partition")
      import; type(c_ptr), value :: c; integer(c_int), value :: first, last
   end function
   integer(c_int) function xtb_set_block_layout(c, parts) bind(C, name="xtb_set_block_layout")
      import; type(c_ptr), value :: c; integer(c_int), value :: parts
   end function
   integer(c_int) function xtb_neighbours(c, rm, nn_size, nn_list, nn_dist, nn_dist_s) bind(C, name="xtb_neighbours")
      import; type(c_ptr), value :: c; real(c_double), value :: rm
      integer(c_int32_t) :: nn_size(*), nn_list(*); real(c_double) :: nn_dist(*), nn_dist_s(*)
   end function
   integer(c_int) function xtb_radial(c, M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, M_x1, M_xrr) bind(C, name="xtb_radial")
      import; type(c_ptr), value :: c
      type(c_ptr), value :: M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, M_x1, M_xrr     ! c_null_ptr = keep on device
   end function
   integer(c_int) function xtb_assemble_HS(c, H_non, Sij) bind(C, name="xtb_assemble_HS")
      import; type(c_ptr), value :: c; type(c_ptr), value :: H_non, Sij
   end function
   integer(c_int) function xtb_set_eigen(c, Ha, Ei, fe) bind(C, name="xtb_set_eigen")
      import; type(c_ptr), value :: c; real(c_double), intent(in) :: Ha(*), Ei(*); type(c_ptr), value :: fe   ! c_null_ptr: occupations follow
   end function
   integer(c_int) function xtb_density(c, Aij, Aij_x_Ei) bind(C, name="xtb_density")
      import; type(c_ptr), value :: c; type(c_ptr), value :: Aij, Aij_x_Ei
   end function
   integer(c_int) function xtb_density_get(c, Aij, Aij_x_Ei) bind(C, name="xtb_density_get")
      import; type(c_ptr), value :: c; type(c_ptr), value :: Aij, Aij_x_Ei
   end function
   integer(c_int) function xtb_forces(c, F_att) bind(C, name="xtb_forces")
      import; type(c_ptr), value :: c; real(c_double) :: F_att(3,*)
   end function
   integer(c_int) function xtb_pressure(c, att) bind(C, name="xtb_pressure")
      import; type(c_ptr), value :: c; real(c_double) :: att(3,3)
   end function
   integer(c_int) function xtb_set_params_dftb(c, n_kinds, n_orb, p) bind(C, name="xtb_set_params_dftb")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds, n_orb; type(xtb_dftb_pair), intent(in) :: p(*)
   end function
   integer(c_int) function xtb_assemble_blocks(c) bind(C, name="xtb_assemble_blocks")
      import; type(c_ptr), value :: c
   end function
   integer(c_int) function xtb_expand_dense(c, first, last, H_non, Sij) bind(C, name="xtb_expand_dense")
      import; type(c_ptr), value :: c; integer(c_int), value :: first, last; type(c_ptr), value :: H_non, Sij
   end function
   integer(c_int) function xtb_density_mirror(c) bind(C, name="xtb_density_mirror")
      import; type(c_ptr), value :: c
   end function
   integer(c_int) function xtb_set_density_cutoff(c, f_cut) bind(C, name="xtb_set_density_cutoff")
      import; type(c_ptr), value :: c; real(c_double), value :: f_cut
   end function
   integer(c_int) function xtb_eigensolve(c, Ei, eigen_S, Ha) bind(C, name="xtb_eigensolve")
      import; type(c_ptr), value :: c; type(c_ptr), value :: Ei, eigen_S, Ha
   end function
   integer(c_int) function xtb_get_eigenvalues(c, Ei) bind(C, name="xtb_get_eigenvalues")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: Ei(*)
   end function
   integer(c_int) function xtb_set_occupations(c, fe) bind(C, name="xtb_set_occupations")
      import; type(c_ptr), value :: c; real(c_double), intent(in) :: fe(*)
   end function
   integer(c_int) function xtb_save_eigen_prev(c) bind(C, name="xtb_save_eigen_prev")
      import; type(c_ptr), value :: c
   end function
   integer(c_int) function xtb_set_eigen_prev(c, Ha0) bind(C, name="xtb_set_eigen_prev")
      import; type(c_ptr), value :: c; real(c_double), intent(in) :: Ha0(*)
   end function
   integer(c_int) function xtb_coupling(c, kind_M, Mij) bind(C, name="xtb_coupling")
      import; type(c_ptr), value :: c; integer(c_int), value :: kind_M; type(c_ptr), value :: Mij
   end function
   integer(c_int) function xtb_sync(c) bind(C, name="xtb_sync")
      import; type(c_ptr), value :: c
   end function
   real(c_double) function xtb_stage_ms(c, stage) bind(C, name="xtb_stage_ms")
      import; type(c_ptr), value :: c; integer(c_int), value :: stage
   end function
   integer(c_int) function xtb_cutoff(c, rm) bind(C, name="xtb_cutoff")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: rm
   end function
   integer(c_int) function xtb_set_params_wall(c, n_kinds, p) bind(C, name="xtb_set_params_wall")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds; type(c_ptr), value :: p      ! c_null_ptr with n_kinds = 0: no wall
   end function
   integer(c_int) function xtb_wall_energy(c, E_wall, Epot) bind(C, name="xtb_wall_energy")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: E_wall; real(c_double), intent(out) :: Epot(*)
   end function
   integer(c_int) function xtb_wall_forces(c, F_wall) bind(C, name="xtb_wall_forces")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: F_wall(3,*)
   end function
   integer(c_int) function xtb_wall_pressure(c, P_wall) bind(C, name="xtb_wall_pressure")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: P_wall(3,3)
   end function
   integer(c_int) function xtb_set_params_short(c, n_kinds, p, Z) bind(C, name="xtb_set_params_short")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds; type(c_ptr), value :: p; real(c_double), intent(in) :: Z(*)
   end function
   integer(c_int) function xtb_short_energy(c, E_short, Epot) bind(C, name="xtb_short_energy")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: E_short; real(c_double), intent(out) :: Epot(*)
   end function
   integer(c_int) function xtb_short_forces(c, F_short) bind(C, name="xtb_short_forces")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: F_short(3,*)
   end function
   integer(c_int) function xtb_short_pressure(c, P_short) bind(C, name="xtb_short_pressure")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: P_short(3,3)
   end function
   integer(c_int) function xtb_set_params_vdw(c, kind, n_kinds, p) bind(C, name="xtb_set_params_vdw")
      import; type(c_ptr), value :: c; integer(c_int), value :: kind, n_kinds; type(c_ptr), value :: p   ! c_null_ptr with n_kinds = 0: no vdW
   end function
   integer(c_int) function xtb_vdw_image_cells(c, N, R_cut) bind(C, name="xtb_vdw_image_cells")
      import; type(c_ptr), value :: c; integer(c_int), intent(out) :: N(3); real(c_double), intent(out) :: R_cut
   end function
   integer(c_int) function xtb_vdw_energy(c, E_vdw, Epot) bind(C, name="xtb_vdw_energy")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: E_vdw; real(c_double), intent(out) :: Epot(*)
   end function
   integer(c_int) function xtb_vdw_forces(c, F_vdw) bind(C, name="xtb_vdw_forces")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: F_vdw(3,*)
   end function
   integer(c_int) function xtb_vdw_pressure(c, P_vdw) bind(C, name="xtb_vdw_pressure")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: P_vdw(3,3)
   end function
   integer(c_int) function xtb_vdw_pressure_noalloc(c, P_vdw) bind(C, name="xtb_vdw_pressure_noalloc")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: P_vdw(3,3)     ! no device allocation inside (fixed scratch, <= 8192 atoms)
   end function
   integer(c_int) function xtb_set_params_coulomb(c, n_kinds, p) bind(C, name="xtb_set_params_coulomb")
      import; type(c_ptr), value :: c; integer(c_int), value :: n_kinds; type(c_ptr), value :: p     ! c_null_ptr with n_kinds = 0: no Coulomb term
   end function
   integer(c_int) function xtb_set_charge(c, Q) bind(C, name="xtb_set_charge")
      import; type(c_ptr), value :: c; real(c_double), value :: Q
   end function
   integer(c_int) function xtb_coulomb_image_cells(c, N) bind(C, name="xtb_coulomb_image_cells")
      import; type(c_ptr), value :: c; integer(c_int), intent(out) :: N(3)
   end function
   integer(c_int) function xtb_coulomb_energy(c, E_coul, Epot) bind(C, name="xtb_coulomb_energy")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: E_coul; real(c_double), intent(out) :: Epot(*)
   end function
   integer(c_int) function xtb_coulomb_forces(c, F_coul) bind(C, name="xtb_coulomb_forces")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: F_coul(3,*)
   end function
   integer(c_int) function xtb_coulomb_pressure(c, P_coul) bind(C, name="xtb_coulomb_pressure")
      import; type(c_ptr), value :: c; real(c_double), intent(out) :: P_coul(3,3)
   end function
   integer(c_int) function xtb_fermi_occupations(c, n_electrons, Te_eV, fe, mu, E_band) bind(C, name="xtb_fermi_occupations")
      import; type(c_ptr), value :: c; real(c_double), value :: n_electrons, Te_eV; type(c_ptr), value :: fe, mu, E_band
   end function
   integer(c_int) function xtb_md_set_velocities(c, SV) bind(C, name="xtb_md_set_velocities")
      import; type(c_ptr), value :: c; real(c_double), intent(in) :: SV(3,*)
   end function
   integer(c_int) function xtb_md_half_step(c, ind, dt_fs, mass_kg, Vsupce, force_terms) bind(C, name="xtb_md_half_step")
      import; type(c_ptr), value :: c; integer(c_int), value :: ind; real(c_double), value :: dt_fs
      real(c_double), intent(in) :: mass_kg(*); type(c_ptr), value :: Vsupce; integer(c_int), value :: force_terms
   end function
   integer(c_int) function xtb_md_get_state(c, S, SV, E_kin, mass_kg) bind(C, name="xtb_md_get_state")
      import; type(c_ptr), value :: c; type(c_ptr), value :: S, SV, E_kin, mass_kg
   end function
   integer(c_int) function xtb_comm_unique_id(id) bind(C, name="xtb_comm_unique_id")
      import; character(kind=c_char), intent(out) :: id(128)
   end function
   integer(c_int) function xtb_comm_init(c, id, rank, nranks) bind(C, name="xtb_comm_init")
      import; type(c_ptr), value :: c; character(kind=c_char), intent(in) :: id(128); integer(c_int), value :: rank, nranks
   end function
   integer(c_int) function xtb_comm_destroy(c) bind(C, name="xtb_comm_destroy")
      import; type(c_ptr), value :: c
   end function
   integer(c_int) function xtb_exchange_blocks(c, which) bind(C, name="xtb_exchange_blocks")
      import; type(c_ptr), value :: c; integer(c_int), value :: which
   end function
   integer(c_int) function xtb_exchange_forces(c, F_att) bind(C, name="xtb_exchange_forces")
      import; type(c_ptr), value :: c; type(c_ptr), value :: F_att
   end function
   integer(c_int) function xtb_allreduce_scforce(c, att) bind(C, name="xtb_allreduce_scforce")
      import; type(c_ptr), value :: c; type(c_ptr), value :: att
   end function
   integer(c_int) function xtb_step(c, flags, io) bind(C, name="xtb_step")
      import; type(c_ptr), value :: c; integer(c_int), value :: flags; type(xtb_step_io), intent(in) :: io
   end function
end interface

contains

! -------------------------------------------------------------------------------------------------
subroutine xtb_b200_init(numpar, Err)
   ! once, after initialize_MPI (XTANT_MAIN_FILE.f90:56-58): rank -> GPU = process_rank mod GPUs per node
   type(Numerics_param), intent(in) :: numpar
   type(Error_handling), intent(inout) :: Err
   integer(c_int) :: rc, dev
   dev = 0
#ifdef MPI_USED
   dev = mod(numpar%MPI_param%process_rank, 8)
#endif
   rc = xtb_create(ctx, dev)
   if (rc /= 0) call Save_error_details(Err, 6, 'xtb_create failed: no CUDA device (there is no CPU fallback)')
#ifdef MPI_USED
   if (rc == 0 .and. numpar%MPI_param%size_of_cluster > 1) call xtb_b200_comm_init(numpar, Err)
#endif
end subroutine xtb_b200_init

! One NCCL communicator over the MPI ranks: rank 0 draws the id, MPI_Bcast carries its 128 bytes (the only MPI call of the
! grafted path; the reference's MPI_Allreduce of dense matrices, MPI_subroutines.f90:3965-3982, is replaced by the exchanges
! inside the library).  xtb_comm_init gives this rank its atom range and the segmented block layout.
subroutine xtb_b200_comm_init(numpar, Err)
   type(Numerics_param), intent(in) :: numpar
   type(Error_handling), intent(inout) :: Err
   character(kind=c_char) :: id(128)
   integer :: ierr
   id = c_null_char
#ifdef MPI_USED
   if (numpar%MPI_param%process_rank == 0) call check(xtb_comm_unique_id(id), 'xtb_comm_unique_id', Err)
   call MPI_Bcast(id, 128, MPI_CHARACTER, 0, MPI_COMM_WORLD, ierr)
   call check(xtb_comm_init(ctx, id, int(numpar%MPI_param%process_rank,c_int), int(numpar%MPI_param%size_of_cluster,c_int)), &
              'xtb_comm_init', Err)
#endif
end subroutine xtb_b200_comm_init

subroutine xtb_b200_finalize()
   integer(c_int) :: rc
   if (c_associated(ctx)) then
      rc = xtb_comm_destroy(ctx)
      rc = xtb_destroy(ctx)
   endif
   ctx = c_null_ptr
   short_pushed = .false.
end subroutine xtb_b200_finalize

! errors raised inside wrappers whose reference signature has no Err argument (Construct_Aij_x_En, the force wrappers):
! call this after get_Hamilonian_and_E (TB.f90:196) to move them into the reference's error record
subroutine xtb_b200_check_pending(Err)
   type(Error_handling), intent(inout) :: Err
   if (len_trim(pending_error) > 0) call Save_error_details(Err, 6, trim(pending_error))
   pending_error = ''
end subroutine xtb_b200_check_pending

subroutine check(rc, where, Err)
   ! the reference records errors and continues (Save_error_details, Objects.f90:1309; e.g. TB_NRL.f90:410-416)
   integer(c_int), intent(in) :: rc
   character(*), intent(in) :: where
   type(Error_handling), intent(inout), optional :: Err
   character(kind=c_char), pointer :: msg(:)
   character(512) :: text
   integer :: i
   if (rc == 0) return
   call c_f_pointer(xtb_last_error(ctx), msg, [512])
   text = ''
   do i = 1, 512
      if (msg(i) == c_null_char) exit
      text(i:i) = msg(i)
   enddo
   if (present(Err)) then
      call Save_error_details(Err, 6, trim(where)//': '//trim(text))
   else
      print*, trim(where)//': '//trim(text)
      pending_error = trim(where)//': '//trim(text)       ! collected by xtb_b200_check_pending
   endif
end subroutine check

subroutine push_system(Scell, NSC)
   ! flatten MDAtoms(:)%KOA, %S and supce (Objects.f90:552-581,692-845)
   type(Super_cell), dimension(:), intent(in) :: Scell
   integer, intent(in) :: NSC
   integer :: na, i
   integer(c_int32_t), allocatable :: koa(:)
   real(c_double), allocatable :: S(:,:)
   na = size(Scell(NSC)%MDAtoms)
   allocate(koa(na), S(3,na))
   do i = 1, na
      koa(i) = Scell(NSC)%MDAtoms(i)%KOA
      S(:,i) = Scell(NSC)%MDAtoms(i)%S(:)
   enddo
   call check(xtb_set_system(ctx, int(na,c_int), koa, S, Scell(NSC)%supce), 'xtb_set_system')
   eigen_current = .false. ; density_current = .false.
end subroutine push_system

subroutine push_params_nrl(TB)
   type(TB_H_NRL), dimension(:,:), intent(in) :: TB
   type(xtb_nrl_pair), allocatable :: p(:)
   integer :: nk, i, j, k
   nk = size(TB,1)
   allocate(p(nk*nk))
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)      ! p[(koa1-1) + n_kinds*(koa2-1)] = TB(koa1,koa2)
         p(k)%ind_split = TB(i,j)%ind_split ; p(k)%ind_overlap = TB(i,j)%ind_overlap
         p(k)%lambd = TB(i,j)%lambd ; p(k)%Rc = TB(i,j)%Rc ; p(k)%lden = TB(i,j)%lden
         p(k)%al = TB(i,j)%al ; p(k)%bl = TB(i,j)%bl ; p(k)%cl = TB(i,j)%cl ; p(k)%dl = TB(i,j)%dl
         p(k)%ellm = TB(i,j)%ellm ; p(k)%fllm = TB(i,j)%fllm ; p(k)%gllm = TB(i,j)%gllm ; p(k)%hllm = TB(i,j)%hllm
         p(k)%pllm = TB(i,j)%pllm ; p(k)%qllm = TB(i,j)%qllm ; p(k)%rllm = TB(i,j)%rllm ; p(k)%sllm = TB(i,j)%sllm
      enddo
   enddo
   call check(xtb_set_params_nrl(ctx, int(nk,c_int), p), 'xtb_set_params_nrl')
end subroutine push_params_nrl

subroutine push_params_pettifor(TB)
   type(TB_H_Pettifor), dimension(:,:), intent(in) :: TB
   type(xtb_pettifor_pair), allocatable :: p(:)
   integer :: nk, i, j, k
   nk = size(TB,1)
   allocate(p(nk*nk))
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)
         p(k)%Es = TB(i,j)%Es ; p(k)%Ep = TB(i,j)%Ep
         p(k)%nc = TB(i,j)%nc ; p(k)%rc = TB(i,j)%rc ; p(k)%c0 = TB(i,j)%c0 ; p(k)%c1 = TB(i,j)%c1
         p(k)%c2 = TB(i,j)%c2 ; p(k)%c3 = TB(i,j)%c3 ; p(k)%V0 = TB(i,j)%V0
         p(k)%r0 = TB(i,j)%r0 ; p(k)%n = TB(i,j)%n ; p(k)%r1 = TB(i,j)%r1 ; p(k)%rm = TB(i,j)%rm
      enddo
   enddo
   call check(xtb_set_params_pettifor(ctx, int(nk,c_int), p), 'xtb_set_params_pettifor')
end subroutine push_params_pettifor

! -------------------------------------------------------------------------------------------------
! S1   Atomic_tools.f90:562
subroutine get_near_neighbours(Scell, numpar, include_vdW, dm)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   type(Numerics_param), intent(in) :: numpar
   logical, intent(in), optional :: include_vdW
   real(8), intent(out), optional :: dm
   integer :: NSC
   real(8) :: rm, r(1), d(1), X
   do NSC = 1, size(Scell)
      ASSOCIATE (ARRAY => Scell(NSC)%TB_Hamil(:,:))
         select type(ARRAY)
         type is (TB_H_Pettifor)
            call push_params_pettifor(ARRAY)
         type is (TB_H_NRL)
            call push_params_nrl(ARRAY)
         type is (TB_H_DFTB)
            call push_params_dftb(ARRAY, numpar%basis_size_ind)
         end select
      END ASSOCIATE
      if (allocated(Scell(NSC)%TB_Repuls)) then
         ASSOCIATE (ARRAY => Scell(NSC)%TB_Repuls(:,:))
            select type(ARRAY)
            type is (TB_Rep_Pettifor)
               call push_params_pettifor_rep(ARRAY)    ! also raises the cut-off to max(dm), Atomic_tools.f90:609-613
            end select
         END ASSOCIATE
      endif
      if (allocated(Scell(NSC)%TB_Expwall)) then
         ASSOCIATE (ARRAY => Scell(NSC)%TB_Expwall(:,:))
            select type(ARRAY)
            type is (TB_Exp_wall_simple)
               call push_params_wall(ARRAY)
            end select
         END ASSOCIATE
      endif
      if (allocated(Scell(NSC)%TB_Waals)) call push_params_vdw(Scell(NSC)%TB_Waals)
      if (allocated(Scell(NSC)%TB_Coul)) then
         ASSOCIATE (ARRAY => Scell(NSC)%TB_Coul(:,:))
            select type(ARRAY)
            type is (TB_Coulomb_cut)
               call push_params_coulomb(ARRAY)
            end select
         END ASSOCIATE
      endif
      ! cut-off exactly as Atomic_tools.f90:570-641: the TB part (attractive, raised to the Pettifor repulsive dm) comes
      ! from the library, the arms it has no parameters for are applied here as the reference does
      call check(xtb_cutoff(ctx, rm), 'xtb_cutoff')
      if (allocated(Scell(NSC)%TB_Repuls)) then
         ASSOCIATE (ARRAY => Scell(NSC)%TB_Repuls(:,:))
            select type(ARRAY)
            type is (TB_Rep_DFTB)                       ! :621-627
               if (ARRAY(1,1)%ToP == 0) then
                  r = maxval(ARRAY(:,:)%rcut)
               else
                  r = maxval(ARRAY(:,:)%rcut_spline)
               endif
               if (rm < r(1)) rm = r(1)
            end select
         END ASSOCIATE
      endif
      if (present(include_vdW) .and. allocated(Scell(NSC)%TB_Waals)) then      ! :633-640
         r = maxval(Scell(NSC)%TB_Waals(:,:)%d0_cut)
         d = maxval(Scell(NSC)%TB_Waals(:,:)%dd_cut)
         X = r(1) + 10.0d0*d(1)
         if (rm < X) rm = X
      endif
      if (present(dm)) then       ! :645-649: a present dm returns the radius and does NOT build the list
         dm = rm
      else
         call push_system(Scell, NSC)
         call check(xtb_neighbours(ctx, rm, Scell(NSC)%Near_neighbor_size, Scell(NSC)%Near_neighbor_list, &
                                   Scell(NSC)%Near_neighbor_dist, Scell(NSC)%Near_neighbor_dist_s), 'xtb_neighbours')
      endif
   enddo
end subroutine get_near_neighbours

! S2a  TB_NRL.f90:1956.  The dense (Na,Na,10) tables are only needed by un-grafted consumers; the grafted
! stages keep everything on the device, so the default forwards NULL pointers and leaves M_* untouched.
subroutine Construct_Vij_NRL(numpar, TB, Scell, NSC, M_Vij, M_dVij, M_SVij, M_dSVij)
   type(Numerics_param), intent(inout) :: numpar
   type(TB_H_NRL), dimension(:,:), intent(in), target :: TB
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   real(8), dimension(:,:,:), allocatable, target :: M_Vij, M_dVij, M_SVij, M_dSVij
   integer :: na
   na = size(Scell(NSC)%MDAtoms)
   if (.not.allocated(M_Vij))   allocate(M_Vij(na,na,10))
   if (.not.allocated(M_dVij))  allocate(M_dVij(na,na,10))
   if (.not.allocated(M_SVij))  allocate(M_SVij(na,na,10))
   if (.not.allocated(M_dSVij)) allocate(M_dSVij(na,na,10))
   call check(xtb_radial(ctx, c_loc(M_Vij), c_loc(M_dVij), c_loc(M_SVij), c_loc(M_dSVij), &
                         c_null_ptr, c_null_ptr, c_null_ptr), 'xtb_radial')
end subroutine Construct_Vij_NRL

! S2b  TB_NRL.f90:74.  Assembly on the GPU; the eigensolve (Loewdin_Orthogonalization + sym_diagonalize,
! TB_NRL.f90:402-421) stays in the reference (outside the graft) and is called on the returned H_non, Sij.
subroutine construct_TB_H_NRL(numpar, matter, TB_Hamil, M_Vij, M_SVij, M_lmn, Scell, NSC, Err)
   use TB_NRL, only : solve_nonorthogonal_NRL => Loewdin_and_diagonalize_from_HS   ! see INTEGRATION.md: 12-line split of Hamil_tot_NRL
   type(Numerics_param), intent(inout) :: numpar
   type(solid), intent(inout) :: matter
   type(TB_H_NRL), dimension(:,:), intent(in) :: TB_Hamil
   real(8), dimension(:,:,:), intent(in) :: M_Vij, M_SVij, M_lmn
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Error_handling), intent(inout) :: Err
   Scell(NSC)%Ha0 = Scell(NSC)%Ha            ! side effects the callers rely on (TB_NRL.f90:88-90)
   Scell(NSC)%Ei0 = Scell(NSC)%Ei
   Scell(NSC)%H_non0 = Scell(NSC)%H_non
   call check(xtb_assemble_HS(ctx, c_loc(Scell(NSC)%H_non), c_loc(Scell(NSC)%Sij)), 'xtb_assemble_HS', Err)
   Scell(NSC)%Hij = Scell(NSC)%H_non
   call solve_nonorthogonal_NRL(numpar, Scell, NSC, Err)      ! fills Ha, Ei, eigen_S, Hij_sol; then find_band_gap (:109)
   call upload_eigen(Scell(NSC)%Ha, Scell(NSC)%Ei)      ! ONE upload of the eigenvectors per step
end subroutine construct_TB_H_NRL

! S3   TB.f90:4328, 4401.  The device pass contracts rho AND rho_E together, once per step: Construct_Aij (called first,
! TB.f90:242) supplies the occupations and triggers it; Construct_Aij_x_En (TB.f90:268) finds it done and only copies rho_E out.
! The eigenvectors and eigenvalues were uploaded once by construct_TB_H_* (upload_eigen); a caller that passes another Ha, or
! energies that are not this step's Ei, is served by a fresh upload and pass (correct, just not free).
subroutine upload_eigen(Ha, Ei)
   real(8), dimension(:,:), intent(in), target :: Ha
   real(8), dimension(:), intent(in) :: Ei
   call check(xtb_set_eigen(ctx, Ha, Ei, c_null_ptr), 'xtb_set_eigen')      ! occupations follow (xtb_set_occupations)
   ha_loc = c_loc(Ha)
   eigen_current = .true. ; density_current = .false.
end subroutine upload_eigen

subroutine ensure_density(Ha, distre, En)
   real(8), dimension(:,:), intent(in), target :: Ha
   real(8), dimension(:), intent(in) :: distre
   real(8), dimension(:), intent(in), optional :: En
   real(c_double), allocatable :: Ei_dev(:)
   logical :: same
   same = eigen_current .and. c_associated(c_loc(Ha), ha_loc)
   if (same .and. present(En)) then                      ! are these the energies that were uploaded with Ha ?
      allocate(Ei_dev(size(En)))
      call check(xtb_get_eigenvalues(ctx, Ei_dev), 'xtb_get_eigenvalues')
      same = all(Ei_dev == En)
   endif
   if (.not. same) then
      if (present(En)) then
         call check(xtb_set_eigen(ctx, Ha, En, c_null_ptr), 'xtb_set_eigen')
         ha_loc = c_loc(Ha) ; eigen_current = .true.
      else
         if (.not. eigen_current) then
            pending_error = 'XTANT_B200_shim: Construct_Aij before construct_TB_H_* of this step (no eigenvalues for rho_E)'
            allocate(Ei_dev(size(distre)), source = 0.0d0)
            call check(xtb_set_eigen(ctx, Ha, Ei_dev, c_null_ptr), 'xtb_set_eigen')
         else
            allocate(Ei_dev(size(distre)))
            call check(xtb_get_eigenvalues(ctx, Ei_dev), 'xtb_get_eigenvalues')
            call check(xtb_set_eigen(ctx, Ha, Ei_dev, c_null_ptr), 'xtb_set_eigen')
         endif
         ha_loc = c_loc(Ha)
      endif
      density_current = .false.
   endif
   if (density_current) then
      if (allocated(fe_used)) then
         if (size(fe_used) == size(distre)) then
            if (all(fe_used == distre)) return           ! same Ha, Ei, fe: the pass of this step stands
         endif
      endif
   endif
   call check(xtb_set_occupations(ctx, distre), 'xtb_set_occupations')
   call check(xtb_density(ctx, c_null_ptr, c_null_ptr), 'xtb_density')     ! ONE contraction: rho and rho_E
   if (allocated(fe_used)) deallocate(fe_used)
   allocate(fe_used, source = distre)
   density_current = .true.
end subroutine ensure_density

subroutine Construct_Aij(numpar, Ha, distre, Aij)
   type(Numerics_param), intent(inout) :: numpar
   real(8), dimension(:,:), intent(in), target :: Ha
   real(8), dimension(:), intent(in) :: distre
   real(8), dimension(:,:), allocatable, intent(inout), target :: Aij
   if (.not. allocated(Aij)) allocate(Aij(size(distre), size(distre)))
   call ensure_density(Ha, distre)
   ! dense copy only for un-grafted consumers of Scell%Aij (TB.f90:250-299); the grafted force stages read the device blocks
   if (xtb_b200_dense_density) call check(xtb_density_get(ctx, c_loc(Aij), c_null_ptr), 'xtb_density(Aij)')
end subroutine Construct_Aij

subroutine Construct_Aij_x_En(numpar, Ha, distre, En, Aij, H_ssc_1_per_S)
   type(Numerics_param), intent(inout) :: numpar
   real(8), dimension(:,:), intent(in), target :: Ha
   real(8), dimension(:), intent(in) :: distre, En
   real(8), dimension(:,:), allocatable, intent(inout), target :: Aij
   real(8), dimension(:,:), intent(in), optional :: H_ssc_1_per_S
   if (.not. allocated(Aij)) allocate(Aij(size(distre), size(distre)))
   if (present(H_ssc_1_per_S)) then
      ! SCC variant fe*(Ei - HperS(i,:)) (TB.f90:4445-4454): outside the graft.  The caller gets a defined (zero) matrix and
      ! the error goes to the reference's error record through xtb_b200_check_pending.
      Aij = 0.0d0
      pending_error = 'XTANT_B200_shim: the SCC variant of Construct_Aij_x_En (TB.f90:4445-4454) is not grafted'
      print*, trim(pending_error)
      return
   endif
   call ensure_density(Ha, distre, En)
   if (xtb_b200_dense_density) call check(xtb_density_get(ctx, c_null_ptr, c_loc(Aij)), 'xtb_density(Aij_x_Ei)')
end subroutine Construct_Aij_x_En

! S4   TB_NRL.f90:2412
subroutine get_dHij_drij_NRL(TB_Hamil, Scell, NSC, numpar, Aij, M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, Aij_x_Ei)
   type(TB_H_NRL), dimension(:,:), intent(in) :: TB_Hamil
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(8), dimension(:,:,:), intent(in) :: M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn
   real(8), dimension(:,:), intent(in) :: Aij, Aij_x_Ei
   real(c_double), allocatable :: F(:,:)
   integer :: k, na
   na = size(Scell(NSC)%MDatoms)
   allocate(F(3,na))
   call check(xtb_forces(ctx, F), 'xtb_forces')        ! uses the device-resident rho, rho_E of this step
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%att(:) = F(:,k)
   enddo
end subroutine get_dHij_drij_NRL

! S5   TB_NRL.f90:2742
subroutine Attract_TB_Forces_Press_NRL(TB_Hamil, Scell, NSC, numpar, Aij, M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, Aij_x_Ei)
   type(TB_H_NRL), dimension(:,:), intent(in) :: TB_Hamil
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(8), dimension(:,:,:), intent(in) :: M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn
   real(8), dimension(:,:), intent(in) :: Aij, Aij_x_Ei
   if (numpar%p_const) then        ! TB_NRL.f90:2758
      call check(xtb_pressure(ctx, Scell(NSC)%SCforce%att), 'xtb_pressure')
   endif
end subroutine Attract_TB_Forces_Press_NRL

! ---- Pettifor arms (TB_Pettifor.f90:1427,141,1134) ----------------------------------------------
subroutine construct_TB_H_Pettifor(numpar, matter, TB_Hamil, Scell, NSC, Ha, Err)
   use Algebra_tools, only : sym_diagonalize
   type(Numerics_param), intent(inout) :: numpar
   type(solid), intent(inout) :: matter
   type(TB_H_Pettifor), dimension(:,:), intent(in) :: TB_Hamil
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   real(8), dimension(:,:), intent(out), target :: Ha
   type(Error_handling), intent(inout) :: Err
   character(200) :: Error_descript
   Error_descript = ''
   Scell(NSC)%Ha0 = Scell(NSC)%Ha
   Scell(NSC)%Ei0 = Scell(NSC)%Ei
   Scell(NSC)%H_non0 = Scell(NSC)%H_non
   call check(xtb_assemble_HS(ctx, c_loc(Ha), c_null_ptr), 'xtb_assemble_HS', Err)
   Scell(NSC)%H_non = Ha                                                  ! TB_Pettifor.f90:1452
   call sym_diagonalize(Ha, Scell(NSC)%Ei, Error_descript, check_M=.true.) ! :1455 (outside the graft)
   call upload_eigen(Ha, Scell(NSC)%Ei)                                    ! ONE upload of the eigenvectors per step
end subroutine construct_TB_H_Pettifor

subroutine dHij_s(TB_Hamil, atoms, Scell, NSC, numpar, Aij, M_x1, M_xrr)
   use Objects, only : Atom
   type(TB_H_Pettifor), dimension(:,:), intent(in) :: TB_Hamil
   type(Atom), dimension(:), intent(inout) :: atoms
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(in) :: numpar
   real(8), dimension(:,:), intent(in) :: Aij
   real(8), dimension(:,:,:), intent(in) :: M_x1, M_xrr
   real(c_double), allocatable :: F(:,:)
   integer :: k, na
   na = size(atoms)
   allocate(F(3,na))
   call check(xtb_forces(ctx, F), 'xtb_forces')
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%att(:) = F(:,k)
   enddo
end subroutine dHij_s

subroutine Attract_TB_Forces_Press(TB_Hamil, atoms, Scell, NSC, numpar, Aij)
   use Objects, only : Atom
   type(TB_H_Pettifor), dimension(:,:), intent(in) :: TB_Hamil
   type(Atom), dimension(:), intent(inout) :: atoms
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(in) :: numpar
   real(8), dimension(:,:), intent(in) :: Aij
   if (numpar%p_const) then        ! TB_Pettifor.f90:1147
      call check(xtb_pressure(ctx, Scell(NSC)%SCforce%att), 'xtb_pressure')
   endif
end subroutine Attract_TB_Forces_Press

! ---- repulsive part of the Pettifor-type branch (TB_Pettifor.f90:2074,2162,2634; called from TB.f90 get_pot_nrg / get_total_forces)
subroutine push_params_pettifor_rep(TB)
   type(TB_Rep_Pettifor), dimension(:,:), intent(in) :: TB
   type(xtb_pettifor_rep_pair), allocatable :: p(:)
   integer :: nk, i, j, k
   nk = size(TB,1)
   allocate(p(nk*nk))
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)
         p(k)%E0_TB = TB(i,j)%E0_TB ; p(k)%a0 = TB(i,j)%a0 ; p(k)%c0 = TB(i,j)%c0
         p(k)%phi0 = TB(i,j)%phi0 ; p(k)%d0 = TB(i,j)%d0 ; p(k)%m = TB(i,j)%m ; p(k)%mc = TB(i,j)%mc
         p(k)%dc = TB(i,j)%dc ; p(k)%d1 = TB(i,j)%d1 ; p(k)%dm = TB(i,j)%dm
      enddo
   enddo
   call check(xtb_set_params_pettifor_rep(ctx, int(nk,c_int), p), 'xtb_set_params_pettifor_rep')
end subroutine push_params_pettifor_rep

subroutine get_Erep_s(TB_Repuls, Scell, NSC, numpar, a)
   type(TB_Rep_Pettifor), dimension(:,:), intent(in) :: TB_Repuls
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(in) :: numpar
   real(8), intent(out) :: a
   real(c_double), allocatable :: Epot(:)
   integer :: i, na
   na = size(Scell(NSC)%MDAtoms)
   allocate(Epot(na))
   call check(xtb_repulsive_energy(ctx, a, Epot), 'xtb_repulsive_energy')
   do i = 1, na
      Scell(NSC)%MDAtoms(i)%Epot = Epot(i)      ! TB_Pettifor.f90:2103
   enddo
end subroutine get_Erep_s

subroutine dErdr_s(TB_Repuls, atoms, Scell, NSC, numpar)
   use Objects, only : Atom
   type(TB_Rep_Pettifor), dimension(:,:), intent(in) :: TB_Repuls
   type(Atom), dimension(:), intent(in) :: atoms
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(c_double), allocatable :: F(:,:)
   integer :: k, na
   na = size(atoms)
   allocate(F(3,na))
   call check(xtb_repulsive_forces(ctx, F), 'xtb_repulsive_forces')
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%rep(:) = F(:,k)     ! TB_Pettifor.f90:2299
   enddo
end subroutine dErdr_s

subroutine dErdr_Pressure_s(TB_Repuls, atoms, Scell, NSC, numpar)
   use Objects, only : Atom
   type(TB_Rep_Pettifor), dimension(:,:), intent(in) :: TB_Repuls
   type(Atom), dimension(:), intent(in) :: atoms
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(in) :: numpar
   if (numpar%p_const) then        ! TB_Pettifor.f90:2646
      call check(xtb_repulsive_pressure(ctx, Scell(NSC)%SCforce%rep), 'xtb_repulsive_pressure')
   endif
end subroutine dErdr_Pressure_s

! ---- DFTB arms (TB_DFTB.f90:56,284,781,1135) ------------------------------------------------------
! The radial tables Rr(:), Vr(:,:), Sr(:,:) of every TB_H_DFTB(KOA1,KOA2) are handed over by address (the library copies
! them to the device once); n_orb follows identify_DFTB_orbitals_per_atom(numpar%basis_size_ind) (TB_DFTB.f90:2237-2248).
subroutine push_params_dftb(TB, basis_size_ind)
   type(TB_H_DFTB), dimension(:,:), intent(in), target :: TB
   integer, intent(in) :: basis_size_ind
   type(xtb_dftb_pair), allocatable :: p(:)
   integer :: nk, i, j, k, n_orb
   nk = size(TB,1)
   allocate(p(nk*nk))
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)      ! p[(koa1-1) + n_kinds*(koa2-1)] = TB(koa1,koa2)
         p(k)%rcut = TB(i,j)%rcut ; p(k)%d = TB(i,j)%d
         p(k)%Ed = TB(i,j)%Ed ; p(k)%Ep = TB(i,j)%Ep ; p(k)%Es = TB(i,j)%Es
         p(k)%n_r = int(size(TB(i,j)%Rr), c_int) ; p(k)%pad_ = 0_c_int
         p(k)%Rr = c_loc(TB(i,j)%Rr) ; p(k)%Vr = c_loc(TB(i,j)%Vr) ; p(k)%Sr = c_loc(TB(i,j)%Sr)   ! Vr(ir,channel): column-major as the header asks
      enddo
   enddo
   select case (basis_size_ind)
   case (0) ; n_orb = 1
   case (1) ; n_orb = 4
   case default ; n_orb = 9
   end select
   call check(xtb_set_params_dftb(ctx, int(nk,c_int), int(n_orb,c_int), p), 'xtb_set_params_dftb')
end subroutine push_params_dftb

! S2a  TB_DFTB.f90:56.  As for NRL: the dense (Na,Na,N_bs) tables are only produced for un-grafted consumers.
subroutine Construct_Vij_DFTB(numpar, TB, Scell, NSC, M_Vij, M_dVij, M_SVij, M_dSVij)
   type(Numerics_param), intent(inout) :: numpar
   type(TB_H_DFTB), dimension(:,:), intent(in) :: TB
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   real(8), dimension(:,:,:), allocatable, target :: M_Vij, M_dVij, M_SVij, M_dSVij
   integer :: na, N_bs
   na = size(Scell(NSC)%MDAtoms)
   select case (numpar%basis_size_ind)      ! identify_DFTB_basis_size, TB_DFTB.f90:2224
   case (0) ; N_bs = 1
   case (1) ; N_bs = 4
   case default ; N_bs = 10
   end select
   if (.not.allocated(M_Vij))   allocate(M_Vij(na,na,N_bs))
   if (.not.allocated(M_dVij))  allocate(M_dVij(na,na,N_bs))
   if (.not.allocated(M_SVij))  allocate(M_SVij(na,na,N_bs))
   if (.not.allocated(M_dSVij)) allocate(M_dSVij(na,na,N_bs))
   call check(xtb_radial(ctx, c_loc(M_Vij), c_loc(M_dVij), c_loc(M_SVij), c_loc(M_dSVij), &
                         c_null_ptr, c_null_ptr, c_null_ptr), 'xtb_radial')
end subroutine Construct_Vij_DFTB

! S2b  TB_DFTB.f90:284.  Non-SCC arm only (the SCC corrections H_scc_0/1 are outside the graft, SURVEY.md 8).
subroutine construct_TB_H_DFTB(numpar, matter, TB_Hamil, M_Vij, M_SVij, M_lmn, Scell, NSC, Err, scc, H_scc_0, H_scc_1)
   use TB_NRL, only : solve_nonorthogonal => Loewdin_and_diagonalize_from_HS   ! the same split as for NRL (INTEGRATION.md)
   type(Numerics_param), intent(inout) :: numpar
   type(solid), intent(inout) :: matter
   type(TB_H_DFTB), dimension(:,:), intent(in) :: TB_Hamil
   real(8), dimension(:,:,:), intent(in) :: M_Vij, M_SVij, M_lmn
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Error_handling), intent(inout) :: Err
   integer, intent(in), optional :: scc
   real(8), dimension(:,:), intent(inout), optional :: H_scc_0, H_scc_1
   if (present(scc) .and. present(H_scc_0) .and. present(H_scc_1)) then
      call Save_error_details(Err, 6, 'XTANT_B200_shim: the SCC arm of construct_TB_H_DFTB (TB_DFTB.f90:303-304) is not grafted')
      return
   endif
   Scell(NSC)%Ha0 = Scell(NSC)%Ha            ! TB_DFTB.f90:298-300
   Scell(NSC)%Ei0 = Scell(NSC)%Ei
   Scell(NSC)%H_non0 = Scell(NSC)%H_non
   call check(xtb_assemble_HS(ctx, c_loc(Scell(NSC)%H_non), c_loc(Scell(NSC)%Sij)), 'xtb_assemble_HS', Err)
   Scell(NSC)%Hij = Scell(NSC)%H_non
   call solve_nonorthogonal(numpar, Scell, NSC, Err)      ! Ha, Ei, eigen_S, Hij_sol; then find_band_gap (:314)
   call upload_eigen(Scell(NSC)%Ha, Scell(NSC)%Ei)      ! ONE upload of the eigenvectors per step
end subroutine construct_TB_H_DFTB

! S4   TB_DFTB.f90:781
subroutine get_dHij_drij_DFTB(numpar, Scell, NSC, Aij, M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, Aij_x_Ei)
   type(Numerics_param), intent(inout) :: numpar
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   real(8), dimension(:,:,:), intent(in) :: M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn
   real(8), dimension(:,:), intent(in) :: Aij, Aij_x_Ei
   real(c_double), allocatable :: F(:,:)
   integer :: k, na
   na = size(Scell(NSC)%MDatoms)
   allocate(F(3,na))
   call check(xtb_forces(ctx, F), 'xtb_forces')        ! uses the device-resident rho, rho_E of this step
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%att(:) = F(:,k)
   enddo
end subroutine get_dHij_drij_DFTB

! S5   TB_DFTB.f90:1135
subroutine Attract_TB_Forces_Press_DFTB(Scell, NSC, numpar, Aij, M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn, Aij_x_Ei)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(8), dimension(:,:,:), intent(in) :: M_Vij, M_dVij, M_SVij, M_dSVij, M_lmn
   real(8), dimension(:,:), intent(in) :: Aij, Aij_x_Ei
   if (numpar%p_const) then
      call check(xtb_pressure(ctx, Scell(NSC)%SCforce%att), 'xtb_pressure')
   endif
end subroutine Attract_TB_Forces_Press_DFTB

! ---- nonadiabatic coupling (Nonadiabatic.f90:782-945) ---------------------------------------------
! Ha and the overlap blocks of this step are on the device already (xtb_set_eigen / xtb_eigensolve, xtb_assemble_HS);
! Ha0 is uploaded unless xtb_save_eigen_prev kept it there.  wr and Sij are part of the reference signature only.
subroutine Electron_ion_coupling_Mij(wr, Ha, Ha0, Mij, kind_M, Sij)
   real(8), dimension(:), intent(in) :: wr
   real(8), dimension(:,:), intent(in) :: Ha, Ha0
   real(8), dimension(:,:), allocatable, intent(inout), target :: Mij
   integer, intent(in), optional :: kind_M
   real(8), dimension(:,:), intent(in), optional :: Sij
   integer(c_int) :: km
   integer :: N
   N = size(wr)
   if (.not. allocated(Mij)) allocate(Mij(N,N))
   km = 0_c_int                                  ! default arm: Tully (Nonadiabatic.f90:876)
   if (present(kind_M)) km = int(kind_M, c_int)
   if (.not. eigen_current) call check(xtb_set_eigen(ctx, Ha, wr, c_null_ptr), 'xtb_set_eigen')   ! occupations are not used by the coupling
   call check(xtb_set_eigen_prev(ctx, Ha0), 'xtb_set_eigen_prev')
   call check(xtb_coupling(ctx, km, c_loc(Mij)), 'xtb_coupling')
end subroutine Electron_ion_coupling_Mij

! ---- short-range wall (Exponential_wall.f90:990,1200,1334; dispatch TB.f90:3533-3556, energy TB.f90:3983) -------------------
subroutine push_params_wall(TB)
   type(TB_Exp_wall_simple), dimension(:,:), intent(in) :: TB
   type(xtb_wall_pair), allocatable, target :: p(:)
   integer :: nk, i, j, k
   nk = size(TB,1)
   allocate(p(nk*nk))
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)
         p(k)%C = TB(i,j)%C ; p(k)%r0 = TB(i,j)%r0 ; p(k)%d0 = TB(i,j)%d0 ; p(k)%dd = TB(i,j)%dd
      enddo
   enddo
   call check(xtb_set_params_wall(ctx, int(nk,c_int), c_loc(p)), 'xtb_set_params_wall')
end subroutine push_params_wall

subroutine get_Exp_wall_s(TB_Expwall, Scell, NSC, numpar, a)
   type(TB_Exp_wall_simple), dimension(:,:), intent(in) :: TB_Expwall
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(8), intent(out) :: a
   real(c_double), allocatable :: Epot(:)
   integer :: i, na
   na = size(Scell(NSC)%MDAtoms)
   allocate(Epot(na))
   call check(xtb_wall_energy(ctx, a, Epot), 'xtb_wall_energy')
   do i = 1, na
      Scell(NSC)%MDAtoms(i)%Epot = Scell(NSC)%MDAtoms(i)%Epot + Epot(i)      ! Exponential_wall.f90:1058 (the half is inside)
   enddo
end subroutine get_Exp_wall_s

subroutine d_Exp_wall_pot_s(Scell, NSC, TB_Expwall, numpar)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(TB_Exp_wall_simple), dimension(:,:), intent(in) :: TB_Expwall
   type(Numerics_param), intent(inout) :: numpar
   real(c_double), allocatable :: F(:,:)
   integer :: k, na
   na = size(Scell(NSC)%MDatoms)
   allocate(F(3,na))
   call check(xtb_wall_forces(ctx, F), 'xtb_wall_forces')
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%rep(:) = Scell(NSC)%MDatoms(k)%forces%rep(:) + F(:,k)    ! :1309 (added to the repulsive part)
   enddo
end subroutine d_Exp_wall_pot_s

subroutine d_Exp_wall_Pressure_s(Scell, NSC, TB_Expwall, numpar)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(TB_Exp_wall_simple), dimension(:,:), intent(in) :: TB_Expwall
   type(Numerics_param), intent(in) :: numpar
   real(c_double) :: P(3,3)
   if (numpar%p_const) then        ! :1346
      call check(xtb_wall_pressure(ctx, P), 'xtb_wall_pressure')
      Scell(NSC)%SCforce%rep = Scell(NSC)%SCforce%rep + P                    ! :1381
   endif
end subroutine d_Exp_wall_Pressure_s

! ---- the 'General' short-range class in the slot of the wall (Exponential_wall.f90:48,519,664; dispatch TB.f90:3533-3556,3975-3993).
!      Its parameters need the atomic numbers (matter%Atoms(:)%Z), which get_near_neighbours does not receive: the three wrappers,
!      which do, push them on their first call.
subroutine ensure_short(TB, matter)
   type(TB_Short_Rep), dimension(:,:), intent(in), target :: TB
   type(solid), intent(in) :: matter
   type(xtb_short_pair), allocatable, target :: p(:)
   real(c_double), allocatable :: Z(:)
   integer :: nk, i, j, k, n
   if (short_pushed) return
   nk = size(TB,1)
   allocate(p(nk*nk), Z(nk))
   do i = 1, nk
      Z(i) = matter%Atoms(i)%Z
   enddo
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)
         p(k)%d0 = TB(i,j)%f_cut%d0 ; p(k)%dd = TB(i,j)%f_cut%dd
         p(k)%use_exp = merge(1, 0, TB(i,j)%f_exp%use_it) ; p(k)%use_inv_exp = merge(1, 0, TB(i,j)%f_inv_exp%use_it)
         p(k)%use_zbl = merge(1, 0, TB(i,j)%f_ZBL%use_it) ; p(k)%use_nlh = merge(1, 0, TB(i,j)%f_NLH%use_it)
         p(k)%exp_Phi = TB(i,j)%f_exp%Phi ; p(k)%exp_r0 = TB(i,j)%f_exp%r0 ; p(k)%exp_a = TB(i,j)%f_exp%a
         p(k)%inv_C = TB(i,j)%f_inv_exp%C ; p(k)%inv_r0 = TB(i,j)%f_inv_exp%r0
         p(k)%pow_Phi = 0.0d0 ; p(k)%pow_r0 = 1.0d0 ; p(k)%pow_m = 0.0d0 ; p(k)%n_pow = 0
         if (allocated(TB(i,j)%f_pow)) then
            n = size(TB(i,j)%f_pow)
            if (n > 8) pending_error = 'TB_Short_Rep: more than 8 power terms (XTB_SHORT_MAX_POW); the extra ones are dropped'
            p(k)%n_pow = min(n, 8)
            p(k)%pow_Phi(1:p(k)%n_pow) = TB(i,j)%f_pow(1:p(k)%n_pow)%Phi
            p(k)%pow_r0(1:p(k)%n_pow) = TB(i,j)%f_pow(1:p(k)%n_pow)%r0
            p(k)%pow_m(1:p(k)%n_pow) = TB(i,j)%f_pow(1:p(k)%n_pow)%m
         endif
         p(k)%nlh_a = TB(i,j)%f_NLH%a ; p(k)%nlh_b = TB(i,j)%f_NLH%b
         p(k)%n_tab = 0 ; p(k)%tab_spline = 0 ; p(k)%pad_ = 0
         p(k)%tab_R = c_null_ptr ; p(k)%tab_E = c_null_ptr
         p(k)%tab_a = c_null_ptr ; p(k)%tab_b = c_null_ptr ; p(k)%tab_c = c_null_ptr ; p(k)%tab_d = c_null_ptr
         if (TB(i,j)%f_tab%use_it) then
            p(k)%n_tab = size(TB(i,j)%f_tab%R)
            p(k)%tab_spline = merge(1, 0, TB(i,j)%f_tab%use_spline)
            p(k)%tab_R = c_loc(TB(i,j)%f_tab%R) ; p(k)%tab_E = c_loc(TB(i,j)%f_tab%E)
            p(k)%tab_a = c_loc(TB(i,j)%f_tab%a) ; p(k)%tab_b = c_loc(TB(i,j)%f_tab%b)
            p(k)%tab_c = c_loc(TB(i,j)%f_tab%c) ; p(k)%tab_d = c_loc(TB(i,j)%f_tab%d)
         endif
      enddo
   enddo
   call check(xtb_set_params_short(ctx, int(nk,c_int), c_loc(p), Z), 'xtb_set_params_short')
   short_pushed = .true.
end subroutine ensure_short

subroutine get_short_range_rep_s(TB_Expwall, Scell, NSC, matter, numpar, a)
   type(TB_Short_Rep), dimension(:,:), intent(in) :: TB_Expwall
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(solid), intent(in), target :: matter
   type(Numerics_param), intent(inout) :: numpar
   real(8), intent(out) :: a
   real(c_double), allocatable :: Epot(:)
   integer :: i, na
   call ensure_short(TB_Expwall, matter)
   na = size(Scell(NSC)%MDAtoms)
   allocate(Epot(na))
   call check(xtb_short_energy(ctx, a, Epot), 'xtb_short_energy')
   do i = 1, na
      Scell(NSC)%MDAtoms(i)%Epot = Scell(NSC)%MDAtoms(i)%Epot + Epot(i)      ! Exponential_wall.f90:120 (the half is inside)
   enddo
end subroutine get_short_range_rep_s

subroutine d_Short_range_pot_s(Scell, NSC, matter, TB_Expwall, numpar)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(solid), intent(in), target :: matter
   type(TB_Short_Rep), dimension(:,:), intent(in) :: TB_Expwall
   type(Numerics_param), intent(inout) :: numpar
   real(c_double), allocatable :: F(:,:)
   integer :: k, na
   call ensure_short(TB_Expwall, matter)
   na = size(Scell(NSC)%MDatoms)
   allocate(F(3,na))
   call check(xtb_short_forces(ctx, F), 'xtb_short_forces')
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%rep(:) = Scell(NSC)%MDatoms(k)%forces%rep(:) + F(:,k)    ! :646
   enddo
end subroutine d_Short_range_pot_s

subroutine d_Short_range_Pressure_s(Scell, NSC, TB_Expwall, matter, numpar)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(TB_Short_Rep), dimension(:,:), intent(in) :: TB_Expwall
   type(solid), intent(in), target :: matter
   type(Numerics_param), intent(in) :: numpar
   real(c_double) :: P(3,3)
   if (numpar%p_const) then        ! :677
      call ensure_short(TB_Expwall, matter)
      call check(xtb_short_pressure(ctx, P), 'xtb_short_pressure')
      Scell(NSC)%SCforce%rep = Scell(NSC)%SCforce%rep + P                    ! :715
   endif
end subroutine d_Short_range_Pressure_s

! ---- van der Waals term (Van_der_Waals.f90:132,158; vdW_forces TB.f90:3599 -> Construct_B, d_Forces_s, d_Forces_Pressure) --------
subroutine push_params_vdw(TB)
   class(TB_vdW), dimension(:,:), allocatable, intent(in) :: TB
   type(xtb_vdw_pair), allocatable, target :: p(:)
   integer :: nk, i, j, k
   integer(c_int) :: kind
   nk = size(TB,1)
   allocate(p(nk*nk))
   kind = 0_c_int
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)
         p(k) = xtb_vdw_pair(0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, &
                             0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0, 0.0d0)
         p(k)%d0_cut = TB(i,j)%d0_cut ; p(k)%dd_cut = TB(i,j)%dd_cut
      enddo
   enddo
   select type (TB)
   type is (TB_vdW_Girifalco)
      kind = 1_c_int
      do j = 1, nk
         do i = 1, nk
            k = i + nk*(j-1)
            p(k)%C12 = TB(i,j)%C12 ; p(k)%C6 = TB(i,j)%C6 ; p(k)%dm = TB(i,j)%dm ; p(k)%d_cut = TB(i,j)%d_cut
            p(k)%a = TB(i,j)%a ; p(k)%b = TB(i,j)%b ; p(k)%c = TB(i,j)%c ; p(k)%d = TB(i,j)%d
            p(k)%dsm = TB(i,j)%dsm ; p(k)%ds_cut = TB(i,j)%ds_cut
            p(k)%as = TB(i,j)%as ; p(k)%bs = TB(i,j)%bs ; p(k)%cs = TB(i,j)%cs
            p(k)%ds = TB(i,j)%ds ; p(k)%es = TB(i,j)%es ; p(k)%fs = TB(i,j)%fs
         enddo
      enddo
   type is (TB_vdW_LJ_cut)
      kind = 2_c_int
      do j = 1, nk
         do i = 1, nk
            k = i + nk*(j-1)
            p(k)%eps = TB(i,j)%eps ; p(k)%r0 = TB(i,j)%r0 ; p(k)%n = TB(i,j)%n
            p(k)%d0_short = TB(i,j)%d0_short ; p(k)%dd_short = TB(i,j)%dd_short
         enddo
      enddo
   type is (TB_vdW_ILJ_cut)
      kind = 3_c_int
      do j = 1, nk
         do i = 1, nk
            k = i + nk*(j-1)
            p(k)%eps = TB(i,j)%eps ; p(k)%r0 = TB(i,j)%r0 ; p(k)%n = TB(i,j)%n ; p(k)%m = TB(i,j)%m
            p(k)%d0_short = TB(i,j)%d0_short ; p(k)%dd_short = TB(i,j)%dd_short
         enddo
      enddo
   end select
   if (kind == 0_c_int) then        ! TB_vdW_Dumitrica: zero energy and zero force in the reference (Van_der_Waals.f90:598,621)
      call check(xtb_set_params_vdw(ctx, 0_c_int, 0_c_int, c_null_ptr), 'xtb_set_params_vdw')
   else
      call check(xtb_set_params_vdw(ctx, kind, int(nk,c_int), c_loc(p)), 'xtb_set_params_vdw')
   endif
   vdw_active = kind /= 0_c_int
end subroutine push_params_vdw

subroutine get_mirror_cell_num(Scell, NSC, numpar, atoms, Nx, Ny, Nz)
   use Objects, only : Atom
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(in) :: numpar
   type(Atom), dimension(:), intent(in) :: atoms
   integer, intent(out) :: Nx, Ny, Nz
   integer(c_int) :: N(3)
   real(c_double) :: R_cut
   call check(xtb_vdw_image_cells(ctx, N, R_cut), 'xtb_vdw_image_cells')
   Nx = N(1) ; Ny = N(2) ; Nz = N(3)
end subroutine get_mirror_cell_num

subroutine get_vdW_s(TB_Waals, Scell, NSC, numpar, a)
   class(TB_vdW), dimension(:,:), allocatable, intent(in) :: TB_Waals
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(8), intent(out) :: a
   real(c_double), allocatable :: Epot(:)
   integer :: i, na
   a = 0.0d0
   if (.not. vdw_active) return     ! Dumitrica class: zero
   na = size(Scell(NSC)%MDAtoms)
   allocate(Epot(na))
   call check(xtb_vdw_energy(ctx, a, Epot), 'xtb_vdw_energy')
   do i = 1, na
      Scell(NSC)%MDAtoms(i)%Epot = Scell(NSC)%MDAtoms(i)%Epot + Epot(i)      ! Van_der_Waals.f90:241 (the half is inside)
   enddo
end subroutine get_vdW_s

subroutine vdW_forces(TB_Waals, Scell, NSC, numpar)      ! TB.f90:3599-3633
   class(TB_vdW), allocatable, dimension(:,:), intent(in) :: TB_Waals
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(c_double), allocatable :: F(:,:)
   real(c_double) :: P(3,3)
   integer :: k, na
   if (.not. allocated(TB_Waals)) return
   if (.not. vdw_active) return
   na = size(Scell(NSC)%MDatoms)
   allocate(F(3,na))
   call check(xtb_vdw_forces(ctx, F), 'xtb_vdw_forces')
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%rep(:) = Scell(NSC)%MDatoms(k)%forces%rep(:) + F(:,k)    ! TB.f90:3777 (added to the repulsive part)
   enddo
   if (numpar%p_const) then         ! TB.f90:3800
      call check(xtb_vdw_pressure(ctx, P), 'xtb_vdw_pressure')
      Scell(NSC)%SCforce%rep = Scell(NSC)%SCforce%rep + P                                   ! TB.f90:3845
   endif
end subroutine vdW_forces

! ---- Coulomb term of a charged supercell (Coulomb.f90:454,715; Coulomb_forces TB.f90:3560 -> d_Forces_s, d_Forces_Pressure) --------
subroutine push_params_coulomb(TB)
   type(TB_Coulomb_cut), dimension(:,:), intent(in) :: TB
   type(xtb_coulomb_pair), allocatable, target :: p(:)
   integer :: nk, i, j, k
   nk = size(TB,1)
   allocate(p(nk*nk))
   do j = 1, nk
      do i = 1, nk
         k = i + nk*(j-1)
         p(k)%k = TB(i,j)%k ; p(k)%dm = TB(i,j)%dm ; p(k)%dd = TB(i,j)%dd
      enddo
   enddo
   call check(xtb_set_params_coulomb(ctx, int(nk,c_int), c_loc(p)), 'xtb_set_params_coulomb')
end subroutine push_params_coulomb

subroutine get_Coulomb_s(TB_Coul, Scell, NSC, numpar, a)
   type(TB_Coulomb_cut), dimension(:,:), intent(in) :: TB_Coul
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(8), intent(out) :: a
   real(c_double), allocatable :: Epot(:)
   integer :: i, na
   na = size(Scell(NSC)%MDAtoms)
   allocate(Epot(na))
   call check(xtb_set_charge(ctx, Scell(NSC)%Q), 'xtb_set_charge')           ! the mean unbalanced charge of THIS step
   call check(xtb_coulomb_energy(ctx, a, Epot), 'xtb_coulomb_energy')
   do i = 1, na
      Scell(NSC)%MDAtoms(i)%Epot = Scell(NSC)%MDAtoms(i)%Epot + Epot(i)      ! Coulomb.f90:543 (not halved, as written)
   enddo
end subroutine get_Coulomb_s

subroutine Coulomb_forces(TB_Coul, Scell, NSC, numpar)      ! TB.f90:3560-3596
   class(TB_Coulomb), allocatable, dimension(:,:), intent(in) :: TB_Coul
   type(Super_cell), dimension(:), intent(inout) :: Scell
   integer, intent(in) :: NSC
   type(Numerics_param), intent(inout) :: numpar
   real(c_double), allocatable :: F(:,:)
   real(c_double) :: P(3,3)
   integer :: k, na
   if (.not. allocated(TB_Coul)) return
   select type (TB_Coul)
   type is (TB_Coulomb_cut)
      na = size(Scell(NSC)%MDatoms)
      allocate(F(3,na))
      call check(xtb_set_charge(ctx, Scell(NSC)%Q), 'xtb_set_charge')
      call check(xtb_coulomb_forces(ctx, F), 'xtb_coulomb_forces')
      do k = 1, na
         Scell(NSC)%MDatoms(k)%forces%rep(:) = Scell(NSC)%MDatoms(k)%forces%rep(:) + F(:,k)    ! TB.f90:3777
      enddo
      if (numpar%p_const) then      ! TB.f90:3800
         call check(xtb_coulomb_pressure(ctx, P), 'xtb_coulomb_pressure')
         Scell(NSC)%SCforce%rep = Scell(NSC)%SCforce%rep + P                                   ! TB.f90:3845
      endif
   end select
end subroutine Coulomb_forces

! ---- Fermi occupations / velocity-Verlet halves on the device (Electron_tools.f90:2660,1913; Atomic_tools.f90:1897) ---------
subroutine xtb_b200_fermi(Netot, Te, fe, mu)
   real(8), intent(in) :: Netot, Te
   real(8), dimension(:), intent(out), target :: fe
   real(8), intent(out), target :: mu
   call check(xtb_fermi_occupations(ctx, Netot, Te, c_loc(fe), c_loc(mu), c_null_ptr), 'xtb_fermi_occupations')
   density_current = .false.
end subroutine xtb_b200_fermi

subroutine xtb_b200_md_set_velocities(SV)
   real(8), dimension(:,:), intent(in) :: SV       ! (3,na) relative velocities
   call check(xtb_md_set_velocities(ctx, SV), 'xtb_md_set_velocities')
end subroutine xtb_b200_md_set_velocities

subroutine xtb_b200_md_half_step(ind, dt, mass_kg, force_terms, Vsupce)
   integer, intent(in) :: ind, force_terms
   real(8), intent(in) :: dt
   real(8), dimension(:), intent(in) :: mass_kg
   real(8), dimension(3,3), intent(in), target, optional :: Vsupce
   if (present(Vsupce)) then
      call check(xtb_md_half_step(ctx, int(ind,c_int), dt, mass_kg, c_loc(Vsupce), int(force_terms,c_int)), 'xtb_md_half_step')
   else
      call check(xtb_md_half_step(ctx, int(ind,c_int), dt, mass_kg, c_null_ptr, int(force_terms,c_int)), 'xtb_md_half_step')
   endif
   if (ind == 2) then
      eigen_current = .false. ; density_current = .false.
   endif
end subroutine xtb_b200_md_half_step

subroutine xtb_b200_md_get_state(S, SV)
   real(8), dimension(:,:), intent(out), target :: S, SV
   call check(xtb_md_get_state(ctx, c_loc(S), c_loc(SV), c_null_ptr, c_null_ptr), 'xtb_md_get_state')
end subroutine xtb_b200_md_get_state

! ---- the whole step as one call (include/xtb_b200.h: xtb_step).  io is filled by the caller with c_loc of its arrays. ----------
subroutine xtb_b200_step(flags, io, Err)
   integer, intent(in) :: flags
   type(xtb_step_io), intent(in) :: io
   type(Error_handling), intent(inout) :: Err
   call check(xtb_step(ctx, int(flags,c_int), io), 'xtb_step', Err)
   eigen_current = .true. ; density_current = .false.
end subroutine xtb_b200_step

! ---- MPI build (TB_NRL.f90:267-327, MPI_subroutines.f90:3965-3982) ---------------------------------
! Each rank owns a contiguous range of atoms (xtb_b200_comm_init, called from xtb_b200_init).  The exchanges between the
! stages run INSIDE the library over NCCL: xtb_b200_exchange_blocks(1) after xtb_b200_assemble_blocks (H, S block segments,
! all-gather), xtb_b200_exchange_blocks(2) after Construct_Aij (rho, rho_E: halo only) followed by xtb_b200_density_mirror,
! xtb_b200_exchange_forces after the force wrapper, xtb_b200_allreduce_scforce after the pressure wrapper; see INTEGRATION.md.
subroutine xtb_b200_exchange_blocks(which)
   integer, intent(in) :: which      ! 1 = H and S blocks, 2 = rho and rho_E blocks
   call check(xtb_exchange_blocks(ctx, int(which,c_int)), 'xtb_exchange_blocks')
end subroutine xtb_b200_exchange_blocks

subroutine xtb_b200_exchange_forces(Scell, NSC)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   real(c_double), allocatable, target :: F(:,:)
   integer :: k, na
   na = size(Scell(NSC)%MDatoms)
   allocate(F(3,na))
   call check(xtb_exchange_forces(ctx, c_loc(F)), 'xtb_exchange_forces')
   do k = 1, na
      Scell(NSC)%MDatoms(k)%forces%att(:) = F(:,k)
   enddo
end subroutine xtb_b200_exchange_forces

subroutine xtb_b200_allreduce_scforce(Scell, NSC)
   type(Super_cell), dimension(:), intent(inout), target :: Scell
   integer, intent(in) :: NSC
   call check(xtb_allreduce_scforce(ctx, c_loc(Scell(NSC)%SCforce%att)), 'xtb_allreduce_scforce')
end subroutine xtb_b200_allreduce_scforce

subroutine xtb_b200_set_partition(first, last, n_ranks)
   integer, intent(in) :: first, last
   integer, intent(in), optional :: n_ranks      ! equal-capacity segments: the exchange is then one in-place all-gather
   if (present(n_ranks)) call check(xtb_set_block_layout(ctx, int(n_ranks,c_int)), 'xtb_set_block_layout')
   call check(xtb_set_partition(ctx, int(first,c_int), int(last,c_int)), 'xtb_set_partition')
end subroutine xtb_b200_set_partition

subroutine xtb_b200_assemble_blocks()
   call check(xtb_assemble_blocks(ctx), 'xtb_assemble_blocks')
end subroutine xtb_b200_assemble_blocks

subroutine xtb_b200_expand_dense(first, last, H_non, Sij)
   integer, intent(in) :: first, last
   real(8), dimension(:,:), intent(inout), target :: H_non, Sij
   call check(xtb_expand_dense(ctx, int(first,c_int), int(last,c_int), c_loc(H_non), c_loc(Sij)), 'xtb_expand_dense')
end subroutine xtb_b200_expand_dense

subroutine xtb_b200_density_mirror()
   call check(xtb_density_mirror(ctx), 'xtb_density_mirror')
end subroutine xtb_b200_density_mirror

! on-device eigensolve (Loewdin + Dsyevd + back-transform as TB_NRL.f90:985-1060,402-421) for device-resident steps
subroutine xtb_b200_eigensolve(Ei, eigen_S, Ha)
   real(8), dimension(:), intent(out), target :: Ei, eigen_S
   real(8), dimension(:,:), intent(out), target, optional :: Ha
   if (present(Ha)) then
      call check(xtb_eigensolve(ctx, c_loc(Ei), c_loc(eigen_S), c_loc(Ha)), 'xtb_eigensolve')
   else
      call check(xtb_eigensolve(ctx, c_loc(Ei), c_loc(eigen_S), c_null_ptr), 'xtb_eigensolve')
   endif
   eigen_current = .true. ; density_current = .false. ; ha_loc = c_null_ptr
end subroutine xtb_b200_eigensolve

end module XTANT_B200_shim
