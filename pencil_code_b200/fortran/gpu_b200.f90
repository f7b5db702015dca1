! $Id$
!
! MODULE_DOC: B200-native GPU module: per-substep RHS + 2N-RK3 update on sm_100a through libpcb200.so.
! MODULE_DOC: Drop-in alternative to nogpu.f90 / gpu_astaroth.f90; selected by GPU=gpu_b200 in src/Makefile.local.
!
! CPARAM logical, parameter :: lgpu = .true.
!
!**************************************************************************
!
!  Same public interface as src/gpu_astaroth.f90 (list in gpu.h).  Fortran talks to CUDA only
!  through ISO_C_BINDING interfaces of the C ABI in include/pcb200.h.  Physics parameters that
!  are private to their modules (nu, eta, chi, cs20, ...) are fetched by the C shim
!  gpu_b200_ansi.c through the existing pushpars2c address tables, because module GPU is used
!  by Sub and Hydro and therefore cannot `use` the physics modules itself.
!
module GPU
!
  use Cdata
  use General, only: keep_compiler_quiet, lpointer, ioptest, loptest
  use Messages
  use iso_c_binding

  implicit none

  include 'gpu.h'

  ! struct pcb200_config of include/pcb200.h (field order must match)
  type, bind(C) :: pcb200_config
    integer(c_int) :: nx, ny, nz
    integer(c_int) :: nprocx, nprocy, nprocz
    integer(c_int) :: ipx, ipy, ipz
    integer(c_int) :: nvar
    integer(c_int) :: iux, ilnrho, iss, iax
    real(c_double) :: dx, dy, dz
    real(c_double) :: x0, y0, z0
    real(c_double) :: cs20, gamma, cp, lnrho0
    real(c_double) :: mu0
    real(c_double) :: nu, nu_hyper3
    integer(c_int) :: lvisc_nu_const, lvisc_hyper3_simplified, lvisc_hyper3_nu_const
    real(c_double) :: diffrho_hyper3
    integer(c_int) :: ldiff_hyper3lnrho
    real(c_double) :: chi
    integer(c_int) :: lheatc_chiconst, lviscosity_heat
    real(c_double) :: eta
    integer(c_int) :: lresi_eta_const, lohmic_heat, llorentzforce
    integer(c_int) :: lupw_uu, lupw_lnrho, lupw_ss, lupw_aa
    real(c_double) :: cdt, cdtv, cdtv2, cdtv3
    real(c_double) :: dtmax
    integer(c_int) :: lperi(3)
    integer(c_int) :: device, kernel_variant
    integer(c_int) :: bcz(8,2)          ! C: bcz[2][8] (bottom/top, variable slot)
    real(c_double) :: gravz, Omega
    integer(c_int) :: lheatc_Kprof
    real(c_double) :: hcond0, hcond1, hcond2, widthss, z1, z2
    real(c_double) :: cool, wcool, cs2cool, ztop
    real(c_double) :: FbotKbot, FtopKtop
    real(c_double) :: cs2bot, cs2top
  endtype pcb200_config

  ! PCB200_BC_* of include/pcb200.h
  integer, parameter :: PCB200_BC_P=0, PCB200_BC_S=1, PCB200_BC_A=2, PCB200_BC_A2=3, PCB200_BC_C1=4, PCB200_BC_CT=5

  interface
    subroutine pcb200_configure(cfg) bind(C,name="pcb200_configure")
      import :: pcb200_config
      type(pcb200_config), intent(in) :: cfg
    endsubroutine
    subroutine pcb200_fill_physics_pars_c(cfg) bind(C,name="pcb200_fill_physics_pars_c")
      import :: pcb200_config
      type(pcb200_config), intent(inout) :: cfg
    endsubroutine
    function pcb200_fill_strat_pars_c(cfg) bind(C,name="pcb200_fill_strat_pars_c") result(ok)
      import :: pcb200_config, c_int
      type(pcb200_config), intent(inout) :: cfg
      integer(c_int) :: ok
    endfunction
    subroutine pcb200_set_dt(dt,ldt) bind(C,name="pcb200_set_dt")
      import :: c_double, c_int
      real(c_double), intent(in) :: dt
      integer(c_int), intent(in) :: ldt
    endsubroutine
    subroutine pcb200_get_dt(dt) bind(C,name="pcb200_get_dt")
      import :: c_double
      real(c_double), intent(out) :: dt
    endsubroutine
    subroutine pcb200_hand_dt1_to_set_dt_c() bind(C,name="pcb200_hand_dt1_to_set_dt_c")
    endsubroutine
    function pcb200_forcing_supported_c(nx_,my_,mz_) bind(C,name="pcb200_forcing_supported_c") result(ok)
      import :: c_int
      integer(c_int), intent(in) :: nx_, my_, mz_
      integer(c_int) :: ok
    endfunction
    subroutine pcb200_reconfigure(cfg) bind(C,name="pcb200_reconfigure")
      import :: pcb200_config
      type(pcb200_config), intent(in) :: cfg
    endsubroutine
    subroutine pcb200_nccl_get_unique_id(id) bind(C,name="pcb200_nccl_get_unique_id")
      import :: c_char
      character(kind=c_char), dimension(128), intent(out) :: id
    endsubroutine
    subroutine pcb200_nccl_init(id,rank,nranks) bind(C,name="pcb200_nccl_init")
      import :: c_char, c_int
      character(kind=c_char), dimension(128), intent(in) :: id
      integer(c_int), intent(in) :: rank, nranks
    endsubroutine
    function pcb200_p2p_blob_bytes() bind(C,name="pcb200_p2p_blob_bytes") result(n)
      import :: c_int
      integer(c_int) :: n
    endfunction
    subroutine pcb200_p2p_export(blob) bind(C,name="pcb200_p2p_export")
      import :: c_char
      character(kind=c_char), dimension(*), intent(out) :: blob
    endsubroutine
    function pcb200_p2p_connect(blobs,nranks) bind(C,name="pcb200_p2p_connect") result(ierr)
      import :: c_char, c_int
      character(kind=c_char), dimension(*), intent(in) :: blobs
      integer(c_int), intent(in) :: nranks
      integer(c_int) :: ierr
    endfunction
    subroutine pcb200_p2p_enable() bind(C,name="pcb200_p2p_enable")
    endsubroutine
    subroutine pcb200_request_diagnostics() bind(C,name="pcb200_request_diagnostics")
    endsubroutine
    subroutine pcb200_push_forcing_c() bind(C,name="pcb200_push_forcing_c")
    endsubroutine
    subroutine initialize_gpu_c(f,comm,t,nt,lread_all,lcpu_dt,lsparse) bind(C,name="initialize_gpu_c")
      import :: c_double, c_int
      real(c_double), dimension(*) :: f
      integer(c_int) :: comm, nt, lread_all, lcpu_dt, lsparse
      real(c_double) :: t
    endsubroutine
    subroutine register_gpu_c() bind(C,name="register_gpu_c")
    endsubroutine
    subroutine finalize_gpu_c() bind(C,name="finalize_gpu_c")
    endsubroutine
    subroutine rhs_gpu_c(isubstep,t) bind(C,name="rhs_gpu_c")
      import :: c_double, c_int
      integer(c_int) :: isubstep
      real(c_double) :: t
    endsubroutine
    subroutine before_boundary_gpu_c(lrmv,isubstep,t,lsubstepping) bind(C,name="before_boundary_gpu_c")
      import :: c_double, c_int
      integer(c_int) :: lrmv, isubstep, lsubstepping
      real(c_double) :: t
    endsubroutine
    subroutine update_after_substep_gpu_c() bind(C,name="update_after_substep_gpu_c")
    endsubroutine
    subroutine load_farray_c() bind(C,name="load_farray_c")
    endsubroutine
    subroutine copy_farray_c(f) bind(C,name="copy_farray_c")
      import :: c_double
      real(c_double), dimension(*) :: f
    endsubroutine
    subroutine gpu_prepare_for_first_substep_c(t) bind(C,name="gpu_prepare_for_first_substep_c")
      import :: c_double
      real(c_double) :: t
    endsubroutine
    subroutine get_farray_ptr_gpu_c(p_in,p_out) bind(C,name="get_farray_ptr_gpu_c")
      import :: c_ptr
      type(c_ptr) :: p_in, p_out
    endsubroutine
    subroutine reload_gpu_config_c() bind(C,name="reload_gpu_config_c")
    endsubroutine
    integer(c_int) function update_on_gpu_scal_by_name_c(name,value) bind(C,name="update_on_gpu_scal_by_name_c")
      import :: c_char, c_double, c_int
      character(kind=c_char), dimension(*) :: name
      real(c_double) :: value
    endfunction
    subroutine update_on_gpu_scal_by_ind_c(index,value) bind(C,name="update_on_gpu_scal_by_ind_c")
      import :: c_double, c_int
      integer(c_int) :: index
      real(c_double) :: value
    endsubroutine
    type(c_ptr) function pos_real_ptr_c(ptr,ind) bind(C,name="pos_real_ptr_c")
      import :: c_ptr, c_int
      type(c_ptr) :: ptr
      integer(c_int) :: ind
    endfunction
    subroutine get_gpu_reduced_vars_c(dst) bind(C,name="get_gpu_reduced_vars_c")
      import :: c_double
      real(c_double), dimension(20) :: dst
    endsubroutine
    subroutine split_update_gpu_c(f) bind(C,name="split_update_gpu_c")
      import :: c_double
      real(c_double), dimension(*) :: f
    endsubroutine
    subroutine test_bcs_c() bind(C,name="test_bcs_c")
    endsubroutine
  endinterface

  type(C_PTR) :: pFarr_GPU_in, pFarr_GPU_out

  logical :: ltest_bcs=.false.     !PAR-DOC: accepted for namelist compatibility with gpu_astaroth
  logical :: ltest_rhs=.false.     !PAR-DOC: run test_rhs_gpu (equ.f90:1974-2143) at step it_test_rhs
  integer :: it_test_rhs=1
  integer :: kernel_variant=0      !PAR-DOC: 0=auto (10 on isotropic grids, else 4; 1 for odd mx or stratified runs), 1=global-load kernel, 4/5/10=TMA z-marching kernels (include/pcb200.h)
  integer :: gpu_device=-1         !PAR-DOC: CUDA device; -1 = local rank modulo device count

  logical :: lp2p_halo=.true.      !PAR-DOC: z halos and y faces as peer-memory puts over NVLink (else NCCL send/recv)
  namelist /gpu_run_pars/ ltest_bcs, ltest_rhs, it_test_rhs, kernel_variant, gpu_device, lp2p_halo

  type(pcb200_config), save :: cfg

contains
!***********************************************************************
    subroutine register_GPU

      call register_gpu_c

    endsubroutine register_GPU
!***********************************************************************
    subroutine initialize_GPU(f)
!
!  Refuses everything the module does not implement (cf. gpu_astaroth.f90:211-228), hands sizes
!  and parameters to the library, bootstraps NCCL over MPI and uploads f.
!
      use Mpicomm, only: mpibcast_char, MPI_COMM_PENCIL

      real, dimension(:,:,:,:), intent(IN) :: f

      character(LEN=512) :: str
      character(kind=c_char), dimension(128) :: nccl_id
      character(LEN=128) :: idbuf
      integer :: i, j, zero
      integer(c_int) :: rank, nranks

      str=''
      if (.not.lhydro)   str=trim(str)//', '//'nohydro'
      if (.not.ldensity) str=trim(str)//', '//'nodensity'
      if (ldensity_nolog) str=trim(str)//', '//'density (non-logarithmic)'
      if (ltemperature)  str=trim(str)//', '//'temperature'
      if (lanelastic)    str=trim(str)//', '//'anelastic'
      if (lboussinesq)   str=trim(str)//', '//'boussinesq'
      if (lshear)        str=trim(str)//', '//'shear'
      if (lshock)        str=trim(str)//', '//'shock'
      if (lradiation)    str=trim(str)//', '//'radiation'
      if (lchemistry)    str=trim(str)//', '//'chemistry'
      if (lpscalar)      str=trim(str)//', '//'pscalar'
      if (ldustvelocity.or.ldustdensity) str=trim(str)//', '//'dust'
      if (lcosmicray)    str=trim(str)//', '//'cosmicray'
      if (ltestscalar)   str=trim(str)//', '//'testscalar'
      if (ltestfield)    str=trim(str)//', '//'testfield'
      if (ltestflow)     str=trim(str)//', '//'testflow'
      if (lparticles)    str=trim(str)//', '//'particles'
      if (lpointmasses)  str=trim(str)//', '//'pointmasses'
      if (lsolid_cells)  str=trim(str)//', '//'solid_cells'
      if (lspecial)      str=trim(str)//', '//'special'
      if (str/='') call fatal_error('initialize_GPU','no gpu_b200 implementation available for module(s) "'// &
                                    trim(str(3:))//'"')
      if (.not.lcartesian_coords) call fatal_error('initialize_GPU','gpu_b200: Cartesian coordinates only')
      if (.not.all(lequidist))    call fatal_error('initialize_GPU','gpu_b200: equidistant grids only')
      if (.not.all(lperi(1:2)))   call fatal_error('initialize_GPU','gpu_b200: x and y must be periodic')
      if (.not.lperi(3).and.(nprocy/=1.or.nprocz/=1)) &
                                  call fatal_error('initialize_GPU','gpu_b200: closed z boundaries need nprocy=nprocz=1')
      if (Omega/=0..and.theta/=0.) call fatal_error('initialize_GPU','gpu_b200: Coriolis force for theta=0 only')
      if (nprocx/=1)              call fatal_error('initialize_GPU','gpu_b200: nprocx must be 1')
      if (itorder/=3)             call fatal_error('initialize_GPU','gpu_b200: itorder=3 only')
      if (ddt>0.)                 call fatal_error('initialize_GPU','gpu_b200: time-step growth limiter (ddt>0) not supported')
      if (loutput_varn_at_exact_tsnap) &
                                  call fatal_error('initialize_GPU','gpu_b200: loutput_varn_at_exact_tsnap (shift_dt) not supported')
!
!  Forcing: run.f90:575 skips addforce when lgpu, so anything but the delta-correlated helical forcing with its
!  default options (no profiles, no second force, no continuous forcing, no momentum forcing) would be dropped
!  silently -> refuse it (the switches are private to Forcing: read through its pushpars2c table by the C shim).
!
      if (lforcing) then
        if (pcb200_forcing_supported_c(nx,my,mz)==0) call fatal_error('initialize_GPU', &
            "gpu_b200: only iforce='helical' without profiles, iforce2, lforcing_cont, lmomentum_ff, iforcing_zsym")
      endif
!
      cfg%nx=nx; cfg%ny=ny; cfg%nz=nz
      cfg%nprocx=nprocx; cfg%nprocy=nprocy; cfg%nprocz=nprocz
      cfg%ipx=ipx; cfg%ipy=ipy; cfg%ipz=ipz
      cfg%nvar=mvar
      cfg%iux=iux-1; cfg%ilnrho=ilnrho-1
      cfg%iss=-1; if (lentropy)  cfg%iss=iss-1
      cfg%iax=-1; if (lmagnetic) cfg%iax=iax-1
      cfg%dx=dx; cfg%dy=dy; cfg%dz=dz
      cfg%x0=x(l1); cfg%y0=y(m1); cfg%z0=z(n1)
      cfg%mu0=mu0
      cfg%cdt=cdt; cfg%cdtv=cdtv; cfg%cdtv2=cdtv2; cfg%cdtv3=cdtv3
      cfg%dtmax=dtmax
      do i=1,3; cfg%lperi(i)=merge(1,0,lperi(i)); enddo
      cfg%device=gpu_device
      cfg%kernel_variant=kernel_variant
!
!  z boundary conditions (bcz12 of Cdata, boundcond.f90:1421-1787) and rotation (Omega, theta of Cdata).
!
      cfg%bcz=PCB200_BC_P
      if (.not.lperi(3)) then
        do j=1,2; do i=1,mvar
          select case (bcz12(i,j))
          case ('s');  cfg%bcz(i,j)=PCB200_BC_S
          case ('a');  cfg%bcz(i,j)=PCB200_BC_A
          case ('a2'); cfg%bcz(i,j)=PCB200_BC_A2
          case ('c1'); cfg%bcz(i,j)=PCB200_BC_C1
          case ('cT'); cfg%bcz(i,j)=PCB200_BC_CT
          case default
            call fatal_error('initialize_GPU','gpu_b200: bcz="'//trim(bcz12(i,j))//'" not supported (s, a, a2, c1, cT)')
          endselect
        enddo; enddo
      endif
      cfg%Omega=Omega
      cfg%ztop=xyz1(3)
!
!  nu, eta, chi, eos constants, upwinding and heating switches, gravity, K-profile conduction and the cooling layer:
!  via pushpars2c tables (C shim; it refuses gravity / heating set-ups other than the ones implemented).
!
      call pcb200_fill_physics_pars_c(cfg)
      if (lgrav) then
        if (pcb200_fill_strat_pars_c(cfg)==0) call fatal_error('initialize_GPU', &
            "gpu_b200: gravity only as gravz_profile='const' (no x, y gravity); heating/cooling only as the 'gaussian' "// &
            "cooling layer (luminosity=0, lcooling_general=F); conduction only 'chi-const' or 'K-profile' (lmultilayer=T)")
      endif
      call pcb200_configure(cfg)
!
!  One rank <-> one GPU; NCCL id is created on root and broadcast over MPI.
!
      if (ncpus>1) then
        if (lroot) call pcb200_nccl_get_unique_id(nccl_id)
        do i=1,128; idbuf(i:i)=nccl_id(i); enddo
        call mpibcast_char(idbuf,root,MPI_COMM_PENCIL)
        do i=1,128; nccl_id(i)=idbuf(i:i); enddo
        rank=iproc; nranks=ncpus
        call pcb200_nccl_init(nccl_id,rank,nranks)
      endif
!
      if (dt<=0.) dt=dtmin
      zero=0
      call initialize_gpu_c(f,MPI_COMM_PENCIL,t,nt,zero,zero,zero)
!
!  z halos as copy-engine puts into the neighbours' ghost planes (NVLink peer memory): every rank exports
!  the CUDA IPC handles of its device buffers, the table of all ranks is all-gathered over MPI.
!  (All ranks of the run share one 8-GPU box; without this block the z exchange stays on NCCL.)
!
      if (ncpus>1 .and. (nprocz>1.or.nprocy>1) .and. lp2p_halo) call connect_p2p_halo
      if (nt>0) call load_farray_to_GPU(f)
      call get_farray_ptr_gpu

    endsubroutine initialize_GPU
!***********************************************************************
    subroutine connect_p2p_halo
!
!  MPI_Allgather of the per-rank IPC blobs (pcb200_p2p_export) -> pcb200_p2p_connect.
!
      use Mpicomm, only: MPI_COMM_PENCIL
      include 'mpif.h'

      character(kind=c_char), dimension(:), allocatable :: blob, table
      integer(c_int) :: nb, nranks
      integer :: ierr, lfail, lfail_any

      nb=pcb200_p2p_blob_bytes()
      allocate(blob(nb),table(nb*ncpus))
      call pcb200_p2p_export(blob)
      call MPI_ALLGATHER(blob,nb,MPI_CHARACTER,table,nb,MPI_CHARACTER,MPI_COMM_PENCIL,ierr)
      if (ierr/=0) call fatal_error('connect_p2p_halo','MPI_ALLGATHER failed')
      nranks=ncpus
      lfail=pcb200_p2p_connect(table,nranks)
!
!  All ranks switch together, or none does (a rank that cannot map its neighbour keeps everybody on NCCL).
!
      call MPI_ALLREDUCE(lfail,lfail_any,1,MPI_INTEGER,MPI_MAX,MPI_COMM_PENCIL,ierr)
      if (lfail_any==0) then
        call pcb200_p2p_enable
      elseif (lroot) then
        call warning('connect_p2p_halo','peer mapping failed on some rank; z halo stays on NCCL')
      endif
      deallocate(blob,table)

    endsubroutine connect_p2p_halo
!***********************************************************************
    subroutine read_gpu_run_pars(iomsg)

      use File_io, only: parallel_unit

      character(LEN=*), intent(out) :: iomsg
      integer :: iostat

      read(parallel_unit, NML=gpu_run_pars, IOSTAT=iostat, IOMSG=iomsg)
      if (iostat==0) iomsg=""

    endsubroutine read_gpu_run_pars
!***********************************************************************
    subroutine write_gpu_run_pars(unit)

      integer, intent(in) :: unit

      write(unit, NML=gpu_run_pars)

    endsubroutine write_gpu_run_pars
!***********************************************************************
    subroutine finalize_GPU

      use Mpicomm, only: mpibarrier
!
!  No rank may free ghost planes that a neighbour is still writing to (p2p halo).
!
      call mpibarrier
      call finalize_gpu_c

    endsubroutine finalize_GPU
!***********************************************************************
    subroutine get_farray_ptr_gpu

      call get_farray_ptr_gpu_c(pFarr_GPU_in,pFarr_GPU_out)

    endsubroutine get_farray_ptr_gpu
!***********************************************************************
    subroutine rhs_GPU(f,isubstep)
!
!  One complete substep on the device: ghost exchange + RHS + 2N-RK update (+ forcing after the
!  last substep).  With ldt the Courant step is reduced on the device inside substep 1 (exact CPU
!  semantics, equ.f90:408) and handed back so that advance_substeps sees it (timestep.f90:150).
!
      real, dimension (mx,my,mz,mfarray), intent(INOUT) :: f
      integer,                            intent(IN)    :: isubstep

      integer(c_int) :: ldt_int, isub
      real(c_double) :: dt_c

      call keep_compiler_quiet(f)

      isub=isubstep
      if (isubstep==1) then
        ldt_int=merge(1,0,ldt)
        dt_c=dt
        call pcb200_set_dt(dt_c,ldt_int)
        if (lout) call pcb200_request_diagnostics
      endif
      if (isubstep==itorder .and. lforcing) call pcb200_push_forcing_c   ! astaroth/forcing.h:1-36
      call rhs_gpu_c(isub,t)
      if (isubstep==1 .and. ldt) then
!
!  The device has reduced dt1 over all cells and all GPUs and used dt=1/dt1 in this substep.  Pencil's own
!  bookkeeping goes through set_dt (sub.f90:4223-4254, as gpu_astaroth.cc:1617-1634 does): with lgpu it skips the
!  MPI reduction and forms dt=1./dt1 -- the same double-precision division as on the device, so both sides
!  integrate with bit-identical dt.  The growth limiter (ddt>0) and shift_dt would change dt on the host only and
!  are refused in initialize_GPU.
!
        call pcb200_hand_dt1_to_set_dt_c   ! C shim: pcb200_get_dt1_max -> Sub::set_dt (module GPU cannot use Sub)
      endif

    endsubroutine rhs_GPU
!***********************************************************************
    subroutine before_boundary_gpu(f,lrmv,isubstep,t)

      real, dimension (mx,my,mz,mfarray), intent(INOUT) :: f
      integer,                            intent(IN)    :: isubstep
      real(KIND=rkind8), intent(IN) :: t
      logical :: lrmv

      integer(c_int) :: lrmv_int, lsub_int, isub
      real(c_double) :: t_c

      call keep_compiler_quiet(f)
      lrmv_int=merge(1,0,lrmv); lsub_int=merge(1,0,lsubstepping_in_time); isub=isubstep; t_c=t
      call before_boundary_gpu_c(lrmv_int,isub,t_c,lsub_int)

    endsubroutine before_boundary_gpu
!***********************************************************************
    subroutine update_after_substep_gpu

      call update_after_substep_gpu_c

    endsubroutine update_after_substep_gpu
!***********************************************************************
    subroutine gpu_prepare_for_first_substep

      call gpu_prepare_for_first_substep_c(t)

    endsubroutine gpu_prepare_for_first_substep
!***********************************************************************
    function get_ptr_GPU(ind1,ind2,lout) result(pFarr)

      integer, optional :: ind1,ind2
      logical, optional :: lout
      real, dimension(:,:,:,:), pointer :: pFarr
      integer :: i1,i2
      integer(c_int) :: ind

      i1=ioptest(ind1,1); i2=ioptest(ind2,i1); ind=i1-1
      if (loptest(lout)) then
        call c_f_pointer(pos_real_ptr_c(pFarr_GPU_out,ind),pFarr,(/mx,my,mz,i2-i1+1/))
      else
        call c_f_pointer(pos_real_ptr_c(pFarr_GPU_in,ind),pFarr,(/mx,my,mz,i2-i1+1/))
      endif

    endfunction get_ptr_GPU
!***********************************************************************
    function get_ptr_GPU_training(ind1,ind2,lout) result(pFarr)

      integer :: ind1
      integer, optional :: ind2
      logical, optional :: lout
      real, dimension(:,:,:,:,:), pointer :: pFarr
      integer :: i2
      integer(c_int) :: ind

      i2=ioptest(ind2,ind1); ind=ind1-1
      if (loptest(lout)) then
        call c_f_pointer(pos_real_ptr_c(pFarr_GPU_out,ind),pFarr,(/mx,my,mz,i2-ind1+1,1/))
      else
        call c_f_pointer(pos_real_ptr_c(pFarr_GPU_in,ind),pFarr,(/mx,my,mz,i2-ind1+1,1/))
      endif

    endfunction get_ptr_GPU_training
!***********************************************************************
    subroutine copy_farray_from_GPU(f,nowait_)
!
!  Same protocol as gpu_astaroth.f90:390-412: with the OpenMP helper thread (MULTITHREADING=openmp) f may still
!  be read by perform_diagnostics, so the copy waits until the helper has finished (lhelper_perf=F) and is
!  done only once per diagnostic step (lfarray_copied); nowait_=T copies unconditionally.
!
!$    use General, only: signal_wait

      real, dimension (mx,my,mz,mfarray), intent(OUT) :: f
      logical, optional :: nowait_

      if (loptest(nowait_)) then
        call copy_farray_c(f)      ! synchronous; all mvar slots incl. valid ghost zones
        return
      endif
!
!$    if (lfarray_copied) return
!
!  Have to wait: the helper thread's diagnostics must not see a half-overwritten f.
!
!$    call signal_wait(lhelper_perf, .false.)
      call copy_farray_c(f)
!$    lfarray_copied = .true.

    endsubroutine copy_farray_from_GPU
!***********************************************************************
    subroutine load_farray_to_GPU(f)

      real, dimension (mx,my,mz,mfarray), intent(IN) :: f

      call keep_compiler_quiet(f)
      call load_farray_c           ! uses the pointer retained by initialize_gpu_c

    endsubroutine load_farray_to_GPU
!***********************************************************************
    subroutine reload_GPU_config
!
!  RELOAD: re-read the physics parameters from the modules and hand them to the library (sizes, decomposition and
!  variable slots must be unchanged: pcb200_reconfigure refuses anything else), then rebuild the device constants.
!
      call pcb200_fill_physics_pars_c(cfg)
      cfg%Omega=Omega
      if (lgrav) then
        if (pcb200_fill_strat_pars_c(cfg)==0) call fatal_error('reload_GPU_config', &
            'gpu_b200: the new gravity / heating / conduction set-up is not supported')
      endif
      cfg%cdt=cdt; cfg%cdtv=cdtv; cfg%cdtv2=cdtv2; cfg%cdtv3=cdtv3; cfg%dtmax=dtmax
      call pcb200_reconfigure(cfg)
      call reload_gpu_config_c

    endsubroutine reload_GPU_config
!***********************************************************************
    subroutine update_on_gpu_vec(index, varname, value)

      integer, intent(inout) :: index
      character(LEN=*), optional :: varname
      real, dimension(3), optional :: value

      call fatal_error('update_on_gpu_vec','gpu_b200 has no vector parameters')

    endsubroutine update_on_gpu_vec
!***********************************************************************
    subroutine update_on_gpu(index, varname, value)

      integer, intent(inout) :: index
      character(LEN=*), optional :: varname
      real, optional :: value

      integer(c_int) :: ind
      real(c_double) :: val

      if (.not.present(value)) call fatal_error('update_on_gpu','gpu_b200 has no array parameters')
      val=value
      if (index>=0) then
        ind=index
        call update_on_gpu_scal_by_ind_c(ind,val)
      else
        index=update_on_gpu_scal_by_name_c(trim(varname)//c_null_char,val)
        if (index<0) call fatal_error('update_on_gpu','variable "'//trim(varname)//'" not found')
      endif

    endsubroutine update_on_gpu
!***********************************************************************
    subroutine radtransfer_gpu

      call fatal_error('radtransfer_gpu','radiation is not supported by gpu_b200')

    endsubroutine radtransfer_gpu
!***********************************************************************
    subroutine get_gpu_reduced_vars(dst)

      real, dimension(20) :: dst

      call get_gpu_reduced_vars_c(dst)

    endsubroutine get_gpu_reduced_vars
!***********************************************************************
    subroutine test_gpu_bcs

      call test_bcs_c

    endsubroutine test_gpu_bcs
!***********************************************************************
    subroutine split_update_gpu(f)

      real, dimension (mx,my,mz,mfarray), intent(INOUT) :: f

      call split_update_gpu_c(f)

    endsubroutine split_update_gpu
!***********************************************************************
    subroutine pushpars2c(p_par)

      use Syscalls, only: copy_addr

      integer, parameter :: n_pars=20
      integer(KIND=ikind8), dimension(n_pars) :: p_par

      call copy_addr(ltest_bcs,p_par(7)) ! bool

    endsubroutine pushpars2c
!***********************************************************************
!  TorchFort hooks of gpu.h: not part of this path.
!***********************************************************************
    subroutine TF_train_gpu(model_name, loss, itsub, t)
      real :: loss
      integer :: itsub
      real(KIND=rkind8), intent(IN) :: t
      character(len=*), intent(in) :: model_name
      call fatal_error('TF_train_gpu','not available in gpu_b200')
    endsubroutine TF_train_gpu
!***********************************************************************
    subroutine TF_infer_gpu(model_name, flag)
      integer :: flag
      character(len=*), intent(in) :: model_name
      call fatal_error('TF_infer_gpu','not available in gpu_b200')
    endsubroutine TF_infer_gpu
!***********************************************************************
    subroutine TF_create_model(model_name, config_file_path, lmpicomm)
      logical :: lmpicomm
      character(len=*), intent(in) :: model_name, config_file_path
      call fatal_error('TF_create_model','not available in gpu_b200')
    endsubroutine TF_create_model
!***********************************************************************
    subroutine tau_snapshots()
    endsubroutine tau_snapshots
!***********************************************************************
    subroutine TF_load_model(model_name, fname)
      character(len=*), intent(in) :: model_name, fname
      call fatal_error('TF_load_model','not available in gpu_b200')
    endsubroutine TF_load_model
!***********************************************************************
    subroutine TF_load_model_checkpoint(model_name, checkpoint_dir)
      character(len=*), intent(in) :: model_name, checkpoint_dir
      call fatal_error('TF_load_model_checkpoint','not available in gpu_b200')
    endsubroutine TF_load_model_checkpoint
!***********************************************************************
    subroutine TF_save_model(model_name, fname)
      character(len=*), intent(in) :: model_name, fname
      call fatal_error('TF_save_model','not available in gpu_b200')
    endsubroutine TF_save_model
!***********************************************************************
    subroutine TF_save_checkpoint(model_name, checkpoint_dir)
      character(len=*), intent(in) :: model_name, checkpoint_dir
      call fatal_error('TF_save_checkpoint','not available in gpu_b200')
    endsubroutine TF_save_checkpoint
!***********************************************************************
endmodule GPU
