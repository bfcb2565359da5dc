!================================================================================!
! crest_cuda_mod.f90 — iso_c_binding shim for libcrest_cuda.so (include/crest_cuda.h)
!
! Drop-in for the hot section of cregen_CRE (crest: src/cregen.f90:1222-1376 and 1421-1469).
! It mirrors the way CREST already binds C code (src/symmetry2.f90:23-39 -> symmetry_i.c:1767):
! an interface block with bind(C), scalars by value, arrays as assumed-size, caller-owned memory.
!
! NOT compiled in the build container (no Fortran compiler there); build line for a CREST tree:
!   gfortran -c fortran/crest_cuda_mod.f90 && link crest with -L<repo>/crest_b200 -lcrest_cuda
! See INTEGRATION.md for the patch to src/cregen.f90.
!================================================================================!
module crest_cuda
  use iso_c_binding
  implicit none
  private
  public :: crest_cuda_thresholds
  public :: crest_cuda_start, crest_cuda_stop
  public :: cregen_dup_cuda, cregen_group_cuda
  public :: compare_blocks_cuda, intraconf_cuda

  !> the scalars of type(systemdata) that cross the boundary (src/classes.f90:332-345)
  type, bind(C) :: crest_cuda_thresholds
    real(c_double) :: ethr      !< env%ethr     kcal/mol
    real(c_double) :: rthr      !< env%rthr     Angstrom
    real(c_double) :: bthr      !< env%bthr2
    real(c_double) :: bthrmax   !< env%bthrmax
    real(c_double) :: bthrshift !< env%bthrshift
  end type crest_cuda_thresholds

  interface
    !> engine of 1-8 GPUs driven by this (single) host thread; ngpu_requested <= 0: all visible GPUs
    function crest_cuda_init(ngpu_requested,ngpu_used) result(rc) bind(C,name="crest_cuda_init")
      import
      integer(c_int),intent(in),value :: ngpu_requested
      integer(c_int),intent(out) :: ngpu_used
      integer(c_int) :: rc
    end function crest_cuda_init

    !> releases every device buffer, stream and pinned staging buffer of the engine
    subroutine crest_cuda_finalize() bind(C,name="crest_cuda_finalize")
    end subroutine crest_cuda_finalize

    !> replaces cregen.f90:1222-1376 (axis + rot, pair RMSD incl. mirror, duplicate sweep)
    function crest_cuda_cregen_dup(nat,nall,at,xyz,er,rmsd_mask,rot_from_mask,enantio, &
        &                          thr,rot,xyz_axis,dup,npairs) result(rc) &
        &    bind(C,name="crest_cuda_cregen_dup")
      import
      integer(c_int),intent(in),value :: nat,nall
      integer(c_int),intent(in) :: at(*)
      real(c_double),intent(in) :: xyz(3,*)
      real(c_double),intent(in) :: er(*)
      type(c_ptr),value :: rmsd_mask          !< int(nat) or c_null_ptr = all atoms
      integer(c_int),intent(in),value :: rot_from_mask,enantio
      type(crest_cuda_thresholds),intent(in) :: thr
      real(c_double),intent(out) :: rot(3,*)
      real(c_double),intent(out) :: xyz_axis(3,*)
      integer(c_int),intent(out) :: dup(*)
      integer(c_long_long),intent(out) :: npairs
      integer(c_int) :: rc
    end function crest_cuda_cregen_dup

    !> replaces cregen.f90:1421-1469 (enuc, greedy rotamer grouping, group ids)
    function crest_cuda_cregen_group(nat,nkept,at,xyz,er,rot,thr,group,enuc) result(rc) &
        &    bind(C,name="crest_cuda_cregen_group")
      import
      integer(c_int),intent(in),value :: nat,nkept
      integer(c_int),intent(in) :: at(*)
      real(c_double),intent(in) :: xyz(3,*)
      real(c_double),intent(in) :: er(*)
      real(c_double),intent(in) :: rot(3,*)
      type(crest_cuda_thresholds),intent(in) :: thr
      integer(c_int),intent(out) :: group(*)     !< group(0:nkept)
      type(c_ptr),value :: enuc                  !< double(nkept) or c_null_ptr
      integer(c_int) :: rc
    end function crest_cuda_cregen_group

    !> replaces the main loop of compare_ensembles (ensemblecomp.f90:213-248)
    function crest_cuda_compare_blocks(nat,nall1,xyz1,nconf1,b1,e1,nall2,xyz2,nconf2,b2,e2,rmat) &
        &    result(rc) bind(C,name="crest_cuda_compare_blocks")
      import
      integer(c_int),intent(in),value :: nat,nall1,nconf1,nall2,nconf2
      real(c_double),intent(in) :: xyz1(3,*),xyz2(3,*)
      integer(c_int),intent(in) :: b1(*),e1(*),b2(*),e2(*)
      real(c_double),intent(out) :: rmat(nconf1,*)
      integer(c_int) :: rc
    end function crest_cuda_compare_blocks

    !> replaces the CONF loop of intraconfRMSD (entropy/entropic.f90:1000-1062)
    function crest_cuda_intraconf(nat,nall,xyz,inc,nconf,low,nrot,symsym,fac,enantiofac) &
        &    result(rc) bind(C,name="crest_cuda_intraconf")
      import
      integer(c_int),intent(in),value :: nat,nall,nconf
      real(c_double),intent(in) :: xyz(3,*)
      integer(c_int),intent(in) :: inc(*),low(*),nrot(*)
      character(kind=c_char),intent(in) :: symsym(3,*)
      real(c_double),intent(out) :: fac(*)
      real(c_double),intent(inout) :: enantiofac(*)
      integer(c_int) :: rc
    end function crest_cuda_intraconf

    function crest_cuda_last_error() result(msg) bind(C,name="crest_cuda_last_error")
      import
      type(c_ptr) :: msg
    end function crest_cuda_last_error
  end interface

contains

  !> call once (e.g. beside new_ompautoset in cregen_setthreads, cregen.f90:2314-2336):
  !> ngpu = 0 takes every visible GPU; returns the number in use
  subroutine crest_cuda_start(ngpu,nused)
    integer,intent(in) :: ngpu
    integer,intent(out) :: nused
    integer(c_int) :: rc,nu
    rc = crest_cuda_init(int(ngpu,c_int),nu)
    if (rc /= 0) error stop 'crest_cuda_init failed (no CUDA device; there is no CPU fallback)'
    nused = int(nu)
  end subroutine crest_cuda_start

  !> call at program end (or never: the driver reclaims everything at exit)
  subroutine crest_cuda_stop()
    call crest_cuda_finalize()
  end subroutine crest_cuda_stop

  !> abort like the reference does on fatal errors (`error stop`); there is no CPU fallback
  subroutine cuda_fail(where)
    character(len=*),intent(in) :: where
    character(kind=c_char),pointer :: cmsg(:)
    character(len=512) :: msg
    type(c_ptr) :: p
    integer :: k
    msg = ''
    p = crest_cuda_last_error()
    if (c_associated(p)) then
      call c_f_pointer(p,cmsg,[512])
      do k = 1,512
        if (cmsg(k) == c_null_char) exit
        msg(k:k) = cmsg(k)
      end do
    end if
    write (*,'(a,a,a,a)') '**ERROR** ',where,': ',trim(msg)
    error stop
  end subroutine cuda_fail

  !> K0-K4 of cregen_CRE.  xyz is replaced by its principal-axis-frame version (what the two
  !> axis() calls leave behind), rot(3,nall) in MHz, double(nall) = 0 or the root it duplicates.
  subroutine cregen_dup_cuda(nat,nall,at,xyz,er,heavy,substruc,includeRMSD,enantio, &
      &                      ethr,rthr,bthr,bthrmax,bthrshift,rot,double)
    integer,intent(in) :: nat,nall
    integer,intent(in) :: at(nat)
    real(c_double),intent(inout) :: xyz(3,nat,nall)
    real(c_double),intent(in) :: er(nall)
    logical,intent(in) :: heavy,substruc,enantio
    integer,intent(in) :: includeRMSD(:)           !< only read if substruc
    real(c_double),intent(in) :: ethr,rthr,bthr,bthrmax,bthrshift
    real(c_double),intent(out) :: rot(3,nall)
    integer,intent(out) :: double(nall)
    type(crest_cuda_thresholds) :: thr
    integer(c_int),allocatable,target :: mask(:)
    integer(c_int),allocatable :: cat(:),cdup(:)
    real(c_double),allocatable :: xa(:,:,:)
    integer(c_long_long) :: npairs
    type(c_ptr) :: pmask
    integer(c_int) :: rc,rfm,ien

    thr = crest_cuda_thresholds(ethr,rthr,bthr,bthrmax,bthrshift)
    allocate (cat(nat),cdup(max(nall,1)),xa(3,nat,max(nall,1)))
    cat = int(at,c_int)
    pmask = c_null_ptr
    rfm = 0
    if (substruc) then                         ! cregen.f90:1199-1202, 1225-1230
      allocate (mask(nat))
      mask = int(includeRMSD(1:nat),c_int)
      pmask = c_loc(mask)
      rfm = 1
    else if (heavy) then                       ! cregen.f90:1293-1299, heavymask: at /= 1
      allocate (mask(nat))
      mask = merge(1_c_int,0_c_int,at /= 1)
      pmask = c_loc(mask)
    end if
    ien = merge(1_c_int,0_c_int,enantio)
    rc = crest_cuda_cregen_dup(int(nat,c_int),int(nall,c_int),cat,xyz,er,pmask,rfm,ien, &
        &                      thr,rot,xa,cdup,npairs)
    if (rc /= 0) call cuda_fail('crest_cuda_cregen_dup')
    xyz = xa(:,:,1:nall)
    double = int(cdup(1:nall))
  end subroutine cregen_dup_cuda

  !> K5-K6 of cregen_CRE on the partitioned ensemble: group(0:nall), group(0) = #groups.
  subroutine cregen_group_cuda(nat,nall,nallout,at,xyz,er,rot,ethr,bthr,bthrmax,bthrshift,group)
    integer,intent(in) :: nat,nall,nallout
    integer,intent(in) :: at(nat)
    real(c_double),intent(in) :: xyz(3,nat,nall)
    real(c_double),intent(in) :: er(nall)
    real(c_double),intent(in) :: rot(3,nall)
    real(c_double),intent(in) :: ethr,bthr,bthrmax,bthrshift
    integer,intent(inout) :: group(0:nall)
    type(crest_cuda_thresholds) :: thr
    integer(c_int),allocatable :: cat(:),cgroup(:)
    integer(c_int) :: rc
    thr = crest_cuda_thresholds(ethr,0.0_c_double,bthr,bthrmax,bthrshift)
    allocate (cat(nat),cgroup(0:nallout))
    cat = int(at,c_int)
    rc = crest_cuda_cregen_group(int(nat,c_int),int(nallout,c_int),cat,xyz,er,rot,thr, &
        &                        cgroup,c_null_ptr)
    if (rc /= 0) call cuda_fail('crest_cuda_cregen_group')
    group = 0
    group(0:nallout) = int(cgroup(0:nallout))
  end subroutine cregen_group_cuda

  !> compare_ensembles: rmat(nconf1,nconf2) = block-wise minimum RMSD (ensemblecomp.f90:213-248)
  subroutine compare_blocks_cuda(nat,nall1,xyz1,nconf1,b1,e1,nall2,xyz2,nconf2,b2,e2,rmat)
    integer,intent(in) :: nat,nall1,nconf1,nall2,nconf2
    real(c_double),intent(in) :: xyz1(3,nat,nall1),xyz2(3,nat,nall2)
    integer,intent(in) :: b1(nconf1),e1(nconf1),b2(nconf2),e2(nconf2)
    real(c_double),intent(out) :: rmat(nconf1,nconf2)
    integer(c_int) :: rc
    rc = crest_cuda_compare_blocks(int(nat,c_int),int(nall1,c_int),xyz1,int(nconf1,c_int), &
        &   int(b1,c_int),int(e1,c_int),int(nall2,c_int),xyz2,int(nconf2,c_int), &
        &   int(b2,c_int),int(e2,c_int),rmat)
    if (rc /= 0) call cuda_fail('crest_cuda_compare_blocks')
  end subroutine compare_blocks_cuda

  !> intraconfRMSD: fac(nconf), enantiofac(nconf) (entropy/entropic.f90:938-1097); the rotamers of
  !> conformer co are zens%xyz(:,:,nmat(co,1):nmat(co,nrot(co)))
  subroutine intraconf_cuda(nat,nall,xyz,inc,nconf,low,nrot,symsym,fac,enantiofac)
    integer,intent(in) :: nat,nall,nconf
    real(c_double),intent(in) :: xyz(3,nat,nall)
    integer,intent(in) :: inc(nat),low(nconf),nrot(nconf)
    character(len=3),intent(in) :: symsym(nconf)
    real(c_double),intent(out) :: fac(nconf)
    real(c_double),intent(inout) :: enantiofac(nconf)
    character(kind=c_char) :: csym(3,max(nconf,1))
    integer(c_int) :: rc
    integer :: c,k
    do c = 1,nconf
      do k = 1,3
        csym(k,c) = symsym(c)(k:k)
      end do
    end do
    rc = crest_cuda_intraconf(int(nat,c_int),int(nall,c_int),xyz,int(inc,c_int),int(nconf,c_int), &
        &   int(low,c_int),int(nrot,c_int),csym,fac,enantiofac)
    if (rc /= 0) call cuda_fail('crest_cuda_intraconf')
  end subroutine intraconf_cuda

end module crest_cuda
