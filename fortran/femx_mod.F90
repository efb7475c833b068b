! femx_mod.F90 -- ISO_C_BINDING interface to libfemx.so (include/femx.h).
!
! This is the reference-side binding a defmod maintainer adds next to src/m_global.F90.
! No Fortran compiler exists in the build image, so this file is not compiled by the test
! suite; the same entry points are exercised through ctypes (fem_b200/_lib.py) with the same
! argument order and types.  See INTEGRATION.md for the call sites it replaces.
module femx
  use iso_c_binding
  implicit none

  integer(c_int), parameter :: FEMX_HOST = 0, FEMX_DEVICE = 1
  integer(c_int), parameter :: FEMX_ROW_MAJOR = 0, FEMX_COL_MAJOR = 1
  integer(c_int), parameter :: FEMX_Q_REFERENCE = 15, FEMX_Q_UEL = 0

  type, bind(C) :: femx_opts
    integer(c_int) :: quirks = FEMX_Q_REFERENCE
    integer(c_int) :: assembly_variant = 0
    integer(c_int) :: reserved(6) = 0
  end type femx_opts

  interface
    integer(c_int) function femx_create(ctx, device, opts) bind(C, name="femx_create")
      import :: c_ptr, c_int, femx_opts
      type(c_ptr), intent(out) :: ctx
      integer(c_int), value :: device
      type(femx_opts), intent(in) :: opts
    end function
    integer(c_int) function femx_destroy(ctx) bind(C, name="femx_destroy")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
    end function
    function femx_last_error(ctx) bind(C, name="femx_last_error") result(msg)
      import :: c_ptr
      type(c_ptr), value :: ctx
      type(c_ptr) :: msg
    end function
    ! local_elements(:) flattened: etype = getElTypeNo, eptr/eind as PartitionBroadcast builds them
    ! (src/m_io.F90:128-135) but with the 1-based LOCAL node ids of src/main.F90:68-73
    integer(c_int) function femx_set_mesh(ctx, pdim, nels, etype, eptr, eind, emat, enlmat, nlnds, coords, &
                                          coords_layout, nl2g, nnds) bind(C, name="femx_set_mesh")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: pdim, nels, nlnds, coords_layout, nnds
      integer(c_int), intent(in) :: etype(*), eptr(*), eind(*), emat(*), enlmat(*), nl2g(*)
      real(c_double), intent(in) :: coords(*)
    end function
    integer(c_int) function femx_set_materials(ctx, nmts, E_nu, ncoh, cohprops) bind(C, name="femx_set_materials")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: nmts, ncoh
      real(c_double), intent(in) :: E_nu(*), cohprops(*)
    end function
    integer(c_int) function femx_set_bcs(ctx, nbcs, bcnode, bcval) bind(C, name="femx_set_bcs")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: nbcs
      integer(c_int), intent(in) :: bcnode(*)
      real(c_double), intent(in) :: bcval(*)
    end function
    integer(c_int) function femx_build_plan(ctx) bind(C, name="femx_build_plan")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
    end function
    integer(c_int) function femx_pattern_size(ctx, nrows, nnz) bind(C, name="femx_pattern_size")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      integer(c_int64_t), intent(out) :: nrows, nnz
    end function
    integer(c_int) function femx_get_pattern(ctx, rowptr, colidx, where) bind(C, name="femx_get_pattern")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, rowptr, colidx      ! int64 rowptr(nrows+1), int32 colidx(nnz); c_null_ptr to skip
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_assemble_K(ctx, vals, where) bind(C, name="femx_assemble_K")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, vals                ! c_loc(host array) or a device pointer from PETSc
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_assemble_K_fint(ctx, u_local, vals, fint, where) bind(C, name="femx_assemble_K_fint")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, u_local, vals, fint
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_apply_kbc(ctx, vals, where) bind(C, name="femx_apply_kbc")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, vals
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_jacobian(ctx, kvals, u_total_local, dt, jvals, where) bind(C, name="femx_jacobian")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx, kvals, u_total_local, jvals
      real(c_double), value :: dt
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_jacobian_update(ctx, kvals, u_total_local, dt, jvals, where) bind(C, name="femx_jacobian_update")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx, kvals, u_total_local, jvals
      real(c_double), value :: dt
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_residual(ctx, jvals, u_total_local, du, f_ext, dt, t_total, r, where) &
        bind(C, name="femx_residual")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx, jvals, u_total_local, du, f_ext, r
      real(c_double), value :: dt, t_total
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_residual_forces(ctx, u_total_local, f_ext, dt, t_total, out, where) &
        bind(C, name="femx_residual_forces")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx, u_total_local, f_ext, out
      real(c_double), value :: dt, t_total
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_stress(ctx, uu_local, stress_node, layout, where) bind(C, name="femx_stress")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, uu_local, stress_node
      integer(c_int), value :: layout, where
    end function
    ! no counterpart in the reference: femx_stress on a static device scratch, no allocation (include/femx.h)
    integer(c_int) function femx_stress_noalloc(ctx, d_uu_local, d_stress_node, layout) bind(C, name="femx_stress_noalloc")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, d_uu_local, d_stress_node
      integer(c_int), value :: layout
    end function
    integer(c_int64_t) function femx_stress_noalloc_max_doubles() bind(C, name="femx_stress_noalloc_max_doubles")
      import :: c_int64_t
    end function
    integer(c_int) function femx_element_ke(ctx, etype, n, ecoords, E_nu, ke) bind(C, name="femx_element_ke")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: etype, n
      real(c_double), intent(in) :: ecoords(*), E_nu(*)
      real(c_double), intent(out) :: ke(*)
    end function
    integer(c_int) function femx_set_ownership(ctx, nranks, rank, row_start, row_end) bind(C, name="femx_set_ownership")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      integer(c_int), value :: nranks, rank
      integer(c_int64_t), value :: row_start, row_end
    end function
    integer(c_int) function femx_gather_sum(ctx, a, na, b, ptr, idx, nout, out) bind(C, name="femx_gather_sum")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx, a, b, ptr, idx, out
      integer(c_int64_t), value :: na, nout
    end function
    ! multi-rank nodal stress: sums of this rank's elements + its count_node (src/main.F90:253-291)
    integer(c_int) function femx_stress_partial(ctx, uu_local, stress_sum, count_node, where) bind(C, name="femx_stress_partial")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, uu_local, stress_sum, count_node
      integer(c_int), value :: where
    end function
    ! FormRHS (src/m_global.F90:373-436): fnode / fval / telsd / tval as ReadDistForces / ReadDistTractions leave them,
    ! TRANSPOSED to row-major (the Fortran arrays are (count, pdim+2) column-major)
    integer(c_int) function femx_set_loads(ctx, nfrcs, fnode, fval, ntrcs, telsd, tval) bind(C, name="femx_set_loads")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: nfrcs, ntrcs
      integer(c_int), intent(in) :: fnode(*), telsd(*)
      real(c_double), intent(in) :: fval(*), tval(*)
    end function
    integer(c_int) function femx_form_rhs(ctx, t_init, dt, t_total, f, where) bind(C, name="femx_form_rhs")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx, f
      real(c_double), value :: t_init, dt, t_total
      integer(c_int), value :: where
    end function
    ! peer-memory interface exchange (include/femx.h): handles are 64-byte cudaIpcMemHandle blobs
    integer(c_int) function femx_peer_arena_create(ctx, n_recv_vals, n_recv_f, handle) bind(C, name="femx_peer_arena_create")
      import :: c_ptr, c_int, c_int64_t, c_char
      type(c_ptr), value :: ctx
      integer(c_int64_t), value :: n_recv_vals, n_recv_f
      character(kind=c_char), intent(out) :: handle(64)
    end function
    integer(c_int) function femx_peer_connect(ctx, nranks, rank, handles, recv_vals, recv_f, val_bounds, val_dst, &
                                              f_bounds, f_dst, recv_from) bind(C, name="femx_peer_connect")
      import :: c_ptr, c_int, c_int64_t, c_char
      type(c_ptr), value :: ctx
      integer(c_int), value :: nranks, rank
      character(kind=c_char), intent(in) :: handles(*)
      integer(c_int64_t), intent(in) :: recv_vals(*), recv_f(*), val_bounds(*), val_dst(*), f_bounds(*), f_dst(*), recv_from(*)
    end function
    ! ranks that are ctxs of one process on one device: peers(q) = the q-th ctx
    integer(c_int) function femx_peer_connect_local(ctx, nranks, rank, peers, recv_vals, recv_f, val_bounds, val_dst, &
                                                    f_bounds, f_dst, recv_from) bind(C, name="femx_peer_connect_local")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      integer(c_int), value :: nranks, rank
      type(c_ptr), intent(in) :: peers(*)
      integer(c_int64_t), intent(in) :: recv_vals(*), recv_f(*), val_bounds(*), val_dst(*), f_bounds(*), f_dst(*), recv_from(*)
    end function
    ! the receiver add of MatAssemblyEnd over the interface rows (src/main.F90:167-168)
    integer(c_int) function femx_peer_set_merge(ctx, n, own_slot, ptr, idx, nf, f_row, f_ptr, f_idx) bind(C, name="femx_peer_set_merge")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      integer(c_int64_t), value :: n, nf
      integer(c_int64_t), intent(in) :: own_slot(*), ptr(*), idx(*), f_row(*), f_ptr(*), f_idx(*)
    end function
    integer(c_int) function femx_peer_merged(ctx, merged, n) bind(C, name="femx_peer_merged")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      type(c_ptr), intent(out) :: merged
      integer(c_int64_t), intent(out) :: n
    end function
    integer(c_int) function femx_peer_complete(ctx) bind(C, name="femx_peer_complete")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
    end function
    integer(c_int) function femx_peer_set_deferred(ctx, on) bind(C, name="femx_peer_set_deferred")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
      integer(c_int), value :: on
    end function
    integer(c_int) function femx_peer_last_error(ctx) bind(C, name="femx_peer_last_error")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
    end function
    ! GetVec_U (src/m_global.F90:549-564): owned slice of Vec_U -> ghosted local vector
    integer(c_int) function femx_peer_set_scatter(ctx, send_ptr, send_rows, u_dst, recv_u) bind(C, name="femx_peer_set_scatter")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      integer(c_int64_t), intent(in) :: send_ptr(*), send_rows(*), u_dst(*), recv_u(*)
    end function
    integer(c_int) function femx_scatter_u(ctx, u_owned, u_local, where) bind(C, name="femx_scatter_u")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, u_owned, u_local     ! c_loc of host arrays or device pointers, by `where`
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_peer_recv(ctx, vals_recv, f_recv) bind(C, name="femx_peer_recv")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
      type(c_ptr), intent(out) :: vals_recv, f_recv
    end function
    integer(c_int) function femx_peer_disconnect(ctx) bind(C, name="femx_peer_disconnect")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
    end function
    ! ---- memory and streams for hosts without a CUDA runtime of their own
    integer(c_int) function femx_host_alloc(ptr, bytes) bind(C, name="femx_host_alloc")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), intent(out) :: ptr
      integer(c_int64_t), value :: bytes
    end function
    integer(c_int) function femx_host_free(ptr) bind(C, name="femx_host_free")
      import :: c_ptr, c_int
      type(c_ptr), value :: ptr
    end function
    integer(c_int) function femx_device_alloc(ctx, ptr, bytes) bind(C, name="femx_device_alloc")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      type(c_ptr), intent(out) :: ptr
      integer(c_int64_t), value :: bytes
    end function
    integer(c_int) function femx_device_free(ctx, ptr) bind(C, name="femx_device_free")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, ptr
    end function
    integer(c_int) function femx_memcpy(ctx, dst, src, bytes, dst_where, src_where) bind(C, name="femx_memcpy")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx, dst, src
      integer(c_int64_t), value :: bytes
      integer(c_int), value :: dst_where, src_where
    end function
    integer(c_int) function femx_synchronize(ctx) bind(C, name="femx_synchronize")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
    end function
    integer(c_int) function femx_set_stream(ctx, cuda_stream) bind(C, name="femx_set_stream")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, cuda_stream          ! e.g. PetscDeviceContextGetStreamHandle
    end function
    integer(c_int) function femx_set_async(ctx, on) bind(C, name="femx_set_async")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx
      integer(c_int), value :: on
    end function
    ! ---- PETSc hand-off: MatSetPreallocationCOO arrays and the MPIAIJ split
    integer(c_int) function femx_get_coo(ctx, coo_i, coo_j, where) bind(C, name="femx_get_coo")
      import :: c_ptr, c_int
      type(c_ptr), value :: ctx, coo_i, coo_j         ! int32 [nnz]
      integer(c_int), value :: where
    end function
    integer(c_int) function femx_mpiaij_sizes(ctx, row_start, row_end, nnz_diag, nnz_offdiag, n_garray) bind(C, name="femx_mpiaij_sizes")
      import :: c_ptr, c_int, c_int64_t
      type(c_ptr), value :: ctx
      integer(c_int64_t), value :: row_start, row_end
      integer(c_int64_t), intent(out) :: nnz_diag, nnz_offdiag, n_garray
    end function
    integer(c_int) function femx_get_pattern_mpiaij(ctx, row_start, row_end, d_rowptr, d_col, o_rowptr, o_col, garray) &
        bind(C, name="femx_get_pattern_mpiaij")
      import :: c_ptr, c_int, c_int64_t, c_int32_t
      type(c_ptr), value :: ctx
      integer(c_int64_t), value :: row_start, row_end
      integer(c_int64_t), intent(out) :: d_rowptr(*), o_rowptr(*)
      integer(c_int32_t), intent(out) :: d_col(*), o_col(*), garray(*)
    end function
    integer(c_int) function femx_get_slot_map(ctx, offsets, blockpos) bind(C, name="femx_get_slot_map")
      import :: c_ptr, c_int, c_int64_t, c_int16_t
      type(c_ptr), value :: ctx
      integer(c_int64_t), intent(out) :: offsets(*)
      integer(c_int16_t), intent(out) :: blockpos(*)   ! uint16 bit patterns
    end function
    ! ---- element-level entry points (unit parity): CalcElStress src/m_local.F90:93, applyTract_1 / applyStiff_1 src/m_global.F90:622, :762
    integer(c_int) function femx_element_stress(ctx, etype, n, ecoords, edisp, E_nu, estress, enodal) bind(C, name="femx_element_stress")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: etype, n
      real(c_double), intent(in) :: ecoords(*), edisp(*), E_nu(*)
      real(c_double), intent(out) :: estress(*), enodal(*)
    end function
    integer(c_int) function femx_element_cohesive(ctx, n, ecoords, nodes, u_local, nu_local, props, dt, force, stiff) &
        bind(C, name="femx_element_cohesive")
      import :: c_ptr, c_int, c_double
      type(c_ptr), value :: ctx
      integer(c_int), value :: n, nu_local
      real(c_double), intent(in) :: ecoords(*), u_local(*), props(*)
      integer(c_int), intent(in) :: nodes(*)
      real(c_double), value :: dt
      real(c_double), intent(out) :: force(*), stiff(*)
    end function
    ! ---- instrumentation
    integer(c_int64_t) function femx_launch_count(ctx) bind(C, name="femx_launch_count")
      import :: c_ptr, c_int64_t
      type(c_ptr), value :: ctx
    end function
    real(c_double) function femx_last_kernel_ms(ctx) bind(C, name="femx_last_kernel_ms")
      import :: c_ptr, c_double
      type(c_ptr), value :: ctx
    end function
    real(c_double) function femx_last_assembly_kernel_ms(ctx) bind(C, name="femx_last_assembly_kernel_ms")
      import :: c_ptr, c_double
      type(c_ptr), value :: ctx
    end function
    real(c_double) function femx_plan_seconds(ctx) bind(C, name="femx_plan_seconds")
      import :: c_ptr, c_double
      type(c_ptr), value :: ctx
    end function
    integer(c_int) function femx_version() bind(C, name="femx_version")
      import :: c_int
    end function
  end interface

contains

  ! Flattens module global's local_elements(:) (src/m_global.F90:40) into the arrays femx_set_mesh takes.
  ! `elems` components used: eltype -> type code, nodes(:), mat, nlMat (src/m_elems.F90:6-12).
  subroutine femx_check(ierr_c, ierr)
    integer(c_int), intent(in) :: ierr_c
    integer, intent(out) :: ierr
    ierr = int(ierr_c)     ! PETSc-style trailing ierr (the reference never checks it; femx keeps the text
                           ! in femx_last_error)
  end subroutine femx_check

end module femx
