/*
 * femx.h -- C ABI of the B200-native defmod assembly hot path.
 *
 * The reference (srinath-chakravarthy/FEM, Fortran 90 + PETSc) has no FFI; the boundary
 * this library sits under is the module-procedure surface that `program main` and
 * PETSc SNES call.  Every entry point below names the reference
 * procedure / loop it replaces (paths relative to the reference tree).  A Fortran
 * host binds these with ISO_C_BINDING (fortran/femx_mod.F90).
 *
 * Conventions kept from the reference, never re-based silently:
 *   - element type codes = getElTypeNo, src/m_elems.F90:25: 1 tri 2 qua 3 tet 4 hex 5 coh
 *     (hex has 8 nodes; the reference's table says 6 -- defect D1, src/m_elems.F90:27)
 *   - element node ids are 1-based LOCAL ids (src/main.F90:68-73), local ids ascend with
 *     the global id (src/main.F90:57-63), nl2g is 1-based global (src/main.F90:76-81)
 *   - global dofs / CSR columns are 0-based: pdim*(g-1)+d (src/main.F90:93-94)
 *   - BC flag 0 = constrained (BC_PRESENT, src/m_global.F90:14)
 *   - element matrices are column-major k(:,:) (src/m_local.F90:43)
 *   - all reals are fp64, all indices 32-bit int except nnz / rowptr (int64)
 *
 * Every function returns 0 on success or a FEMX_ERR_* code; femx_last_error() gives the
 * text.  No exceptions, no exit(), no CPU fallback: without a CUDA device femx_create
 * fails with FEMX_ERR_CUDA.  One ctx per rank / per GPU; a ctx is not re-entrant.
 */
#ifndef FEMX_H
#define FEMX_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct femx_ctx femx_ctx;

enum {
  FEMX_OK = 0,
  FEMX_ERR_ARG = 1,      /* bad argument / call order                     */
  FEMX_ERR_CUDA = 2,     /* CUDA runtime failure (latched in the ctx)      */
  FEMX_ERR_NCCL = 3,     /* NCCL failure                                   */
  FEMX_ERR_ALLOC = 4,
  FEMX_ERR_UNSUPPORTED = 5
};

enum { FEMX_TRI = 1, FEMX_QUA = 2, FEMX_TET = 3, FEMX_HEX = 4, FEMX_COH = 5 };

/* where a caller buffer lives */
enum { FEMX_HOST = 0, FEMX_DEVICE = 1 };

/* coords layout: (node, dim) row-major, or Fortran coords(nlnds, pdim) column-major
 * (src/m_global.F90:22) */
enum { FEMX_ROW_MAJOR = 0, FEMX_COL_MAJOR = 1 };

/* quirk switches: a set bit reproduces the reference as written, a clear
 * bit the corrected Abaqus-UEL form the element was transplanted from. */
enum {
  FEMX_Q_GATHER_LOCAL14 = 1,     /* D3 getCohRels reads local nodes 1..4, src/m_elems.F90:700-702 */
  FEMX_Q_TANGENT_PRECEDENCE = 2, /* D4 lost parentheses, src/m_global.F90:829-831                 */
  FEMX_Q_FORCE_LAST_IP = 4,      /* D8 force assigned per ip, src/m_global.F90:669                */
  FEMX_Q_COH_COUNTS_NODE = 8,    /* D5 coh bumps count_node, src/main.F90:263                     */
  FEMX_Q_REFERENCE = 15,
  FEMX_Q_UEL = 0
};

/* assembly variants: 0 = best available for the mesh (20 for a logical hex8 grid, 11 for a pure
 * tri3 / tet4 mesh, else 9, else 1); 1 = two-pass (element kernel -> packed Ke scratch -> ordered gather; the cross-check);
 * 9 = fused rows kernel for any element mix; 11 = simplex rows kernel (tri3 / tet4 only); 12 = symmetric pair-block hex8
 * tile kernel; 20 = warp-specialised DMMA hex8 tile kernel.  Any other value is refused by femx_create (13 / 21, the
 * clock-accounting builds of 12 / 20, exist only in a library built with `make DIAG=1`). */
typedef struct {
  int quirks;            /* FEMX_Q_* bits                                   */
  int assembly_variant;  /* 0 auto, 1 two-pass, 9 fused rows, 11 simplex rows, 12 / 20 hex8 tile kernels */
  int reserved[6];
} femx_opts;

/* ---- lifetime ------------------------------------------------------------------- */
int femx_create(femx_ctx **ctx, int device, const femx_opts *opts /* NULL = defaults */);
int femx_destroy(femx_ctx *ctx);
const char *femx_last_error(const femx_ctx *ctx /* NULL = last create failure */);
/* pinned host memory for buffers that cross the ABI every step */
int femx_host_alloc(void **ptr, int64_t bytes);
int femx_host_free(void *ptr);
/* raw device memory for callers without a CUDA runtime of their own (tests, Fortran) */
int femx_device_alloc(femx_ctx *ctx, void **ptr, int64_t bytes);
int femx_device_free(femx_ctx *ctx, void *ptr);
int femx_memcpy(femx_ctx *ctx, void *dst, const void *src, int64_t bytes, int dst_where, int src_where);
int femx_synchronize(femx_ctx *ctx);
/* Streams (SURVEY.md 8b threading row).  Default: every call runs on the ctx's own stream and returns synchronised.
 * femx_set_stream puts the ctx on the caller's cudaStream_t (passed as void *; NULL = its own stream again);
 * femx_set_async(ctx, 1) makes calls return once their work is enqueued, so a host can overlap femx_jacobian with
 * anything else; femx_synchronize (or a synchronise of that stream) before results, timings or errors are read. */
int femx_set_stream(femx_ctx *ctx, void *cuda_stream);
int femx_set_async(femx_ctx *ctx, int on);

/* ---- state the path reads (module globals of src/m_global.F90:17-87) ------------- */
/* local_elements(:) as eptr/eind (exactly what PartitionBroadcast builds at
 * src/m_io.F90:128-135, but 1-based local ids), coords, nl2g, nnds */
int femx_set_mesh(femx_ctx *ctx, int pdim, int nels, const int *etype, const int *eptr,
                  const int *eind, const int *emat, const int *enlmat, int nlnds,
                  const double *coords, int coords_layout, const int *nl2g, int nnds);
/* mat(:,1:2) = (E, nu) row-major [nmts*2] (src/m_global.F90:254-255);
 * cohmats(:)%props [ncoh*6] (src/m_seplaw.F90:19-24) */
int femx_set_materials(femx_ctx *ctx, int nmts, const double *E_nu, int ncoh, const double *cohprops);
/* bcnode(nbcs, 1+pdim) row-major: global node, flags; bcval(nbcs, pdim) (src/m_global.F90:33-34) */
int femx_set_bcs(femx_ctx *ctx, int nbcs, const int *bcnode, const double *bcval);

/* ---- one-time integer pass (replaces MatCreateAIJ preallocation guess + the sorted
 *      insert of MatSetValues, src/main.F90:153-171) ------------------------------ */
int femx_build_plan(femx_ctx *ctx);
/* nrows = pdim*nnds (global rows; rows of nodes this rank does not touch are empty) */
int femx_pattern_size(femx_ctx *ctx, int64_t *nrows, int64_t *nnz);
/* CSR of the assembled AIJ: columns ascending, explicit zeros kept (what MatView shows,
 * examples/tmp.txt:15-33).  Either pointer may be NULL. */
int femx_get_pattern(femx_ctx *ctx, int64_t *rowptr, int32_t *colidx, int where);
/* PETSc hand-off (the reference creates its Mat at src/main.F90:153-156 and fills it with MatSetValues):
 * femx_get_coo fills the (i, j) arrays MatSetPreallocationCOO takes -- global row and column of every value, in
 * femx_get_pattern / femx_assemble_K order, so that MatSetValuesCOO(A, vals, INSERT_VALUES) consumes the device array the
 * assembly wrote without any copy or reordering (rows of other ranks may appear: COO assembly ships them, or the caller
 * uses femx_peer_* and sets the interface rows' i to -1, which COO ignores).  32-bit PetscInt.
 * femx_mpiaij_sizes / femx_get_pattern_mpiaij give rows [row_start, row_end) of the rank-local pattern the way
 * Mat_MPIAIJ stores them: the diagonal block (columns inside the range, local numbering) and the off-diagonal block
 * with columns compacted through garray (ascending global columns) -- for MatCreateMPIAIJWithSplitArrays or
 * MatMPIAIJSetPreallocationCSR-style exact preallocation instead of the reference's d_nz = o_nz guess. */
int femx_get_coo(femx_ctx *ctx, int32_t *coo_i, int32_t *coo_j, int where);
int femx_mpiaij_sizes(femx_ctx *ctx, int64_t row_start, int64_t row_end, int64_t *nnz_diag, int64_t *nnz_offdiag,
                      int64_t *n_garray);
int femx_get_pattern_mpiaij(femx_ctx *ctx, int64_t row_start, int64_t row_end, int64_t *d_rowptr, int32_t *d_col,
                            int64_t *o_rowptr, int32_t *o_col, int32_t *garray);
/* element -> nonzero-slot map at node-block granularity: for element e, local nodes a, b
 * slot(e,a,i,b,j) = rowstart(row(a,i)) + pdim*blockpos[e][a*nen+b] + j.
 * blockpos is [sum nen^2] uint16, offsets [nels+1] (int64). */
int femx_get_slot_map(femx_ctx *ctx, int64_t *offsets, uint16_t *blockpos);

/* ---- element-level entry points (unit parity; FormElKE src/m_local.F90:21,
 *      CalcElStress :93, applyTract_1 src/m_global.F90:622, applyStiff_1 :762) ------ */
/* n elements of one type; ecoords [n][nen][dim]; ke [n][ndof*ndof] column-major */
int femx_element_ke(femx_ctx *ctx, int type, int n, const double *ecoords, const double *E_nu /*[n*2]*/,
                    double *ke);
/* estress [n][nip][cp]; enodal [n][nen][cp] (RecoverNodalStress src/m_global.F90:488) */
int femx_element_stress(femx_ctx *ctx, int type, int n, const double *ecoords, const double *edisp,
                        const double *E_nu, double *estress, double *enodal);
/* cohesive: nodes [n][4] 1-based local ids into u_local; force [n][8]; stiff [n][8*8] row-major
 * (row = (node1,dof1), col = (node2,dof2), exactly result(:,:) of applyStiff_1) */
int femx_element_cohesive(femx_ctx *ctx, int n, const double *ecoords, const int *nodes,
                          const double *u_local, int nu_local, const double *props /*[n*6]*/,
                          double dt, double *force, double *stiff);

/* ---- global entry points --------------------------------------------------------- */
/* `do i=1,nels: FormLocalK; MatSetValues(ADD)` + MatAssembly, src/main.F90:157-168.
 * vals[nnz] in femx_get_pattern order.  No BC applied (see femx_apply_kbc). */
int femx_assemble_K(femx_ctx *ctx, double *vals, int where);
/* headline path: K values and the elastic internal force f = sum_e Ke u_e in one pass.
 * u_local [pdim*nlnds] (what GetVec_U yields, src/m_global.F90:549); fint [pdim*nnds] global */
int femx_assemble_K_fint(femx_ctx *ctx, const double *u_local, double *vals, double *fint, int where);
/* ApplyKBC / UpdateKMatBcs, src/m_global.F90:264-299: diag <- 1e30 (INSERT) */
int femx_apply_kbc(femx_ctx *ctx, double *vals, int where);
/* CalcJacobian, src/m_global.F90:98-158: J = K + sum coh tangents, BC diagonal inserted.
 * kvals = values from femx_assemble_K (+apply_kbc), same memory space as jvals */
int femx_jacobian(femx_ctx *ctx, const double *kvals, const double *u_total_local, double dt,
                  double *jvals, int where);
/* Newton iterations after the first: jvals already holds a Jacobian that femx_jacobian / femx_jacobian_update built
 * from the SAME kvals, so only the blocks a cohesive element touches are rebuilt from K (J = K + sum) and the BC
 * diagonals re-inserted -- O(interface) instead of the reference's MatCopy of all nnz values per call
 * (src/m_global.F90:123).  Device-resident only (where = FEMX_DEVICE, jvals != kvals); same result bit for bit. */
int femx_jacobian_update(femx_ctx *ctx, const double *kvals, const double *u_total_local, double dt,
                         double *jvals, int where);
/* CalcResidual, src/m_global.F90:161-235 (single rank: vectors are global [pdim*nnds]):
 * r = J*du - (f_ext + sum f_coh(u_total), BC rows <- 1e30*bcval*dt/t_total) */
int femx_residual(femx_ctx *ctx, const double *jvals, const double *u_total_local, const double *du,
                  const double *f_ext, double dt, double t_total, double *r, int where);
/* the bracket alone, for multi-rank hosts that keep MatMultAdd in PETSc */
int femx_residual_forces(femx_ctx *ctx, const double *u_total_local, const double *f_ext, double dt,
                         double t_total, double *out, int where);
/* RecoverStress + RecoverNodalStress + nodal average, src/main.F90:253-279.
 * stress_node [nlnds][cp] row-major (or column-major stress_node(nlnds,cpdim) by layout) */
/* Multi-rank form of the same: the SUM of the nodal values of this rank's own elements per local node
 * (stress_sum[nlnds*cp] row-major, ascending element order) and the rank's count_node[nlnds] (may be NULL), without
 * the division -- the host adds the partials of the ranks sharing a node and divides by the global count, what the
 * reference does through Mat_Stress and MPI_AllReduce (src/main.F90:270-289).  fem_b200/parallel.py: DistributedStress. */
int femx_stress_partial(femx_ctx *ctx, const double *uu_local, double *stress_sum, int32_t *count_node, int where);
/* Loads.  femx_set_loads takes the arrays ReadDistForces / ReadDistTractions leave on a rank
 * (src/m_io.F90:330-372, :414-475): fnode[nfrcs] GLOBAL 1-based node, fval[nfrcs][pdim+2] = components, t1, t2;
 * telsd[ntrcs][2] = LOCAL 1-based element (0 = skip, src/m_global.F90:427), 1-based side; tval[ntrcs][pdim+2].
 * Row-major.  Call after femx_build_plan (and again when the loads change).
 * femx_form_rhs = FormRHS (src/m_global.F90:373-436): f[pdim*nnds] <- 0, time-ramped nodal forces, tractions spread
 * by ApplyTraction (:343-368) with EdgeAreaNodes* (src/m_elems.F90:282-599) as vvec*area/nps per side node (traction
 * t1, t2 are divided by dt as the reference does, nodal ones are not), contributions per dof added in the reference's
 * call order, then EnforceBCForce (:439-455) INSERTs 1e30*bcval*dt/t on constrained dofs.  nceqs is unsupported
 * (the reference never allocates cval, src/main.F90:108-109). */
int femx_set_loads(femx_ctx *ctx, int nfrcs, const int *fnode, const double *fval, int ntrcs, const int *telsd,
                   const double *tval);
int femx_form_rhs(femx_ctx *ctx, double t_init, double dt, double t_total, double *f, int where);
int femx_stress(femx_ctx *ctx, const double *uu_local, double *stress_node, int layout, int where);
/* No counterpart in the reference.  femx_stress for callers that may not
 * allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): same two kernels, same bits,
 * but the per-(element, node) stress scratch is a file-scope device array of FEMX_STRESS_NOALLOC_MAX_DOUBLES doubles
 * (16 MiB, held by every GPU that loads the library until the process exits; csrc/stress_static.cu) instead of a context
 * buffer grown with cudaMalloc, and the call makes no cudaMalloc / cudaFree / allocating library call.  d_uu_local
 * [nlnds*pdim] and d_stress_node [nlnds*cp] are DEVICE buffers of the caller (pageable host memory -> FEMX_ERR_ARG).
 * A mesh with (sum of element node counts) * cp > FEMX_STRESS_NOALLOC_MAX_DOUBLES (e.g. more than 43690 hex8) is
 * refused with FEMX_ERR_UNSUPPORTED and a message naming both numbers.  All calls on a device share the one array, which
 * keeps the previous call's contents (every slot read is written first within the call, so nothing needs clearing);
 * calls that overlap on the GPU -- two contexts or streams of one device running concurrently -- race: serialise them. */
#define FEMX_STRESS_NOALLOC_MAX_DOUBLES 2097152
int femx_stress_noalloc(femx_ctx *ctx, const double *d_uu_local, double *d_stress_node, int layout);
int64_t femx_stress_noalloc_max_doubles(void);

/* ---- multi-GPU (one ctx per rank) ----------------------------------- */
/* rows [row_start,row_end) of the global matrix are owned by this rank: the PETSC_DECIDE
 * split of MatCreateAIJ, src/main.F90:154.  Recorded and cross-checked by femx_peer_connect; the value ranges per
 * owner are passed there. */
int femx_set_ownership(femx_ctx *ctx, int nranks, int rank, int64_t row_start, int64_t row_end);
/* Ordered merge used after the interface exchange: out[s] = sum_{q in [ptr[s], ptr[s+1])} src(idx[q]) in list
 * order (deterministic), src = a for idx < na else b[idx - na].  All pointers are DEVICE pointers.
 * a = this rank's own values of its owned rows, b = the values received from the other ranks. */
int femx_gather_sum(femx_ctx *ctx, const double *a, int64_t na, const double *b, const int64_t *ptr,
                    const int64_t *idx, int64_t nout, double *out);
/* Peer-memory interface exchange (one node, 2..8 ranks; replaces the stash traffic of MatAssemblyBegin/End,
 * src/main.F90:167-168).  The rank-local CSR lists rows in ascending global order, so the values bound for rank q
 * are the contiguous local value range [val_bounds[q], val_bounds[q+1]) and the internal-force entries (numbered
 * pdim*local_node + i) the range [f_bounds[q], f_bounds[q+1]).  After femx_peer_connect, femx_assemble_K /
 * femx_assemble_K_fint push those ranges into rank q's receive arena over NVLink (remote stores of a kernel behind
 * the assembly kernel; with opts.reserved[3] = 1 and variants 0, 12, 9, 11 the assembly kernel stores them there itself
 * and the non-owned ranges of the local arrays are not written), at val_dst[q] / f_dst[q] doubles from the start of
 * q's value / force receive area.  The call then runs a device-side step barrier among NEIGHBOURS (ranks that exchange
 * a non-empty slice in either direction; recv_from[q] = doubles this rank receives from q) and the receiver add
 * (femx_peer_set_merge): at return the rank's own slice of `vals`, everything its peers sent, the merged interface
 * rows and the owned rows of `fint` are complete.  Assemblies are collective: every rank calls them the same number of
 * times.  The arena is double-buffered on the call parity.
 *   femx_peer_arena_create  allocates the arena for n_recv_vals + n_recv_f doubles (per buffer) and returns its
 *                           cudaIpcMemHandle (64 bytes) for the caller to all-gather among the ranks;
 *   femx_peer_connect       handles = nranks x 64 bytes, recv_vals / recv_f = every rank's arena capacities;
 *   femx_peer_connect_local the same for ranks that are ctxs of ONE process on ONE device (peers[q] instead of
 *                           handles).  Kernels of such ranks run one after another, so none may wait for a peer:
 *                           assemblies push and signal only, and femx_peer_complete -- called on every rank once all
 *                           have assembled -- checks the flags (they are set by then) and runs the receiver add;
 *   femx_peer_set_merge     the receiver add, interface rows only: merged[s] = vals[own_slot[s]] (own_slot -1: none)
 *                           + sum of the received values idx[ptr[s] .. ptr[s+1]) in list order = (source rank, sender
 *                           position); fint[f_row[s]] += received force entries f_idx[f_ptr[s] .. f_ptr[s+1]).
 *                           O(interface) work per assembly; replaces the add inside MatAssemblyEnd;
 *   femx_peer_merged        device pointer / length of the merged interface-row values of the last assembly;
 *   femx_peer_recv          device pointers to what the LAST assembly received (source-rank order, as laid out by
 *                           the caller's val_dst / f_dst);
 *   femx_peer_disconnect    unmaps the peers' arenas (call on every rank before any rank destroys its ctx).
 * Behind the hex8 tile kernel (assembly variant 20, auto) the exchange overlaps the assembly: the CTAs that own interface
 * rows run first on a high-priority stream, push, barrier and add follow them there, the interior CTAs run meanwhile on the
 * ctx stream, which joins before the call returns (in stream / async mode: before later work on that stream).
 * Environment (read once per process): FEMX_PEER_OVERLAP=0 keeps the exchange serial behind one launch;
 * FEMX_PEER_PUSH_CTAS=n bounds the overlapped copy kernel's grid (default 64); FEMX_PEER_TIMING=1 prints push / add times. */
int femx_peer_arena_create(femx_ctx *ctx, int64_t n_recv_vals, int64_t n_recv_f, char handle[64]);
int femx_peer_connect(femx_ctx *ctx, int nranks, int rank, const char *handles, const int64_t *recv_vals,
                      const int64_t *recv_f, const int64_t *val_bounds, const int64_t *val_dst,
                      const int64_t *f_bounds, const int64_t *f_dst, const int64_t *recv_from);
int femx_peer_connect_local(femx_ctx *ctx, int nranks, int rank, femx_ctx *const *peers, const int64_t *recv_vals,
                            const int64_t *recv_f, const int64_t *val_bounds, const int64_t *val_dst,
                            const int64_t *f_bounds, const int64_t *f_dst, const int64_t *recv_from);
int femx_peer_set_merge(femx_ctx *ctx, int64_t n, const int64_t *own_slot, const int64_t *ptr, const int64_t *idx,
                        int64_t nf, const int64_t *f_row, const int64_t *f_ptr, const int64_t *f_idx);
int femx_peer_merged(femx_ctx *ctx, const double **merged, int64_t *n);
int femx_peer_complete(femx_ctx *ctx);
/* femx_peer_connect connections whose ranks SHARE a GPU (more MPI ranks than GPUs: their kernels are time-sliced, a
 * device-side step barrier would wait for context switches): on = 1 makes assemblies / femx_scatter_u push and signal
 * only; the host synchronises, meets the other ranks (MPI_Barrier) and calls femx_peer_complete. */
int femx_peer_set_deferred(femx_ctx *ctx, int on);
int femx_peer_last_error(femx_ctx *ctx);   /* reports (and clears) a step-barrier timeout */
/* GetVec_U across ranks (src/m_global.F90:549-564; scatter of src/main.F90:183-190), the reverse of the force exchange:
 * femx_peer_set_scatter names, per peer q, the owned global rows send_rows[send_ptr[q] .. send_ptr[q+1]) whose
 * displacement q holds as a ghost (= the rows q sends force entries for, the transpose of f_bounds), u_dst[q] = the
 * first local dof of that block in q's numbering, recv_u[q] = pdim * nlnds of rank q.  femx_scatter_u then turns this
 * rank's slice u_owned [row_end - row_start] of the global vector into the ghosted local vector u_local [pdim * nlnds]
 * every kernel reads: owners store into the peers' u areas over NVLink, neighbours' step barrier, local gather.
 * Deferred (local) peers: push + signal only, femx_peer_complete finishes (device buffers only). */
int femx_peer_set_scatter(femx_ctx *ctx, const int64_t *send_ptr, const int64_t *send_rows, const int64_t *u_dst,
                          const int64_t *recv_u);
int femx_scatter_u(femx_ctx *ctx, const double *u_owned, double *u_local, int where);
int femx_peer_recv(femx_ctx *ctx, const double **vals_recv, const double **f_recv);
int femx_peer_disconnect(femx_ctx *ctx);

/* ---- instrumentation -------------------------------------------------------------- */
/* kernels launched by this ctx since creation (bench.py "gpu_launches") */
int64_t femx_launch_count(const femx_ctx *ctx);
/* device time of the last assemble / stress / jacobian call's kernels, CUDA events on the ctx
 * stream, ms (0 if none) */
double femx_last_kernel_ms(const femx_ctx *ctx);
/* ... of which the assembly kernel alone, for the last femx_assemble_* on a peer-connected ctx (the rest is the interface
 * push, the neighbours' step barrier -- i.e. waiting for the slowest neighbour -- and the receiver add); equals
 * femx_last_kernel_ms without peers */
double femx_last_assembly_kernel_ms(femx_ctx *ctx);
/* wall time of the integer pass of the last femx_build_plan (device pass incl. its uploads; the figure the reference
 * pays inside MatSetValues' sorted inserts, src/main.F90:157-166), seconds */
double femx_plan_seconds(const femx_ctx *ctx);
int femx_version(void);

#ifdef __cplusplus
}
#endif
#endif
