/* pynama_b200.h -- C ABI of libpynama_b200.so (B200 / sm_100a).
 *
 * The reference (maiterm/Pynama) has no FFI: its hot path calls petsc4py
 * objects directly.  The entry points below are what a ctypes/cffi binding of
 * that path binds instead; each one names the reference call site it replaces
 * (paths relative to the reference root).  All functions return 0 on success
 * and a negative PYN_E_* code on failure; pyn_last_error() gives the message.
 * Handles are opaque and owned by the library; host pointers are borrowed for
 * the duration of the call.  Reals are IEEE fp64, node/column ids int32, row
 * pointers / nnz / contribution counts int64.  There is no CPU fallback: every
 * compute entry point needs a context created on a CUDA device.
 */
#ifndef PYNAMA_B200_H
#define PYNAMA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PYN_OK 0
#define PYN_E_INVALID (-1)   /* bad argument */
#define PYN_E_CUDA (-2)      /* CUDA runtime error (message in pyn_last_error) */
#define PYN_E_NOMEM (-3)
#define PYN_E_UNSUPPORTED (-4)
#define PYN_E_NCCL (-5)

typedef struct pyn_ctx pyn_ctx;
typedef struct pyn_mesh pyn_mesh;
typedef struct pyn_mat pyn_mat;
typedef struct pyn_vec pyn_vec;
typedef struct pyn_ksp pyn_ksp;
typedef struct pyn_halo pyn_halo;

int pyn_version(void);
const char* pyn_last_error(void);

/* ---- host-only helpers (no GPU needed) -------------------------------------
 * Reference-element tables, src/domain/elements/utilities.py:43-92 (1-D rules),
 * element.py:13-44 (Lagrange basis), spectral.py:230-252,356-441 (local order). */
/* kind: 0 = Gauss-Legendre, 1 = Gauss-Lobatto-Legendre; x,w: n doubles. */
int pyn_rule_1d(int kind, int n, double* x, double* w);
/* h, dh: npts*nnodes doubles, row-major [point][node]. */
int pyn_lagrange_1d(int nnodes, const double* nodes, int npts, const double* pts,
                    double* h, double* dh);
/* Tensor index (a1,a2[,a3]) of every local position of an n^dim grid, in the
 * reference's vertex/edge/face/body order: out is n^dim * dim int32. */
int pyn_local_order(int dim, int n, int32_t* out);

/* ---- context ---------------------------------------------------------------*/
int pyn_init(int device, pyn_ctx** out);
int pyn_finalize(pyn_ctx* ctx);
int pyn_sync(pyn_ctx* ctx);
/* CUDA-event stopwatch on the library's stream (bench.py's device timing). */
int pyn_timer_start(pyn_ctx* ctx);
int pyn_timer_stop(pyn_ctx* ctx, double* ms);
/* number of kernels this library has launched on ctx so far */
int pyn_launch_count(pyn_ctx* ctx, int64_t* count);
/* write `bytes` of device scratch (> L2) to evict the cache between timed runs */
int pyn_flush_l2(pyn_ctx* ctx);
/* Per-kernel-class CUDA-event profiler on the library's stream: while enabled
 * every launch of the class is bracketed by an event pair; pyn_prof_get returns
 * the launches seen and their summed device time since pyn_prof_enable(ctx, 1).
 * cls: 0 SpMV+dot of the Krylov loop, 1 other SpMV, 2 CG update, 3 CG direction,
 * 4 element KLE kernel, 5 assembly numeric kernel. */
int pyn_prof_enable(pyn_ctx* ctx, int on);
int pyn_prof_get(pyn_ctx* ctx, int cls, int64_t* count, double* total_ms);
/* process-wide tuning knobs.  "ksp_check_every": Krylov iterations launched between two
 * polls of the device-side convergence flag (default 8).  "force_idx64": non-zero makes the
 * numeric assembly take its 64-bit index instantiation whatever the mesh size (the path that
 * meshes beyond 2^31 contributions / record offsets use; for parity tests on small meshes).
 * "asm_coarse_unfused": non-zero runs the coarse level of -pc_type asm as one restriction and
 * steps+1 Chebyshev launches instead of the single cooperative launch (comparison / fallback).
 * "halo_wait_at_start": 1 (default) makes the fused-halo product wait for the ghost entries at CTA
 * start; 0 multiplies the interior rows first and lets each warp wait before its first boundary row
 * (measured slower at 2 GPUs, profiles/README.md r2f/r2g; kept for experiments at higher rank counts).
 * "ksp_adaptive_chunks": 0 launches CG iterations in fixed chunks of ksp_check_every between status
 * polls (round 1); default 1 sizes the first chunk from the previous solve's iteration count.
 * "ksp_graphs": 0 launches the kernels of a CG iteration one by one; default 1 replays one captured CUDA
 * graph per iteration on a single GPU (same kernels, same arguments, same results). */
int pyn_tune(const char* key, int value);
/* page-locked host memory for the caller's staging buffers (cudaMallocHost) */
int pyn_host_alloc(int64_t bytes, void** out);
int pyn_host_free(void* p);
/* device memory in use by library objects of this context */
int pyn_mem_used(pyn_ctx* ctx, int64_t* bytes);

/* ---- multi-GPU: one process per GPU, row-partitioned mesh ------------------------------
 * Replaces PETSc/MPI under `mpiexec -n N` (makefile:2; DMPlex.distribute, dmplex.py:21): the
 * ghost update PETSc's VecScatter does inside MatMult and the MPI_Allreduce of every
 * VecDot/VecNorm in the KSP.  NCCL is loaded with dlopen (nccl_lib: path of libnccl.so.2, or
 * NULL for the default search), so single-GPU use needs no NCCL at all.
 * Rank 0 makes the 128-byte id; the host layer ships it to the other ranks. */
int pyn_comm_unique_id(const char* nccl_lib, void* id128);
int pyn_comm_init(pyn_ctx* ctx, const char* nccl_lib, int rank, int nranks, const void* id128);
int pyn_comm_info(pyn_ctx* ctx, int* rank, int* nranks, int* nccl_version, int64_t* collectives);
/* small host-value collective for the drivers: op 0 sum, 1 max; vals[n] in/out, n <= 64 */
int pyn_comm_allreduce(pyn_ctx* ctx, double* vals, int n, int op);
int pyn_comm_finalize(pyn_ctx* ctx);
/* Peer-memory collectives over NVLink/NVSwitch (CUDA IPC): every rank exports a small window,
 * the host layer gathers the 64-byte handles, every rank maps its peers.  Once attached, the
 * Krylov inner products are all-reduced INSIDE the SpMV / update kernels (remote stores + flags)
 * and the halo exchange stores straight into the receiver's window -- no NCCL call per
 * iteration.  handles: nranks x 64 bytes (comm) / nneigh x 64 bytes in neighbour order (halo);
 * peer_off[i] / peer_nghost[i]: where this rank's run starts in neighbour i's ghost section and
 * that neighbour's ghost-node count. */
int pyn_comm_window_export(pyn_ctx* ctx, void* handle64);
int pyn_comm_window_attach(pyn_ctx* ctx, const void* handles);
int pyn_halo_window_export(pyn_halo* h, int max_bs, void* handle64);
int pyn_halo_window_attach(pyn_halo* h, const void* handles, const int64_t* peer_off, const int64_t* peer_nghost);
int pyn_comm_p2p_status(pyn_ctx* ctx, int* enabled, int* timed_out);
/* Node layout of this rank: owned nodes [0, n_own), ghost nodes [n_own, n_loc) grouped by
 * owner.  Per neighbour i: ranks[i]; the owned nodes it needs send_nodes[send_ptr[i]:
 * send_ptr[i+1]] (local ids); the run of the ghost section it fills, recv_off[i] (nodes past
 * n_own) and recv_cnt[i]. */
int pyn_halo_create(pyn_ctx* ctx, int64_t n_own, int64_t n_loc, int nneigh, const int32_t* ranks,
                    const int64_t* send_ptr, const int32_t* send_nodes, const int64_t* recv_off,
                    const int64_t* recv_cnt, pyn_halo** out);
int pyn_halo_destroy(pyn_halo* h);
/* ghost entries of v (n_loc*bs) <- the owners' values (pack kernel + grouped ncclSend/ncclRecv) */
int pyn_halo_exchange(pyn_halo* h, pyn_vec* v, int bs);
/* Mat of a partitioned mesh: products refresh the ghosts of x first and compute the owned
 * rows only; a KSP on it sums its inner products over ranks. */
int pyn_mat_set_halo(pyn_mat* m, pyn_halo* h);
/* Matrix-free products for the KLE matrix K.  on != 0: every product with m
 * (MatMult in KleSolver.solve / the KSP, kle_solver.py:33-41) is evaluated element by element
 * with sum-factorised quadrature instead of reading the stored blocks; the stored blocks stay
 * (export, diagonal blocks for the preconditioner).  m must be the K of pyn_assemble_kle on a
 * 3-D mesh, ngl 2..6.  Same result up to summation order. */
int pyn_mat_set_matfree(pyn_mat* m, int on);
/* Vec of a partitioned mesh: entries [0, owned) are this rank's; dot/norm/sum run over them and
 * are summed over ranks.  owned < 0 (default): plain local Vec. */
int pyn_vec_set_owned(pyn_vec* v, int64_t owned);

/* ---- mesh tables (inputs A0: dmplex.py:93-100,193-196; domain.py:233-250) ----
 * corner_xyz: nel * 2^dim * dim, DMPlex closure-vertex order;
 * elem_nodes: nel * ngl^dim global node ids in the reference's local order.
 * The symbolic node graph (sort + unique of the COO keys) is built here;
 * replaces DMPlexDom.getConnectivityNodes (dmplex.py:352-372) and the
 * preallocation of mat_fs.py:21-113. */
int pyn_mesh_create(pyn_ctx* ctx, int dim, int ngl, int64_t nel, int64_t nnodes,
                    const double* corner_xyz, const int32_t* elem_nodes, pyn_mesh** out);
int pyn_mesh_destroy(pyn_mesh* mesh);
/* node-graph statistics: number of (row node, col node) pairs, COO length */
int pyn_mesh_graph_info(pyn_mesh* mesh, int64_t* npairs, int64_t* ncontrib, int* max_row);
/* dof classes of the velocity dofs (nnodes*dim bytes): 0 free, 1 Dirichlet
 * (free-slip BC, mat_fs.py:158-172), 2 no-slip normal, 3 no-slip tangential
 * (mat_ns.py:59-75). */
/* Row-partitioned runs: the element -> node table of the WHOLE mesh in global node ids (the same on
 * every rank) and the global element of every local element.  Needed by the second level of
 * -pc_type asm, whose coarse space (Q1 on the element corners of the whole mesh) is replicated on every
 * rank.  Replaces nothing in the reference (PETSc would distribute the coarse problem of a PCMG). */
int pyn_mesh_set_global_tables(pyn_mesh* mesh, int64_t n_global_nodes, int64_t nel_global,
                               const int32_t* elem_nodes_global, const int64_t* elem_gid);
int pyn_mesh_set_dof_class(pyn_mesh* mesh, const uint8_t* cls);
/* Congruence classes of element shapes.  getElemKLEMatrices / getElemKLEOperators
 * (spectral.py:92-228) depend on an element only through its corner offsets, and the
 * reference re-evaluates them for every cell of a uniform box (mat_fs.py:152-155,196-198);
 * here elements whose corner offsets agree to tol * (largest offset in the mesh) share one
 * set of element blocks.  pyn_mesh_create applies tol = 1e-13; tol <= 0 switches the
 * merging off (one class per element).  *nclass (may be NULL) returns the class count. */
int pyn_mesh_set_congruence(pyn_mesh* mesh, double tol, int64_t* nclass);
int pyn_mesh_get_congruence(pyn_mesh* mesh, double* tol, int64_t* nclass);

/* ---- element kernels (debug / parity export of raw element blocks) ---------
 * Spectral.getElemKLEMatrices (spectral.py:92-160): K_e [3nn x 3nn],
 * Rw_e [dim*nn x dim_w*nn], Rd_e [dim*nn x nn] per element, row-major, dofs
 * interleaved node*ncomp+comp.  Output pointers are HOST buffers (may be NULL). */
int pyn_elem_kle(pyn_mesh* mesh, int64_t e0, int64_t ne, double* K_e, double* Rw_e, double* Rd_e);
/* Spectral.getElemKLEOperators (spectral.py:162-228): SrT_e [dim_s*nn x dim*nn],
 * DivSrT_e [dim*nn x dim_s*nn], Curl_e [dim_w*nn x dim*nn], W_e [nn]. */
int pyn_elem_ops(pyn_mesh* mesh, int64_t e0, int64_t ne, double* SrT_e, double* Div_e,
                 double* Curl_e, double* W_e);

/* ---- assembly ----------------------------------------------------------------
 * mode 0: MatFS.build/buildFS (mat_fs.py:131-192) -> out[0..3] = K, Krhs, Rw, Rd(NULL)
 * mode 1: MatNS.build/buildNS (mat_ns.py:47-166)  -> out[0..7] = K, Krhs, Rw, Rd,
 *         Kfs, Krhsfs, Rwfs, Rdfs.  Needs pyn_mesh_set_dof_class first. */
#define PYN_MAT_K 0
#define PYN_MAT_KRHS 1
#define PYN_MAT_RW 2
#define PYN_MAT_RD 3
#define PYN_MAT_KFS 4
#define PYN_MAT_KRHSFS 5
#define PYN_MAT_RWFS 6
#define PYN_MAT_RDFS 7
int pyn_assemble_kle(pyn_mesh* mesh, int mode, pyn_mat** out8);
/* MatFS.buildOperators + Operators.setValues/assembleAll (mat_fs.py:194-271):
 * Curl, SrT, DivSrT already row-scaled by the reciprocal lumped mass; weights =
 * per-node lumped mass before the reciprocal (the reference's weig* vectors
 * repeat it per component).  compact != 0 stores only the structurally
 * non-zero node blocks (the exported CSR still carries the reference's
 * explicit zeros). */
int pyn_assemble_ops(pyn_mesh* mesh, int compact, pyn_mat** curl, pyn_mat** srt,
                     pyn_mat** divsrt, pyn_vec** weights);
/* CUDA-event time of the last assembly phases on this mesh:
 * 0 symbolic graph, 1 element KLE kernel, 2 KLE numeric, 3 element operator
 * kernel, 4 operator numeric */
int pyn_mesh_phase_ms(pyn_mesh* mesh, int phase, double* ms);

/* ---- immersed boundary, 2-D -------------------------------
 * Replaces ImmersedBoundaryStatic.collectNodes / computeClose / computeDirac /
 * buildIBMMatrix (src/cases/immersed_boundary.py:170-229,255-274; the dynamic
 * computeClose :346-358) and ImmersedBody.getDiracs with the delta functions
 * fourGrid / threeGrid / linear (src/domain/immersed_body.py:261-266,398-424).
 * lag_xy: nlag*2 marker coordinates (HOST); cand_nodes / cand_xy: the Euler nodes
 * of the affected cells, ascending node id, and their coordinates (HOST).
 * D[l][n] = prod_axis phi(|x_n - X_l|/h)/h on the kept nodes of marker l;
 * *H = scale_h*D (nlag*2 x n_euler_nodes*2), *S = scale_s*D^T, *A = H S, each
 * with 2x2 blocks w*I (the reference sets the two dofs of a marker separately).
 * kind: 0 fourGrid, 1 threeGrid, 2 linear; dynamic != 0 selects the dynamic
 * body's node search.  *max_per_lag (may be NULL): longest row of D;
 * affected (may be NULL): n_euler_nodes HOST bytes, 1 where a node has a positive
 * weight.  A marker without Euler node is PYN_E_INVALID ("Lag Node without Euler",
 * immersed_boundary.py:222-223). */
int pyn_ibm_build(pyn_ctx* ctx, int64_t n_euler_nodes, int64_t nlag, const double* lag_xy,
                  int64_t ncand, const int32_t* cand_nodes, const double* cand_xy, double h,
                  double dl, double scale_h, double scale_s, int kind, int dynamic,
                  pyn_mat** H, pyn_mat** S, pyn_mat** A, int64_t* max_per_lag,
                  uint8_t* affected);

/* ---- Mat (PETSc.Mat AIJ surface used by the hot path) --------------------------*/
int pyn_mat_shape(pyn_mat* m, int64_t* nrows, int64_t* ncols, int* br, int* bc);
/* nnz of the reference (scalar AIJ, explicit zeros kept) pattern and stored blocks */
int pyn_mat_nnz(pyn_mat* m, int64_t* nnz_ref, int64_t* nblocks);
/* scalar CSR in the reference pattern; rowptr nrows+1, col/val nnz_ref */
int pyn_mat_export_csr(pyn_mat* m, int64_t* rowptr, int32_t* col, double* val);
/* y = A x   (Mat.mult, base_problem.py:123-136; Mat.__mul__, kle_solver.py:35-41) */
int pyn_mat_mult(pyn_mat* m, pyn_vec* x, pyn_vec* y);
/* y = alpha*A x + beta*z  (z may alias y or be NULL when beta == 0) */
int pyn_mat_mult_axpby(pyn_mat* m, double alpha, pyn_vec* x, double beta, pyn_vec* z, pyn_vec* y);
/* C = A + B on the union pattern (K + Kfs, kle_solver.py:25) */
int pyn_mat_add_new(pyn_mat* a, pyn_mat* b, pyn_mat** out);
int pyn_mat_get_diagonal(pyn_mat* m, pyn_vec* d);
/* A <- diag(l) A  (Mat.diagonalScale(L=), mat_fs.py:257-267) */
int pyn_mat_scale_rows(pyn_mat* m, pyn_vec* l);
int pyn_mat_bytes(pyn_mat* m, int64_t* stored_bytes, int64_t* spmv_bytes);
int pyn_mat_destroy(pyn_mat* m);

/* ---- Vec (PETSc.Vec surface used by the hot path) ---------------------------------*/
int pyn_vec_create(pyn_ctx* ctx, int64_t n, pyn_vec** out);
int pyn_vec_destroy(pyn_vec* v);
int pyn_vec_size(pyn_vec* v, int64_t* n);
int pyn_vec_set(pyn_vec* v, double a);
int pyn_vec_copy(pyn_vec* src, pyn_vec* dst);
int pyn_vec_scale(pyn_vec* v, double a);
int pyn_vec_axpy(pyn_vec* y, double a, pyn_vec* x);              /* y += a x */
int pyn_vec_aypx(pyn_vec* y, double a, pyn_vec* x);              /* y = a y + x */
int pyn_vec_waxpy(pyn_vec* w, double a, pyn_vec* x, pyn_vec* y); /* w = a x + y */
int pyn_vec_axpby(pyn_vec* y, double a, pyn_vec* x, double b);   /* y = a x + b y */
int pyn_vec_pointwise_mult(pyn_vec* w, pyn_vec* x, pyn_vec* y);
int pyn_vec_reciprocal(pyn_vec* v);
int pyn_vec_dot(pyn_vec* x, pyn_vec* y, double* out);
int pyn_vec_norm2(pyn_vec* x, double* out);
int pyn_vec_sum(pyn_vec* x, double* out);
/* y = y + sum_k a[k] x[k]   (VecMAXPY; RK stage combination) */
int pyn_vec_maxpy(pyn_vec* y, int k, const double* a, pyn_vec* const* x);
/* Vec.setValues(idx, vals, addv) (boundary_conditions.py:252-260) */
int pyn_vec_set_values(pyn_vec* v, int64_t k, const int64_t* idx, const double* vals, int add);
int pyn_vec_get_values(pyn_vec* v, int64_t k, const int64_t* idx, double* vals);
/* persistent scatter plan (k10): the index list of the boundary dofs is uploaded once;
 * every call stages `vals` (k doubles, host) through pinned memory and scatters them on the
 * library's stream without a device synchronisation.  Replaces the per-evaluation
 * Vec.setValues of BoundaryConditions.setValuesToVec (boundary_conditions.py:252-260). */
typedef struct pyn_scatter pyn_scatter;
int pyn_scatter_create(pyn_ctx* ctx, int64_t k, const int64_t* idx, pyn_scatter** out);
int pyn_scatter_apply(pyn_scatter* plan, pyn_vec* v, const double* vals, int add);
int pyn_scatter_destroy(pyn_scatter* plan);
int pyn_vec_to_host(pyn_vec* v, double* out);
int pyn_vec_from_host(pyn_vec* v, const double* in);
/* first n entries only (the owned part of a row-partitioned Vec) */
int pyn_vec_to_host_n(pyn_vec* v, double* out, int64_t n);
int pyn_vec_from_host_n(pyn_vec* v, const double* in, int64_t n);
/* raw device pointer (for torch.distributed plumbing) */
int pyn_vec_device_ptr(pyn_vec* v, void** ptr);

/* ---- KSP (solver/kle_solver.py:49-64 -> PETSc KSP/PC) -------------------------------*/
#define PYN_KSP_CG 0
#define PYN_KSP_GMRES 1
#define PYN_KSP_BCGS 2
#define PYN_KSP_PREONLY 3
#define PYN_PC_NONE 0
#define PYN_PC_JACOBI 1
#define PYN_PC_BJACOBI 2 /* node-block (br x br) Jacobi */
/* -pc_type asm: additive Schwarz with one subdomain per spectral element,
 * sum_e R_e^T W inv(R_e K R_e^T) W R_e with W = 1/sqrt(elements sharing the node); elements with
 * bitwise equal blocks share one inverse.  Operators assembled from a mesh, 2-D or 3-D; row-partitioned
 * operators on meshes of congruent cells (one level); refused (error) when the distinct inverses
 * would exceed 1 GiB.  CG drops the single-reduction
 * variant under this preconditioner. */
#define PYN_PC_ASM 3
int pyn_ksp_create(pyn_ctx* ctx, pyn_mat* a, int type, int pc, pyn_ksp** out);
int pyn_ksp_set_tolerances(pyn_ksp* k, double rtol, double atol, int maxit, int restart);
/* KSPSetUseFischerGuess / -ksp_fischer_guess <model>,<size>: model 1 (A-conjugate projection
 * of the new right-hand side on the last <size> <= 32 solution increments) forms the initial
 * guess; size 0 switches it off (default, as in PETSc). */
/* KSPCGUseSingleReduction / -ksp_cg_single_reduction: CG with the inner products of an
 * iteration merged into one reduction (fused into the product kernel) and the rest of the
 * recurrence in one vector kernel: 2 launches and 1 reduction per iteration instead of 3 and 2. */
int pyn_ksp_set_cg_single_reduction(pyn_ksp* k, int on);
/* How many iterations the caller expects the NEXT solve to take (n < 0: no idea).  CG launches its
 * iterations ahead of the convergence test in chunks; the hint sizes the first chunk, so that few
 * iterations are launched after convergence.  No effect on the result.  (No counterpart in PETSc.) */
int pyn_ksp_hint_iterations(pyn_ksp* k, int n);
/* -pc_asm_coarse <steps> (PYN_PC_ASM only): adds a second level to the Schwarz preconditioner, the Q1
 * space on the element corners with the Galerkin operator P^T K P (probed through K's own product,
 * one product per colour and component), solved approximately by <steps> Chebyshev steps with point
 * Jacobi -- a fixed polynomial, so the preconditioner stays linear and symmetric.  0 = one level. */
int pyn_ksp_set_asm_coarse(pyn_ksp* k, int steps);
int pyn_ksp_set_fischer_guess(pyn_ksp* k, int model, int size);
/* x: initial guess in, solution out (KSP.__call__(b, x), kle_solver.py:35).
 * reason > 0 converged (2 rtol, 3 atol), < 0 diverged (-3 max iterations). */
int pyn_ksp_solve(pyn_ksp* k, pyn_vec* b, pyn_vec* x, int* its, int* reason, double* rnorm);
int pyn_ksp_destroy(pyn_ksp* k);

/* ---- vorticity-transport RHS pieces (cases/base_problem.py:111-154) --------------------*/
/* computeVtensV (base_problem.py:139-154): out[node*dim_s + s] packed v (x) v */
int pyn_vtensv(pyn_vec* vel, pyn_vec* out, int dim);
/* aux = 2 mu * aux - rho * vtensv(vel)   (base_problem.py:127-130 fused) */
int pyn_stress_combine(pyn_vec* aux, pyn_vec* vel, double mu, double rho, int dim);

/* measured fp64 FMA throughput of the device in TFLOP/s (2 flop per DFMA; 16 independent chains per
 * thread, best of four timed launches): the denominator of the element kernels' roofline fraction.
 * No counterpart in the reference (measurement support). */
int pyn_fp64_peak(pyn_ctx* ctx, double* tflops);

/* ---- allocation-free entry point ---------------------------------------------------------*/
/* No counterpart in the reference: added for callers that may
 * not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * out[0] = x . y over n entries; x, y, out are DEVICE pointers supplied by the caller; the work is
 * queued on `stream` (a cudaStream_t; NULL = the context's stream) and the call does not wait for it.
 * It is pyn_vec_dot's kernel with the reduction scratch (per-chunk partial sums + arrival ticket) in a
 * file-scope device array of 131 328 bytes that every GPU loading the library holds until exit, so
 * the call path has no cudaMalloc / cudaFree / allocating routine.  n > PYN_NOALLOC_MAX_N is refused
 * (PYN_E_INVALID).  All calls on a device share that array: calls that overlap on the GPU race; calls
 * ordered on one stream (or by events) are safe.  The current device must be the context's. */
#define PYN_NOALLOC_MAX_N 268435456LL /* 16384 chunks x 16384 entries */
int pyn_vec_dot_noalloc(pyn_ctx* ctx, const double* x, const double* y, int64_t n, double* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PYNAMA_B200_H */
