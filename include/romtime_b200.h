/* romtime_b200.h -- C ABI of libromtime_b200.so (B200 / sm_100a).
 *
 * Drop-in boundary for the offline-compression + online-solve hot path of KikeM/romtime
 * (reference @ 7bca28a; file:line citations are relative to the reference's src/romtime/).
 * The reference has no FFI layer: its boundary is the Python API, and the arithmetic goes to
 * LAPACK/BLAS through NumPy/SciPy.  Every entry point below replaces one of those call sites;
 * the Python classes in romtime_b200/ (same names and signatures as the reference) call them
 * through ctypes (see INTEGRATION.md for the reference-side stubs).
 *
 * Conventions
 *  - plain C: pointers + sizes, no torch types.  Every function returns an rt_status
 *    (0 = ok, negative = error); nothing throws across the ABI.  rt_last_error(h) gives text.
 *  - all matrix/vector pointers are DEVICE pointers owned by the caller (allocated by torch and
 *    handed over as data_ptr()); the library never frees or keeps them.  Scratch memory is owned
 *    by the handle.  `stream` is a cudaStream_t passed as void* (0 = legacy default stream); all
 *    work is enqueued on it and the call returns without synchronising unless stated.
 *  - matrices are fp64 with explicit element strides (stride_row, stride_col) so that NumPy
 *    C-order (np.hstack, rom/rom.py:368) and F-order (np.array(list).T, deim/deim.py:384;
 *    LAPACK outputs) arrays are both accepted without a host-side copy.
 *  - indices are int64 (NumPy intp).
 *  - one handle per device; a handle is not thread-safe.
 *
 * Limits (the reference accepts any size; these are the sizes the kernels are built for -- a call outside
 * them returns RT_EINVAL with the limit named in rt_last_error, the Python classes raise ValueError):
 *  - reduced dimension N <= 128 (rt_online_*: the dense solve keeps one system per CTA in shared memory);
 *  - collateral basis size m <= 160 for rt_invert_f64 / rt_deim_greedy_f64 (interpolation matrix in shared
 *    memory), snapshots n <= 16384 for rt_sym_eig_f64, n <= 65535 for rt_pod_gram_f64;
 *  - rt_errors_summary_f64: nt <= 16384;  rt_rom_difference_f64: N <= Nhat <= 128;
 *  - rt_pod_batched_f64: n <= 256 snapshots per problem.
 */
#ifndef ROMTIME_B200_H
#define ROMTIME_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_handle rt_handle;

typedef enum {
    RT_OK = 0,
    RT_EINVAL = -1,    /* bad argument (null pointer, negative size, unsupported shape)          */
    RT_ECUDA = -2,     /* CUDA runtime / driver error; text in rt_last_error                     */
    RT_ESINGULAR = -3, /* exactly singular matrix (mirrors numpy.linalg.LinAlgError)             */
    RT_ENOMEM = -4,    /* scratch allocation failed                                              */
    RT_ENOTCONV = -5,  /* eigensolver did not converge within its sweep limit                    */
    RT_ENODEVICE = -6  /* no sm_100 device: there is NO CPU fallback                             */
} rt_status;

#define RT_NF 12     /* number of (mu, t) features of the piston evaluator, see rt_rom_model */
#define RT_MAX_OPS 8

/* ---------------------------------------------------------------- life cycle -------- */
int rt_version(void);
int rt_create(int device, rt_handle** out);
int rt_destroy(rt_handle* h);
const char* rt_last_error(const rt_handle* h);
/* Enable loads from buffers that live on device `peer` (peer memory over NVLink / NVSwitch): afterwards any S / A / B
 * pointer of the GEMM-shaped calls below may be peer memory (e.g. a neighbour rank's snapshot columns, opened through
 * CUDA IPC); outputs stay local.  Used by the column-sharded POD Gram. */
int rt_enable_peer_access(rt_handle* h, int peer);
/* measured FP64 DMMA peak of this device in TFLOP/s (register-resident mma.sync.m8n8k4.f64 loop);
 * synchronous.  Used as the roofline denominator of the FP64-bound kernels. */
int rt_fp64_peak_tflops(rt_handle* h, double* tflops);

/* ---------------------------------------------------------------- POD (rom/pod.py:7-62) */
/* Snapshot Gram matrix  G = S^T S  (n x n, symmetric, written in full, row-major == col-major)
 * and column norms  colnorm[j] = sqrt(G[j][j])  (rom/pod.py:32-33 computes np.linalg.norm(axis=0)).
 * normalize != 0: G <- D^-1 G D^-1 with D = diag(colnorm), i.e. the Gram matrix of the
 * column-normalised snapshots that rom/pod.py:33 feeds to gesvd (:38).
 * Rows [row0, row0+rows) only are read when rows < Nh is wanted: pass the sub-pointer. */
int rt_pod_gram_f64(rt_handle* h, const double* S, int64_t Nh, int64_t n, int64_t stride_row,
                    int64_t stride_col, int normalize, double* G, double* colnorm, void* stream);

/* Cross Gram block  C = A^T B  (na x nb), A: Nh x na, B: Nh x nb, C row-major with leading
 * dimension ldc.  Used for the column-sharded multi-GPU Gram (block row G[g,:] = S_g^T [S_0 .. S_P-1],
 * followed by an NCCL all-reduce of G issued by the host) and for DEIM.project_basis
 * (V^T U_f, deim/deim.py:508-509). */
int rt_gram_cross_f64(rt_handle* h, const double* A, int64_t na, int64_t a_stride_row,
                      int64_t a_stride_col, const double* B, int64_t nb, int64_t b_stride_row,
                      int64_t b_stride_col, int64_t Nh, double* C, int64_t ldc, void* stream);

/* In-place finalisation of an assembled (e.g. all-reduced) Gram matrix: colnorm = sqrt(diag),
 * optional D^-1 G D^-1 scaling, exact symmetrisation (lower <- upper). */
int rt_gram_finalize_f64(rt_handle* h, double* G, int64_t n, int normalize, double* colnorm,
                         void* stream);

/* Symmetric eigen-decomposition  G = W diag(w) W^T  (replaces the right-singular side of
 * scipy.linalg.svd(..., lapack_driver="gesvd"), rom/pod.py:38: s = sqrt(w), VT = W^T).
 * G (n x n) is overwritten; W is written row-major n x n with eigenvector k in COLUMN k;
 * w is sorted descending (gesvd ordering).  Parallel-order block Jacobi.  `info` (device int,
 * may be NULL): number of sweeps used, or -1 if not converged. */
int rt_sym_eig_f64(rt_handle* h, double* G, int64_t n, double* w, double* W, int* info,
                   void* stream);

/* Symmetric eigen-decomposition by PLAIN block Jacobi (no pivoted-Cholesky preconditioner): all n eigenpairs, W a
 * product of plane rotations, rotation criterion |a_pq| > eps sqrt(a_pp a_qq).  On the Gram matrix H = Y^T Y of nearly
 * orthogonal, arbitrarily graded columns every eigenvalue has a small RELATIVE error -- the rotation step of the
 * one-sided (Hestenes) refinement that brings the method-of-snapshots POD to gesvd-level accuracy on every retained
 * mode (rom/pod.py:38 resolves sigma_k to eps sigma_1; the Gram route alone only to eps sigma_1^2 / sigma_k).
 * A (n x n row-major, ld n) is not modified; w descending; W row-major n x n, eigenvector k in column k;
 * info (device int, may be NULL): sweeps used; defect (device double, may be NULL): max_{i != j} |a_ij| / sqrt(a_ii a_jj)
 * of the input, i.e. how far the columns behind H are from orthogonal. */
int rt_sym_eig_jacobi_f64(rt_handle* h, const double* A, int64_t n, double* w, double* W, int* info,
                          double* defect, void* stream);

/* rt_sym_eig_jacobi_static_f64: rt_sym_eig_jacobi_f64 for callers that may not allocate.  No counterpart in the reference:
 * added (callers under CUDA-graph capture set-up, tight serving loops and harnesses
 * that forbid cudaMalloc).  Same arguments, same results; the work space is a fixed device array inside the library
 * (static __device__, about 322 KB, held by every GPU that loads the library until exit), so n <= 128
 * (RT_STATIC_EIG_MAX_N) -- a larger n returns RT_EINVAL with a message naming the limit.  The call makes no cudaMalloc /
 * cudaFree and calls no allocating library routine.  All calls on a device share the array: calls that overlap on the GPU
 * (two streams, two host threads) race and must be serialised by the caller; nothing in it survives between calls. */
#define RT_STATIC_EIG_MAX_N 128
int rt_sym_eig_jacobi_static_f64(rt_handle* h, const double* A, int64_t n, double* w, double* W, int* info,
                                 double* defect, void* stream);

/* Basis lift  Q = S Wr  (Nh x r): left singular vectors from the Gram eigenvectors,
 * Wr (n x r, row-major, leading dimension ldw) already scaled by 1/(sigma_k * colnorm_j).
 * Q is written with strides (q_stride_row, q_stride_col). */
int rt_pod_lift_f64(rt_handle* h, const double* S, int64_t Nh, int64_t n, int64_t stride_row,
                    int64_t stride_col, const double* Wr, int64_t ldw, int64_t r, double* Q,
                    int64_t q_stride_row, int64_t q_stride_col, void* stream);

/* Deflation step of the POD's second level: P (Nh x n row-major) holds Q (Q^T S D^-1) on entry (rt_gram_cross_f64 +
 * rt_pod_lift_f64) and  S D^-1 - Q (Q^T S D^-1)  on exit -- what the delivered modes Q leave of the (column-normalised,
 * colnorm != NULL) snapshots.  Its POD resolves the singular values the first Gram matrix cannot see
 * (sigma_k < 4e-8 sigma_1; rom/pod.py:56 keeps every s > 1e-7 ABSOLUTE). */
int rt_deflate_f64(rt_handle* h, const double* S, int64_t Nh, int64_t n, int64_t stride_row, int64_t stride_col,
                   const double* colnorm, double* P, void* stream);

/* ---------------------------------------------- DEIM greedy (deim/deim.py:517-561) ---- */
/* Classic DEIM point selection on the collateral basis U (Nh x m):
 *   p_0 = argmax |u_0| ; c = (P^T U_j)^-1 P^T u_j ; res = u_j - U_j c ; p_j = argmax |res|
 * with numpy.argmax semantics (FIRST maximum wins; NaN counts as maximum).
 * Outputs (device): dofs[m] (int64), PT_U[m*m] row-major = U[dofs, :] (deim/deim.py:159, :212),
 * info[1]: 0 ok, j>0 if the interpolation matrix became exactly singular at step j. */
int rt_deim_greedy_f64(rt_handle* h, const double* U, int64_t Nh, int64_t m, int64_t stride_row,
                       int64_t stride_col, int64_t* dofs, double* PT_U, int* info, void* stream);

/* ---------------------------------------------- projections ---------------------------- */
/* DEIM.project_basis (deim/deim.py:495-515): out (N x m, row-major) = V^T U_f. */
int rt_project_vec_f64(rt_handle* h, const double* V, int64_t v_stride_row, int64_t v_stride_col,
                       const double* U, int64_t u_stride_row, int64_t u_stride_col, int64_t Nh,
                       int64_t N, int64_t m, double* out, void* stream);

/* MDEIM.project_basis (deim/mdeim.py:153-192 + utils.py:96-149): for every collateral mode q,
 * A_q = csr(U[:, q]; rows, cols), out[(i*N + j), q] = (V^T A_q V)[i, j]; duplicates (rows[k],
 * cols[k]) are summed like scipy's COO->CSR conversion.  out is (N*N) x m with element strides
 * (out_stride_row, out_stride_col). */
int rt_project_mat_f64(rt_handle* h, const double* U, int64_t u_stride_row, int64_t u_stride_col,
                       const int64_t* rows, const int64_t* cols, int64_t nnz, const double* V,
                       int64_t v_stride_row, int64_t v_stride_col, int64_t Nh, int64_t N,
                       int64_t m, double* out, int64_t out_stride_row, int64_t out_stride_col,
                       void* stream);

/* ---------------------------------------------- theta solve + assembly ----------------- */
/* Inverse of a small dense matrix by Gauss-Jordan with partial pivoting ("factor once",
 * replaces the per-call LU of np.linalg.solve, deim/deim.py:491-492).  A, Ainv: m x m row-major.
 * info (device int): 0 ok, k>0 zero pivot at step k. */
int rt_invert_f64(rt_handle* h, const double* A, int64_t m, double* Ainv, int* info, void* stream);

/* compute_thetas for a batch (deim/deim.py:477-493): theta[b, :] = PT_U^-1 F[b, :].
 * PT_Uinv from rt_invert_f64; F, theta: B x m row-major. */
int rt_theta_f64(rt_handle* h, const double* PT_Uinv, int64_t m, const double* F, int64_t B,
                 double* theta, void* stream);

/* _interpolate for a batch (deim/deim.py:445): out[b, :] = sum_q theta[b, q] * basis[:, q].
 * basis: R x m with element strides; out: B x R row-major (R = N*N for MDEIM which="rom",
 * reshaped (N, N) row-major per deim/mdeim.py:250-251). */
int rt_assemble_batch_f64(rt_handle* h, const double* basis, int64_t R, int64_t m,
                          int64_t b_stride_row, int64_t b_stride_col, const double* theta,
                          int64_t B, double* out, void* stream);

/* ---------------------------------------------- online sweep (rom/rom.py:430-555) ------ */
/* One hyper-reduced operator of the ROM (a DEIM/MDEIM/N-MDEIM reductor after project_basis). */
typedef struct {
    int32_t m;               /* collateral basis size (deim N); 0 = operator absent                */
    int32_t is_vector;       /* 0: matrix operator (basis_rom is N*N x m), 1: vector (N x m)       */
    const double* basis_rom; /* column q contiguous: element (r, q) at basis_rom[q * ld + r]       */
    int64_t ld;              /* leading dimension of basis_rom (>= N*N or N)                       */
    const double* PT_Uinv;   /* m x m row-major, from rt_invert_f64                                */
    const double* C;         /* m x RT_NF row-major: sampled entries f = C phi(mu,t) (+ b0 T u*)   */
    const double* T;         /* m x N row-major or NULL: state-dependent part (trilinear operator) */
} rt_rom_operator;

/* RomConstructorNonlinear model (rom/rom.py:877-929).  ops order: mass, stiffness, convection,
 * nonlinear_lifting, trilinear, rhs.  evaluator (the closed-form features phi(mu, t) of the device sampled-entry
 * evaluator, entries = C phi + b0 T u*; tables from romtime_b200/evaluators.py):
 *   0 = piston "rest", 1 = piston "sudden" (problems/piston.py:18-60), mu rows (a0, omega, delta):
 *       phi = [1, L, 1/L, L', a0, b0 bL, bL' L, b0 bL^2, a0 bL, L' bL, bL/L, b0]   (fom/nonlinear.py:374-581)
 *   2 = MFP1 heat problem on L(t) = 1 - sin(omega t) (problems/mfp1.py:27-75; omega = 0: fixed domain), mu rows
 *       (delta, beta, alpha_0, omega), E = exp(-beta t):
 *       phi = [1, L, alpha_0/L, L', beta E L, beta E delta^2 L^3, alpha_0 (1-E) delta^2 L, (1-E) delta^2 L' L^2, 0, 0, 0, 0]
 *       (fom/heat.py:80-297, fom/base.py:377-495)
 * n_mu: doubles per row of the sweep's mu array (0 means 3). */
typedef struct {
    int32_t N;
    int32_t n_ops; /* 6 */
    int32_t evaluator;
    int32_t generation; /* 0: stage the model on every call.  Non-zero: a caller-chosen id, unique per set of device
                         * tables; consecutive calls with a bitwise-identical rt_rom_model (same id, same pointers)
                         * reuse the staged, PT_U^-1-folded operators kept in the handle.  Change the id whenever
                         * the CONTENTS behind the pointers change. */
    int32_t n_mu;
    int32_t reserved;
    rt_rom_operator ops[RT_MAX_OPS];
} rt_rom_model;

/* Batched RomConstructorNonlinear.solve: for every parameter b (mu[b, 0:3] = a0, omega, delta)
 * (n_mu doubles per row for the other evaluators) run nt implicit BDF steps from a zero initial condition (rom/rom.py:451-452):
 *   sampled entries -> theta -> K = bdf M + dt (A + C + N(u*) + N^) -> b = M w + dt f -> K u = b
 * (dense LU with partial pivoting replaces GMRES, rom/rom.py:36, :492).
 * bdf: 1 or 2.  uN_out: B x nt x N (row-major; [b][n][:] = reduced solution after step n).
 * info: B ints, 0 ok, n>0 zero pivot at step n.  All pointers device. */
int rt_online_sweep_f64(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B,
                        int64_t nt, double dt, int bdf, double* uN_out, int* info, void* stream);

/* Same sweep that only keeps the final state (uN_last: B x N) -- used for long benchmark sweeps
 * where the B x nt x N history (33.5 GB at C5) is not wanted on the device. */
int rt_online_sweep_last_f64(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B,
                             int64_t nt, double dt, int bdf, double* uN_last, int* info,
                             void* stream);

/* Steps n0 .. n1 - 1 of the same nt-step sweep, with the state carried by the caller (Un, Unm1: B x N row-major, device;
 * zeroed by the call when n0 == 0, updated in place).  Consecutive segments [0, n1), [n1, n2), ... reproduce the
 * uninterrupted sweep bit for bit (same time accumulation t += dt, rom/rom.py:476-479; same BDF start-up, :481-483).
 * Every segment writes its rows [b][n0 .. n1 - 1][:] of uN_out (B x nt x N): the host-delivered sweep copies a finished
 * time slab to pinned memory (rt_memcpy2d_async) while the next segment is computed. */
int rt_online_sweep_segment_f64(rt_handle* h, const rt_rom_model* model, const double* mu, int64_t B,
                                int64_t nt, int64_t n0, int64_t n1, double dt, int bdf, double* Un,
                                double* Unm1, double* uN_out, int* info, void* stream);

/* cudaMemcpy2DAsync on `stream`: `rows` rows of `width_bytes`, pitches in bytes; to_host != 0: device -> host,
 * else host -> device.  (RomSolutionsStorage histories leave the device through this call, base.py:70-79.) */
int rt_memcpy2d_async(rt_handle* h, void* dst, int64_t dst_pitch, const void* src, int64_t src_pitch,
                      int64_t width_bytes, int64_t rows, int to_host, void* stream);

/* ONE implicit step for a batch of B parameters with caller-supplied sampled entries -- the form the
 * reference's own time loop needs when the assemble(mu, t, entries[, u_n]) callback stays on the host
 * (FEniCS; rom/rom.py:486-492 with deim/deim.py:437-445 inside every assemble_* call):
 *   F      : B x Qsum row-major, Qsum = sum of ops[i].m; row b = the sampled entries f_local of every
 *            reductor side by side in model order (an operator with m == 0 is absent and skipped);
 *   theta_i = PT_U_i^-1 f_i ;  M = sum theta_mass A_q ;  K = bdf M + dt sum_(other matrix ops) ;
 *   b = M w + dt f_rhs  with w = u_n (bdf 1) or 2 u_n - 0.5 u_nm1 (bdf 2, rom/rom.py:919-925) ;  K u = b.
 *   bdf coefficient: 1.5 when bdf == 2 and step > 0, else 1.0 (rom/rom.py:481-483).
 * Optional dense terms for operators WITHOUT a reductor, projected by the caller (rom/rom.py:135-158):
 *   K += Kadd[b], M += Madd[b] (B x N x N row-major), b += fadd[b] (B x N); pass NULL when unused.
 *   (Kadd must already contain bdf * Madd + dt * (...): it is added to K as is.)
 * State (device, caller-owned, B x N row-major): Unm1 <- Un, Un <- solution (rom/rom.py:499-503).
 * model->ops[i].C / .T / model->evaluator are not read.  info: B ints, zero-initialised by the caller
 * before the first step; set to step+1 where a zero pivot was met. */
int rt_online_step_f64(rt_handle* h, const rt_rom_model* model, const double* F, int64_t B, double dt,
                       int bdf, int64_t step, const double* Kadd, const double* Madd,
                       const double* fadd, double* Un, double* Unm1, int* info, void* stream);

/* Host-only helper (no device work): parameters per chunk the sweep uses for a batch of B parameters,
 * reduced dimension N, on a device with sm_count SMs -- a multiple of 128 chosen so that the assembly
 * GEMM's output tiles fill whole waves of sm_count CTAs.  Exposed for tests and for callers that want
 * to size parameter shards (multi-GPU: romtime_b200/distributed.py). */
int64_t rt_sweep_chunk(int64_t B, int N, int sm_count);

/* Per-kernel-class device timing of the online sweep (CUDA events on the caller's stream around every
 * launch of the first parameter chunk).  Classes: 0 state/features, 1 T u* GEMM, 2 sampled entries,
 * 3 theta GEMM, 4 assembly GEMM (K and M), 5 forcing GEMM, 6 LU solve, 7 the fused pre-step kernel (classes 0, 1, 2 and 5
 * in one launch: N <= 64, trilinear and forcing bases of at most 128 modes).
 * rt_profile_read synchronises, returns accumulated milliseconds and launch counts, and resets. */
#define RT_PROFILE_CLASSES 8
int rt_profile_enable(rt_handle* h, int on);
int rt_profile_read(rt_handle* h, double* ms, int64_t* launches);

/* number of kernels launched by this handle since creation (bench.py "gpu_launches") */
int64_t rt_launch_count(const rt_handle* h);

/* ---------------------------------------------- batched small POD (tree walk) ------------- */
/* The reference consumes orth as many small per-parameter time-PODs followed by one stacked mu-POD (rom/rom.py:313-406,
 * deim/deim.py:279-397, deim/nonlinear.py:435-466).  These two calls do the eigen-work of a whole BATCH of such PODs in
 * one launch each (one CTA per problem, everything in shared memory, no host round trip inside); the Gram matrices and
 * the lifts around them are the ordinary rt_pod_gram_f64 / rt_pod_lift_f64 launches, queued without synchronisation.
 *
 * rt_small_gram_eig_f64: problem p has the n_p x n_p Gram matrix G_all + p * stride_g (row-major, leading dimension
 * n_p; n_of[p] = n_p <= nmax <= 256, device ints).  Out: w_all[p * nmax + k] eigenvalues descending (0 beyond the
 * numerical rank: pivoted-Cholesky cut at 8 ulp of the largest diagonal entry, as rt_sym_eig_f64), W_all + p * stride_w
 * the n_p x n_p row-major eigenvector matrix (vector k in column k, zero beyond the rank), rank[p], info[p]:
 * 0 ok, 1 = numerical rank above 96 (use rt_sym_eig_f64 for this problem), 2 = Jacobi did not converge. */
int rt_small_gram_eig_f64(rt_handle* h, const double* G_all, int64_t stride_g, const int* n_of, int64_t P,
                          int64_t nmax, double* w_all, double* W_all, int64_t stride_w, int* rank, int* info,
                          void* stream);
/* The Gram matrices / lifts of a batch, queued by one call: d[p] (HOST array) holds the arguments of rt_pod_gram_f64 /
 * rt_pod_lift_f64 for problem p (device pointers). */
typedef struct {
    const double* S;
    int64_t Nh, n, stride_row, stride_col;
    double* G;       /* n x n, leading dimension n */
    double* colnorm; /* n */
} rt_gram_desc;
typedef struct {
    const double* S;
    int64_t Nh, n, stride_row, stride_col;
    const double* Wr;
    int64_t ldw, r;
    double* Q;
    int64_t q_stride_row, q_stride_col;
} rt_lift_desc;
int rt_pod_gram_batch_f64(rt_handle* h, const rt_gram_desc* d, int64_t P, int normalize, void* stream);
int rt_pod_lift_batch_f64(rt_handle* h, const rt_lift_desc* d, int64_t P, void* stream);

/* rt_small_refine_f64: the Hestenes refinement step for a batch: H_p = Y_p^T Y_p (c_p x c_p row-major, leading dimension
 * c_p, c_of[p] = c_p <= cmax <= 96) of lifted, nearly orthogonal graded columns -> all eigenpairs with small RELATIVE
 * errors (Cholesky + one-sided Jacobi).  Out: lam_all[p * cmax + k] descending, X_all + p * stride_x (c_p x c_p row-major,
 * vector k in column k), defect[p] = max |h_ij| / sqrt(h_ii h_jj) of the input, info[p]: 0 ok, 1 = H not positive
 * definite (use rt_sym_eig_jacobi_f64), 2 = not converged. */
int rt_small_refine_f64(rt_handle* h, const double* H_all, int64_t stride_h, const int* c_of, int64_t P,
                        int64_t cmax, double* lam_all, double* X_all, int64_t stride_x, double* defect, int* info,
                        void* stream);

/* ---------------------------------------------- certification (rom/hrom.py:504-626) -------- */
/* S-ROM error estimator of utils.compute_rom_difference (utils.py:173-212) for every (parameter, step) of a pair
 * of sweeps:  err[b*nt + n] = || V_srom (uN_srom[b, n, :] - pad(uN[b, n, :])) ||_2 / sqrt(Nh).
 * uN (B x nt x N) and uN_srom (B x nt x Nhat, N <= Nhat <= 128) are the device histories of rt_online_sweep_f64;
 * G (Nhat x Nhat row-major) = V_srom^T V_srom, e.g. from rt_pod_gram_f64 with normalize = 0; Nh = rows of V_srom. */
int rt_rom_difference_f64(rt_handle* h, const double* uN, int64_t B, int64_t nt, int64_t N,
                          const double* uN_srom, int64_t Nhat, const double* G, int64_t Nh, double* err,
                          void* stream);

/* Reductor.create_errors_summary (rom/base.py:154-163) on the device: err (B x nt) ->
 * summary (B x 4 row-major) = {mean, median, max, min} over the nt steps of every parameter; nt <= 16384. */
int rt_errors_summary_f64(rt_handle* h, const double* err, int64_t B, int64_t nt, double* summary, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ROMTIME_B200_H */
