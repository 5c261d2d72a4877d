/* rails_cuda.h -- C ABI of the B200 (sm_100a) backend for the RAILS hot path.
 *
 * This is the drop-in boundary below the reference's plugin surface
 * (SURVEY.md §8(b)).  The reference's solver loop
 * (/root/reference/src/LyapunovSolver.hpp:100-346) talks only to three
 * duck-typed C++ classes -- Operator, MultiVector, DenseMatrix
 * (/root/reference/src/LyapunovSolverDecl.hpp:9-51).  The C++ wrapper classes
 * in rails_b200/host/ implement those concepts for the GPU and call nothing
 * but the functions declared here; each declaration cites the reference
 * method whose arithmetic it replaces (StlWrapper = /root/reference/src/
 * StlWrapper.cpp, EpetraMV = src/Epetra_MultiVectorWrapper.cpp, EpetraOp =
 * src/Epetra_OperatorWrapper.cpp).
 *
 * Conventions
 *   - plain pointers and sizes only; every function returns 0 on success and
 *     a non-zero code on failure (the message is kept for
 *     rails_cuda_last_error()).  No exceptions cross this boundary, matching
 *     the reference's int-code error style (src/LyapunovSolver.hpp:89-95).
 *   - a rails_mv_t is a column-major FP64 block resident in HBM: `rows` local
 *     rows, leading dimension padded to a multiple of 16 doubles so that every
 *     column starts on a 128-byte line, `cols` columns of capacity.
 *     Operations address a column range [col0, col0+ncols).
 *   - one process drives one GPU.  With a communicator
 *     (rails_cuda_comm_init) a DISTRIBUTED multivector holds this rank's
 *     contiguous row block of an Epetra-style linear map and inner products
 *     are summed over ranks; a REPLICATED one (Epetra_LocalMap analogue,
 *     src/Epetra_MultiVectorWrapper.cpp:226,451) is identical on all ranks.
 *   - work is enqueued on the library's stream; a call synchronises only when
 *     it returns a value to host memory.
 *   - there is no CPU fallback: without a usable CUDA device every compute
 *     entry point fails with RAILS_ERR_CUDA.
 */
#ifndef RAILS_CUDA_H
#define RAILS_CUDA_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rails_mv_s *rails_mv_t;
typedef struct rails_csr_s *rails_csr_t;
typedef struct rails_lu_s *rails_lu_t;
typedef struct rails_idx_s *rails_idx_t;

enum {
    RAILS_OK = 0,
    RAILS_ERR_CUDA = 1,      /* CUDA runtime / NCCL failure or no device */
    RAILS_ERR_ARG = 2,       /* bad handle, range or shape               */
    RAILS_ERR_STATE = 3      /* library not initialised                  */
};

enum {
    RAILS_MV_DISTRIBUTED = 0,
    RAILS_MV_REPLICATED = 1
};

#define RAILS_NCCL_ID_BYTES 128

/* ---- context --------------------------------------------------------- */
int rails_cuda_init(int device);
int rails_cuda_finalize(void);
int rails_cuda_is_initialized(void);
int rails_cuda_sync(void);
const char *rails_cuda_last_error(void);
/* number of failed calls so far: lets a caller that drives the wrappers in the
 * reference's "print and carry on" style (src/StlWrapper.cpp:173-179) stop a
 * solve cleanly after, e.g., an out-of-memory allocation */
long long rails_cuda_error_count(void);
/* number of this library's kernels launched since init (bench.py's
 * gpu_launches) */
long long rails_cuda_kernel_launches(void);
/* stream synchronisations made by the library so far: how often the host waited for
 * the device (bench.py reports the count per outer iteration) */
long long rails_cuda_host_syncs(void);
/* Sustained FP64 throughput of this GPU in TFLOP/s, measured with a register-only
 * kernel on every SM: kind 0 = DFMA (vector pipe), 1 = DMMA (tensor cores).  The
 * denominator of the FLOP-bound rooflines bench.py reports (restart compression,
 * warm-start projection: SURVEY F13); MEASURED_PEAKS.json has no FP64 figure. */
int rails_cuda_fp64_peak(int kind, double *tflops);
/* device memory currently held by multivectors and operators, bytes */
long long rails_cuda_bytes_in_use(void);
/* CUDA-event timing on the library stream (bench.py, Timer sections):
 * rails_cuda_event_record(slot) marks a point, rails_cuda_event_elapsed_ms
 * synchronises on `stop` and returns the time between two marks. */
int rails_cuda_event_record(int slot);
int rails_cuda_event_elapsed_ms(int start_slot, int stop_slot, float *ms);

/* ---- communicator (row partition over GPUs; Epetra_MpiComm analogue,
 *      src/main.cpp:47) ------------------------------------------------- */
int rails_cuda_comm_unique_id(char id[RAILS_NCCL_ID_BYTES]);
int rails_cuda_comm_init(int rank, int nranks, const char id[RAILS_NCCL_ID_BYTES]);
/* 1 when small reductions go through the peer-memory mailboxes (NVLink/NVSwitch,
 * CUDA IPC), 0 when they use NCCL all-reduce */
int rails_cuda_comm_p2p(void);
int rails_cuda_comm_rank(void);
int rails_cuda_comm_size(void);
/* Epetra_Map(m, 0, comm) linear map: rows [begin, end) of rank r of P */
int rails_partition_rows(long long m, int nranks, int rank,
                         long long *begin, long long *end);

/* ---- multivector storage (StlWrapper ctors/resize/view/push_back,
 *      src/StlWrapper.cpp:19-121,219-263,323-392) ----------------------- */
int rails_mv_alloc(long long rows, int cols, int flags, rails_mv_t *out);
int rails_mv_free(rails_mv_t mv);
long long rails_mv_rows(rails_mv_t mv);
long long rails_mv_ld(rails_mv_t mv);
int rails_mv_cols(rails_mv_t mv);
int rails_mv_flags(rails_mv_t mv);
/* dst[:, dcol..] = src[:, scol..]   (memcpy in push_back :367-374, view
 * assignment :65-121, copy ctor :31-44) */
int rails_mv_copy_cols(rails_mv_t dst, int dcol, rails_mv_t src, int scol, int ncols);
/* host <-> device, host block is column-major with leading dimension ldh */
int rails_mv_upload(rails_mv_t mv, int col0, int ncols, const double *host, long long ldh);
int rails_mv_download(rails_mv_t mv, int col0, int ncols, double *host, long long ldh);
/* test-only element access (the (m,n) ctor + operator()(i,j) of
 * test/GenericMultiVectorWrapper_test.cpp:23-29) */
int rails_mv_get(rails_mv_t mv, long long row, int col, double *value);
int rails_mv_set(rails_mv_t mv, long long row, int col, double value);

/* ---- BLAS-1 family (StlWrapper :123-166) ------------------------------ */
int rails_mv_fill(rails_mv_t mv, int col0, int ncols, double value);     /* = double :123-129 */
int rails_mv_scal(rails_mv_t mv, int col0, int ncols, double alpha);     /* *=, /=   :131-145 */
/* y[:, yc0..] += alpha * x[:, xc0..]                    (+=, -= :147-161) */
int rails_mv_axpy(rails_mv_t y, int yc0, double alpha, rails_mv_t x, int xc0, int ncols);
/* dst[:, dcol..] = alpha * src[:, scol..]: the scaled copy that
 * `double * MultiVector` returns (:481-487) */
int rails_mv_scaled_copy(rails_mv_t dst, int dcol, double alpha, rails_mv_t src, int scol, int ncols);
/* y += sign * round(alpha * x), sign = +1 or -1: `y -= alpha * x` of
 * resid_lanczos (src/LyapunovSolver.hpp:411-413) in one pass, with the
 * product rounded before the add exactly as the scaled temporary would be */
int rails_mv_add_scaled(rails_mv_t y, int yc0, double sign, double alpha, rails_mv_t x, int xc0, int ncols);
/* max_i sum_j |x_ij| over this rank's rows, then max over ranks (:291-303) */
int rails_mv_norm_inf(rails_mv_t mv, int col0, int ncols, double *out);

/* ---- inner products: out = X' Y  (a x b, host, column-major, ld = ldo)
 *      StlWrapper::dot :394-412, EpetraMV::dot :418-438.  Includes the
 *      cross-rank sum when both operands are distributed. ---------------- */
int rails_mv_dot(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b,
                 double *out, int ldo);
/* `count` such products with one host synchronisation (and one all-reduce):
 * out[i] (a[i] x b[i], ld = ldo[i]) = X[i][:, xc0[i]..]' Y[i][:, yc0[i]..].  The
 * projection update (src/LyapunovSolver.hpp:173-195) issues its dots this way. */
int rails_mv_dot_batch(int count, const rails_mv_t *X, const int *xc0, const int *a,
                       const rails_mv_t *Y, const int *yc0, const int *b,
                       double *const *out, const int *ldo);
/* same product left on the device in a REPLICATED multivector (a rows);
 * used for Y = B' q (src/LyapunovSolver.hpp:150,389) without a host sync */
int rails_mv_dot_dev(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b,
                     rails_mv_t Out, int oc0);
/* Allocation-free variant of rails_mv_dot_dev (local product, no sum over ranks).
 * NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot allocate (code under CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  Calls no cudaMalloc / cudaFree / allocating
 * routine, does not synchronise; the per-CTA partials live in a static device array
 * of RAILS_NOALLOC_SCRATCH_DOUBLES doubles (1.16 MiB, held by every GPU that loads
 * the library until exit).  A product that needs more -- CTAs (<= 4 per SM) x
 * streamed columns x held columns per launch (<= 8) -- is rejected with
 * RAILS_ERR_ARG, as are row-partitioned operands with more than one rank.  All calls
 * share the array: calls that overlap on the GPU race; calls on the library stream
 * are ordered.  Same kernels and summation order as rails_mv_dot_dev when neither
 * block is wider than 8 columns (bit-identical results). */
#define RAILS_NOALLOC_SCRATCH_DOUBLES 151552
int rails_mv_dot_dev_noalloc(rails_mv_t X, int xc0, int a, rails_mv_t Y, int yc0, int b,
                             rails_mv_t Out, int oc0);

/* ---- tall-skinny times small: Out[:, oc0..oc0+c) = alpha * V[:, vc0..vc0+k) * C
 *      + beta * Out      (StlWrapper::operator* :168-187, EpetraMV :194-241)
 *      C is k x c column-major on the host (ldc) or a replicated multivector
 *      on the device.  Out must not overlap the V columns read. ---------- */
int rails_mv_gemm(rails_mv_t V, int vc0, int k, const double *C, int ldc, int c,
                  double alpha, double beta, rails_mv_t Out, int oc0);
int rails_mv_gemm_dev(rails_mv_t V, int vc0, int k, rails_mv_t C, int cc0, int c,
                      double alpha, double beta, rails_mv_t Out, int oc0);

/* ---- fused residual-Lanczos step (Solver::resid_lanczos,
 *      src/LyapunovSolver.hpp:384-434).  Up to 4 column blocks
 *      X_b[:, xc0_b : xc0_b + xn_b) -- in the solver [V, AV, B] -- are treated
 *      as one wide multivector so that each is streamed once per kernel. ----- */
/* out (host, sum_b xn_b doubles, block after block) = [X_0 X_1 ..]' (qscale * q[:, qc]);
 * when qscale != 1 the scaled column is written back in the same pass (the
 * `q /= beta` of :431, deferred from the previous step).  Replaces the three
 * dots of :389,394,400. */
int rails_mv_multi_dot(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                       rails_mv_t q, int qc, double qscale, double *out);
/* out[:, oc] = sum_b X_b * coef_b  (coef: the blocks' coefficient vectors
 * concatenated, host); with q != NULL also *alpha = out[:, oc]' q[:, qc].
 * Replaces `next = B*Y; next += AV*Z; next += V*Z; alpha = next.dot(q)`
 * (:390,396,402,406). */
int rails_mv_multi_gemv(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                        const double *coef, rails_mv_t out, int oc,
                        rails_mv_t q, int qc, double *alpha);
/* next[:, nc] -= alpha * q[:, qc]; next[:, nc] -= beta * p[:, pc] (p may be
 * NULL); *norm = |next[:, nc]|_2.  Each product is rounded before it is
 * subtracted, as the scaled temporaries of :411-413 are; the norm is :417. */
int rails_mv_lanczos_update(rails_mv_t next, int nc, double alpha, rails_mv_t q, int qc,
                            double beta, rails_mv_t p, int pc, double *norm);

/* The whole L-step recurrence (:375-434) as one asynchronous chain: blocks 0 and
 * 1 (k columns each; V or MV, and AV) exchange their projections through the
 * host matrix T (k x k, ldt), further blocks (B) use theirs directly.
 * Q[:, qc0] holds the un-normalised start vector and Q has room for L + 1
 * columns from qc0.  On return alpha[i] = H(i,i), beta[i] = H(i+1,i) for
 * i < *steps; *steps < L after a breakdown (beta < 1e-14, :419-427).  One host
 * synchronisation for the whole call; per step one kernel that reads the blocks
 * once (TMA-staged tiles) and one single-CTA kernel that finishes the reduction,
 * exchanges it with the other ranks and forms the k x k products and alpha. */
int rails_mv_lanczos_run(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                         const double *T, int ldt, rails_mv_t Q, int qc0, int L,
                         double *alpha, double *beta, int *steps);

/* Widest [X_0 X_1 ..] (sum of xn) rails_mv_lanczos_run, _begin, _project and the
 * multi_* entry points accept (2040: the reduced vector plus the squared norm fits one
 * peer-memory exchange); a caller with a wider space uses the plain sequence. */
int rails_mv_lanczos_max_columns(void);

/* Host-only (no CUDA call, usable before rails_cuda_init): the launch shape the
 * one-kernel Lanczos step takes for blocks of xn[0..nblk) columns.  shape6 = {tile
 * rows (0: the two-kernel chain runs instead), pipeline stages, column groups, dynamic
 * shared memory in bytes, TMA requests per 128-row slab, consumer warps}. */
int rails_lanczos_step_shape(int nblk, const int *xn, int *shape6);

/* Optional head start: the first projection of the chain does not depend on T, so
 * it can be enqueued before the host solves the projected problem
 * (src/LyapunovSolver.hpp:209); the next rails_mv_lanczos_run with the same Q,
 * qc0 and blocks continues from it. */
int rails_mv_lanczos_begin(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                           rails_mv_t Q, int qc0);
/* First kernel of a step of that chain on its own (unit tests, roofline timing):
 * with prev_col >= 0, Q[:, col] -= alpha * q_prev + |Q[:, prev_col]| * Q[:, prev2_col]
 * (q_prev = Q[:, prev_col] normalised, written back; prev2_col < 0: no last term),
 * i.e. the update of :411-413 and the `/= beta` of :431 fused into the
 * projection; then out[0..total) = [X_0 X_1 ..]' Q[:, col], out[total] = |Q[:, col]|^2. */
int rails_mv_lanczos_project(int nblk, const rails_mv_t *X, const int *xc0, const int *xn,
                             rails_mv_t Q, int col, int prev_col, int prev2_col,
                             double alpha, double *out);

/* Measurement aid (bench.py `roofline`; no reference counterpart): with enable != 0
 * every launch of the one-kernel Lanczos step inside rails_mv_lanczos_run is bracketed
 * by CUDA events on the library's stream.  _read synchronises, returns the summed
 * kernel time in ms, the number of launches and the shape of the last launch
 * {block columns, tile rows, pipeline stages, column groups, CTAs}, and resets. */
int rails_lanczos_step_timing(int enable);
int rails_lanczos_step_timing_read(double *ms_total, int *launches, int *shape5);

/* ---- Gram-Schmidt: column-wise CGS2 of columns [first, last] against all
 *      earlier columns (StlWrapper::orthogonalize :305-321), executed as one
 *      device-side chain with no host round trip. ------------------------ */
int rails_mv_orthogonalize(rails_mv_t V, int col0, int first, int last);

/* ---- random(): bit-exact restatement of StlWrapper::random :414-423 with
 *      libstdc++'s default_random_engine (= minstd_rand0) seeded by `seed`
 *      (the caller passes std::rand()), filling the GLOBAL rows x ncols
 *      block row-major; this rank writes rows [row_offset, row_offset+rows).
 *      Values in (-1, 1). ------------------------------------------------ */
int rails_mv_random(rails_mv_t mv, int col0, int ncols, long long row_offset,
                    unsigned int seed);

/* ---- operator: CSR x tall-skinny (EpetraOp::operator* :75-91 ->
 *      Epetra_CrsMatrix::Apply).  rowptr has rows+1 entries; colidx holds
 *      GLOBAL column indices; this rank owns global rows
 *      [row_offset, row_offset+rows) of a global_rows x global_cols matrix. */
int rails_csr_create(long long rows, long long global_rows, long long global_cols,
                     long long row_offset, const int *rowptr, const int *colidx,
                     const double *vals, rails_csr_t *out);
int rails_csr_free(rails_csr_t A);
/* Host-only halo plan of a row-partitioned operator: what Epetra_CrsMatrix's
 * column map + Epetra_Import hold for the ghost import of
 * src/Epetra_OperatorWrapper.cpp:87.  local_col[nnz] = column indices
 * renumbered [owned | ghost]; the *nghost ghost rows in ascending global order
 * (ghost_global), as row ids in their owner's numbering (ghost_owner_row), and
 * need_from[nranks] = how many come from each rank.  No CUDA call is made: this
 * is the specification the CPU tests pin (tests/test_partition_gloo.py). */
int rails_csr_halo_plan(long long rows, long long global_rows, long long row_offset,
                        int nranks, const int *rowptr, const int *colidx,
                        int *local_col, int *nghost, int *ghost_global,
                        int *ghost_owner_row, int *need_from);
/* The same plan built on the device, as rails_csr_create builds it (ghost columns
 * marked in a byte map of the global index range, compacted in ascending order,
 * indices renumbered in place), then exchanged over NCCL there.  Same arguments and
 * results as rails_csr_halo_plan; host arrays in and out. */
int rails_csr_halo_plan_device(long long rows, long long global_rows, long long row_offset,
                               int nranks, const int *rowptr, const int *colidx,
                               int *local_col, int *nghost, int *ghost_global,
                               int *ghost_owner_row, int *need_from);
/* New values on the same sparsity pattern (same order as the arrays given to
 * rails_csr_create): a continuation step or a re-assembled Jacobian keeps the row
 * blocks, the renumbered columns and the halo plan.  The reference replaces the
 * Epetra_CrsMatrix behind the operator (src/main.cpp:66-98 per run). */
int rails_csr_update_values(rails_csr_t A, const double *vals);
/* test-only element access, entries of the sparsity pattern only (the (m, n)
 * constructor + operator()(i, j) of src/Epetra_OperatorWrapper.cpp:224-282,
 * test/GenericOperatorWrapper_test.cpp:36-46) */
int rails_csr_get_value(rails_csr_t A, long long row, long long col, double *value);
int rails_csr_set_value(rails_csr_t A, long long row, long long col, double value);
/* the operator as a dense rows x cols block in Out[:, oc0 ..) (small operators:
 * Operator::norm() forms A'A explicitly, src/Epetra_OperatorWrapper.cpp:115-145) */
int rails_csr_to_dense(rails_csr_t A, rails_mv_t Out, int oc0);
long long rails_csr_rows(rails_csr_t A);
long long rails_csr_nnz(rails_csr_t A);
/* Y[:, yc0..yc0+ncols) = op(A) * X[:, xc0..xc0+ncols);  trans != 0 applies A'
 * (transpose flag of EpetraOp :50-55,81; single-rank only) */
int rails_csr_spmm(rails_csr_t A, int trans, rails_mv_t X, int xc0, int ncols,
                   rails_mv_t Y, int yc0);
/* sqrt(lambda_max(A'A)) by power iteration on the device (EpetraOp::norm
 * :115-145; reached only when B is an Operator) */
int rails_csr_norm2(rails_csr_t A, double *out);

/* ---- sparse LU solves: the `A11 \ x` inside every apply of the reference's
 *      Schur-complement operator for a singular mass matrix
 *      (src/SchurOperator.cpp:177-186 factorise with Amesos_Klu, :204-208,
 *      :252-256,:262-266 solve; matlab/RAILSschur.m:31-36 lu() and the two
 *      solve closures) and the `A \ x` behind the inverse projection methods
 *      (matlab/RAILSschur.m:60-66, RAILSsolver.m:294-298,520-524).
 *      The factors P A Q = L U are computed once on the host (as KLU's are):
 *      L unit lower triangular with only its strict part stored, U upper
 *      triangular with its diagonal, both column-compressed in pivot numbering;
 *      pinv[i] = pivot position of row i, q[j] = column eliminated at step j.
 *      They are kept in HBM with a level schedule per triangle; a solve is a few
 *      launches whatever the number of levels.  Single rank only (the reference
 *      solver is serial too). ------------------------------------------------- */
int rails_lu_create(int n, const int *Lp, const int *Li, const double *Lx,
                    const int *Up, const int *Ui, const double *Ux,
                    const int *pinv, const int *q, rails_lu_t *out);
int rails_lu_free(rails_lu_t F);
/* X[:, xc0..xc0+nc) = op(A)^-1 B[:, bc0..bc0+nc);  trans != 0: A'.  X may be B. */
int rails_lu_solve(rails_lu_t F, int trans, rails_mv_t B, int bc0, int nc,
                   rails_mv_t X, int xc0);
/* stored entries of L + U (unit diagonal of L not counted) */
long long rails_lu_nnz(rails_lu_t F);
/* levels of the schedule of triangle `which` (0: L, 1: U, 2: U', 3: L') and the
 * kernels one solve launches */
int rails_lu_levels(rails_lu_t F, int which);
int rails_lu_launches_per_solve(rails_lu_t F, int trans);
/* Host-only check of the device schedule (no CUDA call, like rails_csr_halo_plan):
 * builds the level schedules, chains, shared-memory slots and windows exactly as
 * rails_lu_create does and runs one solve over x (n x nrhs, ld, in place) in the
 * order and from the arrays the kernels use.  stats[12], per triangle of the
 * solve: levels, launches segments, chains, windows, rows in chains, entries
 * between rows of a chain. */
int rails_lu_plan_emulate(int n, const int *Lp, const int *Li, const double *Lx,
                          const int *Up, const int *Ui, const double *Ux,
                          const int *pinv, const int *q, int trans, double *x, int nrhs,
                          long long ld, long long *stats);

/* ---- row index lists: the idx1 / idx2 splitting of the Schur complement
 *      (Epetra_Import between the maps of src/SchurOperator.cpp:93-94,131-132,
 *      228-238,280-286; x(idx2, :) and Xreorder of matlab/RAILSschur.m:52-60) --- */
int rails_idx_create(long long n, const int *host_idx, rails_idx_t *out);
int rails_idx_free(rails_idx_t I);
/* dst[i, :] = scale * src[idx[i], :]        (dst has n rows) */
int rails_mv_gather_rows(rails_mv_t dst, int dc0, rails_mv_t src, int sc0, int nc,
                         rails_idx_t I, double scale);
/* dst[idx[i], :] = scale * src[i, :]; the other rows of dst are left alone */
int rails_mv_scatter_rows(rails_mv_t dst, int dc0, rails_mv_t src, int sc0, int nc,
                          rails_idx_t I, double scale);

#ifdef __cplusplus
}
#endif

#endif
