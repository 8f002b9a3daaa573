/*
 * include/lp_b200.h -- C ABI of liblp_b200.so, the B200 (sm_100a) dense LP engine.
 *
 * This is the drop-in boundary for the ONE hot path of aakarsh/linear_programming:
 *   - the two-phase tableau simplex   week-2/diet/simplex.py:163-660
 *     (the solver the course stubs week-2/ad_allocation/ad_allocation.cpp:9-19 and
 *      week-2/diet/diet.cc:10-15 are meant to contain), and
 *   - the Gauss-Jordan elimination    week-2/energy_values/energy_values.cc:85-197.
 *
 * Conventions (all entry points)
 *   - extern "C", plain pointers and sizes, no C++/torch types; never throws.
 *   - Return value: 0 = ok, negative = error (LPB200_E_*); lpb200_last_error() returns a
 *     thread-local message for the last failing call on this thread.
 *   - Naming follows the reference: n = number of constraints (rows of A), m = number of
 *     variables (columns of A); A row-major [n][m], b[n], c[m]; maximise c.x s.t. Ax <= b, x >= 0.
 *   - "_device" entry points take DEVICE pointers on the current CUDA device and enqueue
 *     on `stream` (a cudaStream_t passed as void*, NULL = default stream) without
 *     synchronising; the caller owns every buffer.  Entry points without that suffix take
 *     HOST pointers, copy in, run, copy out and synchronise (the end-to-end path).
 *   - Domain of the bit-exact claim (status, pivot sequence, x, objective identical to the reference): integer
 *     inputs (what the reference's parser produces, simplex.py:712-717), and floating-point inputs in which no
 *     non-zero entry of A has magnitude <= tol.  The engine does not distinguish "coefficient is None" from
 *     "coefficient is 0.0"; the reference's substitute() (simplex.py:203-205) keeps a sub-tolerance coefficient in
 *     place where the engine zeroes it.  On such inputs status and pivot sequence still match and x / objective agree
 *     within 1e-9 relative (tests/test_subtolerance_inputs.py, vectors produced by the reference).
 *   - max_pivots is the one deliberate deviation from the drop-in behaviour: the reference has no cap (and can
 *     cycle); the Python mirror defaults to max(10000, 50(n+m)), the CLI takes --max-pivots=N.
 *   - Per-problem outcome codes (status[]): the reference's anst values plus the classes in
 *     which the reference itself does not return:
 */
#ifndef LP_B200_H
#define LP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LPB200_STATUS_BOUNDED 0              /* simplex.py:608,626  "Bounded solution" */
#define LPB200_STATUS_NO_SOLUTION (-1)       /* simplex.py:614,630  "No solution"      */
#define LPB200_STATUS_INFINITY 1             /* simplex.py:611,628  "Infinity"         */
#define LPB200_STATUS_ITERATION_LIMIT 2      /* max_pivots reached (the reference has no cap) */
#define LPB200_STATUS_REF_ATTRIBUTE_ERROR 3  /* reference raises AttributeError, simplex.py:626 */
#define LPB200_STATUS_REF_INDEX_ERROR 4      /* reference raises IndexError, simplex.py:541->183 */
#define LPB200_STATUS_REF_TYPE_ERROR 5       /* reference raises TypeError, simplex.py:409 */

#define LPB200_E_INVALID (-1)     /* bad argument */
#define LPB200_E_CUDA (-2)        /* CUDA runtime error (message has the cudaError string) */
#define LPB200_E_NO_DEVICE (-3)   /* no CUDA device / driver: there is NO CPU fallback */
#define LPB200_E_TOO_LARGE (-4)   /* problem does not fit the requested mode */

/* Library / device probes (never touch the GPU unless stated). */
const char *lpb200_version(void);
const char *lpb200_last_error(void);
int lpb200_device_count(void); /* cudaGetDeviceCount; <= 0 means the engine cannot run */

/* Largest problem the one-CTA-per-LP kernel can hold in shared memory: returns the dynamic
 * shared memory (bytes) the kernel needs for an n x m LP; > 232448 (227 KB) does not fit. */
int64_t lpb200_simplex_batch_smem_bytes(int n, int m);

/*
 * Batched simplex: B independent LPs of one shape, one CTA per LP, tableau resident in
 * shared memory, persistent CTAs pulling problems off a device work queue.
 * Replaces, per problem, Simplex(A,b,c,n,m).solve()            simplex.py:496-505, 586-630
 * i.e. the body of allocate_ads(n,m,A,b,c)                     ad_allocation.cpp:9-19
 *      and solve_diet_problem(n,m,A,b,c)                       diet.cc:10-15.
 *   A[B][n][m], b[B][n], c[B][m]    inputs
 *   tol                              global_tolerance, simplex.py:16 (default 1e-9)
 *   max_pivots                       cap on pivots per LP (incl. the forced phase-1 pivot)
 *   status[B]                        outcome code
 *   x[B][m]                          assignment (zeros unless status == 0), simplex.py:603,622
 *   objective[B]                     optimum (simplex.py:600,620), 0 unless status == 0
 *   npivots[B]                       pivots performed
 *   pivot_log[B][log_cap][2]         nullable; (entering, leaving) variable ids per pivot
 */
int lpb200_simplex_batch_device(int B, int n, int m, const double *A, const double *b, const double *c,
                                double tol, int max_pivots, int32_t *status, double *x,
                                double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                                void *stream);
int lpb200_simplex_batch(int B, int n, int m, const double *A, const double *b, const double *c,
                         double tol, int max_pivots, int32_t *status, double *x, double *objective,
                         int32_t *npivots, int32_t *pivot_log, int log_cap);

/*
 * The same two calls, extended: a typed constraint matrix and a choice of pivoting rule.
 *
 * a_dtype -- the reference reads its input with int() (simplex.py:712-717) and the course bounds
 * the coefficients to |A[i][j]| <= 100, so a caller that parsed a problem holds small integers:
 * handing them over as int8/int16/int32 (or float32) cuts the PCIe/HBM bytes of A, which is 98 %
 * of a problem, by 8x/4x/2x.  Elements are widened to fp64 exactly when the tableau is built
 * (SlackForm.__init__, simplex.py:230-302); from there on everything is bit-identical to the
 * fp64 call on the same values.  b and c stay fp64.
 *
 * rule -- LPB200_RULE_REFERENCE is the reference's behaviour, bugs included (this is what every
 * other entry point does).  LPB200_RULE_BLAND is the corrected mode (NOT
 * reference behaviour -- the bugs it removes are catalogued in SURVEY Appendix A.8): entering =
 * lowest variable id with a positive coefficient (Bland, terminates), ratio test that keeps
 * zero ratios (ties -> lowest id), phase dispatch on the true minimum of b (no AttributeError
 * class, simplex.py:626), x_aux pivoted out of the basis after phase 1 instead of IndexError
 * (simplex.py:541), no None-tracking (TypeError, simplex.py:409).  Same tableau arithmetic
 * (two roundings, snap to zero at tol).  status is then one of 0, -1, 1, 2.
 * LPB200_RULE_DANTZIG is the same corrected mode with largest-coefficient pricing (the textbook
 * rule the reference's pick_entering_idx, simplex.py:390-392, does not implement): the column with
 * the largest objective coefficient enters, ties -> lowest id; after more than 8 degenerate pivots
 * in a row (leaving ratio 0) Bland's choice is used until a pivot makes progress, so it cannot
 * cycle.  Typically 5-10x fewer pivots than either other rule.
 * LPB200_RULE_DEVEX prices by Devex reference-framework weights (Forrest & Goldfarb's approximation of
 * steepest edge): the column with the largest c*c/w among c > tol enters; a pivot with entering position p,
 * new pivot row nr[] and wq = w[p] sets w[k] = max(w[k], (nr[k]*nr[k])*wq), w[p] = max((nr[p]*nr[p])*wq, 1);
 * weights restart at 1 wherever a pivot loop is entered; same Bland fallback.  Element-wise updates, so the
 * batch kernel, the grid-wide kernel, any number of GPUs and the CPU restatement agree bit for bit.  About a
 * third fewer pivots than Dantzig on the mixed-sign 100x100 batch.
 */
#define LPB200_DTYPE_F64 0
#define LPB200_DTYPE_F32 1
#define LPB200_DTYPE_I32 2
#define LPB200_DTYPE_I16 3
#define LPB200_DTYPE_I8 4
#define LPB200_RULE_REFERENCE 0
#define LPB200_RULE_BLAND 1
#define LPB200_RULE_DANTZIG 2
#define LPB200_RULE_DEVEX 3
int lpb200_simplex_batch_ex_device(int B, int n, int m, const void *A, int a_dtype, const double *b,
                                   const double *c, double tol, int max_pivots, int rule, int32_t *status,
                                   double *x, double *objective, int32_t *npivots, int32_t *pivot_log,
                                   int log_cap, void *stream);
int lpb200_simplex_batch_ex(int B, int n, int m, const void *A, int a_dtype, const double *b, const double *c,
                            double tol, int max_pivots, int rule, int32_t *status, double *x,
                            double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap);

/*
 * Single large LP, grid-wide: the condensed tableau lives in HBM and every pivot is one
 * streaming sweep over it by all SMs (cooperative persistent kernel).  Same semantics and
 * outputs as one problem of the batch call.  `workspace` is a device buffer of at least
 * lpb200_simplex_single_workspace_bytes(n, m) bytes: TWO copies of the condensed tableau (2 x 8(n+1)(m+1) bytes,
 * 8.6 GB at 16384 x 32768 -- the pivot loop sweeps from one into the other so that the next pivot can be selected
 * while the current one is applied) plus small vectors.
 * stats[0] = kernel milliseconds (CUDA events), stats[1] = pivots (nullable, host memory).
 */
int64_t lpb200_simplex_single_workspace_bytes(int n, int m);
int lpb200_simplex_single_device(int n, int m, const double *A, const double *b, const double *c,
                                 double tol, int max_pivots, int32_t *status, double *x,
                                 double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                                 void *workspace, int64_t workspace_bytes, void *stream);
int lpb200_simplex_single(int n, int m, const double *A, const double *b, const double *c, double tol,
                          int max_pivots, int32_t *status, double *x, double *objective,
                          int32_t *npivots, int32_t *pivot_log, int log_cap, double *stats);

/*
 * Export of the tableau a grid-wide (or column-sharded) solve stopped on, indexed by VARIABLE id (slack i = id i,
 * x_j = id n + j, x_aux = id n + m), read from the `workspace` of that solve on the current device:
 *   obj_by_var[n+m+2]    objective-row coefficient (the row pick_entering_idx scans, simplex.py:390-392) of every
 *                        non-basic variable, 0 for basic ones; the last element is the row's constant (the value
 *                        `v` of simplex.py:480).  After a sharded solve a rank reports the coefficients of the
 *                        columns IT owns (0 elsewhere): the sum over ranks of elements [0, n+m+1) is the full row.
 *   const_by_var[n+m+1]  constant-column entry (the b_i of the ratio test, simplex.py:414) of every basic
 *                        variable's row, 0 for non-basic ones (replicated on every rank).
 * Either pointer may be NULL.  With max_pivots = K this is the state after the first K pivots: how an LP of
 * config-5 size, which no solver reaches the end of within a test, is compared with the CPU oracle.
 */
int lpb200_simplex_single_state_device(int n, int m, const void *workspace, double *obj_by_var,
                                       double *const_by_var, void *stream);

/*
 * Batched Simplex.verify_bounds (simplex.py:663-671; stress.py:21-87 runs it after every "Bounded solution"):
 * for every LP with status[p] == LPB200_STATUS_BOUNDED (status == NULL: every LP) and every constraint r, in
 * order: sum = 0; sum += A[r][i] * x[i] for i = 0..m-1 (product and sum rounded separately, left to right);
 * the reference asserts sum <= b[r] + 1e-3 (it ignores its `tolerance` argument; pass 1e-3 to reproduce it).
 *   first_bad_row[B]   the first r whose assertion would fire, -1 when the answer passes (or was skipped)
 * One CTA per LP, A read from HBM once (8*n*m bytes per LP: HBM-bound).
 */
int lpb200_verify_bounds_batch_device(int B, int n, int m, const double *A, const double *b, const double *x,
                                      const int32_t *status, double tolerance, int32_t *first_bad_row,
                                      void *stream);
int lpb200_verify_bounds_batch(int B, int n, int m, const double *A, const double *b, const double *x,
                               const int32_t *status, double tolerance, int32_t *first_bad_row);

/*
 * Host-core helpers of the end-to-end path (no GPU involved).  lpb200_simplex_batch[_ex] called with HOST
 * pointers and an fp64 A runs two lanes over the chunks of a large batch: the copy engine ships chunks as
 * they are while the host cores narrow other chunks to int8 -- only chunks in which every element is an
 * integer in [-127, 127] (and no -0.0), so that the kernel's exact widening reproduces the caller's bits --
 * and ship 1/8 of the bytes; answers are bit-identical either way.  LPB_HOST_THREADS sets the pool size
 * (default = the cores the process may run on, at most 32); the second lane runs when the pool has at least
 * 10 threads (with fewer -- many GPU ranks sharing one host -- the cores' reads only take host memory
 * bandwidth from the copy engines), LPB_HOST_NARROW=0 / 1 forces it off / on.
 *   lpb200_narrow_f64_to_i8     the same exact narrowing for callers that want to hold A as int8 themselves
 *                               (then pass LPB200_DTYPE_I8): 1 = dst is valid, 0 = some element is not
 *                               representable (dst unspecified), < 0 = bad argument
 *   lpb200_host_threads         size of the host pool
 *   lpb200_simplex_batch_last_lanes  out4[0..3] = chunks served by the copy-engine lane / by the narrowing
 *                               lane / narrowed to int8, and the bytes of A that crossed PCIe, in the calling
 *                               thread's last chunked host-pointer batch call
 */
int lpb200_narrow_f64_to_i8(const double *src, int8_t *dst, int64_t count);
int lpb200_host_threads(void);
int lpb200_simplex_batch_last_lanes(int64_t *out4);

/*
 * The same two calls with a choice of pivoting rule (LPB200_RULE_* above; LPB200_RULE_REFERENCE is what
 * the calls without _ex do).  The corrected rules make a large LP solvable to optimality: the
 * reference's set-order rule needs thousands of pivots where Dantzig pricing needs tens.  Decisions
 * and arithmetic are those of the batch kernel's corrected modes (same CPU restatement, bit-exact).
 */
int lpb200_simplex_single_ex_device(int n, int m, const double *A, const double *b, const double *c,
                                    double tol, int max_pivots, int rule, int32_t *status, double *x,
                                    double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                                    void *workspace, int64_t workspace_bytes, void *stream);
int lpb200_simplex_single_ex(int n, int m, const double *A, const double *b, const double *c, double tol,
                             int max_pivots, int rule, int32_t *status, double *x, double *objective,
                             int32_t *npivots, int32_t *pivot_log, int log_cap, double *stats);

/*
 * Single large LP, COLUMN-SHARDED over the GPUs of one NVLink/NVSwitch box (SURVEY.md 8(e),
 * BASELINE.json configs[4]).  One process per GPU; rank g owns column positions
 * [g*W, (g+1)*W) of every tableau row (W from lpb200_simplex_sharded_columns); per pivot the
 * ranks exchange an 8-byte entering key (all-to-all, min) and the owner of the entering
 * column broadcasts it (n+1 doubles) -- both from inside the persistent kernel, as flagged
 * packets written straight into peer-mapped exchange buffers.  Same semantics and outputs as
 * lpb200_simplex_single (every rank receives the full status / x / objective / pivot log).
 *
 * Communicator: lpb200_comm_create allocates this rank's exchange buffer on the current
 * device and returns an opaque lpb200_comm_handle_bytes()-byte handle; the caller gathers
 * the handles of all ranks (any transport, e.g. torch.distributed.all_gather) and passes
 * the concatenation, in rank order, to lpb200_comm_connect (CUDA IPC + peer access).  All
 * ranks must call lpb200_simplex_sharded_device the same number of times with the same
 * (n, m, tol, max_pivots).  A rank that receives no packet for timeout_s seconds (<= 0: 20 s)
 * aborts its kernel (the CUDA context reports a launch failure) instead of hanging the GPU.
 *   A_cols   this rank's columns of A: A_cols[i*lda + j] = A[i][c_lo + j], j < c_count
 *   b, c     full vectors (replicated on every rank)
 */
typedef struct lpb200_comm lpb200_comm_t;
int64_t lpb200_comm_handle_bytes(void);
int lpb200_comm_create(int rank, int nranks, int max_n, lpb200_comm_t **comm, void *handle_out);
int lpb200_comm_connect(lpb200_comm_t *comm, const void *all_handles);
int lpb200_comm_destroy(lpb200_comm_t *comm);
int lpb200_simplex_sharded_columns(int m, int nranks, int rank, int *c_lo, int *c_count);
int64_t lpb200_simplex_sharded_workspace_bytes(int n, int m, int nranks);
int lpb200_simplex_sharded_device(lpb200_comm_t *comm, int n, int m, const double *A_cols, int64_t lda,
                                  const double *b, const double *c, double tol, int max_pivots,
                                  int32_t *status, double *x, double *objective, int32_t *npivots,
                                  int32_t *pivot_log, int log_cap, void *workspace, int64_t workspace_bytes,
                                  double timeout_s, void *stream);
/* The same nranks-way protocol with every rank run as a CTA group of ONE cooperative launch
 * on the current device (HOST pointers; outputs are per rank: status[nranks], x[nranks][m],
 * objective[nranks], npivots[nranks], pivot_log[nranks][log_cap][2]).  For parity tests and
 * for boxes with fewer GPUs than ranks. */
int lpb200_simplex_sharded_emulated(int nranks, int n, int m, const double *A, const double *b,
                                    const double *c, double tol, int max_pivots, int32_t *status, double *x,
                                    double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap);

/* The sharded calls with a pivoting rule.  Under LPB200_RULE_DANTZIG / LPB200_RULE_DEVEX the entering
 * exchange carries (~bits(value), variable id) as two packets per rank and the minimum is lexicographic;
 * Devex weights stay with the rank that owns the column, the weight of the entering column rides along
 * with the pivot-column broadcast; the x_aux drive-out of the corrected modes costs one more key
 * exchange after phase 1. */
int lpb200_simplex_sharded_ex_device(lpb200_comm_t *comm, int n, int m, const double *A_cols, int64_t lda,
                                     const double *b, const double *c, double tol, int max_pivots, int rule,
                                     int32_t *status, double *x, double *objective, int32_t *npivots,
                                     int32_t *pivot_log, int log_cap, void *workspace, int64_t workspace_bytes,
                                     double timeout_s, void *stream);
int lpb200_simplex_sharded_emulated_ex(int nranks, int n, int m, const double *A, const double *b,
                                       const double *c, double tol, int max_pivots, int rule, int32_t *status,
                                       double *x, double *objective, int32_t *npivots, int32_t *pivot_log,
                                       int log_cap);
/* The emulated launch plus, per rank, the tableau export of lpb200_simplex_single_state_device (host memory):
 * obj_by_var[nranks][n+m+2] (a rank reports the columns it owns; the sum over ranks of the first n+m+1 elements
 * is the full objective row), const_by_var[nranks][n+m+1] (replicated).  Either may be NULL. */
int lpb200_simplex_sharded_emulated_state(int nranks, int n, int m, const double *A, const double *b,
                                          const double *c, double tol, int max_pivots, int rule, int32_t *status,
                                          double *x, double *objective, int32_t *npivots, int32_t *pivot_log,
                                          int log_cap, double *obj_by_var, double *const_by_var);

/*
 * Batched Gauss-Jordan: B independent n x n systems a.x = b.
 * Replaces, per system, SolveEquation(Equation{a,b})           energy_values.cc:166-197.
 *   a[B][n][n], b[B][n]  inputs (not modified);  x[B][n]  solution (b after the last step).
 * `workspace`: device buffer of lpb200_gauss_jordan_workspace_bytes(B, n) bytes (may be 0).
 */
int64_t lpb200_gauss_jordan_workspace_bytes(int B, int n);
int lpb200_gauss_jordan_batch_device(int B, int n, const double *a, const double *b, double *x,
                                     void *workspace, int64_t workspace_bytes, void *stream);
int lpb200_gauss_jordan_batch(int B, int n, const double *a, const double *b, double *x);

/*
 * The same two calls with a choice of pivot selection.  LPB200_PIVOT_REFERENCE is the reference's
 * SelectPivotElement (energy_values.cc:85-114: the diagonal element, a row swap only on an exact
 * zero).  LPB200_PIVOT_PARTIAL (NOT reference behaviour) pivots step k on the
 * row i >= k with the largest |a[i][k]| (first maximum); the elimination arithmetic
 * (ProcessPivotElement, :132-159) is unchanged.
 */
#define LPB200_PIVOT_REFERENCE 0
#define LPB200_PIVOT_PARTIAL 1
int lpb200_gauss_jordan_batch_ex_device(int B, int n, const double *a, const double *b, double *x, int pivoting,
                                        void *workspace, int64_t workspace_bytes, void *stream);
int lpb200_gauss_jordan_batch_ex(int B, int n, const double *a, const double *b, double *x, int pivoting);

/*
 * The same two calls with a per-system outcome.  The reference has none: when SelectPivotElement
 * (energy_values.cc:100-110) finds no row to swap a zero pivot with, it divides by zero and prints inf/nan; x[]
 * reproduces that, and a batch caller would have to scan it.  status[B] (device / host memory like x):
 *   -1   every step found a non-zero pivot
 *   k    the first step k whose column holds no non-zero at or below the diagonal (the matrix is singular to the
 *        arithmetic performed); x of that system is the reference's inf/nan output
 */
int lpb200_gauss_jordan_batch_status_device(int B, int n, const double *a, const double *b, double *x, int pivoting,
                                            int32_t *status, void *workspace, int64_t workspace_bytes, void *stream);
int lpb200_gauss_jordan_batch_status(int B, int n, const double *a, const double *b, double *x, int pivoting,
                                     int32_t *status);

/*
 * Batched Gauss-Jordan that ALLOCATES NOTHING and needs no caller workspace.  This entry point has no counterpart in
 * the reference: it was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels, same arithmetic and the
 * same outputs as lpb200_gauss_jordan_batch_status_device (x bit-identical, status as described there; `status` may
 * be NULL); device pointers on the current device, enqueued on `stream`, no synchronisation, no cudaMalloc / cudaFree
 * / pool allocation on any path.
 *   The working copy of `a` (which the elimination destroys) lives in a file-scope __device__ array of
 *   LPB200_GJ_STATIC_SCRATCH_BYTES inside the library -- every GPU that loads liblp_b200.so holds these bytes until
 *   the process exits.  Limit: B * n * n * 8 <= LPB200_GJ_STATIC_SCRATCH_BYTES (16 systems of 512 x 512, 4096 of
 *   32 x 32, 10485 of the course's 20 x 20); a larger batch is rejected with LPB200_E_TOO_LARGE and a message that
 *   states both sizes.
 *   All calls share that one array per GPU.  It keeps the last call's contents (nothing needs clearing: a call
 *   overwrites what it uses before reading it).  Calls that OVERLAP on the GPU -- different streams, or several host
 *   threads on one device -- race on it and corrupt each other's answers; calls on one stream are ordered and safe.
 */
#define LPB200_GJ_STATIC_SCRATCH_BYTES (32 * 1024 * 1024)
int lpb200_gauss_jordan_batch_static_device(int B, int n, const double *a, const double *b, double *x, int pivoting,
                                            int32_t *status, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* LP_B200_H */
