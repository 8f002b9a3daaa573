/*
 * oracle/lp_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, scalar, no FMA contraction) of the reference's hot path:
 *   - two-phase slack-form simplex      /root/reference/week-2/diet/simplex.py:163-660
 *   - Gauss-Jordan elimination          /root/reference/week-2/energy_values/energy_values.cc:85-197
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker / the timed CPU baseline.
 * The product (linear_programming_b200) never links or imports it.
 *
 * Parity pin: tests/test_oracle_vs_golden.py checks this restatement against golden
 * vectors produced by the unmodified reference (tests/golden/make_golden.py) -- status,
 * full pivot sequence and the bit pattern of every x -- and, when /root/reference is
 * present, against the live reference on fresh random inputs.
 */
#ifndef LP_ORACLE_H
#define LP_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- outcome codes (same values as include/lp_b200.h) ---- */
#define LP_STATUS_BOUNDED 0
#define LP_STATUS_NO_SOLUTION (-1)
#define LP_STATUS_INFINITY 1
#define LP_STATUS_ITERATION_LIMIT 2
#define LP_STATUS_REF_ATTRIBUTE_ERROR 3 /* simplex.py:626 */
#define LP_STATUS_REF_INDEX_ERROR 4     /* simplex.py:541 -> :183 */
#define LP_STATUS_REF_TYPE_ERROR 5      /* simplex.py:409 */

/* ---- CPython set model (pyset_model.c) ---- */
#define PYSET_UNUSED (-1)
#define PYSET_DUMMY (-2)
typedef struct {
    int32_t *table; /* key >= 0, PYSET_UNUSED or PYSET_DUMMY */
    int size;       /* power of two */
    int fill;       /* active + dummy */
    int used;       /* active */
} pyset_t;

void pyset_init(pyset_t *s);
void pyset_free(pyset_t *s);
void pyset_init_range(pyset_t *s, int32_t lo, int32_t hi);
int pyset_add(pyset_t *s, int32_t key);
int pyset_remove(pyset_t *s, int32_t key);
int pyset_order(const pyset_t *s, int32_t *out);
int lp_oracle_pyset_replay(int32_t lo, int32_t hi, int nops, const int32_t *ops, int32_t *order_out,
                           int32_t *table_size_out);

/* ---- host-thread pool for the timed CPU baseline (parallel_for.c) ----
 * Runs fn(arg, i) for i in [0, count) on `threads` pthreads pulling indices off a shared
 * counter (threads <= 0: one per online core).  Returns the number of threads used. */
int lp_oracle_parallel_for(int count, int threads, void (*fn)(void *, int), void *arg);

/* ---- simplex (simplex_oracle.c) ----
 * A is row-major [n][m], b[n], c[m] (the reference's naming: n constraints, m variables).
 * x[m] receives the assignment when status == 0 (zeros otherwise); *objective the optimum.
 * pivot_log (nullable) receives (entering, leaving) variable ids, up to log_cap pairs;
 * *npivots counts every pivot including the forced phase-1 pivot (simplex.py:524).
 * Returns the status code. */
int lp_oracle_simplex(int n, int m, const double *A, const double *b, const double *c, double tol,
                      int max_pivots, double *x, double *objective, int32_t *npivots,
                      int32_t *pivot_log, int log_cap);

/* The same with a pivoting rule: LP_RULE_REFERENCE is lp_oracle_simplex; LP_RULE_BLAND is the
 * corrected mode described at the top of simplex_oracle.c (not reference behaviour). */
#define LP_RULE_REFERENCE 0
#define LP_RULE_BLAND 1
#define LP_RULE_DANTZIG 2 /* largest coefficient, Bland after LP_DEGENERATE_STREAK degenerate pivots in a row */
#define LP_RULE_DEVEX 3   /* Devex reference-framework weights (Forrest & Goldfarb), same Bland fallback */
#define LP_DEGENERATE_STREAK 8
int lp_oracle_simplex_rule(int n, int m, const double *A, const double *b, const double *c, double tol,
                           int max_pivots, int rule, double *x, double *objective, int32_t *npivots,
                           int32_t *pivot_log, int log_cap);
/* The same with the rows of each pivot swept on `threads` host threads (bit-identical for any count) and an
 * export of the final tableau by variable id: obj_by_var[n+m+2], const_by_var[n+m+1] (see simplex_oracle.c). */
int lp_oracle_simplex_state(int n, int m, const double *A, const double *b, const double *c, double tol,
                            int max_pivots, int rule, int threads, double *x, double *objective,
                            int32_t *npivots, int32_t *pivot_log, int log_cap, double *obj_by_var,
                            double *const_by_var);
int lp_oracle_simplex_batch_rule(int B, int n, int m, const double *A, const double *b, const double *c,
                                 double tol, int max_pivots, int rule, int32_t *status, double *x,
                                 double *objective, int32_t *npivots, int threads);

/* Simplex.verify_bounds (simplex.py:663-671): first row r with sum_i A[r][i]*x[i] > b[r] + tolerance, else -1 */
int lp_oracle_verify_bounds(int n, int m, const double *A, const double *b, const double *x, double tolerance);
int lp_oracle_driveouts(int reset); /* LP_RULE_BLAND: x_aux drive-out pivots so far (test instrumentation) */

/* Batch driver used for the timed CPU baseline: pthreads over problems
 * (threads <= 0 means all cores).  Returns the number of threads used. */
int lp_oracle_simplex_batch(int B, int n, int m, const double *A, const double *b, const double *c,
                            double tol, int max_pivots, int32_t *status, double *x, double *objective,
                            int32_t *npivots, int threads);

/* ---- Gauss-Jordan (gauss_jordan_oracle.c) ----
 * a is row-major [n][n], b[n]; x[n] receives the solution (b after the last step).
 * full_sweep != 0 updates every column like the reference (energy_values.cc:146-158);
 * full_sweep == 0 restricts the column range to [k, n) (bit-identical x). */
void lp_oracle_gauss_jordan(int n, const double *a, const double *b, double *x, int full_sweep);
int lp_oracle_gauss_jordan_batch(int B, int n, const double *a, const double *b, double *x,
                                 int full_sweep, int threads);
/* The same with a pivot rule: LP_PIVOT_REFERENCE as above, LP_PIVOT_PARTIAL = largest |a[i][k]|
 * of the column (not reference behaviour; see gauss_jordan_oracle.c). */
#define LP_PIVOT_REFERENCE 0
#define LP_PIVOT_PARTIAL 1
void lp_oracle_gauss_jordan_pivot(int n, const double *a, const double *b, double *x, int full_sweep,
                                  int pivoting);
int lp_oracle_gauss_jordan_batch_pivot(int B, int n, const double *a, const double *b, double *x,
                                       int full_sweep, int pivoting, int threads);

#ifdef __cplusplus
}
#endif
#endif
