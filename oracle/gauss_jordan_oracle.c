/*
 * oracle/gauss_jordan_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU oracle, never shipped).
 *
 * Scalar C restatement of /root/reference/week-2/energy_values/energy_values.cc:
 *   select_pivot_row()  SelectPivotElement   energy_values.cc:85-114
 *                       (SwapLines :116-130 is a no-op: row == column == step)
 *   eliminate()         ProcessPivotElement  energy_values.cc:132-159
 *   lp_oracle_gauss_jordan()  SolveEquation  energy_values.cc:166-197
 *
 * Build with -ffp-contract=off: the reference is compiled for baseline x86-64
 * (no FMA), so `a[i][j] += scale * a[r][j]` rounds the product and the sum separately.
 * The row scale is a true division per element (:143-145).
 * It is pinned against the compiled reference itself (oracle/_ref/energy_values, built
 * by oracle/Makefile from the sources where they lie) in tests/test_oracle_gauss_jordan.py
 * and against the committed golden vectors in tests/golden/.
 *
 * pivoting = LP_PIVOT_PARTIAL is NOT reference behaviour and has nothing
 * to be pinned to ("parity unpinned" for this mode): step k pivots on the row i >= k with the
 * largest |a[i][k]| (first maximum, compared as the bit pattern of fabs), everything else as
 * above.  tests/test_oracle_gauss_jordan.py checks it against numpy.linalg.solve.
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdint.h>
#include "lp_oracle.h"

void lp_oracle_gauss_jordan(int n, const double *a_in, const double *b_in, double *x, int full_sweep)
{
    lp_oracle_gauss_jordan_pivot(n, a_in, b_in, x, full_sweep, LP_PIVOT_REFERENCE);
}

static void swap_rows(double *a, double *b, int n, int k, int sw)
{
    for (int j = 0; j < n; ++j) {
        const double t = a[(size_t)k * n + j];
        a[(size_t)k * n + j] = a[(size_t)sw * n + j];
        a[(size_t)sw * n + j] = t;
    }
    const double t = b[k];
    b[k] = b[sw];
    b[sw] = t;
}

void lp_oracle_gauss_jordan_pivot(int n, const double *a_in, const double *b_in, double *x, int full_sweep,
                                  int pivoting)
{
    if (n <= 0) return;
    double *a = (double *)malloc(sizeof(double) * (size_t)n * (size_t)n);
    double *b = x;
    memcpy(a, a_in, sizeof(double) * (size_t)n * (size_t)n);
    memmove(b, b_in, sizeof(double) * (size_t)n);
    for (int k = 0; k < n; ++k) {
        /* SelectPivotElement: exact-zero test, first row >= k with a non-zero in column k */
        if (pivoting == LP_PIVOT_PARTIAL) {
            int r = k;
            uint64_t best = 0;
            for (int i = k; i < n; ++i) {
                const double m = fabs(a[(size_t)i * n + k]);
                uint64_t bits;
                memcpy(&bits, &m, sizeof(bits));
                if (bits > best) {
                    best = bits;
                    r = i;
                }
            }
            if (r != k) swap_rows(a, b, n, k, r);
        } else if (a[(size_t)k * n + k] == 0.0) {
            int sw = k;
            while (sw < n && a[(size_t)sw * n + k] == 0.0) ++sw;
            if (sw < n) swap_rows(a, b, n, k, sw);
            /* else: the reference carries on and divides by zero */
        }
        /* ProcessPivotElement */
        const int j0 = full_sweep ? 0 : k;
        double *ak = a + (size_t)k * n;
        const double value = ak[k];
        for (int j = j0; j < n; ++j) ak[j] /= value;
        b[k] /= value;
        for (int i = 0; i < n; ++i) {
            if (i == k) continue;
            double *ai = a + (size_t)i * n;
            const double scale = -1 * ai[k];
            for (int j = j0; j < n; ++j) {
                const double prod = scale * ak[j];
                ai[j] += prod;
            }
            const double pb = scale * b[k];
            b[i] += pb;
        }
    }
    free(a);
}

typedef struct {
    int n;
    const double *a, *b;
    double *x;
    int full_sweep, pivoting;
} gj_job_t;

static void gj_job(void *arg, int s)
{
    gj_job_t *j = (gj_job_t *)arg;
    const int n = j->n;
    lp_oracle_gauss_jordan_pivot(n, j->a + (size_t)s * n * n, j->b + (size_t)s * n, j->x + (size_t)s * n,
                                 j->full_sweep, j->pivoting);
}

int lp_oracle_gauss_jordan_batch(int B, int n, const double *a, const double *b, double *x,
                                 int full_sweep, int threads)
{
    return lp_oracle_gauss_jordan_batch_pivot(B, n, a, b, x, full_sweep, LP_PIVOT_REFERENCE, threads);
}

int lp_oracle_gauss_jordan_batch_pivot(int B, int n, const double *a, const double *b, double *x,
                                       int full_sweep, int pivoting, int threads)
{
    gj_job_t j = {n, a, b, x, full_sweep, pivoting};
    return lp_oracle_parallel_for(B, threads, gj_job, &j);
}
