// oracle/ref_shim_gauss_jordan.cc -- TEST INFRASTRUCTURE ONLY.
//
// Exposes the UNMODIFIED reference Gauss-Jordan (energy_values.cc:166 SolveEquation)
// through a C entry point so tests can call it in-process and bench.py can time it as
// the CPU baseline (cpu_baseline.kind == "reference").  The reference translation unit
// is #included from where it lies under /root/reference (path passed as -DREF_SOURCE by
// oracle/Makefile); no reference source is copied into this repository.  Its main() is
// renamed so the object can live in a shared library.
#define main lp_reference_energy_values_main
#include REF_SOURCE
#undef main

extern "C" void lp_ref_gauss_jordan(int n, const double *a, const double *b, double *x)
{
    Matrix A(n, Row(n, 0.0));
    Column B(n, 0.0);
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) A[i][j] = a[(size_t)i * n + j];
        B[i] = b[i];
    }
    Column sol = SolveEquation(Equation(A, B));
    for (int i = 0; i < n; ++i) x[i] = sol[i];
}
