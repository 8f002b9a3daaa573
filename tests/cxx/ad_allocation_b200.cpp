// tests/cxx/ad_allocation_b200.cpp -- the C++ side of the drop-in boundary, compiled.
//
// allocate_ads() below is the body INTEGRATION.md section 2 tells a maintainer to put into the reference's unsolved
// stubs (week-2/ad_allocation/ad_allocation.cpp:9-19; solve_diet_problem, week-2/diet/diet.cc:10-15, is the same
// call).  main() speaks the text protocol of the reference's own main() (ad_allocation.cpp:21-56): "n m", n rows of
// A, b, c on stdin; "No solution" / "Bounded solution" + x with 18 decimals / "Infinity" on stdout.  Outcomes the
// reference's switch has no case for (include/lp_b200.h status 2..5) print nothing and become the exit code 10 + status.
//   g++ -pipe -O2 -std=c++11 ad_allocation_b200.cpp -Iinclude -Llinear_programming_b200 -llp_b200
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <utility>
#include <vector>

#include "lp_b200.h"

typedef std::vector<std::vector<double> > matrix;

std::pair<int, std::vector<double> > allocate_ads(int n, int m, matrix A, std::vector<double> b, std::vector<double> c)
{
    std::vector<double> flat;
    flat.reserve((size_t)n * m);
    for (size_t i = 0; i < A.size(); ++i) flat.insert(flat.end(), A[i].begin(), A[i].end());
    int32_t status = 0, npivots = 0;
    double objective = 0.0;
    std::vector<double> x(m, 0.0);
    const int rc = lpb200_simplex_batch(1, n, m, flat.data(), b.data(), c.data(), 1e-9, 1000000, &status, x.data(),
                                        &objective, &npivots, NULL, 0);
    if (rc != 0) {
        std::fprintf(stderr, "lp_b200: %s\n", lpb200_last_error());
        std::exit(2);
    }
    return std::make_pair((int)status, x);
}

int main()
{
    int n = 0, m = 0;
    if (!(std::cin >> n >> m)) return 1;
    matrix A(n, std::vector<double>(m));
    std::vector<double> b(n), c(m);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < m; ++j) std::cin >> A[i][j];
    for (int i = 0; i < n; ++i) std::cin >> b[i];
    for (int j = 0; j < m; ++j) std::cin >> c[j];
    const std::pair<int, std::vector<double> > ans = allocate_ads(n, m, A, b, c);
    if (ans.first == LPB200_STATUS_NO_SOLUTION) {
        std::printf("No solution\n");
    } else if (ans.first == LPB200_STATUS_BOUNDED) {
        std::printf("Bounded solution\n");
        for (int j = 0; j < m; ++j) std::printf("%.18f%c", ans.second[j], j + 1 == m ? '\n' : ' ');
    } else if (ans.first == LPB200_STATUS_INFINITY) {
        std::printf("Infinity\n");
    } else {
        return 10 + ans.first;
    }
    return 0;
}
