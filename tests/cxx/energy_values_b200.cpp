// tests/cxx/energy_values_b200.cpp -- SolveEquation (week-2/energy_values/energy_values.cc:166-197) replaced by the
// engine, as INTEGRATION.md section 3 shows, behind the reference's text protocol: "size" then size rows of size + 1
// numbers (a | b) on stdin (ReadEquation, :66-80), the solution one value per line with precision 20 on stdout
// (PrintColumn, :199-205).
//   g++ -pipe -O2 -std=c++0x energy_values_b200.cpp -Iinclude -Llinear_programming_b200 -llp_b200
#include <cstdlib>
#include <iostream>
#include <vector>

#include "lp_b200.h"

typedef std::vector<double> Column;
typedef std::vector<Column> Matrix;

struct Equation {
    Matrix a;
    Column b;
};

Column SolveEquation(Equation equation)
{
    const int n = (int)equation.a.size();
    std::vector<double> flat;
    for (int i = 0; i < n; ++i) flat.insert(flat.end(), equation.a[i].begin(), equation.a[i].end());
    Column x(n);
    if (lpb200_gauss_jordan_batch(1, n, flat.data(), equation.b.data(), x.data()) != 0) {
        std::cerr << "lp_b200: " << lpb200_last_error() << std::endl;
        std::exit(2);
    }
    return x;
}

int main()
{
    int size = 0;
    std::cin >> size;
    Equation eq;
    eq.a.assign(size, Column(size, 0.0));
    eq.b.assign(size, 0.0);
    for (int i = 0; i < size; ++i) {
        for (int j = 0; j < size; ++j) std::cin >> eq.a[i][j];
        std::cin >> eq.b[i];
    }
    const Column x = SolveEquation(eq);
    std::cout.precision(20);
    for (int i = 0; i < size; ++i) std::cout << x[i] << std::endl;
    return 0;
}
