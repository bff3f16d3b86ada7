// Host build of wxfactory_b200/csrc/fastmath.cuh against libm in long double (tests/test_fastmath.py compiles and runs it).
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <random>
#include "fastmath.cuh"
int main() {
    double tab[WXB_FASTTAB_DOUBLES];
    wxbfm::fill_tables(tab);
    std::mt19937_64 g(1);
    double worst_log = 0, worst_exp = 0, worst_log_abs = 0;
    for (int t = 0; t < 4000000; ++t) {
        std::uniform_real_distribution<double> u(-12.0, 12.0);
        double lx = u(g);
        double x = std::exp(lx);
        if (t % 4 == 0) x = 0.5 + 1.0 * (double)(g() % 1000000) / 1e6;  // near 1
        long double want = logl((long double)x);
        double got = wxbfm::fast_log(x, tab);
        double ulp = std::fabs(std::nextafter((double)want, INFINITY) - (double)want); if (ulp < 0x1p-54) ulp = 0x1p-54;
        double e = std::fabs((double)((long double)got - want)) / ulp;
        if (e > worst_log) worst_log = e;
        double ea = std::fabs((double)((long double)got - want));
        if (std::fabs((double)want) < 1 && ea > worst_log_abs) worst_log_abs = ea;
        double y = u(g) * 2;
        long double wante = expl((long double)y);
        double gote = wxbfm::fast_exp(y, tab + WXB_EXPTAB_OFF);
        double ulpe = std::fabs(std::nextafter((double)wante, INFINITY) - (double)wante);
        double ee = std::fabs((double)((long double)gote - wante)) / ulpe;
        if (ee > worst_exp) worst_exp = ee;
    }
    printf("log_ulp %.4f log_abs %.4e exp_ulp %.4f\n", worst_log, worst_log_abs, worst_exp);
    const double* et = tab + WXB_EXPTAB_OFF;
    const bool special = std::isnan(wxbfm::fast_log(0.0, tab)) && std::isnan(wxbfm::fast_log(-1.0, tab)) && std::isnan(wxbfm::fast_log(NAN, tab)) &&
                         std::isnan(wxbfm::fast_log(INFINITY, tab)) && std::isnan(wxbfm::fast_exp(NAN, et)) && wxbfm::fast_exp(0.0, et) == 1.0 &&
                         std::fabs(wxbfm::fast_log(1.0, tab)) < 1e-17;
    printf("special %d\n", (int)special);
    return 0;
}
