// Prints values of nori_b200/host/nori/hypothesis.h for tests/test_host_tests.py (compared with scipy).
#include <nori/hypothesis.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
int main(int argc, char **argv) {
    if (argc < 2) return 2;
    if (!strcmp(argv[1], "chi2") && argc == 4) printf("%.17g\n", hypothesis::chi2_cdf(atof(argv[2]), atoi(argv[3])));
    else if (!strcmp(argv[1], "t") && argc == 4) printf("%.17g\n", hypothesis::students_t_cdf(atof(argv[2]), atoi(argv[3])));
    else if (!strcmp(argv[1], "simpson")) {
        printf("%.17g\n", hypothesis::adaptiveSimpson([](double x) { return std::exp(-x * x); }, -1.0, 2.0));
        printf("%.17g\n", hypothesis::adaptiveSimpson2D([](double x, double y) { return std::sin(x) * std::cos(y) + x * y; }, 0.0, 0.0, 1.5, 0.75));
        std::vector<double> w = hypothesis::simpsonWeights(8, 0.0, 2.0);
        double s = 0; for (int i = 0; i <= 8; ++i) { const double x = 2.0 * i / 8; s += w[i] * x * x * x; }
        printf("%.17g\n", s);
    } else if (!strcmp(argv[1], "chi2test")) {
        // 6 cells, two of them below the minimum expected frequency -> pooled
        const double obs[6] = {48, 35, 15, 3, 2, 1}, expd[6] = {50, 30, 18, 3.5, 1.5, 1.0};
        auto r = hypothesis::chi2_test(6, obs, expd, 104, 5, 0.01, 3);
        printf("%d\n%s", r.first ? 1 : 0, r.second.c_str());
    } else if (!strcmp(argv[1], "ttest") && argc == 5) {
        auto r = hypothesis::students_t_test(atof(argv[2]), atof(argv[3]), atof(argv[4]), 100000, 0.01, 5);
        printf("%d\n%s", r.first ? 1 : 0, r.second.c_str());
    } else if (!strcmp(argv[1], "chi2file") && argc == 3) {
        // file: nCells sampleCount minExp significance numTests, then nCells obs, then nCells exp
        FILE *f = fopen(argv[2], "r"); if (!f) return 3;
        int n, sc, nt; double me, sig;
        if (fscanf(f, "%d %d %lf %lf %d", &n, &sc, &me, &sig, &nt) != 5) return 3;
        std::vector<double> obs(n), expd(n);
        for (double &v : obs) if (fscanf(f, "%lf", &v) != 1) return 3;
        for (double &v : expd) if (fscanf(f, "%lf", &v) != 1) return 3;
        fclose(f);
        auto r = hypothesis::chi2_test(n, obs.data(), expd.data(), sc, me, sig, nt);
        printf("%d\n%s", r.first ? 1 : 0, r.second.c_str());
    } else if (!strcmp(argv[1], "ttestn") && argc == 8) {
        auto r = hypothesis::students_t_test(atof(argv[2]), atof(argv[3]), atof(argv[4]), atoi(argv[5]), atof(argv[6]), atoi(argv[7]));
        printf("%d\n%s", r.first ? 1 : 0, r.second.c_str());
    } else return 2;
    return 0;
}
