// nori/hypothesis.h — the statistics behind the <test> plugins (chi2test, ttest, warptest).
//
// Stands in for the reference's third-party header ext/hypothesis/hypothesis.h (+ cephes.h, pinned in
// .SUBMODULES.json) with the same decision rules, written from the published definitions:
//   * chi^2 goodness of fit over a contingency table with pooling of low-expectation cells
//     (hypothesis.h:143-243): cells visited in order of increasing expected frequency; a cell joins
//     the pool while its own expectation or the pool's is below the minimum; dof = terms - 1;
//   * two-sided one-sample Student t-test (hypothesis.h:321-353): t = |mean - ref| sqrt(n / max(var, 1e-5));
//   * Sidak-corrected significance level 1 - (1 - alpha)^(1 / numTests) in both;
//   * adaptive Simpson quadrature with the |S2 - S1| < 15 eps acceptance rule and Richardson
//     extrapolation, nested for rectangles (hypothesis.h:65-109).
// The distribution functions are the regularised incomplete gamma and beta functions, evaluated by
// series / continued fractions (modified Lentz). Pinned: verdicts, pooled counts, statistics, p-values and CDF values
// equal the reference header's (compiled into oracle/_ref) on the seeded cases of tests/golden/hypothesis_ref.json,
// and the CDFs agree with scipy.stats (tests/test_host_tests.py).
#pragma once
#include <algorithm>
#include <cmath>
#include <functional>
#include <limits>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

namespace hypothesis {

/// Regularised lower incomplete gamma function P(a, x), a > 0, x >= 0.
inline double gammaP(double a, double x) {
    if (!(x > 0.0) || !(a > 0.0)) return 0.0;
    const double logPrefix = a * std::log(x) - x - std::lgamma(a);
    if (x < a + 1.0) {                                  // power series  x^a e^-x / Gamma(a) * sum x^k / (a)_(k+1)
        double term = 1.0 / a, sum = term, ak = a;
        for (int k = 0; k < 100000; ++k) {
            ak += 1.0; term *= x / ak; sum += term;
            if (std::fabs(term) < std::fabs(sum) * 1e-16) break;
        }
        return std::min(1.0, sum * std::exp(logPrefix));
    }
    // continued fraction for Q(a, x) = 1 - P (modified Lentz)
    const double tiny = 1e-300;
    double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, h = d;
    for (int k = 1; k < 100000; ++k) {
        const double an = -k * (k - a);
        b += 2.0;
        d = an * d + b; if (std::fabs(d) < tiny) d = tiny;
        c = b + an / c; if (std::fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        const double delta = d * c;
        h *= delta;
        if (std::fabs(delta - 1.0) < 1e-16) break;
    }
    return std::max(0.0, 1.0 - std::exp(logPrefix) * h);
}

/// Regularised incomplete beta function I_x(a, b), a, b > 0, 0 <= x <= 1.
inline double betaI(double a, double b, double x) {
    if (!(x > 0.0)) return 0.0;
    if (!(x < 1.0)) return 1.0;
    const double logFront = std::lgamma(a + b) - std::lgamma(a) - std::lgamma(b) + a * std::log(x) + b * std::log1p(-x);
    auto cf = [](double a, double b, double x) {        // continued fraction of I_x(a, b) * a / front (Lentz)
        const double tiny = 1e-300;
        double c = 1.0, d = 1.0 - (a + b) * x / (a + 1.0);
        if (std::fabs(d) < tiny) d = tiny;
        d = 1.0 / d;
        double h = d;
        for (int m = 1; m < 200000; ++m) {
            const double m2 = 2.0 * m;
            double num = m * (b - m) * x / ((a + m2 - 1.0) * (a + m2));
            d = 1.0 + num * d; if (std::fabs(d) < tiny) d = tiny;
            c = 1.0 + num / c; if (std::fabs(c) < tiny) c = tiny;
            d = 1.0 / d; h *= d * c;
            num = -(a + m) * (a + b + m) * x / ((a + m2) * (a + m2 + 1.0));
            d = 1.0 + num * d; if (std::fabs(d) < tiny) d = tiny;
            c = 1.0 + num / c; if (std::fabs(c) < tiny) c = tiny;
            d = 1.0 / d;
            const double delta = d * c;
            h *= delta;
            if (std::fabs(delta - 1.0) < 1e-16) break;
        }
        return h;
    };
    if (x < (a + 1.0) / (a + b + 2.0)) return std::exp(logFront) * cf(a, b, x) / a;
    return 1.0 - std::exp(logFront) * cf(b, a, 1.0 - x) / b;
}

/// CDF of the chi^2 distribution with `dof` degrees of freedom.
inline double chi2_cdf(double x, int dof) {
    if (dof < 1 || x < 0) return 0.0;
    return gammaP(0.5 * dof, 0.5 * x);
}

/// CDF of Student's t distribution with `dof` degrees of freedom.
inline double students_t_cdf(double x, int dof) {
    const double tail = 0.5 * betaI(0.5 * dof, 0.5, dof / (x * x + dof));
    return x > 0 ? 1.0 - tail : tail;
}

/// Adaptive Simpson quadrature of f over [x0, x1].
inline double adaptiveSimpson(const std::function<double(double)> &f, double x0, double x1, double eps = 1e-6, int depth = 6) {
    struct Rec {
        const std::function<double(double)> &f;
        // [a, c] with midpoint b and the three-point estimate `whole` of its integral
        double run(double a, double b, double c, double fa, double fb, double fc, double whole, double eps, int depth) const {
            const double l = 0.5 * (a + b), r = 0.5 * (b + c), fl = f(l), fr = f(r), w = c - a;
            const double left = w * (fa + 4.0 * fl + fb) / 12.0, right = w * (fb + 4.0 * fr + fc) / 12.0, both = left + right;
            if (depth <= 0 || std::fabs(both - whole) < 15.0 * eps) return both + (both - whole) / 15.0;
            return run(a, l, b, fa, fl, fb, left, 0.5 * eps, depth - 1) + run(b, r, c, fb, fr, fc, right, 0.5 * eps, depth - 1);
        }
    } rec = {f};
    const double mid = 0.5 * (x0 + x1), f0 = f(x0), fm = f(mid), f1 = f(x1);
    return rec.run(x0, mid, x1, f0, fm, f1, (x1 - x0) * (f0 + 4.0 * fm + f1) / 6.0, eps, depth);
}

/// Nested adaptive Simpson quadrature of f(x, y) over [x0, x1] x [y0, y1].
inline double adaptiveSimpson2D(const std::function<double(double, double)> &f, double x0, double y0, double x1, double y1,
                                double eps = 1e-6, int depth = 6) {
    return adaptiveSimpson([&](double y) { return adaptiveSimpson([&](double x) { return f(x, y); }, x0, x1, eps, depth); }, y0, y1, eps, depth);
}

/// Composite Simpson weights for n (even) intervals of [a, b]: n + 1 nodes a + i (b - a) / n.
inline std::vector<double> simpsonWeights(int n, double a, double b) {
    std::vector<double> w((size_t) n + 1);
    const double h = (b - a) / n;
    for (int i = 0; i <= n; ++i) w[(size_t) i] = (i == 0 || i == n ? 1.0 : (i & 1 ? 4.0 : 2.0)) * h / 3.0;
    return w;
}

inline double sidak(double significanceLevel, int numTests) { return 1.0 - std::pow(1.0 - significanceLevel, 1.0 / numTests); }

/// Chi^2 test of observed against expected cell frequencies. Returns (accepted, report).
inline std::pair<bool, std::string> chi2_test(int nCells, const double *obs, const double *expd, int sampleCount, double minExpFrequency,
                                              double significanceLevel, int numTests = 1) {
    std::vector<int> order((size_t) nCells);
    for (int i = 0; i < nCells; ++i) order[(size_t) i] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return expd[a] < expd[b]; });
    std::ostringstream out;
    double poolObs = 0, poolExp = 0, chsq = 0;
    int pooled = 0, terms = 0;
    for (int i : order) {
        const double e = expd[i], o = obs[i];
        if (e < 0) {
            out << "Encountered a negative expected number of samples (" << e << "). Rejecting the null hypothesis!" << std::endl;
            return std::make_pair(false, out.str());
        }
        if (e == 0) {                                   // a few stray samples are tolerated (rounding), more are not
            if (o > sampleCount * 1e-5) {
                out << "Encountered " << o << " samples in a cell with expected frequency 0. Rejecting the null hypothesis!" << std::endl;
                return std::make_pair(false, out.str());
            }
        } else if (e < minExpFrequency || (poolExp > 0 && poolExp < minExpFrequency)) {
            poolObs += o; poolExp += e; ++pooled;
        } else {
            chsq += (o - e) * (o - e) / e; ++terms;
        }
    }
    if (poolExp > 0 || poolObs > 0) {
        out << "Pooled " << pooled << " to ensure sufficiently high expected cell frequencies (>" << minExpFrequency << ")" << std::endl;
        chsq += (poolObs - poolExp) * (poolObs - poolExp) / poolExp; ++terms;
    }
    const int dof = terms - 1;
    if (dof <= 0) {
        out << "The number of degrees of freedom (" << dof << ") is too low!" << std::endl;
        return std::make_pair(false, out.str());
    }
    out << "Chi^2 statistic = " << chsq << " (d.o.f. = " << dof << ")" << std::endl;
    const double pval = 1.0 - chi2_cdf(chsq, dof), alpha = sidak(significanceLevel, numTests);
    const bool ok = !(pval < alpha) && std::isfinite(pval);
    out << (ok ? "Accepted" : "***** Rejected *****") << " the null hypothesis (p-value = " << pval << ", significance level = " << alpha << ")" << std::endl;
    return std::make_pair(ok, out.str());
}

/// Two-sided t-test of a sample mean against a reference value. Returns (accepted, report).
inline std::pair<bool, std::string> students_t_test(double mean, double variance, double reference, int sampleCount,
                                                    double significanceLevel, int numTests) {
    std::ostringstream out;
    const double t = std::fabs(mean - reference) * std::sqrt(sampleCount / std::max(variance, 1e-5));
    const int dof = sampleCount - 1;
    out << "Sample mean = " << mean << " (reference value = " << reference << ")" << std::endl;
    out << "Sample variance = " << variance << std::endl;
    out << "t-statistic = " << t << " (d.o.f. = " << dof << ")" << std::endl;
    const double pval = 2.0 * (1.0 - students_t_cdf(t, dof)), alpha = sidak(significanceLevel, numTests);
    const bool ok = !(pval < alpha);
    out << (ok ? "Accepted" : "***** Rejected *****") << " the null hypothesis (p-value = " << pval << ", significance level = " << alpha << ")" << std::endl;
    return std::make_pair(ok, out.str());
}

} // namespace hypothesis
