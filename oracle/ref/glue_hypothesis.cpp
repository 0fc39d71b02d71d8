// glue_hypothesis.cpp — C access to the reference's statistics header (ext/hypothesis/hypothesis.h + cephes.h,
// compiled as they are): TEST INFRASTRUCTURE. Used by oracle/make_hypothesis_golden.py to pin the host layer's own
// nori_b200/host/nori/hypothesis.h (tests/golden/hypothesis_ref.json).
#include <hypothesis.h>
#include <cstring>
#include <string>

static int copyOut(const std::pair<bool, std::string> &r, char *msg, int cap) {
    if (msg && cap > 0) { std::strncpy(msg, r.second.c_str(), (size_t) cap - 1); msg[cap - 1] = 0; }
    return r.first ? 1 : 0;
}

extern "C" {
double ref_chi2_cdf(double x, int dof) { return hypothesis::chi2_cdf(x, dof); }
double ref_students_t_cdf(double x, int dof) { return hypothesis::students_t_cdf(x, dof); }
int ref_chi2_test(int nCells, const double *obs, const double *expd, int sampleCount, double minExpFrequency, double significanceLevel,
                  int numTests, char *msg, int cap) {
    return copyOut(hypothesis::chi2_test(nCells, obs, expd, sampleCount, minExpFrequency, significanceLevel, numTests), msg, cap);
}
int ref_students_t_test(double mean, double variance, double reference, int sampleCount, double significanceLevel, int numTests, char *msg, int cap) {
    return copyOut(hypothesis::students_t_test(mean, variance, reference, sampleCount, significanceLevel, numTests), msg, cap);
}
}
