// common.cpp — string helpers and transforms of the host layer (see nori/common.h, nori/transform.h).
#include <nori/transform.h>
#include <algorithm>
#include <cstdlib>
#include <cstring>

NORI_NAMESPACE_BEGIN

std::string toLower(const std::string &value) {
    std::string r(value);
    std::transform(r.begin(), r.end(), r.begin(), [](unsigned char c) { return (char) std::tolower(c); });
    return r;
}

bool toBool(const std::string &str) {
    const std::string v = toLower(str);
    if (v == "true") return true;
    if (v == "false") return false;
    throw NoriException("Could not parse boolean value \"%s\"", str);
}

int toInt(const std::string &str) {
    char *end = nullptr;
    const long r = strtol(str.c_str(), &end, 10);
    if (*end != '\0') throw NoriException("Could not parse integer value \"%s\"", str);
    return (int) r;
}

float toFloat(const std::string &str) {
    char *end = nullptr;
    const float r = strtof(str.c_str(), &end);
    if (*end != '\0') throw NoriException("Could not parse floating point value \"%s\"", str);
    return r;
}

std::vector<std::string> tokenize(const std::string &s, const std::string &delim, bool includeEmpty) {
    // src/common.cpp:116-131: empty tokens between delimiters are dropped, the segment after the
    // last delimiter is always kept (so "1, 2, 3," has four tokens and is rejected by toVector3f)
    std::vector<std::string> out;
    size_t start = 0;
    for (;;) {
        const size_t pos = s.find_first_of(delim, start);
        if (pos == std::string::npos) { out.push_back(s.substr(start)); break; }
        if (pos != start || includeEmpty) out.push_back(s.substr(start, pos - start));
        start = pos + 1;
    }
    return out;
}

Vector3f toVector3f(const std::string &str) {
    const std::vector<std::string> t = tokenize(str);
    if (t.size() != 3) throw NoriException("Expected 3 values");
    return Vector3f(toFloat(t[0]), toFloat(t[1]), toFloat(t[2]));
}

std::string indent(const std::string &s, int amount) {
    std::string out, pad((size_t) amount, ' ');
    for (size_t i = 0; i < s.size(); ++i) {
        out += s[i];
        if (s[i] == '\n' && i + 1 < s.size()) out += pad;
    }
    return out;
}

std::string timeString(double t, bool precise) {
    if (std::isnan(t) || std::isinf(t)) return "inf";
    const char *suffix = "ms";
    if (t > 1000) { t /= 1000; suffix = "s";
        if (t > 60) { t /= 60; suffix = "m";
            if (t > 60) { t /= 60; suffix = "h";
                if (t > 12) { t /= 12; suffix = "d"; } } } }
    char buf[64];
    snprintf(buf, sizeof(buf), precise ? "%.4f%s" : "%.1f%s", t, suffix);
    return buf;
}

std::string memString(size_t size, bool precise) {
    static const char *suffixes[] = {"B", "KiB", "MiB", "GiB", "TiB", "PiB"};
    double v = (double) size; int s = 0;
    while (s < 5 && v > 1024.0) { v /= 1024.0; ++s; }
    char buf[64];
    snprintf(buf, sizeof(buf), s == 0 ? "%.0f %s" : (precise ? "%.4f %s" : "%.1f %s"), v, suffixes[s]);
    return buf;
}

// ---- matrices -------------------------------------------------------------------------------
// Inverse by 2x2 blocks (the "divide and conquer" scheme of Intel's AP-928 note, which is what Eigen's
// single-precision 4x4 inverse runs on x86-64): with M = [P Q; R S],
//   det M   = |P||S| + |Q||R| - tr(adj(P) Q adj(S) R)
//   M^-1    = 1/det [ P|S| - Q adj(S) R      ...
// evaluated in float, one rounding per product / sum, in the order below. Scene files reach the
// device through this routine twice (normals use the inverse transpose of toWorld, the camera the
// inverse of its projection), so its rounding is part of "same inputs as the reference".
// Block vectors hold a 2x2 matrix as (x0 x1; x2 x3); the blocks are taken from the TRANSPOSE of m
// (Eigen stores column-major), which changes nothing mathematically: inv(M^T) = inv(M)^T.
Matrix4f Matrix4f::inverse() const {
    const float A[4] = {m[0][0], m[1][0], m[0][1], m[1][1]}, B[4] = {m[2][0], m[3][0], m[2][1], m[3][1]};
    const float C[4] = {m[0][2], m[1][2], m[0][3], m[1][3]}, D[4] = {m[2][2], m[3][2], m[2][3], m[3][3]};
    // adj(A) B and adj(D) C
    const float AB[4] = {A[3] * B[0] - A[1] * B[2], A[3] * B[1] - A[1] * B[3], A[0] * B[2] - A[2] * B[0], A[0] * B[3] - A[2] * B[1]};
    const float DC[4] = {D[3] * C[0] - D[1] * C[2], D[3] * C[1] - D[1] * C[3], D[0] * C[2] - D[2] * C[0], D[0] * C[3] - D[2] * C[1]};
    const float dA = A[3] * A[0] - A[1] * A[2], dB = B[3] * B[0] - B[1] * B[2];
    const float dC = C[3] * C[0] - C[1] * C[2], dD = D[3] * D[0] - D[1] * D[2];
    // trace(adj(A) B adj(D) C)
    const float t0 = DC[0] * AB[0], t1 = DC[2] * AB[1], t2 = DC[1] * AB[2], t3 = DC[3] * AB[3];
    const float tr = (t0 + t2) + (t1 + t3);
    float iD[4] = {C[0] * AB[0] + C[1] * AB[2], C[0] * AB[1] + C[1] * AB[3], C[2] * AB[0] + C[3] * AB[2], C[2] * AB[1] + C[3] * AB[3]};
    float iA[4] = {B[0] * DC[0] + B[1] * DC[2], B[0] * DC[1] + B[1] * DC[3], B[2] * DC[0] + B[3] * DC[2], B[2] * DC[1] + B[3] * DC[3]};
    const float d1 = dA * dD, d2 = dB * dC;
    for (int i = 0; i < 4; ++i) { iD[i] = D[i] * dA - iD[i]; iA[i] = A[i] * dD - iA[i]; }
    const float det = (d1 + d2) - tr;
    const float rd = 1.0f / det;
    float iB[4] = {D[0] * AB[3] - D[1] * AB[2], D[1] * AB[0] - D[0] * AB[1], D[2] * AB[3] - D[3] * AB[2], D[3] * AB[0] - D[2] * AB[1]};
    float iC[4] = {A[0] * DC[3] - A[1] * DC[2], A[1] * DC[0] - A[0] * DC[1], A[2] * DC[3] - A[3] * DC[2], A[3] * DC[0] - A[2] * DC[1]};
    for (int i = 0; i < 4; ++i) { iB[i] = C[i] * dB - iB[i]; iC[i] = B[i] * dC - iC[i]; }
    const float sgn[4] = {rd, -rd, -rd, rd};
    for (int i = 0; i < 4; ++i) { iA[i] = sgn[i] * iA[i]; iB[i] = sgn[i] * iB[i]; iC[i] = sgn[i] * iC[i]; iD[i] = sgn[i] * iD[i]; }
    Matrix4f r;                                               // columns of the result
    const float col0[4] = {iA[3], iA[1], iB[3], iB[1]}, col1[4] = {iA[2], iA[0], iB[2], iB[0]};
    const float col2[4] = {iC[3], iC[1], iD[3], iD[1]}, col3[4] = {iC[2], iC[0], iD[2], iD[0]};
    for (int i = 0; i < 4; ++i) { r.m[i][0] = col0[i]; r.m[i][1] = col1[i]; r.m[i][2] = col2[i]; r.m[i][3] = col3[i]; }
    return r;
}

std::string Transform::toString() const {
    std::string s = "[";
    char buf[32];
    for (int i = 0; i < 4; ++i) {
        for (int j = 0; j < 4; ++j) {
            snprintf(buf, sizeof(buf), "%g", matrix(i, j));
            s += buf;
            if (j < 3) s += ", ";
        }
        if (i < 3) s += ";\n ";
    }
    return s + "]";
}

// Affine accumulation: `a` has last row (0,0,0,1). The product keeps the accumulated last row and
// combines the upper three rows column-wise, like the matrix product above.
void TransformBuilder::leftMultiplyAffine(const Matrix4f &a) { m_cur = a * m_cur; }

void TransformBuilder::translate(const Vector3f &v) {
    // Translation * T only adds to the translation column (for an affine T whose last row is 0 0 0 1)
    m_cur(0, 3) += v.x; m_cur(1, 3) += v.y; m_cur(2, 3) += v.z;
}

void TransformBuilder::scale(const Vector3f &v) {
    for (int j = 0; j < 4; ++j) { m_cur(0, j) *= v.x; m_cur(1, j) *= v.y; m_cur(2, j) *= v.z; }
}

void TransformBuilder::rotate(float angleDeg, const Vector3f &axis) {
    // Rodrigues' formula in the arrangement of an angle-axis -> matrix conversion:
    // R = c I + s [a]x + (1 - c) a a^T, with the axis used as given (the reference does not normalise it)
    const float angle = degToRad(angleDeg);
    const float s = std::sin(angle), c = std::cos(angle);
    const Vector3f sinAxis = axis * s, cos1Axis = axis * (1.0f - c);
    Matrix4f r = Matrix4f::identity();
    float tmp;
    tmp = cos1Axis.x * axis.y; r(0, 1) = tmp - sinAxis.z; r(1, 0) = tmp + sinAxis.z;
    tmp = cos1Axis.x * axis.z; r(0, 2) = tmp + sinAxis.y; r(2, 0) = tmp - sinAxis.y;
    tmp = cos1Axis.y * axis.z; r(1, 2) = tmp - sinAxis.x; r(2, 1) = tmp + sinAxis.x;
    r(0, 0) = cos1Axis.x * axis.x + c; r(1, 1) = cos1Axis.y * axis.y + c; r(2, 2) = cos1Axis.z * axis.z + c;
    leftMultiplyAffine(r);
}

void TransformBuilder::matrix(const float v[16]) {
    Matrix4f a;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) a(i, j) = v[4 * i + j];
    leftMultiplyAffine(a);
}

void TransformBuilder::lookAt(const Vector3f &origin, const Vector3f &target, const Vector3f &up) {
    const Vector3f dir = (target - origin).normalized();
    const Vector3f left = up.normalized().cross(dir).normalized();
    const Vector3f newUp = dir.cross(left).normalized();
    Matrix4f a = Matrix4f::identity();
    for (int i = 0; i < 3; ++i) { a(i, 0) = left[i]; a(i, 1) = newUp[i]; a(i, 2) = dir[i]; a(i, 3) = origin[i]; }
    leftMultiplyAffine(a);
}

NORI_NAMESPACE_END
