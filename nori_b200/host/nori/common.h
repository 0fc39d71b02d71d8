// nori/common.h — basic types of the C++ host layer of nori_b200.
//
// The host layer keeps the reference's plugin surface (NoriObject factory, PropertyList, XML scene
// description, Scene/Mesh/BSDF/Camera/... classes registered under the reference's names) and hands
// the render path to libnori_b200.so through the C ABI of include/nori_gpu.h. It has no CPU
// renderer: every operator that computes something (rayIntersect, BSDF::sample/eval/pdf, render)
// runs on the device or throws.
//
// Reference: include/nori/common.h (NoriException :147-152, string helpers src/common.cpp:63-140).
#pragma once
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

#define NORI_NAMESPACE_BEGIN namespace nori {
#define NORI_NAMESPACE_END }

NORI_NAMESPACE_BEGIN

/// Error type of the whole host layer (reference: NoriException, include/nori/common.h:147-152).
class NoriException : public std::runtime_error {
public:
    explicit NoriException(const std::string &msg) : std::runtime_error(msg) {}
    /// printf-style constructor (the reference formats with tinyformat; same call sites, %s/%d/%f)
    template <typename... Args> NoriException(const char *fmt, const Args &...args) : std::runtime_error(format(fmt, pass(args)...)) {}

private:
    static const char *pass(const std::string &s) { return s.c_str(); }
    template <typename T> static T pass(const T &v) { return v; }
    static std::string format(const char *fmt, ...) {
        va_list ap, ap2;
        va_start(ap, fmt); va_copy(ap2, ap);
        const int n = vsnprintf(nullptr, 0, fmt, ap);
        va_end(ap);
        std::string out((size_t) (n > 0 ? n : 0), '\0');
        if (n > 0) vsnprintf(&out[0], (size_t) n + 1, fmt, ap2);
        va_end(ap2);
        return out;
    }
};

// ---- small fixed-size vectors (plain float, no SIMD: every operation rounds once, in the order written)
struct Vector2f { float x = 0, y = 0; Vector2f() {} Vector2f(float x_, float y_) : x(x_), y(y_) {} };
struct Vector2i { int x = 0, y = 0; Vector2i() {} Vector2i(int x_, int y_) : x(x_), y(y_) {} };
typedef Vector2i Point2i;
struct Vector3f {
    float x = 0, y = 0, z = 0;
    Vector3f() {}
    explicit Vector3f(float v) : x(v), y(v), z(v) {}
    Vector3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    Vector3f operator+(const Vector3f &b) const { return Vector3f(x + b.x, y + b.y, z + b.z); }
    Vector3f operator-(const Vector3f &b) const { return Vector3f(x - b.x, y - b.y, z - b.z); }
    Vector3f operator*(float s) const { return Vector3f(x * s, y * s, z * s); }
    Vector3f operator/(float s) const { return Vector3f(x / s, y / s, z / s); }
    Vector3f operator-() const { return Vector3f(-x, -y, -z); }
    /// Eigen's 3-vector reductions add as x0 + (x1 + x2) (see oracle/noracle_math.h)
    float dot(const Vector3f &b) const { return x * b.x + (y * b.y + z * b.z); }
    float squaredNorm() const { return x * x + (y * y + z * z); }
    float norm() const { return std::sqrt(squaredNorm()); }
    /// Eigen normalized(): unchanged if the squared norm is not positive
    Vector3f normalized() const { const float n2 = squaredNorm(); return n2 > 0.0f ? *this / std::sqrt(n2) : *this; }
    Vector3f cross(const Vector3f &b) const { return Vector3f(y * b.z - z * b.y, z * b.x - x * b.z, x * b.y - y * b.x); }
};
typedef Vector3f Point3f;
typedef Vector3f Normal3f;
typedef Vector2f Point2f;

/// RGB triple (reference: Color3f, include/nori/color.h)
struct Color3f {
    float r = 0, g = 0, b = 0;
    Color3f() {}
    explicit Color3f(float v) : r(v), g(v), b(v) {}
    Color3f(float r_, float g_, float b_) : r(r_), g(g_), b(b_) {}
    float maxCoeff() const { return std::fmax(r, std::fmax(g, b)); }
    /// src/common.cpp:210-217: no NaN/Inf/negative component
    bool isValid() const { return r >= 0 && g >= 0 && b >= 0 && std::isfinite(r) && std::isfinite(g) && std::isfinite(b); }
    /// src/common.cpp:219-221
    float getLuminance() const { return r * 0.212671f + g * 0.715160f + b * 0.072169f; }
    /// src/common.cpp:178-192
    Color3f toSRGB() const {
        auto f = [](float v) { return v <= 0.0031308f ? 12.92f * v : (1.0f + 0.055f) * std::pow(v, 1.0f / 2.4f) - 0.055f; };
        return Color3f(f(r), f(g), f(b));
    }
};

/// Ray segment (reference: TRay, include/nori/ray.h)
struct Ray3f {
    Point3f o; Vector3f d; float mint = 1e-4f, maxt = INFINITY;
    Ray3f() {}
    Ray3f(const Point3f &o_, const Vector3f &d_, float mint_ = 1e-4f, float maxt_ = INFINITY) : o(o_), d(d_), mint(mint_), maxt(maxt_) {}
};

// ---- string helpers (src/common.cpp:63-140: same accept/reject behaviour)
std::string toLower(const std::string &value);
bool toBool(const std::string &str);
int toInt(const std::string &str);
float toFloat(const std::string &str);
Vector3f toVector3f(const std::string &str);
std::vector<std::string> tokenize(const std::string &s, const std::string &delim = ", ", bool includeEmpty = false);
std::string indent(const std::string &s, int amount = 2);
std::string timeString(double ms, bool precise = false);
std::string memString(size_t size, bool precise = false);
/// Unit vector from spherical coordinates (src/common.cpp:244-252)
inline Vector3f sphericalDirection(float theta, float phi) {
    float sinTheta, cosTheta, sinPhi, cosPhi;
    sincosf(theta, &sinTheta, &cosTheta); sincosf(phi, &sinPhi, &cosPhi);
    return Vector3f(sinTheta * cosPhi, sinTheta * sinPhi, cosTheta);
}

/// PCG32 (XSH-RR 64/32, O'Neill 2014), the generator behind the reference's `independent` sampler and
/// its test plugins (ext/pcg32/pcg32.h:40-112): default stream, nextUInt, nextFloat by the
/// exponent-bits trick. The device has its own copy (csrc/nmath.cuh); this one drives the <test> plugins.
struct Pcg32 {
    uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
    Pcg32() {}
    Pcg32(uint64_t initstate, uint64_t initseq) { seed(initstate, initseq); }
    void seed(uint64_t initstate, uint64_t initseq) {
        state = 0; inc = (initseq << 1) | 1u;
        nextUInt(); state += initstate; nextUInt();
    }
    uint32_t nextUInt() {
        const uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        const uint32_t xorshifted = (uint32_t) (((old >> 18u) ^ old) >> 27u), rot = (uint32_t) (old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    float nextFloat() {
        union { uint32_t u; float f; } x;
        x.u = (nextUInt() >> 9) | 0x3f800000u;
        return x.f - 1.0f;
    }
};

inline float degToRad(float v) { return v * (3.14159265358979323846f / 180.0f); }

NORI_NAMESPACE_END
