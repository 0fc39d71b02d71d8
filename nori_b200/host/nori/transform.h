// nori/transform.h — 4x4 homogeneous transforms of the host layer.
// Reference: include/nori/transform.h:30-98 (Transform keeps the matrix and its inverse),
// src/parser.cpp:249-298 (how <translate>/<scale>/<rotate>/<matrix>/<lookat> compose),
// src/perspective.cpp:53-86 (the camera's sample-to-camera matrix).
//
// The reference does this arithmetic with Eigen; the matrices reach the device as 16 floats each
// (NoriGpuCameraDesc), so what matters is that they carry the same values. The routines below keep
// Eigen's evaluation orders where the result depends on them (column-combination products).
#pragma once
#include "common.h"

NORI_NAMESPACE_BEGIN

struct Matrix4f {
    float m[4][4];                                   // m[row][col]
    static Matrix4f identity() {
        Matrix4f r;
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i][j] = i == j ? 1.0f : 0.0f;
        return r;
    }
    float operator()(int i, int j) const { return m[i][j]; }
    float &operator()(int i, int j) { return m[i][j]; }
    /// Product as a combination of the left operand's columns: ((c0 b0 + c1 b1) + c2 b2) + c3 b3.
    Matrix4f operator*(const Matrix4f &b) const {
        Matrix4f r;
        for (int j = 0; j < 4; ++j)
            for (int i = 0; i < 4; ++i)
                r.m[i][j] = ((m[i][0] * b.m[0][j] + m[i][1] * b.m[1][j]) + m[i][2] * b.m[2][j]) + m[i][3] * b.m[3][j];
        return r;
    }
    /// General inverse by 2x2 blocks in single precision (see common.cpp for why this scheme).
    Matrix4f inverse() const;
    void toRowMajor(float out[16]) const { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[4 * i + j] = m[i][j]; }
};

struct Transform {
    Matrix4f matrix, inv;
    Transform() : matrix(Matrix4f::identity()), inv(Matrix4f::identity()) {}
    explicit Transform(const Matrix4f &t) : matrix(t), inv(t.inverse()) {}
    Transform(const Matrix4f &t, const Matrix4f &i) : matrix(t), inv(i) {}
    const Matrix4f &getMatrix() const { return matrix; }
    const Matrix4f &getInverseMatrix() const { return inv; }
    Transform inverse() const { return Transform(inv, matrix); }
    Transform operator*(const Transform &t) const { return Transform(matrix * t.matrix, t.inv * inv); }
    /// direction: upper-left 3x3 times v, rows reduced as x0 + (x1 + x2)
    Vector3f applyVector(const Vector3f &v) const {
        return Vector3f(matrix(0, 0) * v.x + (matrix(0, 1) * v.y + matrix(0, 2) * v.z),
                        matrix(1, 0) * v.x + (matrix(1, 1) * v.y + matrix(1, 2) * v.z),
                        matrix(2, 0) * v.x + (matrix(2, 1) * v.y + matrix(2, 2) * v.z));
    }
    /// normal: transpose of the inverse's upper-left 3x3
    Normal3f applyNormal(const Normal3f &n) const {
        return Normal3f(inv(0, 0) * n.x + (inv(1, 0) * n.y + inv(2, 0) * n.z),
                        inv(0, 1) * n.x + (inv(1, 1) * n.y + inv(2, 1) * n.z),
                        inv(0, 2) * n.x + (inv(1, 2) * n.y + inv(2, 2) * n.z));
    }
    /// point: homogeneous product (columns combined left to right) and divide by w
    Point3f applyPoint(const Point3f &p) const {
        float r[4];
        for (int i = 0; i < 4; ++i) r[i] = ((matrix(i, 0) * p.x + matrix(i, 1) * p.y) + matrix(i, 2) * p.z) + matrix(i, 3);
        return Point3f(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
    }
    std::string toString() const;
};

/// Incremental builder for the contents of a <transform> element: every operation is applied on the
/// LEFT of what has been accumulated so far (src/parser.cpp:252-296).
class TransformBuilder {
public:
    TransformBuilder() : m_cur(Matrix4f::identity()) {}
    void translate(const Vector3f &v);
    void scale(const Vector3f &v);
    void rotate(float angleDeg, const Vector3f &axis);
    void matrix(const float rowMajor[16]);
    void lookAt(const Vector3f &origin, const Vector3f &target, const Vector3f &up);
    const Matrix4f &result() const { return m_cur; }
private:
    void leftMultiplyAffine(const Matrix4f &a);
    Matrix4f m_cur;
};

NORI_NAMESPACE_END
