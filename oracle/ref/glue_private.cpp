// glue_private.cpp — compiles two UNMODIFIED reference translation units inside this one so the
// glue can read their private parameters exactly (TEST INFRASTRUCTURE; oracle/_ref only).
// src/perspective.cpp and src/diffuse.cpp define their classes inside the .cpp, so there is no
// header to reach m_sampleToCamera / m_albedo through. Nothing is copied: the files are
// #included from /root/reference where they lie (the Makefile passes -I$(REF)/src).
#include <nori/camera.h>
#include <nori/rfilter.h>
#include <nori/warp.h>
#include <nori/bsdf.h>
#include <nori/frame.h>
#include <Eigen/Geometry>
#include <sstream>
#include <string>

#define private public
#define protected public
#include <perspective.cpp>
#include <diffuse.cpp>
#undef private
#undef protected

namespace nori {

/// Row-major copies of PerspectiveCamera's matrices after activate() (perspective.cpp:53-86).
bool refCameraState(const Camera *cam, float *s2c, float *c2w, float *nearClip, float *farClip, float *fov) {
    const PerspectiveCamera *pc = dynamic_cast<const PerspectiveCamera *>(cam);
    if (!pc) return false;
    const Eigen::Matrix4f &a = pc->m_sampleToCamera.getMatrix(), &b = pc->m_cameraToWorld.getMatrix();
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { s2c[4 * i + j] = a(i, j); c2w[4 * i + j] = b(i, j); }
    *nearClip = pc->m_nearClip; *farClip = pc->m_farClip; *fov = pc->m_fov;
    return true;
}

bool refDiffuseAlbedo(const BSDF *bsdf, float *rgb) {
    const Diffuse *d = dynamic_cast<const Diffuse *>(bsdf);
    if (!d) return false;
    rgb[0] = d->m_albedo.r(); rgb[1] = d->m_albedo.g(); rgb[2] = d->m_albedo.b();
    return true;
}

} // namespace nori
