// camera_host.h — Camera::new (camera.rs:17-31) evaluated on the host (it is host code upstream too): the ray matrix
// matrix * inv_projection and the eye position handed to the raygen device code (shading.h: camera_ray).
#pragma once
#include <cmath>

#include "exact_math.h"

namespace ptb {

using ex::V3;

// Mat4::inverse by cofactors (column-major 4x4), the general form glam uses for Mat4
inline void mat4_inverse(const float P[16], float IP[16])
{
    const float a00 = P[0], a01 = P[1], a02 = P[2], a03 = P[3], a10 = P[4], a11 = P[5], a12 = P[6], a13 = P[7];
    const float a20 = P[8], a21 = P[9], a22 = P[10], a23 = P[11], a30 = P[12], a31 = P[13], a32 = P[14], a33 = P[15];
    const float b00 = a00 * a11 - a01 * a10, b01 = a00 * a12 - a02 * a10, b02 = a00 * a13 - a03 * a10, b03 = a01 * a12 - a02 * a11;
    const float b04 = a01 * a13 - a03 * a11, b05 = a02 * a13 - a03 * a12, b06 = a20 * a31 - a21 * a30, b07 = a20 * a32 - a22 * a30;
    const float b08 = a20 * a33 - a23 * a30, b09 = a21 * a32 - a22 * a31, b10 = a21 * a33 - a23 * a31, b11 = a22 * a33 - a23 * a32;
    const float det = b00 * b11 - b01 * b10 + b02 * b09 + b03 * b08 - b04 * b07 + b05 * b06;
    const float id = 1.0f / det;
    IP[0] = (a11 * b11 - a12 * b10 + a13 * b09) * id; IP[1] = (a02 * b10 - a01 * b11 - a03 * b09) * id; IP[2] = (a31 * b05 - a32 * b04 + a33 * b03) * id; IP[3] = (a22 * b04 - a21 * b05 - a23 * b03) * id;
    IP[4] = (a12 * b08 - a10 * b11 - a13 * b07) * id; IP[5] = (a00 * b11 - a02 * b08 + a03 * b07) * id; IP[6] = (a32 * b02 - a30 * b05 - a33 * b01) * id; IP[7] = (a20 * b05 - a22 * b02 + a23 * b01) * id;
    IP[8] = (a10 * b10 - a11 * b08 + a13 * b06) * id; IP[9] = (a01 * b08 - a00 * b10 - a03 * b06) * id; IP[10] = (a30 * b04 - a31 * b02 + a33 * b00) * id; IP[11] = (a21 * b02 - a20 * b04 - a23 * b00) * id;
    IP[12] = (a11 * b07 - a10 * b09 - a12 * b06) * id; IP[13] = (a00 * b09 - a01 * b07 + a02 * b06) * id; IP[14] = (a31 * b01 - a30 * b03 - a32 * b00) * id; IP[15] = (a20 * b03 - a21 * b01 + a22 * b00) * id;
}

// Camera::new (camera.rs:17-31) on the host: look_at_rh(...).inverse(), perspective_infinite_rh(fov, aspect, 1).inverse(),
// ray matrix = matrix * inv_projection.
inline void camera_setup(const float origin[3], const float target[3], float fov_deg, float aspect, float m_out[16], float eye_out[3])
{
    const V3 eye = ex::v3(origin[0], origin[1], origin[2]), tgt = ex::v3(target[0], target[1], target[2]);
    const V3 up = ex::v3(0.0f, 1.0f, 0.0f);
    const V3 f = ex::normalize(ex::sub(tgt, eye));
    const V3 s = ex::normalize(ex::cross(f, up));
    const V3 u = ex::cross(s, f);
    // view = rows (s, u, -f), translation (-eye.s, -eye.u, eye.f); invert as glam Affine3A::inverse
    const V3 c0 = ex::v3(s.x, u.x, -f.x), c1 = ex::v3(s.y, u.y, -f.y), c2 = ex::v3(s.z, u.z, -f.z);
    const V3 tr = ex::v3(-ex::dot3(eye, s), -ex::dot3(eye, u), ex::dot3(eye, f));
    const V3 t0 = ex::cross(c1, c2), t1 = ex::cross(c2, c0), t2 = ex::cross(c0, c1);
    const float idet = ex::div(1.0f, ex::dot3(c2, t2));
    const V3 r0 = ex::mul(t0, idet), r1 = ex::mul(t1, idet), r2 = ex::mul(t2, idet);
    const V3 i0 = ex::v3(r0.x, r1.x, r2.x), i1 = ex::v3(r0.y, r1.y, r2.y), i2 = ex::v3(r0.z, r1.z, r2.z);
    const V3 it = ex::neg(ex::m3_mul(i0, i1, i2, tr));
    // inverse of perspective_infinite_rh(fov, aspect, 1) = cols (f/a,0,0,0) (0,f,0,0) (0,0,-1,-1) (0,0,-1,0) by cofactors
    const float fov = ex::mul(fov_deg, ex::div(3.14159265358979323846f, 180.0f));
    const float ff = ex::div(1.0f, std::tan(ex::mul(0.5f, fov)));
    float P[16] = {ex::div(ff, aspect), 0, 0, 0, 0, ff, 0, 0, 0, 0, -1.0f, -1.0f, 0, 0, -1.0f, 0};
    float IP[16];
    mat4_inverse(P, IP);
    // Mat4::from(affine) * inv_projection: column k = ((A0*x + A1*y) + A2*z) + A3*w
    const float A[16] = {i0.x, i0.y, i0.z, 0.0f, i1.x, i1.y, i1.z, 0.0f, i2.x, i2.y, i2.z, 0.0f, it.x, it.y, it.z, 1.0f};
    for (int k = 0; k < 4; ++k)
        for (int r = 0; r < 4; ++r)
        {
            const float* v = IP + k * 4;
            m_out[k * 4 + r] = ((A[r] * v[0] + A[4 + r] * v[1]) + A[8 + r] * v[2]) + A[12 + r] * v[3];
        }
    eye_out[0] = it.x; eye_out[1] = it.y; eye_out[2] = it.z;
}


}  // namespace ptb
