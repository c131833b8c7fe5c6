// exact_math.h — the floating-point contract of the parity-critical tier, usable from host and device.
//
// The reference is Rust + glam 0.23 (SSE2): a*b+c is never contracted except at explicit mul_add sites
// (ray.rs:20).  Device code is compiled with the default -fmad=true (the shading tier and the
// conservative wide-node tests want FMA), so every parity-critical operation goes through the ex::mul /
// ex::add / ex::sub / ex::div / ex::fma wrappers below, which map to the non-contractable __f*_rn intrinsics
// on the device and to plain operators on the host (host objects are built with -ffp-contract=off).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define PTB_HD __host__ __device__ __forceinline__
#else
#define PTB_HD inline
#endif

namespace ex {

constexpr float EPS = 5e-04f;  // utility.rs:4

#if defined(__CUDA_ARCH__)
PTB_HD float mul(float a, float b) { return __fmul_rn(a, b); }
PTB_HD float add(float a, float b) { return __fadd_rn(a, b); }
PTB_HD float sub(float a, float b) { return __fsub_rn(a, b); }
PTB_HD float div(float a, float b) { return __fdiv_rn(a, b); }
PTB_HD float fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
PTB_HD float sqrt(float a) { return __fsqrt_rn(a); }
PTB_HD float rcp(float a) { return __frcp_rn(a); }  // correctly rounded 1/a == __fdiv_rn(1, a), without the general division
// min/max for the slab test: one FMNMX each.  They differ from SSE minps/maxps only when the SECOND operand is NaN or for
// the sign of a zero result; in AABB::intersect_t the second operand is EPSILON, t_max or a value already clamped by one
// of them (never NaN), and a zero's sign cannot change the comparison or the returned t >= EPSILON.
PTB_HD float mn_fast(float a, float b) { return fminf(a, b); }
PTB_HD float mx_fast(float a, float b) { return fmaxf(a, b); }
PTB_HD uint32_t bits(float f) { return __float_as_uint(f); }
PTB_HD float from_bits(uint32_t u) { return __uint_as_float(u); }
#else
PTB_HD float mul(float a, float b) { return a * b; }
PTB_HD float add(float a, float b) { return a + b; }
PTB_HD float sub(float a, float b) { return a - b; }
PTB_HD float div(float a, float b) { return a / b; }
PTB_HD float fma(float a, float b, float c) { return std::fmaf(a, b, c); }
PTB_HD float sqrt(float a) { return std::sqrt(a); }
PTB_HD float rcp(float a) { return 1.0f / a; }
PTB_HD float mn_fast(float a, float b) { return (a < b) ? a : b; }
PTB_HD float mx_fast(float a, float b) { return (a > b) ? a : b; }
PTB_HD uint32_t bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
PTB_HD float from_bits(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
#endif

PTB_HD float inf() { return from_bits(0x7f800000u); }
PTB_HD bool is_nan(float a) { return a != a; }
// SSE minps/maxps semantics (second operand on NaN / equality): glam Vec3A::min/max
PTB_HD float mn(float a, float b) { return (a < b) ? a : b; }
PTB_HD float mx(float a, float b) { return (a > b) ? a : b; }
// f32::signum equality (primitive.rs:122,131,138): NaN never equal, otherwise compare sign bits
PTB_HD bool same_sign(float a, float b) { return !(is_nan(a) || is_nan(b)) && (((bits(a) ^ bits(b)) >> 31) == 0u); }
// f32::total_cmp key (boundingbox.rs:67, blas_bvh.rs:109)
PTB_HD int32_t total_key(float f)
{
    int32_t b = (int32_t)bits(f);
    return b ^ (int32_t)(((uint32_t)(b >> 31)) >> 1);
}
PTB_HD bool total_lt(float a, float b) { return total_key(a) < total_key(b); }

struct V3
{
    float x, y, z;
};
PTB_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PTB_HD V3 add(V3 a, V3 b) { return v3(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
PTB_HD V3 sub(V3 a, V3 b) { return v3(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
PTB_HD V3 mul(V3 a, V3 b) { return v3(mul(a.x, b.x), mul(a.y, b.y), mul(a.z, b.z)); }
PTB_HD V3 mul(V3 a, float s) { return v3(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
PTB_HD V3 div(V3 a, float s) { return v3(div(a.x, s), div(a.y, s), div(a.z, s)); }
PTB_HD V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
PTB_HD V3 vmin(V3 a, V3 b) { return v3(mn(a.x, b.x), mn(a.y, b.y), mn(a.z, b.z)); }
PTB_HD V3 vmax(V3 a, V3 b) { return v3(mx(a.x, b.x), mx(a.y, b.y), mx(a.z, b.z)); }
PTB_HD float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
// glam sse2 horizontal sums: dot3 = (x+y)+z ; dot4 = (x+z)+(y+w)
PTB_HD float dot3(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
PTB_HD float dot4(float ax, float ay, float az, float aw, float bx, float by, float bz, float bw)
{
    return add(add(mul(ax, bx), mul(az, bz)), add(mul(ay, by), mul(aw, bw)));
}
PTB_HD V3 cross(V3 a, V3 b)
{
    return v3(sub(mul(a.y, b.z), mul(b.y, a.z)), sub(mul(a.z, b.x), mul(b.z, a.x)), sub(mul(a.x, b.y), mul(b.x, a.y)));
}
PTB_HD float length(V3 a) { return sqrt(dot3(a, a)); }
PTB_HD V3 normalize(V3 a) { return div(a, sqrt(dot3(a, a))); }
PTB_HD V3 recip(V3 a) { return v3(rcp(a.x), rcp(a.y), rcp(a.z)); }
// Mat3A * Vec3A = ((c0*x + c1*y) + c2*z)
PTB_HD V3 m3_mul(V3 c0, V3 c1, V3 c2, V3 v) { return add(add(mul(c0, v.x), mul(c1, v.y)), mul(c2, v.z)); }
// Ray::at (ray.rs:20): direction.mul_add(splat(t), origin), fused
PTB_HD V3 ray_at(V3 o, V3 d, float t) { return v3(fma(d.x, t, o.x), fma(d.y, t, o.y), fma(d.z, t, o.z)); }

struct Box
{
    V3 lo, hi;
};
// AABB::intersect_t (boundingbox.rs:115-131): returns hit and the entry distance (>= EPS, independent of t_max)
PTB_HD bool slab(V3 lo, V3 hi, V3 o, V3 inv, float t_max, float* t_enter)
{
    V3 t0 = mul(sub(lo, o), inv);
    V3 t1 = mul(sub(hi, o), inv);
    float lx = mn_fast(mx_fast(t0.x, EPS), mx_fast(t1.x, EPS)), ly = mn_fast(mx_fast(t0.y, EPS), mx_fast(t1.y, EPS)), lz = mn_fast(mx_fast(t0.z, EPS), mx_fast(t1.z, EPS));
    float hx = mx_fast(mn_fast(t0.x, t_max), mn_fast(t1.x, t_max)), hy = mx_fast(mn_fast(t0.y, t_max), mn_fast(t1.y, t_max)), hz = mx_fast(mn_fast(t0.z, t_max), mn_fast(t1.z, t_max));
    float t = mx_fast(mx_fast(lx, ly), lz);
    *t_enter = t;
    return t <= mn_fast(mn_fast(hx, hy), hz);
}
PTB_HD float surface_area(V3 lo, V3 hi)  // boundingbox.rs:90-95
{
    V3 v = sub(hi, lo);
    return mul(2.0f, dot3(v, v3(v.z, v.x, v.y)));
}

// Triangle::intersect_naive on a pre-translated origin (primitive.rs:117-144).  n0/n1/n2 = planes with d in w.
// Outputs the undivided (td, ud, vd, det).
PTB_HD bool tri_naive(const float n0[4], const float n1[4], const float n2[4], V3 o, V3 d, float t_min, float t_max, float out[4])
{
    float det = dot3(d, v3(n0[0], n0[1], n0[2]));
    float td = -dot4(o.x, o.y, o.z, -1.0f, n0[0], n0[1], n0[2], n0[3]);
    if (!same_sign(sub(td, mul(det, t_min)), sub(mul(det, t_max), td))) return false;
    V3 p = add(mul(o, det), mul(d, td));
    float ud = dot4(p.x, p.y, p.z, det, n1[0], n1[1], n1[2], n1[3]);
    if (!same_sign(ud, sub(det, ud))) return false;
    float vd = dot4(p.x, p.y, p.z, det, n2[0], n2[1], n2[2], n2[3]);
    if (!same_sign(vd, sub(sub(det, ud), vd))) return false;
    out[0] = td; out[1] = ud; out[2] = vd; out[3] = det;
    return true;
}
// Triangle::intersect (primitive.rs:147-178) up to (t,u,v); the normal is resolved once per ray afterwards.
PTB_HD bool tri_hit(const float n0[4], const float n1[4], const float n2[4], V3 o, V3 d, float t_max, float t_est, float* t, float* u, float* v)
{
    float r[4];
    if (!tri_naive(n0, n1, n2, ray_at(o, d, t_est), d, sub(EPS, t_est), sub(t_max, t_est), r)) return false;
    *t = add(div(r[0], r[3]), t_est);
    *u = div(r[1], r[3]);
    *v = div(r[2], r[3]);
    return true;
}
// the same with the barycentric divisions left to the caller: t as above, raw = (ud, vd, det); u = ud / det, v = vd / det.
// A closest-hit search replaces its candidate several times per ray and only the last one's (u, v) is ever used.
PTB_HD bool tri_hit_raw(const float n0[4], const float n1[4], const float n2[4], V3 o, V3 d, float t_max, float t_est, float* t, float raw[3])
{
    float r[4];
    if (!tri_naive(n0, n1, n2, ray_at(o, d, t_est), d, sub(EPS, t_est), sub(t_max, t_est), r)) return false;
    *t = add(div(r[0], r[3]), t_est);
    raw[0] = r[1]; raw[1] = r[2]; raw[2] = r[3];
    return true;
}
PTB_HD bool tri_hit_bool(const float n0[4], const float n1[4], const float n2[4], V3 o, V3 d, float t_max, float t_est)  // primitive.rs:181-189
{
    float r[4];
    return tri_naive(n0, n1, n2, ray_at(o, d, t_est), d, sub(EPS, t_est), sub(t_max, t_est), r);
}
// Triangle::get_normal (primitive.rs:57-63)
PTB_HD V3 tri_normal(V3 na, V3 nb, V3 nc, float u, float v)
{
    float w = sub(sub(1.0f, u), v);
    return normalize(m3_mul(na, nb, nc, v3(w, u, v)));
}

}  // namespace ex
