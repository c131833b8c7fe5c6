// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under path_tracer_b200/ may include, link or call this.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// PARITY UNPINNED: the reference (CouncilmanJeremyJamm/path_tracer) ships no tests, golden vectors or
// fixtures and cannot be compiled here (no Rust toolchain, un-vendored crates), so this restatement is
// pinned only by its own self-checks (tests/test_oracle_*.py).  See DESIGN.md "Oracle".
//
// po_math.hpp — the arithmetic contract of glam 0.23.0 (SSE2 backend; Cargo.lock:486-487) and
// nanorand 0.7.0 (Cargo.lock:900-901) as used by the reference.  Neither crate is vendored under
// /root/reference; their published algorithms are restated here.
// Compile with -ffp-contract=off and without -ffast-math: Rust never contracts a*b+c, only the explicit
// mul_add call sites (ray.rs:20, material.rs:205,487, integrator.rs:211,258-259) are fused.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

namespace po {

constexpr float EPSILON = 5e-04f;                                   // utility.rs:4
constexpr float INF = std::numeric_limits<float>::infinity();       // utility.rs:5
constexpr float PI_F = 3.14159265358979323846264338327950288f;      // std::f32::consts::PI
constexpr float TAU_F = 6.28318530717958647692528676655900577f;     // std::f32::consts::TAU
constexpr float FRAC_1_PI_F = 0.318309886183790671537767526745028724f;

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// SSE minps/maxps: "if either operand is NaN return the second" == (a<b)?a:b / (a>b)?a:b
inline float sse_min(float a, float b) { return (a < b) ? a : b; }
inline float sse_max(float a, float b) { return (a > b) ? a : b; }

// f32::total_cmp key (IEEE totalOrder): bits ^ (((bits as i32) >> 31) as u32 >> 1), compared as i32
inline int32_t total_key(float f)
{
    int32_t b = (int32_t)f2u(f);
    b ^= (int32_t)(((uint32_t)(b >> 31)) >> 1);
    return b;
}
inline bool total_lt(float a, float b) { return total_key(a) < total_key(b); }

// f32::signum: NaN -> NaN, else copysign(1, x).  Equality of two signums (primitive.rs:122,131,138):
inline bool same_signum(float a, float b)
{
    if (std::isnan(a) || std::isnan(b)) return false;  // NaN != NaN
    return (f2u(a) >> 31) == (f2u(b) >> 31);
}
inline float signum(float a) { return std::isnan(a) ? a : std::copysign(1.0f, a); }

struct V3
{
    float x, y, z;
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct V4
{
    float x, y, z, w;
};

inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 splat(float s) { return V3{s, s, s}; }
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }  // true division per lane
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
inline V3& operator*=(V3& a, V3 b) { a = a * b; return a; }
inline V3 vmin(V3 a, V3 b) { return {sse_min(a.x, b.x), sse_min(a.y, b.y), sse_min(a.z, b.z)}; }
inline V3 vmax(V3 a, V3 b) { return {sse_max(a.x, b.x), sse_max(a.y, b.y), sse_max(a.z, b.z)}; }
inline float max_element(V3 a) { return sse_max(sse_max(a.x, a.y), a.z); }
inline float min_element(V3 a) { return sse_min(sse_min(a.x, a.y), a.z); }

// glam sse2 dot3_in_x: (x*x' + y*y') + z*z'
inline float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// glam sse2 dot4_in_x: (x*x' + z*z') + (y*y' + w*w')
inline float dot4(V4 a, V4 b) { return (a.x * b.x + a.z * b.z) + (a.y * b.y + a.w * b.w); }
// glam sse2 cross: two rounded products and one subtract per lane
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline float length_squared(V3 a) { return dot3(a, a); }
inline float length(V3 a) { return std::sqrt(dot3(a, a)); }
// glam sse2 Vec3A::normalize: _mm_div_ps(v, sqrt(dot))
inline V3 normalize(V3 a) { return a / std::sqrt(dot3(a, a)); }
inline V3 recip(V3 a) { return {1.0f / a.x, 1.0f / a.y, 1.0f / a.z}; }
// Vec3A::mul_add(self, a, b) = self*a + b, single rounding (glam fast-math + target-cpu=native -> FMA)
inline V3 mul_add(V3 s, V3 a, V3 b) { return {std::fmaf(s.x, a.x, b.x), std::fmaf(s.y, a.y, b.y), std::fmaf(s.z, a.z, b.z)}; }
inline bool is_finite(V3 a) { return std::isfinite(a.x) && std::isfinite(a.y) && std::isfinite(a.z); }
inline bool is_nan(V3 a) { return std::isnan(a.x) || std::isnan(a.y) || std::isnan(a.z); }

// Mat3A: three columns.  mul_vec3a = ((c0*x + c1*y) + c2*z)
struct M3
{
    V3 c0, c1, c2;
};
inline V3 mul(const M3& m, V3 v) { return (m.c0 * v.x + m.c1 * v.y) + m.c2 * v.z; }
inline M3 transpose(const M3& m) { return {{m.c0.x, m.c1.x, m.c2.x}, {m.c0.y, m.c1.y, m.c2.y}, {m.c0.z, m.c1.z, m.c2.z}}; }
// glam Mat3A::inverse: cross products, det = z_axis . (x_axis x y_axis), times 1/det, transposed
inline M3 inverse(const M3& m)
{
    V3 t0 = cross(m.c1, m.c2);
    V3 t1 = cross(m.c2, m.c0);
    V3 t2 = cross(m.c0, m.c1);
    float det = dot3(m.c2, t2);
    V3 inv_det = splat(1.0f / det);
    return transpose(M3{t0 * inv_det, t1 * inv_det, t2 * inv_det});
}

// Affine3A: matrix3 + translation
struct Aff3
{
    M3 m;
    V3 t;
};
inline V3 transform_point(const Aff3& a, V3 p) { return mul(a.m, p) + a.t; }
inline V3 transform_vector(const Aff3& a, V3 v) { return mul(a.m, v); }
inline Aff3 inverse(const Aff3& a)
{
    M3 mi = inverse(a.m);
    return {mi, -(mul(mi, a.t))};
}
inline Aff3 aff_from_cols12(const float* f) { return {{{f[0], f[1], f[2]}, {f[3], f[4], f[5]}, {f[6], f[7], f[8]}}, {f[9], f[10], f[11]}}; }
inline void aff_to_cols12(const Aff3& a, float* f)
{
    f[0] = a.m.c0.x; f[1] = a.m.c0.y; f[2] = a.m.c0.z;
    f[3] = a.m.c1.x; f[4] = a.m.c1.y; f[5] = a.m.c1.z;
    f[6] = a.m.c2.x; f[7] = a.m.c2.y; f[8] = a.m.c2.z;
    f[9] = a.t.x; f[10] = a.t.y; f[11] = a.t.z;
}

// ---- ray.rs:1-28 -------------------------------------------------------------------------------
struct Ray
{
    V3 o, d, inv;
};
inline Ray make_ray(V3 o, V3 d) { return {o, d, recip(d)}; }                                  // ray.rs:10-18
inline V3 ray_at(const Ray& r, float t) { return mul_add(r.d, splat(t), r.o); }               // ray.rs:20
inline Ray ray_transform(const Ray& r, const Aff3& inv_m)                                     // ray.rs:22-28
{
    return make_ray(transform_point(inv_m, r.o), transform_vector(inv_m, r.d));
}

// ---- nanorand 0.7.0 WyRand (recalled; parity unpinned, streams are entropy-seeded upstream) ------
// One generator per (seed, pixel, sample) instead of one per OS thread (main.rs:190): the reference's
// render is non-deterministic, ours is a counter-keyed stream so images do not depend on scheduling.
struct WyRand
{
    uint64_t s;
    uint64_t next()
    {
        s += 0xa0761d6478bd642fULL;
        __uint128_t t = (__uint128_t)s * (__uint128_t)(s ^ 0xe7037ed1a0b428dbULL);
        return (uint64_t)(t >> 64) ^ (uint64_t)t;
    }
    uint32_t u32() { return (uint32_t)next(); }                     // low 4 bytes of one output
    float f32() { return (float)u32() / 4294967296.0f; }            // u32 as f32 / (u32::MAX as f32)
};
inline uint64_t splitmix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
// stream key: (global seed, pixel index, global sample index)
inline WyRand stream_for(uint64_t seed, uint32_t pixel, uint32_t sample)
{
    uint64_t k = splitmix64(seed ^ splitmix64(((uint64_t)pixel << 32) | (uint64_t)sample));
    return WyRand{k};
}

}  // namespace po
