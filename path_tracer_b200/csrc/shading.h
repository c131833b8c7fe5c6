// shading.h — statistical-tier device math (host+device): RNG, Sobol jitter, camera rays, ONBs, the five materials,
// homogeneous media, light sampling.  FMA contraction is allowed here (parity is statistical);
// the integer parts (WyRand, Sobol hashes) are bit-exact.
//   utility.rs, sampling.rs, camera.rs:94-105, material/onb.rs, material.rs, material/volume.rs, light_sampler.rs
#pragma once
#include "../../include/ptb.h"
#include "exact_math.h"
#include "layout.h"

// Development switch (A/B builds only): the two large material functions as real calls instead of three inlined copies each.
#if defined(PTB_SHADE_NOINLINE) && defined(__CUDACC__)
#define PTB_SHADE_FN __host__ __device__ __noinline__
#else
#define PTB_SHADE_FN PTB_HD
#endif

namespace ptb {

using ex::V3;

constexpr float PI_F = 3.14159265358979323846f;
constexpr float INV_PI_F = 0.318309886183790671538f;

// ---- small float3 helpers (contractable) ------------------------------------------------------------------------------
PTB_HD V3 operator+(V3 a, V3 b) { return ex::v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PTB_HD V3 operator-(V3 a, V3 b) { return ex::v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PTB_HD V3 operator*(V3 a, V3 b) { return ex::v3(a.x * b.x, a.y * b.y, a.z * b.z); }
PTB_HD V3 operator*(V3 a, float s) { return ex::v3(a.x * s, a.y * s, a.z * s); }
PTB_HD V3 operator*(float s, V3 a) { return ex::v3(a.x * s, a.y * s, a.z * s); }
PTB_HD V3 operator/(V3 a, float s) { return ex::v3(a.x / s, a.y / s, a.z / s); }
PTB_HD V3 operator-(V3 a) { return ex::v3(-a.x, -a.y, -a.z); }
PTB_HD float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
PTB_HD V3 crossf(V3 a, V3 b) { return ex::v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
PTB_HD V3 normalizef(V3 a) { return a / sqrtf(dot(a, a)); }
PTB_HD float max3(V3 a) { return fmaxf(fmaxf(a.x, a.y), a.z); }
PTB_HD bool finite1(float x) { return fabsf(x) <= 3.4028234664e38f; }  // false for NaN and +-inf
PTB_HD bool finite3(V3 a) { return finite1(a.x) && finite1(a.y) && finite1(a.z); }
PTB_HD bool nan3(V3 a) { return (a.x != a.x) || (a.y != a.y) || (a.z != a.z); }
struct M3
{
    V3 c0, c1, c2;
};
PTB_HD V3 mulm(const M3& m, V3 v) { return (m.c0 * v.x + m.c1 * v.y) + m.c2 * v.z; }
PTB_HD V3 mulmt(const M3& m, V3 v) { return ex::v3(dot(m.c0, v), dot(m.c1, v), dot(m.c2, v)); }  // transpose(m) * v (same sums, dot3 order)

// ---- nanorand WyRand, one stream per (seed, pixel, sample) — bit-exact integer arithmetic ------------------------
struct Rng
{
    unsigned long long s;
};
PTB_HD unsigned long long mulhi64(unsigned long long a, unsigned long long b)
{
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (unsigned long long)(((unsigned __int128)a * (unsigned __int128)b) >> 64);
#endif
}
PTB_HD uint32_t rng_u32(Rng& r)
{
    r.s += 0xa0761d6478bd642fULL;
    const unsigned long long b = r.s ^ 0xe7037ed1a0b428dbULL;
    return (uint32_t)(mulhi64(r.s, b) ^ (r.s * b));
}
PTB_HD float rng_f32(Rng& r) { return (float)rng_u32(r) * 2.3283064365386963e-10f; }  // (u32 as f32) / 2^32, exact scaling
PTB_HD unsigned long long splitmix64(unsigned long long z)
{
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
PTB_HD Rng rng_stream(unsigned long long seed, uint32_t pixel, uint32_t sample)
{
    Rng r;
    r.s = splitmix64(seed ^ splitmix64(((unsigned long long)pixel << 32) | (unsigned long long)sample));
    return r;
}

// ---- sampling.rs: scrambled Sobol jitter (integer part bit-exact) ----------------------------------------------------------------
PTB_HD uint32_t rev32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __brev(x);
#else
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
    x = ((x >> 8) & 0x00ff00ffu) | ((x & 0x00ff00ffu) << 8);
    return (x >> 16) | (x << 16);
#endif
}
PTB_HD uint32_t lk_hash(uint32_t x, uint32_t seed)  // sampling.rs:53-68
{
    x ^= x * 0x3d20adeau;
    x += seed;
    x *= (seed >> 16) | 1u;
    x ^= x * 0x05526c56u;
    x ^= x * 0x53a22864u;
    return x;
}
PTB_HD uint32_t scramble_base2(uint32_t x, uint32_t seed) { return rev32(lk_hash(rev32(x), seed)); }  // :71
PTB_HD uint32_t low_bias_hash(uint32_t x)                                                            // :76-91
{
    x ^= x >> 16;
    x *= 0x21f0aaadu;
    x ^= x >> 15;
    x *= 0xd35a2d97u;
    x ^= x >> 15;
    return x;
}
// dimension 1 of the table (sampling.rs:24-30): XOR of direction numbers over the set bits of the index.
// DIRECTIONS[b] (sampling.rs:4-8) is the Sobol dimension-2 generator v_b = v_{b-1} ^ (v_{b-1} >> 1), v_0 = 1<<31.
// BITS: the index is known to be below 2^BITS (get_ss_sobol reduces it modulo NUM_POINTS = 512 first: 9 of the 32 rounds do the work).
template <int BITS = 32>
PTB_HD uint32_t sobol_dim1(uint32_t index)
{
    uint32_t x = 0, v = 0x80000000u;
    for (int b = 0; b < BITS; ++b)
    {
        if ((index >> b) & 1u) x ^= v;
        v ^= v >> 1;
    }
    return x;
}
constexpr uint32_t SOBOL_POINTS = 512;  // NUM_POINTS, main.rs:47-48
// get_ss_sobol (sampling.rs:97-114); the 512-entry table is recomputed on the fly (dimension 0 = bit reversal)
PTB_HD void ss_sobol(uint32_t index, uint32_t seed, float* px, float* py)
{
    const uint32_t x_seed = low_bias_hash(seed), y_seed = low_bias_hash(seed + 1u), shuffle_seed = low_bias_hash(seed + 2u);
    const uint32_t i = scramble_base2(index, shuffle_seed) % SOBOL_POINTS;
    const uint32_t x = scramble_base2(rev32(i), x_seed), y = scramble_base2(sobol_dim1<9>(i), y_seed);  // i < 512
    // x as f32 / 2^32 (sampling.rs:112-113): a division by a power of two is exact, so is the product with 2^-32
    *px = ex::mul((float)x, 2.3283064365386963e-10f);
    *py = ex::mul((float)y, 2.3283064365386963e-10f);
}

// ---- camera.rs:94-105 ------------------------------------------------------------------------------------------------------
struct CameraDev
{
    float m[16];  // matrix * inv_projection, column-major
    float eye[3];
};
PTB_HD void camera_ray(const CameraDev& c, float s, float t, V3* o, V3* d)
{
    const float nx = ex::sub(ex::mul(s, 2.0f), 1.0f), ny = ex::sub(ex::mul(t, 2.0f), 1.0f), nz = 0.0f;
    float r[4];
    for (int k = 0; k < 4; ++k)
    {
        float a = ex::mul(c.m[k], nx);
        a = ex::add(ex::mul(c.m[4 + k], ny), a);
        a = ex::add(ex::mul(c.m[8 + k], nz), a);
        r[k] = ex::add(c.m[12 + k], a);
    }
    const float rw = ex::div(1.0f, r[3]);
    const V3 eye = ex::v3(c.eye[0], c.eye[1], c.eye[2]);
    const V3 p = ex::v3(ex::mul(r[0], rw), ex::mul(r[1], rw), ex::mul(r[2], rw));
    *o = eye;
    *d = ex::normalize(ex::sub(p, eye));
}

// ---- utility.rs ---------------------------------------------------------------------------------------------------------------
PTB_HD void sincos_(float x, float* s, float* c)
{
#if defined(__CUDA_ARCH__)
    sincosf(x, s, c);
#else
    *s = sinf(x);
    *c = cosf(x);
#endif
}
PTB_HD V3 random_cosine_vector(Rng& rng)  // :7-19
{
    const float r = sqrtf(rng_f32(rng));
    const float z = sqrtf(1.0f - r * r);
    const float phi = 6.28318530717958647692f * rng_f32(rng);
    float sn, cs;
    sincos_(phi, &sn, &cs);
    return ex::v3(cs * r, sn * r, z);
}
PTB_HD V3 reflect(V3 i, V3 n) { return i - (2.0f * dot(n, i)) * n; }  // :21
PTB_HD V3 refract(V3 i, V3 n, float eta)                                // :23-36, NaN = total internal reflection
{
    const float ndi = dot(n, i);
    const float k = 1.0f - eta * eta * (1.0f - ndi * ndi);
    if (k <= 0.0f) { const float q = ex::from_bits(0x7fc00000u); return ex::v3(q, q, q); }
    return eta * i - (eta * ndi + sqrtf(k)) * n;
}

// ---- onb.rs --------------------------------------------------------------------------------------------------------------------
PTB_HD M3 onb(V3 n)  // glam any_orthonormal_pair (Duff et al. 2017), onb.rs:1-7
{
    const float sign = copysignf(1.0f, n.z);
    const float a = -1.0f / (sign + n.z);
    const float b = n.x * n.y * a;
    M3 m;
    m.c0 = ex::v3(1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    m.c1 = ex::v3(b, sign + n.y * n.y * a, -n.y);
    m.c2 = n;
    return m;
}
PTB_HD M3 onb_ggx(V3 v)  // onb.rs:9-27
{
    M3 m;
    if (v.z > 0.99999f) { m.c0 = ex::v3(1, 0, 0); m.c1 = ex::v3(0, -1, 0); m.c2 = ex::v3(0, 0, 1); return m; }
    m.c0 = normalizef(crossf(v, ex::v3(0, 0, 1)));
    m.c1 = crossf(m.c0, v);
    m.c2 = v;
    return m;
}

// ---- material.rs ---------------------------------------------------------------------------------------------------------------
struct Mat
{
    uint32_t kind;
    V3 colour;
    float a, ior;
    int volume;
};
PTB_HD Mat load_mat(const MaterialRow* rows, uint32_t blas)
{
    const MaterialRow& r = rows[blas];
    Mat m;
    m.kind = r.kind;
    m.colour = ex::v3(r.colour[0], r.colour[1], r.colour[2]);
    m.a = r.a;
    m.ior = r.ior;
    m.volume = r.volume;
    return m;
}
PTB_HD bool is_delta(uint32_t kind) { return kind == PTB_SPECULAR || kind == PTB_DIELECTRIC; }
PTB_HD float weakening(uint32_t kind, V3 wo, V3 n) { return is_delta(kind) ? 1.0f : fabsf(dot(wo, n)); }  // :67-77
PTB_HD float pow5(float x) { const float x2 = x * x; return x * (x2 * x2); }

PTB_HD float ggx_d(float a, V3 h)  // :187-200 (the sqrt(1-cos^2)/cos^2 "tan_sq" quirk is the reference's, H3)
{
    if (h.z <= 0.0f) return 0.0f;
    const float c2 = h.z * h.z;
    const float tan_sq = sqrtf(1.0f - c2) / c2;
    const float x = a * a + tan_sq;
    return a * a / (PI_F * c2 * c2 * x * x);
}
PTB_HD float schlick(float vdh, float f0) { return fmaf(pow5(1.0f - vdh), 1.0f - f0, f0); }  // :205
PTB_HD float ggx_g1(float a, V3 v, V3 h)                                                      // :210-221
{
    if (v.z * dot(h, v) <= 0.0f) return 0.0f;
    const float t2 = 1.0f / (v.z * v.z) - 1.0f;
    return 2.0f / (1.0f + sqrtf(1.0f + a * a * t2));
}
PTB_HD float ggx_g_uncorrelated(float a, V3 wi, V3 wo)  // :227-244
{
    if (wi.z <= 0.0f || wo.z <= 0.0f) return 0.0f;
    const float x = 2.0f * wi.z * wo.z, y = 1.0f - a * a, sy = sqrtf(y);
    const float z = wo.z * hypotf(a, wi.z * sy), w = wi.z * hypotf(a, wo.z * sy);
    return x / (z + w);
}
PTB_HD V3 ggx_half_vector(float a, Rng& rng, V3 incoming, V3 normal)  // :248-284 (VNDF, Dupuy & Benyoub)
{
    const M3 A = onb(normal);
    const V3 v_ = mulmt(A, -incoming);
    const V3 v = normalizef(v_ * ex::v3(a, a, 1.0f));
    const M3 B = onb_ggx(v);
    const float u1 = rng_f32(rng), u2 = rng_f32(rng);
    const float a_ = 1.0f / (1.0f + v.z);
    const bool cond = u2 < a_;
    const float r = fminf(sqrtf(u1), 0.9999f);
    const float phi = cond ? (PI_F * u2 / a_) : (PI_F + ((u2 - a_) / (1.0f - a_)) * PI_F);
    float sn, cs;
    sincos_(phi, &sn, &cs);
    const float p1 = r * cs, p2 = r * sn * (cond ? 1.0f : v.z);
    const V3 h_ = mulm(B, ex::v3(p1, p2, sqrtf(1.0f - p1 * p1 - p2 * p2)));
    return mulm(A, normalizef(h_ * ex::v3(a, a, 1.0f)));
}
PTB_HD float dielectric_f(float cosine, float eta)  // :477-488
{
    if (eta * eta * (1.0f - cosine * cosine) > 1.0f) return 1.0f;
    const float r = (eta - 1.0f) / (eta + 1.0f);
    return schlick(cosine, r * r);
}

PTB_SHADE_FN V3 scatter_direction(const Mat& m, Rng& rng, V3 incoming, V3 normal, bool front)
{
    switch (m.kind)
    {
    case PTB_LAMBERTIAN: return mulm(onb(normal), random_cosine_vector(rng));  // :104-107
    case PTB_EMISSIVE: return ex::v3(0, 0, 0);                                 // :133
    case PTB_SPECULAR: return reflect(incoming, normal);                       // :151
    case PTB_GGX_METAL: return reflect(incoming, ggx_half_vector(m.a, rng, incoming, normal));  // :317-325
    case PTB_GGX_DIELECTRIC:                                                                   // :326-346
    {
        const V3 h = ggx_half_vector(m.a, rng, incoming, normal);
        const float eta = front ? (1.0f / m.ior) : m.ior;
        const float r = (eta - 1.0f) / (eta + 1.0f);
        const float f = schlick(-dot(incoming, h), r * r);
        const V3 refracted = refract(incoming, h, eta);
        const bool reflected = nan3(refracted) || rng_f32(rng) < f;  // short-circuit: no draw on TIR (:335)
        return reflected ? reflect(incoming, h) : refracted;
    }
    default:  // PTB_DIELECTRIC :493-506
    {
        const float eta = front ? (1.0f / m.ior) : m.ior;
        const float cosine = -dot(incoming, normal);
        if (rng_f32(rng) < dielectric_f(cosine, eta)) return reflect(incoming, normal);
        return refract(incoming, normal, eta);
    }
    }
}

struct BsdfPdf
{
    V3 bsdf;
    float pdf;
};
PTB_SHADE_FN BsdfPdf bsdf_pdf(const Mat& m, V3 incoming, V3 outgoing, V3 normal, bool front)
{
    BsdfPdf o;
    switch (m.kind)
    {
    case PTB_LAMBERTIAN:  // :109-115
        o.bsdf = m.colour * INV_PI_F;
        o.pdf = dot(outgoing, normal) * INV_PI_F;
        return o;
    case PTB_EMISSIVE:  // :134
    case PTB_SPECULAR:  // :155
        o.bsdf = m.colour;
        o.pdf = 1.0f;
        return o;
    case PTB_GGX_METAL:
    case PTB_GGX_DIELECTRIC:  // :349-450
    {
        const bool tr = m.kind == PTB_GGX_DIELECTRIC;
        const M3 A = onb(normal);
        const V3 wi = mulmt(A, outgoing), wo = mulmt(A, incoming);
        const bool transmitted = wi.z < 0.0f;
        const float eta = front ? m.ior : (1.0f / m.ior);
        V3 h;
        if (tr && transmitted)
        {
            const V3 h_ = normalizef(eta * wi + wo);
            h = h_ * ((h_.z != h_.z) ? h_.z : copysignf(1.0f, h_.z));
        }
        else h = normalizef(wi + wo);
        const float idh = dot(wi, h), odh = dot(wo, h);
        const float d = ggx_d(m.a, h);
        float f, g;
        if (!tr) { f = 1.0f; g = ggx_g_uncorrelated(m.a, wi, wo); }
        else
        {
            const float r = (eta - 1.0f) / (eta + 1.0f);
            f = schlick(fabsf(idh), r * r);
            g = ggx_g1(m.a, wi, h) * ggx_g1(m.a, wo, h);
        }
        if (transmitted)
        {
            if (!tr) { o.bsdf = ex::v3(0, 0, 0); o.pdf = 0.0f; return o; }  // BsdfPdf::invalid
            const float x = fabsf(idh * odh), y = fabsf(wi.z * wo.z);
            const float z = (1.0f - f) * g * d, w = eta * idh + odh;
            const float btdf = (x * z) / (y * w * w);
            const float jac = fabsf(odh) / (w * w);
            o.bsdf = m.colour * btdf * eta * eta;
            o.pdf = d * (1.0f - f) * fabsf(h.z) * jac;
            return o;
        }
        const float brdf = f * g * d / (4.0f * fabsf(wi.z * wo.z));
        const float jac = 1.0f / (4.0f * fabsf(odh));
        o.pdf = d * h.z * f * jac;
        const V3 one = ex::v3(1, 1, 1);
        const V3 tint = tr ? one : (m.colour + ((one - m.colour) * pow5(1.0f - fabsf(idh))));  // f_vector :207
        o.bsdf = brdf * tint;
        return o;
    }
    default:  // PTB_DIELECTRIC :508-527 (Fresnel cosine -dot(wi,wo) is the reference's own quirk, H3)
    {
        const float cosine = -dot(incoming, outgoing);
        const float eta = front ? (1.0f / m.ior) : m.ior;
        const float f = dielectric_f(cosine, eta);
        if (dot(outgoing, normal) > 0.0f) { o.bsdf = ex::v3(f, f, f); o.pdf = f; return o; }
        o.bsdf = m.colour * ((1.0f - f) / (eta * eta));
        o.pdf = 1.0f - f;
        return o;
    }
    }
}

// ---- volume.rs -------------------------------------------------------------------------------------------------------------------
PTB_HD V3 volume_transmission(const VolumeRow& v, float dist)  // :112-113
{
    return ex::v3(expf(-v.absorption[0] * dist), expf(-v.absorption[1] * dist), expf(-v.absorption[2] * dist));
}
PTB_HD bool volume_scatter(const VolumeRow& v, Rng& rng, V3 incoming, float t_max, float* t_out, V3* dir)  // :68-83, :25-50
{
    const float t = -logf(rng_f32(rng)) / v.c;
    if (t > t_max) return false;
    const float u0 = rng_f32(rng), u1 = rng_f32(rng);
    const float phi = 2.0f * PI_F * u0;
    float z;
    if (v.g == 0.0f) z = 1.0f - 2.0f * u1;
    else
    {
        const float x = (1.0f - v.g * v.g) / (1.0f + v.g * (1.0f - 2.0f * u1));
        z = (1.0f + v.g * v.g - x * x) / (2.0f * v.g);
    }
    float sn, cs;
    sincos_(phi, &sn, &cs);
    const float r = sqrtf(1.0f - z * z);
    *dir = mulm(onb(-incoming), ex::v3(r * cs, r * sn, z));
    *t_out = t;
    return true;
}

// ---- light_sampler.rs:31-37: first cdf entry >= x under total order, clamped to the last light (H3) -------------------------------
PTB_HD uint32_t light_pick(const LightRow* lights, uint32_t n, float x)
{
    uint32_t lo = 0, hi = n;
    while (lo < hi)
    {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (ex::total_lt(lights[mid].cdf, x)) lo = mid + 1;
        else hi = mid;
    }
    return lo >= n ? n - 1 : lo;
}

// ---- image_helper.rs:61-88 ----------------------------------------------------------------------------------------------------------
PTB_HD V3 env_pixel(const float* env, uint32_t w, uint32_t h, uint32_t x, uint32_t y)
{
    const float* p = env + ((size_t)(y % h) * w + (x % w)) * 3;
    return ex::v3(p[0], p[1], p[2]);
}
PTB_HD uint32_t sat_u32(float f) { return !(f > 0.0f) ? 0u : (f >= 4294967296.0f ? 0xffffffffu : (uint32_t)f); }  // Rust `as u32`
PTB_HD V3 env_bilinear(const float* env, uint32_t w, uint32_t h, float u, float v)
{
    const float x = (float)w * u, y = (float)h * v;
    const uint32_t x0 = sat_u32(x), y0 = sat_u32(y);
    const float xf = x - truncf(x), yf = y - truncf(y);
    const V3 c00 = env_pixel(env, w, h, x0, y0), c01 = env_pixel(env, w, h, x0, y0 + 1), c10 = env_pixel(env, w, h, x0 + 1, y0), c11 = env_pixel(env, w, h, x0 + 1, y0 + 1);
    return (1.0f - xf) * (1.0f - yf) * c00 + (1.0f - xf) * yf * c01 + xf * (1.0f - yf) * c10 + xf * yf * c11;
}

}  // namespace ptb
