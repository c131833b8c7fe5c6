// ORACLE — TEST INFRASTRUCTURE ONLY (see po_math.hpp header).  PARITY UNPINNED.
// po_shading.hpp — sampling, camera, materials, media, lights, the path integrator and the frame driver:
//   SobolSampler    src/sampling.rs
//   Camera          src/camera.rs:17-31,94-105
//   utility         src/utility.rs
//   onb             src/tlas/tlas_bvh/blas/primitive/material/onb.rs
//   materials       src/tlas/tlas_bvh/blas/primitive/material.rs
//   media           src/tlas/tlas_bvh/blas/primitive/material/volume.rs
//   lights/scene    src/scene.rs, src/scene/light_sampler.rs
//   integrator      src/integrator.rs
//   env lookup      src/image_helper.rs:61-88
//   frame driver    src/main.rs:181-207 ; accumulate src/shaders/accumulate.wgsl:20-23
#pragma once
#include <array>
#include <memory>

#include "po_accel.hpp"

namespace po {

// ---- sampling.rs ------------------------------------------------------------------------------
constexpr uint32_t SOBOL_DIRECTIONS[32] = {                                                          // :4-8
    0x80000000, 0xc0000000, 0xa0000000, 0xf0000000, 0x88000000, 0xcc000000, 0xaa000000, 0xff000000, 0x80800000, 0xc0c00000, 0xa0a00000,
    0xf0f00000, 0x88880000, 0xcccc0000, 0xaaaa0000, 0xffff0000, 0x80008000, 0xc000c000, 0xa000a000, 0xf000f000, 0x88008800, 0xcc00cc00,
    0xaa00aa00, 0xff00ff00, 0x80808080, 0xc0c0c0c0, 0xa0a0a0a0, 0xf0f0f0f0, 0x88888888, 0xcccccccc, 0xaaaaaaaa, 0xffffffff,
};
constexpr uint32_t SOBOL_N = 512;  // NUM_POINTS = SAMPLES_PER_PIXEL*2, main.rs:47-48

inline uint32_t reverse_bits(uint32_t x)
{
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
    x = ((x >> 8) & 0x00ff00ffu) | ((x & 0x00ff00ffu) << 8);
    return (x >> 16) | (x << 16);
}
inline uint32_t sobol_dim1(uint32_t index)                                                           // :24-30
{
    uint32_t x = 0;
    for (int bit = 0; bit < 32; ++bit) x ^= ((index >> bit) & 1u) * SOBOL_DIRECTIONS[bit];
    return x;
}
inline uint32_t lk_hash(uint32_t x, uint32_t seed)                                                   // :53-68
{
    x ^= x * 0x3d20adeau;
    x += seed;
    x *= (seed >> 16) | 1u;
    x ^= x * 0x05526c56u;
    x ^= x * 0x53a22864u;
    return x;
}
inline uint32_t scramble_base2(uint32_t x, uint32_t seed) { return reverse_bits(lk_hash(reverse_bits(x), seed)); }  // :71
inline uint32_t low_bias_hash(uint32_t x)                                                            // :76-91
{
    x ^= x >> 16;
    x *= 0x21f0aaadu;
    x ^= x >> 15;
    x *= 0xd35a2d97u;
    x ^= x >> 15;
    return x;
}
struct SobolSampler
{
    uint32_t pts[SOBOL_N][2];
    SobolSampler()                                                                                   // :33-48
    {
        for (uint32_t i = 0; i < SOBOL_N; ++i) { pts[i][0] = reverse_bits(i); pts[i][1] = sobol_dim1(i); }
    }
    void get_ss_sobol(uint32_t index, uint32_t seed, float* px, float* py) const                     // :97-114
    {
        uint32_t x_seed = low_bias_hash(seed);
        uint32_t y_seed = low_bias_hash(seed + 1u);
        uint32_t shuffle_seed = low_bias_hash(seed + 2u);
        uint32_t shuffled = scramble_base2(index, shuffle_seed);
        const uint32_t* pt = pts[shuffled % SOBOL_N];
        uint32_t x = scramble_base2(pt[0], x_seed);
        uint32_t y = scramble_base2(pt[1], y_seed);
        *px = (float)x / 4294967296.0f;
        *py = (float)y / 4294967296.0f;
    }
};

// ---- camera.rs --------------------------------------------------------------------------------
// glam Mat4 helpers (column-major).  mul_vec4 = ((x*vx + y*vy) + z*vz) + w*vw.
struct M4
{
    V4 c[4];
};
inline V4 v4_scale(V4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline V4 v4_add(V4 a, V4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline V4 m4_mul_v4(const M4& m, V4 v) { return v4_add(v4_add(v4_add(v4_scale(m.c[0], v.x), v4_scale(m.c[1], v.y)), v4_scale(m.c[2], v.z)), v4_scale(m.c[3], v.w)); }
inline M4 m4_mul(const M4& a, const M4& b) { return {{m4_mul_v4(a, b.c[0]), m4_mul_v4(a, b.c[1]), m4_mul_v4(a, b.c[2]), m4_mul_v4(a, b.c[3])}}; }
inline M4 m4_from_affine(const Aff3& a)
{
    return {{{a.m.c0.x, a.m.c0.y, a.m.c0.z, 0.0f}, {a.m.c1.x, a.m.c1.y, a.m.c1.z, 0.0f}, {a.m.c2.x, a.m.c2.y, a.m.c2.z, 0.0f}, {a.t.x, a.t.y, a.t.z, 1.0f}}};
}
// General cofactor inverse (glam's exact SIMD evaluation order is not available offline; camera parity is
// tolerance-tier, SURVEY §8a a3 "negligible").
inline M4 m4_inverse(const M4& m)
{
    const float a00 = m.c[0].x, a01 = m.c[0].y, a02 = m.c[0].z, a03 = m.c[0].w;
    const float a10 = m.c[1].x, a11 = m.c[1].y, a12 = m.c[1].z, a13 = m.c[1].w;
    const float a20 = m.c[2].x, a21 = m.c[2].y, a22 = m.c[2].z, a23 = m.c[2].w;
    const float a30 = m.c[3].x, a31 = m.c[3].y, a32 = m.c[3].z, a33 = m.c[3].w;
    const float b00 = a00 * a11 - a01 * a10, b01 = a00 * a12 - a02 * a10, b02 = a00 * a13 - a03 * a10;
    const float b03 = a01 * a12 - a02 * a11, b04 = a01 * a13 - a03 * a11, b05 = a02 * a13 - a03 * a12;
    const float b06 = a20 * a31 - a21 * a30, b07 = a20 * a32 - a22 * a30, b08 = a20 * a33 - a23 * a30;
    const float b09 = a21 * a32 - a22 * a31, b10 = a21 * a33 - a23 * a31, b11 = a22 * a33 - a23 * a32;
    const float det = b00 * b11 - b01 * b10 + b02 * b09 + b03 * b08 - b04 * b07 + b05 * b06;
    const float id = 1.0f / det;
    M4 r;
    r.c[0] = {(a11 * b11 - a12 * b10 + a13 * b09) * id, (a02 * b10 - a01 * b11 - a03 * b09) * id, (a31 * b05 - a32 * b04 + a33 * b03) * id, (a22 * b04 - a21 * b05 - a23 * b03) * id};
    r.c[1] = {(a12 * b08 - a10 * b11 - a13 * b07) * id, (a00 * b11 - a02 * b08 + a03 * b07) * id, (a32 * b02 - a30 * b05 - a33 * b01) * id, (a20 * b05 - a22 * b02 + a23 * b01) * id};
    r.c[2] = {(a10 * b10 - a11 * b08 + a13 * b06) * id, (a01 * b08 - a00 * b10 - a03 * b06) * id, (a30 * b04 - a31 * b02 + a33 * b00) * id, (a21 * b02 - a20 * b04 - a23 * b00) * id};
    r.c[3] = {(a11 * b07 - a10 * b09 - a12 * b06) * id, (a00 * b09 - a01 * b07 + a02 * b06) * id, (a31 * b01 - a30 * b03 - a32 * b00) * id, (a20 * b03 - a21 * b01 + a22 * b00) * id};
    return r;
}
struct Camera
{
    Aff3 matrix;        // camera-to-world (look_at_rh(...).inverse()), :19
    M4 inv_projection;  // :20-21
    M4 ray_matrix;      // matrix * inv_projection (recomputed per create_ray call upstream, :98)
};
inline Camera camera_new(V3 origin, V3 target, float fov_deg, float aspect)                          // :17-31
{
    // glam look_at_rh(eye, center, up) = look_to_rh(eye, center-eye, up): f = normalize(dir); s = normalize(f x up); u = s x f
    V3 up = {0.0f, 1.0f, 0.0f};
    V3 f = normalize(target - origin);
    V3 s = normalize(cross(f, up));
    V3 u = cross(s, f);
    Aff3 view = {{{s.x, u.x, -f.x}, {s.y, u.y, -f.y}, {s.z, u.z, -f.z}}, {-dot3(origin, s), -dot3(origin, u), dot3(origin, f)}};
    Camera c;
    c.matrix = inverse(view);
    // perspective_infinite_rh(fov_y, aspect, z_near = 1): f = 1/tan(fov/2)
    float fov = fov_deg * (PI_F / 180.0f);  // f32::to_radians
    float ff = 1.0f / std::tan(0.5f * fov);
    M4 proj = {{{ff / aspect, 0, 0, 0}, {0, ff, 0, 0}, {0, 0, -1.0f, -1.0f}, {0, 0, -1.0f, 0}}};
    c.inv_projection = m4_inverse(proj);
    c.ray_matrix = m4_mul(m4_from_affine(c.matrix), c.inv_projection);
    return c;
}
inline Ray camera_create_ray(const Camera& c, float s, float t)                                      // :94-105
{
    V3 ndc = {s * 2.0f - 1.0f, t * 2.0f - 1.0f, 0.0f};
    // Mat4::project_point3: x_axis*x + y_axis*y + z_axis*z + w_axis, then * (1/w)
    const M4& m = c.ray_matrix;
    V4 r = v4_scale(m.c[0], ndc.x);
    r = v4_add(v4_scale(m.c[1], ndc.y), r);
    r = v4_add(v4_scale(m.c[2], ndc.z), r);
    r = v4_add(m.c[3], r);
    float rw = 1.0f / r.w;
    V3 point = {r.x * rw, r.y * rw, r.z * rw};
    V3 dir = normalize(point - c.matrix.t);
    return make_ray(c.matrix.t, dir);
}

// ---- utility.rs -------------------------------------------------------------------------------
inline V3 random_cosine_vector(WyRand& rng)                                                          // :7-19
{
    float r = std::sqrt(rng.f32());
    float z = std::sqrt(1.0f - r * r);
    float phi = TAU_F * rng.f32();
    float sn = std::sin(phi), cs = std::cos(phi);
    return {cs * r, sn * r, z};
}
inline V3 reflect(V3 i, V3 n) { return i - 2.0f * dot3(n, i) * n; }                                  // :21
inline V3 refract(V3 i, V3 n, float eta)                                                             // :23-36 (NaN = TIR)
{
    float n_dot_i = dot3(n, i);
    float k = 1.0f - eta * eta * (1.0f - n_dot_i * n_dot_i);
    if (k <= 0.0f) return splat(std::numeric_limits<float>::quiet_NaN());
    return eta * i - (eta * n_dot_i + std::sqrt(k)) * n;
}

// ---- onb.rs -----------------------------------------------------------------------------------
inline M3 generate_onb(V3 n)                                                                         // :1-7 (glam any_orthonormal_pair, Duff et al.)
{
    float sign = std::copysign(1.0f, n.z);
    float a = -1.0f / (sign + n.z);
    float b = n.x * n.y * a;
    V3 c0 = {1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x};
    V3 c1 = {b, sign + n.y * n.y * a, -n.y};
    return {c0, c1, n};
}
inline M3 generate_onb_ggx(V3 v)                                                                     // :9-27
{
    if (v.z > 0.99999f) return {{1, 0, 0}, {0, -1, 0}, {0, 0, 1}};
    V3 t1 = normalize(cross(v, V3{0, 0, 1}));
    V3 t2 = cross(t1, v);
    return {t1, t2, v};
}

// ---- volume.rs --------------------------------------------------------------------------------
struct Volume
{
    bool has_absorption = false, has_scatter = false;
    V3 absorption = {0, 0, 0};  // absorption * k (:109)
    float c = 0, g = 0;
};
inline Volume volume_new(V3 absorption, float k, float c, float g)                                   // :136-142
{
    Volume v;
    v.has_absorption = !(k == 0.0f);
    v.absorption = absorption * k;
    v.has_scatter = !(c == 0.0f);
    v.c = c;
    v.g = std::fmin(std::fmax(g, -0.999f), 0.999f);                                                  // :21
    return v;
}
inline V3 volume_transmission(const Volume& v, float dist)                                           // :112-113
{
    V3 a = -v.absorption * dist;
    return {std::exp(a.x), std::exp(a.y), std::exp(a.z)};
}
inline V3 volume_scatter_direction(const Volume& v, WyRand& rng, V3 incoming)                        // :25-50
{
    float u0 = rng.f32();
    float u1 = rng.f32();
    float phi = 2.0f * PI_F * u0;
    float z;
    if (v.g == 0.0f) z = 1.0f - 2.0f * u1;
    else
    {
        float x = (1.0f - v.g * v.g) / (1.0f + v.g * (1.0f - 2.0f * u1));
        z = (1.0f + v.g * v.g - x * x) / (2.0f * v.g);
    }
    float sn = std::sin(phi), cs = std::cos(phi);
    float r = std::sqrt(1.0f - z * z);
    return mul(generate_onb(-incoming), V3{r * cs, r * sn, z});
}
// :68-83.  Returns true with (t, direction) when a scattering event happens before t_max.
inline bool volume_scatter(const Volume& v, WyRand& rng, V3 incoming, float t_max, float* t_out, V3* dir_out)
{
    float t = -std::log(rng.f32()) / v.c;
    if (t > t_max) return false;
    *dir_out = volume_scatter_direction(v, rng, incoming);
    *t_out = t;
    return true;  // scatter_pdf (:52-66) is computed upstream but never used by the integrator
}

// ---- material.rs ------------------------------------------------------------------------------
enum : uint32_t { MAT_LAMBERTIAN = 0, MAT_EMISSIVE = 1, MAT_SPECULAR = 2, MAT_GGX_METAL = 3, MAT_GGX_DIELECTRIC = 4, MAT_DIELECTRIC = 5 };
struct Material
{
    uint32_t kind = MAT_LAMBERTIAN;
    V3 colour = {0, 0, 0};  // albedo | emitted | colour
    float a = 0;            // GGX: clamp(roughness^2, 1e-4, 0.9999) (:294,:309)
    float ior = 1;
    int volume = -1;        // index into Scene::volumes, or -1
};
struct BsdfPdf                                                                                       // :15-36
{
    V3 bsdf;
    float pdf;
};
inline bool mat_is_delta(const Material& m) { return m.kind == MAT_SPECULAR || m.kind == MAT_DIELECTRIC; }   // DELTA consts :140,:491
inline bool mat_is_emissive(const Material& m) { return m.kind == MAT_EMISSIVE; }                           // :130
inline float mat_weakening(const Material& m, V3 wo, V3 n) { return mat_is_delta(m) ? 1.0f : std::fabs(dot3(wo, n)); }  // :67-77

inline float powi5(float x) { float x2 = x * x; return x * (x2 * x2); }  // compiler-rt __powisf2 order for n=5

inline float ggx_d(float a, V3 h)                                                                    // :187-200 (tan_sq quirk kept, H3)
{
    if (h.z <= 0.0f) return 0.0f;
    float cosine_sq = h.z * h.z;
    float tan_sq = std::sqrt(1.0f - cosine_sq) / cosine_sq;
    float x = (a * a) + tan_sq;
    return a * a / (PI_F * cosine_sq * cosine_sq * x * x);
}
inline float ggx_f(float v_dot_h, float f0) { return std::fmaf(powi5(1.0f - v_dot_h), 1.0f - f0, f0); }   // :205
inline V3 ggx_f_vector(float v_dot_h, V3 f0) { return f0 + ((splat(1.0f) - f0) * powi5(1.0f - v_dot_h)); }   // :207
inline float ggx_g1(float a, V3 v, V3 h)                                                             // :210-221
{
    if (v.z * dot3(h, v) <= 0.0f) return 0.0f;
    float tan_squared = 1.0f / (v.z * v.z) - 1.0f;
    return 2.0f / (1.0f + std::sqrt(1.0f + a * a * tan_squared));
}
inline float ggx_g(float a, V3 wi, V3 wo, V3 h) { return ggx_g1(a, wi, h) * ggx_g1(a, wo, h); }      // :224
inline float ggx_g_uncorrelated(float a, V3 wi, V3 wo)                                               // :227-244
{
    if (wi.z <= 0.0f || wo.z <= 0.0f) return 0.0f;
    float a_squared = a * a;
    float x = 2.0f * wi.z * wo.z;
    float y = 1.0f - a_squared;
    float z = wo.z * std::hypot(a, wi.z * std::sqrt(y));
    float w = wi.z * std::hypot(a, wo.z * std::sqrt(y));
    return x / (z + w);
}
inline V3 ggx_half_vector(float a, WyRand& rng, V3 incoming, V3 normal)                              // :248-284
{
    M3 onb_a = generate_onb(normal);
    V3 _v = mul(transpose(onb_a), -incoming);
    V3 v = normalize(_v * V3{a, a, 1.0f});
    M3 onb_b = generate_onb_ggx(v);
    float u1 = rng.f32();
    float u2 = rng.f32();
    float _a = 1.0f / (1.0f + v.z);
    bool condition = u2 < _a;
    float r = std::fmin(std::sqrt(u1), 0.9999f);
    float phi = condition ? (PI_F * u2 / _a) : (PI_F + ((u2 - _a) / (1.0f - _a)) * PI_F);
    float sn = std::sin(phi), cs = std::cos(phi);
    float p1 = r * cs;
    float p2 = r * sn * (condition ? 1.0f : v.z);
    V3 _h = mul(onb_b, V3{p1, p2, std::sqrt(1.0f - p1 * p1 - p2 * p2)});
    return mul(onb_a, normalize(_h * V3{a, a, 1.0f}));
}
inline float dielectric_f(float cosine, float eta)                                                   // :477-488
{
    if (eta * eta * (1.0f - cosine * cosine) > 1.0f) return 1.0f;
    float r = (eta - 1.0f) / (eta + 1.0f);
    float f0 = r * r;
    return std::fmaf(powi5(1.0f - cosine), 1.0f - f0, f0);
}

inline V3 mat_scatter_direction(const Material& m, WyRand& rng, V3 incoming, V3 normal, bool front_facing)
{
    switch (m.kind)
    {
    case MAT_LAMBERTIAN: return mul(generate_onb(normal), random_cosine_vector(rng));               // :104-107
    case MAT_EMISSIVE: return {0, 0, 0};                                                             // :133
    case MAT_SPECULAR: return reflect(incoming, normal);                                             // :151
    case MAT_GGX_METAL: return reflect(incoming, ggx_half_vector(m.a, rng, incoming, normal));        // :317-325
    case MAT_GGX_DIELECTRIC:                                                                         // :326-346
    {
        V3 h = ggx_half_vector(m.a, rng, incoming, normal);
        float eta = front_facing ? (1.0f / m.ior) : m.ior;
        float r = (eta - 1.0f) / (eta + 1.0f);
        float f0 = r * r;
        float f = ggx_f(-dot3(incoming, h), f0);
        V3 refracted = refract(incoming, h, eta);
        bool reflected = is_nan(refracted) || rng.f32() < f;  // short-circuit: no draw on TIR
        return reflected ? reflect(incoming, h) : refracted;
    }
    default:                                                                                         // MAT_DIELECTRIC :493-506
    {
        float eta = front_facing ? (1.0f / m.ior) : m.ior;
        float cosine = -dot3(incoming, normal);
        if (rng.f32() < dielectric_f(cosine, eta)) return reflect(incoming, normal);
        return refract(incoming, normal, eta);
    }
    }
}

inline BsdfPdf mat_bsdf_pdf(const Material& m, V3 incoming, V3 outgoing, V3 normal, bool front_facing)
{
    switch (m.kind)
    {
    case MAT_LAMBERTIAN:                                                                             // :109-115
    {
        float cosine = dot3(outgoing, normal);
        return {m.colour * FRAC_1_PI_F, cosine * FRAC_1_PI_F};
    }
    case MAT_EMISSIVE: return {m.colour, 1.0f};                                                      // :134
    case MAT_SPECULAR: return {m.colour, 1.0f};                                                      // :155
    case MAT_GGX_METAL:
    case MAT_GGX_DIELECTRIC:                                                                         // :349-450
    {
        const bool transmissive = m.kind == MAT_GGX_DIELECTRIC;
        M3 onb_inv = transpose(generate_onb(normal));
        V3 wi = mul(onb_inv, outgoing);
        V3 wo = mul(onb_inv, incoming);
        bool ray_transmitted = wi.z < 0.0f;
        V3 h;
        if (transmissive && ray_transmitted)
        {
            float eta = front_facing ? m.ior : (1.0f / m.ior);
            V3 _h = normalize(eta * wi + wo);
            h = _h * signum(_h.z);
        }
        else h = normalize(wi + wo);
        float i_dot_h = dot3(wi, h);
        float o_dot_h = dot3(wo, h);
        float d = ggx_d(m.a, h);
        float f, g;
        if (!transmissive) { f = 1.0f; g = ggx_g_uncorrelated(m.a, wi, wo); }
        else
        {
            float eta = front_facing ? m.ior : (1.0f / m.ior);
            float r = (eta - 1.0f) / (eta + 1.0f);
            f = ggx_f(std::fabs(i_dot_h), r * r);
            g = ggx_g(m.a, wi, wo, h);
        }
        if (ray_transmitted)
        {
            if (!transmissive) return {{0, 0, 0}, 0.0f};                                             // BsdfPdf::invalid
            float eta = front_facing ? m.ior : (1.0f / m.ior);
            float x = std::fabs(i_dot_h * o_dot_h);
            float y = std::fabs(wi.z * wo.z);
            float z = (1.0f - f) * g * d;
            float w = (eta * i_dot_h) + o_dot_h;
            float btdf = (x * z) / (y * w * w);
            float ja = std::fabs(o_dot_h);
            float jacobian = ja / (w * w);
            float pdf = d * (1.0f - f) * std::fabs(h.z) * jacobian;
            return {m.colour * btdf * eta * eta, pdf};
        }
        float brdf = f * g * d / (4.0f * std::fabs(wi.z * wo.z));
        float jacobian = 1.0f / (4.0f * std::fabs(o_dot_h));
        float pdf = d * h.z * f * jacobian;
        V3 tint = transmissive ? splat(1.0f) : ggx_f_vector(std::fabs(i_dot_h), m.colour);
        return {brdf * tint, pdf};
    }
    default:                                                                                         // MAT_DIELECTRIC :508-527 (Fresnel-cosine quirk kept)
    {
        float cosine = -dot3(incoming, outgoing);
        float eta = front_facing ? (1.0f / m.ior) : m.ior;
        float f = dielectric_f(cosine, eta);
        if (dot3(outgoing, normal) > 0.0f) return {splat(f), f};
        float bsdf = (1.0f - f) / (eta * eta);
        return {m.colour * bsdf, 1.0f - f};
    }
    }
}

// ---- light_sampler.rs / scene.rs --------------------------------------------------------------
struct LightItem                                                                                     // :9-15
{
    uint32_t blas;  // lights-TLAS-local blas index
    uint32_t prim;
    float pdf;
};
struct LightSampler
{
    std::vector<LightItem> lights;
    std::vector<float> cdf;
    float max = 0;
};
struct EnvMap
{
    uint32_t w = 0, h = 0;
    std::vector<V3> data;  // linear RGB
};
struct Scene                                                                                         // scene.rs:12-17
{
    std::vector<std::unique_ptr<Blas>> blases;  // Model order (tlas.rs:31-33)
    std::vector<Material> materials;            // one per BLAS
    std::vector<Volume> volumes;
    Tlas world, lights;
    std::vector<const Blas*> world_blases, light_blases;
    LightSampler sampler;
    EnvMap env;
    SobolSampler sobol;
};
inline void light_sampler_new(Scene& s)                                                              // tlas.rs:55-64, blas.rs:203-212, light_sampler.rs:41-61
{
    std::vector<LightItem> data;
    for (uint32_t lb = 0; lb < s.light_blases.size(); ++lb)
    {
        const Blas& b = *s.light_blases[lb];
        float le_len = length(s.materials[b.material].colour);
        for (uint32_t p = 0; p < b.prims.size(); ++p) data.push_back({lb, p, tri_area(b.prims[p]) * le_len});
    }
    float mx = 0.0f;
    for (auto& d : data) mx = mx + d.pdf;  // Iterator::sum = sequential left fold from 0.0
    s.sampler.max = mx;
    float state = 0.0f;
    for (auto& d : data)
    {
        LightItem li{d.blas, d.prim, d.pdf / mx};
        s.sampler.lights.push_back(li);
        state += li.pdf;
        s.sampler.cdf.push_back(state);
    }
}
// light_sampler.rs:31-37.  binary_search_by(total_cmp): index of the first cdf entry >= x.  Upstream can index
// one past the end when x exceeds the last entry (H3); the restatement clamps (and so does the GPU path).
inline const LightItem& light_sample(const LightSampler& ls, WyRand& rng)
{
    float x = rng.f32();
    size_t lo = 0, hi = ls.cdf.size();
    while (lo < hi)
    {
        size_t mid = lo + (hi - lo) / 2;
        if (total_lt(ls.cdf[mid], x)) lo = mid + 1;
        else hi = mid;
    }
    if (lo >= ls.lights.size()) lo = ls.lights.size() - 1;
    return ls.lights[lo];
}

// ---- image_helper.rs:61-88 --------------------------------------------------------------------
inline V3 env_pixel(const EnvMap& e, uint32_t x, uint32_t y) { return e.data[(size_t)(y % e.h) * e.w + (x % e.w)]; }
inline V3 env_bilinear(const EnvMap& e, float u, float v)
{
    float x = (float)e.w * u;
    float y = (float)e.h * v;
    // Rust `as u32` saturates (negative -> 0, NaN -> 0)
    auto to_u32 = [](float f) -> uint32_t { if (!(f > 0.0f)) return 0u; if (f >= 4294967296.0f) return 0xffffffffu; return (uint32_t)f; };
    uint32_t x0 = to_u32(x), y0 = to_u32(y);
    float xf = x - std::trunc(x), yf = y - std::trunc(y);
    V3 c00 = env_pixel(e, x0, y0), c01 = env_pixel(e, x0, y0 + 1), c10 = env_pixel(e, x0 + 1, y0), c11 = env_pixel(e, x0 + 1, y0 + 1);
    return (1.0f - xf) * (1.0f - yf) * c00 + (1.0f - xf) * yf * c01 + xf * (1.0f - yf) * c10 + xf * yf * c11;
}

// ---- integrator.rs ----------------------------------------------------------------------------
struct RayCounters
{
    uint64_t extend = 0, shadow = 0, probe = 0, paths = 0, bounces = 0;
    Counters world;   // work of world-TLAS closest-hit traversals (extend)
    Counters shadow_w; // work of world-TLAS any-hit traversals (shadow)
    Counters lights;  // work of lights-TLAS probes
};
inline float mis_heuristic2(float f, float g) { return (f * f) / ((f * f) + (g * g)); }              // :22, HEURISTIC_POWER = 2

inline V3 estimate_direct_explicit(WyRand& rng, const Ray& in_ray, const HitInfo& hi, const Material& mat, const Scene& sc, RayCounters& rc)  // :25-74
{
    V3 incoming = -in_ray.d;
    const LightItem& li = light_sample(sc.sampler, rng);                                             // scene.rs:37-45
    const Blas& lb = *sc.light_blases[li.blas];
    const Material& lmat = sc.materials[lb.material];
    const Triangle& light = lb.prims[li.prim];
    // primitive.rs:77-91 random_point (BLAS-local space; instance matrices ignored upstream, H3)
    float u = rng.f32();
    float v = rng.f32();
    if (u + v > 1.0f) { u = 1.0f - u; v = 1.0f - v; }
    V3 point = tri_position(light, u, v);
    V3 light_normal = tri_normal(light, u, v);
    V3 o = ray_at(in_ray, hi.t);
    V3 d = point - o;
    float distance_squared = length_squared(d);
    float distance = std::sqrt(distance_squared);
    Ray out = make_ray(o, normalize(d));
    if (dot3(out.d, hi.normal) > 0.0f)
    {
        rc.shadow += 1;
        if (!tlas_any_intersect(sc.world, sc.world_blases, out, (1.0f - EPSILON) * distance, &rc.shadow_w))
        {
            BsdfPdf bp = mat_bsdf_pdf(mat, incoming, out.d, hi.normal, hi.front_facing);
            float sample_pdf = li.pdf / tri_area(light);
            float cosine = std::fabs(dot3(out.d, light_normal));
            float light_pdf = sample_pdf * (distance_squared / cosine);
            float weight = mis_heuristic2(light_pdf, bp.pdf);
            return lmat.colour * weight * mat_weakening(mat, out.d, hi.normal) * bp.bsdf / light_pdf;
        }
    }
    return {0, 0, 0};
}
inline V3 estimate_direct_bsdf(WyRand& rng, const Ray& in_ray, const HitInfo& hi, const Material& mat, const Scene& sc, RayCounters& rc)  // :77-130
{
    V3 incoming = -in_ray.d;
    Ray out = make_ray(ray_at(in_ray, hi.t), mat_scatter_direction(mat, rng, in_ray.d, hi.normal, hi.front_facing));
    if (dot3(out.d, hi.normal) > 0.0f)
    {
        TlasHit lh;
        rc.probe += 1;
        if (tlas_intersect(sc.lights, sc.light_blases, out, INF, &lh, &rc.lights))
        {
            rc.shadow += 1;
            if (!tlas_any_intersect(sc.world, sc.world_blases, out, lh.hi.t * (1.0f - EPSILON), &rc.shadow_w))
            {
                BsdfPdf bp = mat_bsdf_pdf(mat, incoming, out.d, hi.normal, hi.front_facing);
                if (bp.pdf > 0.0f)  // MIN_PDF = 0 (:10)
                {
                    const Blas& lb = *sc.light_blases[lh.blas];
                    const Material& lmat = sc.materials[lb.material];
                    const Triangle& light = lb.prims[lh.prim];
                    float sample_pdf = (tri_area(light) * length(lmat.colour) / sc.sampler.max) / tri_area(light);  // light_sampler.rs:39
                    float cosine = std::fabs(dot3(out.d, lh.hi.normal));
                    float light_pdf = sample_pdf * (lh.hi.t * lh.hi.t / cosine);
                    float weight = mis_heuristic2(bp.pdf, light_pdf);
                    return lmat.colour * weight * mat_weakening(mat, out.d, hi.normal) * bp.bsdf / bp.pdf;
                }
            }
        }
    }
    return {0, 0, 0};
}

struct PathResult
{
    V4 colour;    // (rgb, 1)
    V4 position;  // (first hit point, t)
    uint8_t first_id;
};
inline PathResult integrate(Ray r, const Scene& sc, WyRand& rng, uint32_t max_bounces, bool enable_nee, RayCounters& rc)  // :143-281
{
    V3 accumulated = {0, 0, 0};
    V3 path_weight = {1, 1, 1};
    V3 p0 = ray_at(r, 1e5f);
    V4 position = {p0.x, p0.y, p0.z, 1e5f};
    uint8_t first_id = 255;
    bool last_delta = false;
    uint64_t volume_set = 0;  // IntSet<&Volume> (:161) as a bitmask over Scene::volumes; iterated in index order
    rc.paths += 1;
    for (uint64_t b = 0; b <= (uint64_t)max_bounces; ++b)
    {
        if (b > 3)                                                                                   // :166-177
        {
            float survive = std::fmin(max_element(path_weight), 0.9999f);
            if (rng.f32() > survive) break;
            path_weight = path_weight / survive;
        }
        TlasHit th;
        rc.extend += 1;
        rc.bounces += 1;
        if (tlas_intersect(sc.world, sc.world_blases, r, INF, &th, &rc.world))                       // :179
        {
            const HitInfo& hi = th.hi;
            const Material& mat = sc.materials[sc.world_blases[th.blas]->material];
            if (b == 0)
            {
                V3 p = ray_at(r, hi.t);
                position = {p.x, p.y, p.z, hi.t};
                first_id = (uint8_t)th.blas;                                                         // id.index() as u8 (:184)
            }
            V3 wi = -r.d;
            // media (:189-205)
            bool scattered = false;
            float best_t = 0;
            V3 best_dir = {0, 0, 0};
            for (uint32_t vi = 0; vi < sc.volumes.size(); ++vi)
            {
                if (!((volume_set >> vi) & 1) || !sc.volumes[vi].has_scatter) continue;
                float t;
                V3 dir;
                if (volume_scatter(sc.volumes[vi], rng, r.d, hi.t, &t, &dir))
                {
                    if (!scattered || total_lt(t, best_t)) { best_t = t; best_dir = dir; scattered = true; }
                }
            }
            V3 trans = {1, 1, 1};
            for (uint32_t vi = 0; vi < sc.volumes.size(); ++vi)
                if (((volume_set >> vi) & 1) && sc.volumes[vi].has_absorption) trans = trans * volume_transmission(sc.volumes[vi], scattered ? best_t : hi.t);
            path_weight *= trans;
            if (scattered)
            {
                last_delta = true;
                r = make_ray(ray_at(r, best_t), best_dir);
                continue;
            }
            if (mat_is_emissive(mat))                                                                // :207-214
            {
                if (!enable_nee || last_delta || b == 0) accumulated = mul_add(mat.colour, path_weight, accumulated);
                break;
            }
            if (mat.volume >= 0)                                                                     // :217-227
            {
                if (hi.front_facing) volume_set |= (1ull << mat.volume);
                else volume_set &= ~(1ull << mat.volume);
            }
            bool is_delta = mat_is_delta(mat);
            if (enable_nee && !is_delta)                                                             // :231-234
            {
                V3 e = estimate_direct_explicit(rng, r, hi, mat, sc, rc);
                V3 s = estimate_direct_bsdf(rng, r, hi, mat, sc, rc);
                accumulated += path_weight * (e + s);
            }
            Ray in_ray = r;
            r = make_ray(ray_at(in_ray, hi.t), mat_scatter_direction(mat, rng, in_ray.d, hi.normal, hi.front_facing));  // :236-239
            BsdfPdf mi = mat_bsdf_pdf(mat, wi, r.d, hi.normal, hi.front_facing);
            if (mi.pdf < 0.0f) break;                                                                // :243-247
            path_weight *= mat_weakening(mat, r.d, hi.normal) * mi.bsdf / mi.pdf;                    // :249
            last_delta = is_delta;
        }
        else
        {
            if (sc.env.w > 0)                                                                        // :256-262
            {
                float u = std::fmaf(std::atan2(r.d.x, r.d.z), FRAC_1_PI_F * 0.5f, 0.5f);
                float v = std::fmaf(std::asin(r.d.y), -FRAC_1_PI_F, 0.5f);
                accumulated += env_bilinear(sc.env, u, v) * path_weight;
            }
            else accumulated += V3{0.006f, 0.006f, 0.006f} * path_weight;                            // :265
            break;
        }
    }
    PathResult out;
    out.position = position;
    out.first_id = first_id;
    if (is_finite(accumulated))                                                                      // :272-280
    {
        // glam clamp_length_max(100): if len^2 > max^2 { v * (len^2.sqrt().recip() * max) }
        float l2 = length_squared(accumulated);
        V3 c = accumulated;
        if (l2 > 100.0f * 100.0f) c = accumulated * ((1.0f / std::sqrt(l2)) * 100.0f);
        out.colour = {c.x, c.y, c.z, 1.0f};
    }
    else out.colour = {0, 0, 0, 1.0f};
    return out;
}

}  // namespace po
