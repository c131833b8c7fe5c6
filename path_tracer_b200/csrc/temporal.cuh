// temporal.cuh — per-pixel bodies of the reference's three post-frame compute passes, restated as CUDA device code
// (no graphics API involved):
//   accumulate.wgsl:20-23   static camera: accumulation += (rgb, 1)
//   velocity.wgsl:16-40     screen-space velocity from the first-hit world position and last frame's view-projection
//   compute.wgsl:103-213    TAA resolve: 3x3 YCoCg statistics, id-validated reprojection, Catmull-Rom history fetch,
//                           variance clip, 15 % blend
// Arithmetic: every operation is a separately rounded IEEE f32 op in the order the shader source writes it (ex:: wrappers,
// no contraction), matrix * vector = ((c0*x + c1*y) + c2*z) + c3*w, mix(a,b,t) = a*(1-t) + b*t (WGSL definition),
// min/max = minNum/maxNum (a NaN operand is ignored, -0 < +0), f32 -> i32 saturating with NaN -> 0.  WGSL leaves these
// free; they are fixed here so that the CUDA path and the CPU restatement (oracle/po_temporal.hpp) agree bit for bit.
// textureSampleLevel (linear, clamp-to-edge, state.rs:169-177) is evaluated in full f32 (the hardware unit the reference
// would use quantises the weights to 8 fractional bits; that path is unavailable on this box).
#pragma once
#include <cmath>
#include <cstdint>

#include "exact_math.h"

namespace ptb {
namespace tp {

using ex::V3;

struct F4 { float x, y, z, w; };
struct F2 { float x, y; };

PTB_HD float fmx(float a, float b)
{
#if defined(__CUDA_ARCH__)
    return fmaxf(a, b);
#else
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) return (ex::bits(a) >> 31) ? b : a;
    return a > b ? a : b;
#endif
}
PTB_HD float ffloor(float a)
{
#if defined(__CUDA_ARCH__)
    return floorf(a);
#else
    return std::floor(a);
#endif
}
PTB_HD int to_i32(float a)  // saturating, NaN -> 0
{
#if defined(__CUDA_ARCH__)
    return __float2int_rz(a);
#else
    if (a != a) return 0;
    if (a >= 2147483648.0f) return 2147483647;
    if (a <= -2147483648.0f) return (int)0x80000000;
    return (int)a;
#endif
}
PTB_HD int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

PTB_HD F4 f4(float x, float y, float z, float w) { F4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
PTB_HD F4 add4(F4 a, F4 b) { return f4(ex::add(a.x, b.x), ex::add(a.y, b.y), ex::add(a.z, b.z), ex::add(a.w, b.w)); }
PTB_HD F4 mul4(F4 a, float s) { return f4(ex::mul(a.x, s), ex::mul(a.y, s), ex::mul(a.z, s), ex::mul(a.w, s)); }
PTB_HD F4 div4(F4 a, float s) { return f4(ex::div(a.x, s), ex::div(a.y, s), ex::div(a.z, s), ex::div(a.w, s)); }
PTB_HD V3 xyz(F4 a) { return ex::v3(a.x, a.y, a.z); }

// w_divide (velocity.wgsl:11-14, compute.wgsl:11-14): xyz / max(w, 1)
// (x / 1 == x exactly: the common case of a history texel written by the resolve pass, w = 1, skips the three IEEE divisions)
PTB_HD V3 w_divide(F4 v)
{
    const float d = fmx(v.w, 1.0f);
    return d == 1.0f ? xyz(v) : ex::div(xyz(v), d);
}

// column-major mat4 * vec4
PTB_HD F4 mat4_mul(const float* m, F4 v)
{
    F4 r;
    r.x = ex::add(ex::add(ex::add(ex::mul(m[0], v.x), ex::mul(m[4], v.y)), ex::mul(m[8], v.z)), ex::mul(m[12], v.w));
    r.y = ex::add(ex::add(ex::add(ex::mul(m[1], v.x), ex::mul(m[5], v.y)), ex::mul(m[9], v.z)), ex::mul(m[13], v.w));
    r.z = ex::add(ex::add(ex::add(ex::mul(m[2], v.x), ex::mul(m[6], v.y)), ex::mul(m[10], v.z)), ex::mul(m[14], v.w));
    r.w = ex::add(ex::add(ex::add(ex::mul(m[3], v.x), ex::mul(m[7], v.y)), ex::mul(m[11], v.z)), ex::mul(m[15], v.w));
    return r;
}

// compute.wgsl:62-69 — the constructor arguments are COLUMNS in WGSL, so this is the transpose of the usual YCoCg matrix
PTB_HD V3 rgb_to_ycocg(V3 c)
{
    return ex::v3(ex::add(ex::add(ex::mul(0.25f, c.x), ex::mul(0.5f, c.y)), ex::mul(-0.25f, c.z)),
                  ex::add(ex::add(ex::mul(0.5f, c.x), ex::mul(0.0f, c.y)), ex::mul(0.5f, c.z)),
                  ex::add(ex::add(ex::mul(0.25f, c.x), ex::mul(-0.5f, c.y)), ex::mul(-0.25f, c.z)));
}
// compute.wgsl:71-78 (columns (1,1,-1) (1,0,1) (1,-1,-1))
PTB_HD V3 ycocg_to_rgb(V3 c)
{
    return ex::v3(ex::add(ex::add(ex::mul(1.0f, c.x), ex::mul(1.0f, c.y)), ex::mul(1.0f, c.z)),
                  ex::add(ex::add(ex::mul(1.0f, c.x), ex::mul(0.0f, c.y)), ex::mul(-1.0f, c.z)),
                  ex::add(ex::add(ex::mul(-1.0f, c.x), ex::mul(1.0f, c.y)), ex::mul(-1.0f, c.z)));
}

// compute.wgsl:80-99
PTB_HD V3 clip_aabb(V3 aabb_min, V3 aabb_max, V3 q)
{
    const V3 p_clip = ex::mul(ex::add(aabb_max, aabb_min), 0.5f);
    const V3 e_clip = ex::mul(ex::sub(aabb_max, aabb_min), 0.5f);
    const V3 v_clip = ex::sub(q, p_clip);
    const V3 v_unit = ex::v3(ex::div(v_clip.x, e_clip.x), ex::div(v_clip.y, e_clip.y), ex::div(v_clip.z, e_clip.z));
    const float ma_unit = fmx(fabsf(v_unit.x), fmx(fabsf(v_unit.y), fabsf(v_unit.z)));
    if (ma_unit > 1.0f) return ex::add(p_clip, ex::div(v_clip, ma_unit));
    return q;
}

// textureLoad on a W x H rgba32float image stored row-major (texel (x,y) = element y*W + x, the layout of the reference's
// `data` slices, state.rs:519-528)
// `mean`: the buffer holds (sum of n samples, n) — the library's own accumulation buffer used as the frame input — and is read
// as (sum / n, 1); with the reference's one sample per frame (n = 1) that is the stored value bit for bit.
struct Image
{
    const F4* px;
    int W, H;
    int mean;
};
PTB_HD F4 load(const Image& t, int x, int y)
{
#if defined(__CUDA_ARCH__)
    const float4 q = __ldg(reinterpret_cast<const float4*>(t.px) + (size_t)y * t.W + x);
    const F4 v = f4(q.x, q.y, q.z, q.w);
#else
    const F4 v = t.px[(size_t)y * t.W + x];
#endif
    if (t.mean && v.w > 0.0f && v.w != 1.0f) return f4(ex::div(v.x, v.w), ex::div(v.y, v.w), ex::div(v.z, v.w), 1.0f);
    return v;
}

// textureSampleLevel(tex, linear clamp-to-edge sampler, uv, 0)
PTB_HD F4 sample_linear(const Image& t, float u, float v)
{
    const float x = ex::sub(ex::mul(u, (float)t.W), 0.5f), y = ex::sub(ex::mul(v, (float)t.H), 0.5f);
    const float x0 = ffloor(x), y0 = ffloor(y);
    const float fx = ex::sub(x, x0), fy = ex::sub(y, y0);
    const int ix = to_i32(x0), iy = to_i32(y0);
    const int xa = iclamp(ix, 0, t.W - 1), xb = iclamp(ix < 2147483647 ? ix + 1 : ix, 0, t.W - 1);
    const int ya = iclamp(iy, 0, t.H - 1), yb = iclamp(iy < 2147483647 ? iy + 1 : iy, 0, t.H - 1);
    const float gx = ex::sub(1.0f, fx), gy = ex::sub(1.0f, fy);
    const F4 top = add4(mul4(load(t, xa, ya), gx), mul4(load(t, xb, ya), fx));
    const F4 bot = add4(mul4(load(t, xa, yb), gx), mul4(load(t, xb, yb), fx));
    return add4(mul4(top, gy), mul4(bot, fy));
}

// compute.wgsl:16-60
PTB_HD V3 sample_catmull_rom(const Image& t, F2 tex_size, F2 uv)
{
    const F2 sp = {ex::add(ex::mul(uv.x, tex_size.x), 0.5f), ex::add(ex::mul(uv.y, tex_size.y), 0.5f)};
    const F2 p1 = {ex::add(ffloor(ex::sub(sp.x, 0.5f)), 0.5f), ex::add(ffloor(ex::sub(sp.y, 0.5f)), 0.5f)};
    const float f[2] = {ex::sub(sp.x, p1.x), ex::sub(sp.y, p1.y)};
    float w0[2], w3[2], w12[2], o12[2];
#pragma unroll
    for (int k = 0; k < 2; ++k)
    {
        const float a = f[k];
        w0[k] = ex::mul(a, ex::add(-0.5f, ex::mul(a, ex::sub(1.0f, ex::mul(0.5f, a)))));
        const float w1 = ex::add(1.0f, ex::mul(ex::mul(a, a), ex::add(-2.5f, ex::mul(1.5f, a))));
        const float w2 = ex::mul(a, ex::add(0.5f, ex::mul(a, ex::sub(2.0f, ex::mul(1.5f, a)))));
        w3[k] = ex::mul(ex::mul(a, a), ex::add(-0.5f, ex::mul(0.5f, a)));
        w12[k] = ex::add(w1, w2);
        o12[k] = ex::div(w2, ex::add(w1, w2));
    }
    const float px[3] = {ex::div(ex::sub(p1.x, 1.0f), tex_size.x), ex::div(ex::add(p1.x, o12[0]), tex_size.x), ex::div(ex::add(p1.x, 2.0f), tex_size.x)};
    const float py[3] = {ex::div(ex::sub(p1.y, 1.0f), tex_size.y), ex::div(ex::add(p1.y, o12[1]), tex_size.y), ex::div(ex::add(p1.y, 2.0f), tex_size.y)};
    const float wx[3] = {w0[0], w12[0], w3[0]}, wy[3] = {w0[1], w12[1], w3[1]};
    V3 c = ex::v3(0.0f, 0.0f, 0.0f);
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
        for (int i = 0; i < 3; ++i) c = ex::add(c, ex::mul(ex::mul(w_divide(sample_linear(t, px[i], py[j])), wx[i]), wy[j]));
    return c;
}

// velocity.wgsl:19-39 for texel (x, y): current_uv - previous_uv
PTB_HD F2 velocity_at(const Image& position, const float* last_inv_projection, F2 dims, int x, int y)
{
    const F2 cur = {ex::div(ex::add((float)x, 0.5f), dims.x), ex::div(ex::add((float)y, 0.5f), dims.y)};
    const F4 p = load(position, x, y);
    const V3 ndc = w_divide(mat4_mul(last_inv_projection, f4(p.x, p.y, p.z, 1.0f)));
    const F2 prev = {ex::add(ex::mul(ndc.x, 0.5f), 0.5f), ex::add(ex::mul(ndc.y, 0.5f), 0.5f)};
    return F2{ex::sub(cur.x, prev.x), ex::sub(cur.y, prev.y)};
}

// compute.wgsl:103-213 for texel (x, y)
PTB_HD F4 resolve_at(const Image& input, const Image& history, const Image& position, const uint32_t* id, const float* last_inv_projection, int x, int y)
{
    const int W = input.W, H = input.H;
    const F2 dims = {(float)W, (float)H};
    const V3 current = xyz(load(input, x, y));
    V3 m1 = ex::v3(0.0f, 0.0f, 0.0f), m2 = m1;
    float closest_depth = 1e20f;
    int cx = 0, cy = 0;
    const int x_lo = x - 1 > 0 ? x - 1 : 0, y_lo = y - 1 > 0 ? y - 1 : 0;
    const int x_hi = x + 1 < W - 1 ? x + 1 : W - 1, y_hi = y + 1 < H - 1 ? y + 1 : H - 1;
    const int n = (x_hi + 1 - x_lo) * (y_hi + 1 - y_lo);
    for (int xx = x_lo; xx <= x_hi; ++xx)
        for (int yy = y_lo; yy <= y_hi; ++yy)
        {
            const F4 dd = load(input, xx, yy);
            const V3 d = rgb_to_ycocg(xyz(dd));
            m1 = ex::add(m1, d);
            m2 = ex::add(m2, ex::mul(d, d));
            if (dd.w < closest_depth) { closest_depth = dd.w; cx = xx; cy = yy; }
        }
    const F2 cur = {ex::div(ex::add((float)x, 0.5f), dims.x), ex::div(ex::add((float)y, 0.5f), dims.y)};
    const F2 vel = velocity_at(position, last_inv_projection, dims, cx, cy);
    const F2 prev = {ex::sub(cur.x, vel.x), ex::sub(cur.y, vel.y)};
    const int pcx = to_i32(ffloor(ex::mul(prev.x, dims.x))), pcy = to_i32(ffloor(ex::mul(prev.y, dims.y)));
    const bool outside = pcx < 0 || pcy < 0 || pcx >= W || pcy >= H;
    const uint32_t current_id = id[(size_t)y * W + x] & 0xffffu;
    const uint32_t old_id = outside ? 0u : ((id[(size_t)pcy * W + pcx] >> 16) & 0xffffu);
    if (current_id != old_id || outside)
    {
        const float c0x = ex::div((float)x, dims.x), c0y = ex::div((float)y, dims.y);
        const float c1x = ex::add(c0x, ex::div(1.0f, dims.x)), c1y = ex::add(c0y, ex::div(1.0f, dims.y));
        const F4 a = sample_linear(input, c0x, c0y), b = sample_linear(input, c0x, c1y);
        const F4 c = sample_linear(input, c1x, c0y), d = sample_linear(input, c1x, c1y);
        return mul4(add4(add4(add4(a, b), c), d), 0.25f);  // (a + b + c + d) / 4.0: a power-of-two divisor, the product rounds identically
    }
    const float fn = (float)n;
    const V3 mu = ex::div(m1, fn);
    const V3 var = ex::sub(ex::div(m2, fn), ex::mul(mu, mu));
    const V3 sigma = ex::v3(ex::sqrt(var.x), ex::sqrt(var.y), ex::sqrt(var.z));
    const float gamma = 1.0f;
    const V3 lo = ex::sub(mu, ex::mul(sigma, gamma)), hi = ex::add(mu, ex::mul(sigma, gamma));
    const V3 previous = sample_catmull_rom(history, dims, prev);
    const V3 clamped = ycocg_to_rgb(clip_aabb(lo, hi, rgb_to_ycocg(previous)));
    const V3 out = ex::add(ex::mul(clamped, ex::sub(1.0f, 0.15f)), ex::mul(current, 0.15f));
    return f4(out.x, out.y, out.z, 1.0f);
}

}  // namespace tp
}  // namespace ptb
