// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under path_tracer_b200/ may include, link or call this.
//
// PARITY UNPINNED: the reference's WGSL passes need a wgpu device (unavailable on this box) and the reference ships no
// expected images; this is a plain C++ reading of the shader text.
//
// po_temporal.hpp — State::update (state.rs:505-579) on host arrays: accumulate.wgsl, velocity.wgsl, compute.wgsl executed
// texel by texel with WGSL-like value types.  Conventions WGSL leaves open, fixed here (and in the CUDA path):
//   * each operator is one IEEE f32 operation (file compiled with -ffp-contract=off); mat * vec sums columns left to right;
//   * mix(a, b, t) = a * (1 - t) + b * t;  max(a, b) ignores a NaN operand;  i32(f32) saturates, NaN -> 0;
//   * textureSampleLevel with the linear clamp-to-edge sampler (state.rs:169-177) is an exact f32 bilinear blend:
//     lerp along x for both rows, then along y, weights (1 - f) and f;
//   * textureLoad outside the texture (possible for t_id at previous_coords, compute.wgsl:159) is never used: the same
//     condition forces the fallback branch.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace po {
namespace wgsl {

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct ivec2 { int x, y; };

inline vec2 operator+(vec2 a, vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline vec2 operator-(vec2 a, vec2 b) { return {a.x - b.x, a.y - b.y}; }
inline vec2 operator*(vec2 a, vec2 b) { return {a.x * b.x, a.y * b.y}; }
inline vec2 operator/(vec2 a, vec2 b) { return {a.x / b.x, a.y / b.y}; }
inline vec2 operator+(vec2 a, float s) { return {a.x + s, a.y + s}; }
inline vec2 operator-(vec2 a, float s) { return {a.x - s, a.y - s}; }
inline vec2 operator*(vec2 a, float s) { return {a.x * s, a.y * s}; }
inline vec2 operator*(float s, vec2 a) { return {s * a.x, s * a.y}; }
inline vec2 operator+(float s, vec2 a) { return {s + a.x, s + a.y}; }
inline vec2 operator-(float s, vec2 a) { return {s - a.x, s - a.y}; }
inline vec2 floor(vec2 a) { return {std::floor(a.x), std::floor(a.y)}; }

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 sqrt(vec3 a) { return {std::sqrt(a.x), std::sqrt(a.y), std::sqrt(a.z)}; }
inline vec3 abs(vec3 a) { return {std::fabs(a.x), std::fabs(a.y), std::fabs(a.z)}; }

inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline vec4 operator/(vec4 a, float s) { return {a.x / s, a.y / s, a.z / s, a.w / s}; }

inline float max_(float a, float b)
{
    if (std::isnan(a)) return b;
    if (std::isnan(b)) return a;
    if (a == b) return std::signbit(a) ? b : a;
    return a > b ? a : b;
}
inline int to_i32(float f)
{
    if (std::isnan(f)) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return -2147483647 - 1;
    return (int)f;
}
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }

struct mat3 { vec3 c0, c1, c2; };  // mat3x3<f32>(nine scalars): consecutive triples are columns
inline vec3 operator*(const mat3& m, vec3 v) { return m.c0 * v.x + m.c1 * v.y + m.c2 * v.z; }
struct mat4 { vec4 c0, c1, c2, c3; };
inline vec4 operator*(const mat4& m, vec4 v) { return m.c0 * v.x + m.c1 * v.y + m.c2 * v.z + m.c3 * v.w; }

// a W x H rgba32float texture over a row-major array (write_texture with bytes_per_row = 16 W, state.rs:519-528)
struct Texture
{
    const float* data;
    int W, H;
    vec4 load(ivec2 c) const
    {
        const float* p = data + ((size_t)c.y * W + c.x) * 4;
        return {p[0], p[1], p[2], p[3]};
    }
    vec4 sample(vec2 uv) const  // textureSampleLevel(.., linear clamp-to-edge, uv, 0.0)
    {
        const float x = uv.x * (float)W - 0.5f, y = uv.y * (float)H - 0.5f;
        const float x0 = std::floor(x), y0 = std::floor(y);
        const float fx = x - x0, fy = y - y0;
        auto cl = [](long long v, int hi) { return (int)(v < 0 ? 0 : (v > hi ? hi : v)); };
        const long long ix = to_i32(x0), iy = to_i32(y0);
        const int xa = cl(ix, W - 1), xb = cl(ix == 2147483647 ? ix : ix + 1, W - 1), ya = cl(iy, H - 1), yb = cl(iy == 2147483647 ? iy : iy + 1, H - 1);
        const vec4 top = load({xa, ya}) * (1.0f - fx) + load({xb, ya}) * fx;
        const vec4 bot = load({xa, yb}) * (1.0f - fx) + load({xb, yb}) * fx;
        return top * (1.0f - fy) + bot * fy;
    }
};

inline vec3 w_divide(vec4 v) { return vec3{v.x, v.y, v.z} / max_(v.w, 1.0f); }  // velocity.wgsl:11-14, compute.wgsl:11-14

inline vec3 sample_catmull_rom(const Texture& tex, vec2 tex_size, vec2 uv)  // compute.wgsl:16-60
{
    const vec2 samplePos = (uv * tex_size) + 0.5f;
    const vec2 texPos1 = floor(samplePos - 0.5f) + 0.5f;
    const vec2 f = samplePos - texPos1;
    const vec2 w0 = f * (-0.5f + f * (1.0f - 0.5f * f));
    const vec2 w1 = 1.0f + f * f * (-2.5f + 1.5f * f);
    const vec2 w2 = f * (0.5f + f * (2.0f - 1.5f * f));
    const vec2 w3 = f * f * (-0.5f + 0.5f * f);
    const vec2 w12 = w1 + w2;
    const vec2 offset12 = w2 / (w1 + w2);
    const vec2 texPos0 = (texPos1 - 1.0f) / tex_size;
    const vec2 texPos3 = (texPos1 + 2.0f) / tex_size;
    const vec2 texPos12 = (texPos1 + offset12) / tex_size;
    vec3 c = {0.0f, 0.0f, 0.0f};
    c = c + w_divide(tex.sample({texPos0.x, texPos0.y})) * w0.x * w0.y;
    c = c + w_divide(tex.sample({texPos12.x, texPos0.y})) * w12.x * w0.y;
    c = c + w_divide(tex.sample({texPos3.x, texPos0.y})) * w3.x * w0.y;
    c = c + w_divide(tex.sample({texPos0.x, texPos12.y})) * w0.x * w12.y;
    c = c + w_divide(tex.sample({texPos12.x, texPos12.y})) * w12.x * w12.y;
    c = c + w_divide(tex.sample({texPos3.x, texPos12.y})) * w3.x * w12.y;
    c = c + w_divide(tex.sample({texPos0.x, texPos3.y})) * w0.x * w3.y;
    c = c + w_divide(tex.sample({texPos12.x, texPos3.y})) * w12.x * w3.y;
    c = c + w_divide(tex.sample({texPos3.x, texPos3.y})) * w3.x * w3.y;
    return c;
}

inline vec3 rgb_to_ycocg(vec3 in) { return mat3{{0.25f, 0.5f, 0.25f}, {0.5f, 0.0f, -0.5f}, {-0.25f, 0.5f, -0.25f}} * in; }  // compute.wgsl:62-69
inline vec3 ycocg_to_rgb(vec3 in) { return mat3{{1.0f, 1.0f, -1.0f}, {1.0f, 0.0f, 1.0f}, {1.0f, -1.0f, -1.0f}} * in; }     // compute.wgsl:71-78

inline vec3 clip_aabb(vec3 aabb_min, vec3 aabb_max, vec3 q)  // compute.wgsl:80-99
{
    const vec3 p_clip = 0.5f * (aabb_max + aabb_min);
    const vec3 e_clip = 0.5f * (aabb_max - aabb_min);
    const vec3 v_clip = q - p_clip;
    const vec3 v_unit = v_clip / e_clip;
    const vec3 a_unit = abs(v_unit);
    const float ma_unit = max_(a_unit.x, max_(a_unit.y, a_unit.z));
    if (ma_unit > 1.0f) return p_clip + v_clip / ma_unit;
    return q;
}

}  // namespace wgsl

// One State::update.  accumulation[W*H*4] is the accumulation texture (in: previous, out: new, state.rs:577);
// velocity[W*H*2] is the rg32float texture (written only when the camera moved).  Returns 1 if the TAA path ran.
inline int state_update(int W, int H, const float* data, const float* position, const uint32_t* id, const float* inv_projection, const float* last_inv_projection,
                        float* accumulation, float* velocity)
{
    using namespace wgsl;
    bool same = true;
    for (int k = 0; k < 16; ++k) same = same && (inv_projection[k] == last_inv_projection[k]);  // glam Mat4 PartialEq, state.rs:561
    const size_t npix = (size_t)W * H;
    std::vector<float> output(npix * 4);
    const Texture t_input{data, W, H}, t_accumulation{accumulation, W, H}, t_position{position, W, H};
    if (same)
    {
        for (size_t i = 0; i < npix; ++i)  // accumulate.wgsl:20-23
        {
            output[i * 4 + 0] = accumulation[i * 4 + 0] + data[i * 4 + 0];
            output[i * 4 + 1] = accumulation[i * 4 + 1] + data[i * 4 + 1];
            output[i * 4 + 2] = accumulation[i * 4 + 2] + data[i * 4 + 2];
            output[i * 4 + 3] = accumulation[i * 4 + 3] + 1.0f;
        }
    }
    else
    {
        const float* m = last_inv_projection;
        const mat4 last{{m[0], m[1], m[2], m[3]}, {m[4], m[5], m[6], m[7]}, {m[8], m[9], m[10], m[11]}, {m[12], m[13], m[14], m[15]}};
        const vec2 dimensions = {(float)W, (float)H};
        // velocity.wgsl:16-40
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
            {
                const vec2 current_uv = (vec2{(float)x, (float)y} + 0.5f) / dimensions;
                const vec4 p = t_position.load({x, y});
                const vec3 nd = w_divide(last * vec4{p.x, p.y, p.z, 1.0f});
                const vec2 previous_uv = vec2{nd.x, nd.y} * 0.5f + 0.5f;
                const vec2 v = current_uv - previous_uv;
                velocity[((size_t)y * W + x) * 2] = v.x;
                velocity[((size_t)y * W + x) * 2 + 1] = v.y;
            }
        // compute.wgsl:103-213
        for (int cy = 0; cy < H; ++cy)
            for (int cx = 0; cx < W; ++cx)
            {
                const vec4 cur4 = t_input.load({cx, cy});
                const vec3 current_diffuse = {cur4.x, cur4.y, cur4.z};
                vec3 m1 = {0, 0, 0}, m2 = {0, 0, 0};
                float closest_depth = 1e20f;
                ivec2 closest = {0, 0};
                const ivec2 min_bounds = {cx - 1 > 0 ? cx - 1 : 0, cy - 1 > 0 ? cy - 1 : 0};
                const ivec2 max_bounds = {cx + 1 < W - 1 ? cx + 1 : W - 1, cy + 1 < H - 1 ? cy + 1 : H - 1};
                const int n = (max_bounds.x + 1 - min_bounds.x) * (max_bounds.y + 1 - min_bounds.y);
                for (int x = min_bounds.x; x <= max_bounds.x; x++)
                    for (int y = min_bounds.y; y <= max_bounds.y; y++)
                    {
                        const vec4 dd = t_input.load({x, y});
                        const vec3 diffuse = rgb_to_ycocg({dd.x, dd.y, dd.z});
                        m1 = m1 + diffuse;
                        m2 = m2 + diffuse * diffuse;
                        if (dd.w < closest_depth) { closest_depth = dd.w; closest = {x, y}; }
                    }
                const vec2 current_uv = (vec2{(float)cx, (float)cy} + 0.5f) / dimensions;
                const vec2 vel = {velocity[((size_t)closest.y * W + closest.x) * 2], velocity[((size_t)closest.y * W + closest.x) * 2 + 1]};
                const vec2 previous_uv = current_uv - vel;
                const vec2 pf = floor(previous_uv * dimensions);
                const ivec2 previous_coords = {to_i32(pf.x), to_i32(pf.y)};
                const bool outside = previous_coords.x < 0 || previous_coords.y < 0 || previous_coords.x >= W || previous_coords.y >= H;
                const uint32_t current_id = id[(size_t)cy * W + cx] & 0xFFFFu;
                const uint32_t old_id = outside ? 0u : ((id[(size_t)previous_coords.y * W + previous_coords.x] >> 16) & 0xFFFFu);
                float* o = &output[((size_t)cy * W + cx) * 4];
                if (current_id != old_id || outside)
                {
                    const vec2 c0 = vec2{(float)cx, (float)cy} / dimensions;
                    const vec2 c1 = c0 + (vec2{1.0f, 1.0f} / dimensions);
                    const vec4 a = t_input.sample(c0), b = t_input.sample({c0.x, c1.y}), c = t_input.sample({c1.x, c0.y}), d = t_input.sample(c1);
                    const vec4 r = (a + b + c + d) / 4.0f;
                    o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = r.w;
                }
                else
                {
                    const vec3 mu = m1 / (float)n;
                    const vec3 sigma = sqrt(m2 / (float)n - mu * mu);
                    const float gamma = 1.0f;
                    const vec3 minimum = mu - gamma * sigma, maximum = mu + gamma * sigma;
                    const vec3 previous_diffuse = sample_catmull_rom(t_accumulation, dimensions, previous_uv);
                    const vec3 clamped = ycocg_to_rgb(clip_aabb(minimum, maximum, rgb_to_ycocg(previous_diffuse)));
                    const vec3 out = mix(clamped, current_diffuse, 0.15f);
                    o[0] = out.x; o[1] = out.y; o[2] = out.z; o[3] = 1.0f;
                }
            }
    }
    for (size_t i = 0; i < npix * 4; ++i) accumulation[i] = output[i];  // copy_texture_to_texture, state.rs:577
    return same ? 0 : 1;
}

}  // namespace po
