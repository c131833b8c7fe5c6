// wavefront.h — the per-slot bodies of the wavefront stages (host+device), restating integrate() (integrator.rs:143-281)
// and the frame loop (main.rs:181-207) as a state machine over per-slot path state:
//
//   regen  : finish the previous sample of a slot (integrator.rs:272-280, accumulate.wgsl:23) and start the next one
//            (main.rs:187-199: seed draw, Sobol jitter, Camera::create_ray)
//   extend : scene.world.intersect (integrator.rs:179)
//   shade  : everything between two intersect calls: b==0 bookkeeping, media, emission, volume set, NEE set-up,
//            continuation, path weight, Russian roulette of the next bounce
//   shadow : explicit light sample visibility + contribution (integrator.rs:56-70)
//   probe  : BSDF-sampled direct light: lights-TLAS closest hit, world any-hit, contribution (integrator.rs:100-123)
//
// A pixel owns K slots (K = 1 in the host emulation); slot k of a pixel runs the samples k, k+K, k+2K, ... of that pixel back
// to back (path regeneration) and sums them into its own accumulator, so every accumulator has a single writer and needs
// no atomics; the K partial sums are folded in a fixed order afterwards (deterministic).  Queues between stages are handled by the callers (render.cu on the device, warp-aggregated;
// tests/hostemu sequentially on the host): the bodies only return flags.
//
// Path state lives in HBM as 32-byte records (one DRAM sector each, two 128-bit accesses): a stage that touches a record
// uses the whole sector, whatever order the queue presents the slots in.
#pragma once
#include "../../include/ptb.h"
#include "shading.h"
#include "traverse.h"

namespace ptb {

struct alignas(32) RayRec     // current ray of the path (ray.rs:1-8 without inv_direction, which is recomputed)
{
    float4 o;  // origin.xyz, w = bits: bounce (low 24) | last_delta << 24 | has_sample << 25
    float4 d;  // direction.xyz, w = bits: samples started so far in this ptb_render call
};
struct alignas(32) PathRec    // integrator.rs:153-154
{
    float4 weight;    // path_weight.xyz
    float4 radiance;  // accumulated.xyz
};
struct alignas(32) HitRec     // HitInfo (model.rs:5-11) as the shade stage needs it + the path's RNG state
{
    float4 tn;        // t, world-space ray-facing normal.xyz
    uint32_t info;    // InstanceRow index | front_facing << 31 ; 0xffffffff = miss
    uint32_t rng_lo, rng_hi;
    uint32_t kind;    // shade kind of this hit (material kind 0..5, Q_MISS), written by k_extend for k_shade_ordered's span-local sort
};
struct alignas(32) ShadowRec  // explicit light sample awaiting its visibility test (integrator.rs:56)
{
    float4 d;  // direction.xyz, w = t_max
    float4 c;  // contribution if unoccluded
};
struct alignas(32) ProbeRec   // BSDF-sampled light direction awaiting lights-TLAS probe + visibility (integrator.rs:100-103)
{
    float4 d;  // direction.xyz, w = bsdf pdf
    float4 f;  // path_weight * weakening * bsdf / pdf
};

struct RenderView
{
    SceneDev S;
    CameraDev cam;
    uint32_t W, H;
    uint32_t npix;           // W * H
    uint32_t K;              // slots per pixel: slot = k * npix + j renders the samples k, k + K, k + 2K, ... of pixel slot_pixel(j)
    uint32_t tile_nodes;     // tile lists also name the BLAS nodes a tile's frustum stays within (option tile_nodes)
    uint32_t tile_settle;    // camera rays of a tile whose list of TLAS leaves is empty are settled in the regeneration loop (option tile_settle)
    uint32_t tiles_x;        // > 0: slots are laid over the frame in tiles of 8 x 4 pixels (W / 8 tiles per row), so that the 32 rays of a warp
                             // cover a compact patch of the image instead of a 32 x 1 line; 0: j = pixel (frame sizes that do not divide)
    uint32_t sample_begin, sample_count;
    unsigned long long seed;
    uint32_t max_bounces;
    uint32_t enable_nee;
    uint32_t extend_kernel;  // 0 wide, 1 reference-order
    // per slot: partial outputs of the slot's own samples, folded into the per-pixel arrays by the combine step
    float4* accum;     // (sum rgb, count)                                   accumulate.wgsl:23
    float4* accum_sq;  // optional
    float4* position;  // of the slot's latest sample                        main.rs:205
    uint32_t* id;      // (id << 16) | first_id over the slot's samples      main.rs:206
    // per slot: path state
    RayRec* ray;
    PathRec* path;
    HitRec* hit;
    ShadowRec* shadow;  // origin of shadow and probe rays = ray[slot].o (the hit point, written by shade)
    ProbeRec* probe;
    unsigned long long* volset;
};

enum : uint32_t { WF_NEXT = 1u, WF_TERM = 2u, WF_SHADOW = 4u, WF_PROBE = 8u,
                  WF_PROBE_OUT = 16u };  // a BSDF-sampled light direction that cannot reach the bounds of the lights: counted as a probe, nothing to trace
constexpr uint32_t ST_LAST_DELTA = 1u << 24, ST_HAS_SAMPLE = 1u << 25, ST_BOUNCE_MASK = (1u << 24) - 1u;
constexpr uint32_t NEE_SHADOW_BIT = 0x80000000u, NEE_PROBE_BIT = 0x40000000u, NEE_SLOT_MASK = 0x3fffffffu;  // nee queue entry = slot | flags
constexpr uint32_t Q_MISS = 6;  // shade queue kinds: 0..5 = material kinds, 6 = miss

// slot index within a layer -> pixel, and back (bijections; identity when tiles_x == 0)
PTB_HD uint32_t slot_pixel(const RenderView& V, uint32_t j)
{
    if (V.tiles_x == 0u) return j;
    const uint32_t tile = j >> 5, in = j & 31u;
    const uint32_t x = (tile % V.tiles_x) * 8u + (in & 7u), y = (tile / V.tiles_x) * 4u + (in >> 3);
    return y * V.W + x;
}
PTB_HD uint32_t pixel_slot(const RenderView& V, uint32_t p)
{
    if (V.tiles_x == 0u) return p;
    const uint32_t x = p % V.W, y = p / V.W;
    return (((y >> 2) * V.tiles_x + (x >> 3)) << 5) | ((y & 3u) << 3) | (x & 7u);
}

PTB_HD float4 f4(float x, float y, float z, float w)
{
    float4 r;
    r.x = x; r.y = y; r.z = z; r.w = w;
    return r;
}

// ---- candidate list of one 8 x 4 pixel tile (k_tile_lists, render.cu; host-callable so that the culling can be tested on the CPU) -------------
// Walks the world TLAS with the tile's frustum — the four rays through the corners of its pixel area, jitter and a margin included — and
// writes the TLAS leaves the frustum touches, nearest first, each with a lower bound of the distance at which a ray can enter it.  Inside
// an instance the walk goes on in BLAS space while the frustum stays within one to three nodes: those nodes follow the instance's entry,
// and a camera ray of the tile starts its BLAS walk there instead of at the BLAS root (an instance none of whose nodes the frustum
// touches is dropped).
// out: TILE_LIST_STRIDE entries, [0] = (number of entries that follow | TILE_LIST_NONE when the tile sees more than the list holds, 0);
// an instance entry = (leaf-header index | k << TILE_NODES_SHIFT, distance bound), followed by its k node entries (wide node index, 0),
// nearest first; k = 0: the BLAS root.
constexpr uint32_t TILE_LIST_CAP = TILE_LIST_STRIDE - 1, TILE_NODES_SHIFT = 28, TILE_NODES_MAX = 3;
struct TileFrustum
{
    V3 e;     // apex
    V3 n[4];  // inward unit normals of the four side planes
};
PTB_HD TileFrustum tile_frustum(V3 e, const V3* c)  // c: the four corner directions in order around the tile
{
    TileFrustum F;
    F.e = e;
    for (int k = 0; k < 4; ++k)
    {
        V3 m = crossf(c[k], c[(k + 1) & 3]);
        m = m * (1.0f / sqrtf(dot(m, m)));
        if (dot(m, c[(k + 2) & 3]) < 0.0f) m = -m;
        F.n[k] = m;
    }
    return F;
}
// false: every corner of the box lies outside one side plane (margin: 1e-4 of the box's distance scale).  *dmin: lower bound of the
// distance from the apex to the box
PTB_HD bool frustum_touches(const TileFrustum& F, V3 lo, V3 hi, float* dmin)
{
    const V3 a = lo - F.e, b = hi - F.e;
    const float R = fmaxf(fmaxf(fmaxf(fabsf(a.x), fabsf(b.x)), fmaxf(fabsf(a.y), fabsf(b.y))), fmaxf(fabsf(a.z), fabsf(b.z)));
    for (int k = 0; k < 4; ++k)
    {
        const float s = fmaxf(F.n[k].x * a.x, F.n[k].x * b.x) + fmaxf(F.n[k].y * a.y, F.n[k].y * b.y) + fmaxf(F.n[k].z * a.z, F.n[k].z * b.z);
        if (s < -1e-4f * R) return false;
    }
    const float dx = fmaxf(fmaxf(a.x, -b.x), 0.0f), dy = fmaxf(fmaxf(a.y, -b.y), 0.0f), dz = fmaxf(fmaxf(a.z, -b.z), 0.0f);
    *dmin = fmaxf(sqrtf(dx * dx + dy * dy + dz * dz) * 0.9999f - 1e-4f * R, 0.0f);
    return true;
}
// one 8-wide node: its words, and the quantised box of child slot s (a superset of the child's own) with one more step of slack
struct WideWords
{
    uint4 q0, q1, q2, q3, q4;
};
PTB_HD WideWords wide_words(const SceneDev& S, uint32_t node)
{
    const uint4* np = S.wide + (size_t)node * 5;
    WideWords w;
    w.q0 = PTB_LDG4U(np); w.q1 = PTB_LDG4U(np + 1); w.q2 = PTB_LDG4U(np + 2); w.q3 = PTB_LDG4U(np + 3); w.q4 = PTB_LDG4U(np + 4);
    return w;
}
PTB_HD void wide_child_box(const WideWords& w, uint32_t s, V3* lo, V3* hi)
{
    const V3 p = ex::v3(ex::from_bits(w.q0.x), ex::from_bits(w.q0.y), ex::from_bits(w.q0.z));
    const V3 sc = ex::v3(ex::from_bits((w.q0.w & 0xffu) << 23), ex::from_bits(((w.q0.w >> 8) & 0xffu) << 23), ex::from_bits(((w.q0.w >> 16) & 0xffu) << 23));
    const uint32_t lo_w[6] = {w.q2.x, w.q2.y, w.q2.z, w.q2.w, w.q3.x, w.q3.y}, hi_w[6] = {w.q3.z, w.q3.w, w.q4.x, w.q4.y, w.q4.z, w.q4.w};  // x, y, z planes of the slots 0-3 | 4-7
    const uint32_t sh = (s & 3u) * 8u, h = s >> 2;
    *lo = ex::v3(p.x + ((float)((lo_w[0 + h] >> sh) & 0xffu) - 1.0f) * sc.x, p.y + ((float)((lo_w[2 + h] >> sh) & 0xffu) - 1.0f) * sc.y,
                 p.z + ((float)((lo_w[4 + h] >> sh) & 0xffu) - 1.0f) * sc.z);
    *hi = ex::v3(p.x + ((float)((hi_w[0 + h] >> sh) & 0xffu) + 1.0f) * sc.x, p.y + ((float)((hi_w[2 + h] >> sh) & 0xffu) + 1.0f) * sc.y,
                 p.z + ((float)((hi_w[4 + h] >> sh) & 0xffu) + 1.0f) * sc.z);
}
// the frustum's walk inside one BLAS: from the root down while it stays within TILE_NODES_MAX nodes and touches no triangle leaf of the
// node it would leave.  Returns the number of nodes (0: the frustum touches nothing in this BLAS; 1 with nodes[0] == root: no gain).
PTB_HD uint32_t tile_blas_frontier(const SceneDev& S, const TileFrustum& F, uint32_t root, uint32_t* nodes, float* dist)
{
    uint32_t n = 1;
    bool final_[TILE_NODES_MAX] = {};
    nodes[0] = root;
    dist[0] = 0.0f;
    for (int iter = 0; iter < 24 && n > 0; ++iter)
    {
        uint32_t i = 0;
        while (i < n && final_[i]) ++i;
        if (i == n) break;
        const WideWords w = wide_words(S, nodes[i]);
        const uint32_t imask = w.q0.w >> 24, lmask = w.q1.z & 0xffu;
        uint32_t ch[8], nch = 0;
        float cd[8];
        bool leaf_touched = false;
        for (uint32_t s = 0; s < 8u; ++s)
        {
            if (!(((imask | lmask) >> s) & 1u)) continue;
            V3 lo, hi;
            float d;
            wide_child_box(w, s, &lo, &hi);
            if (!frustum_touches(F, lo, hi, &d)) continue;
            if ((lmask >> s) & 1u) { leaf_touched = true; break; }
            ch[nch] = w.q1.x + popc32(imask & ((1u << s) - 1u));
            cd[nch++] = d;
        }
        if (leaf_touched || n - 1u + nch > TILE_NODES_MAX) { final_[i] = true; continue; }
        // replace node i by its touched inner children (none: the node drops out)
        nodes[i] = nodes[n - 1]; dist[i] = dist[n - 1]; final_[i] = final_[n - 1];
        --n;
        for (uint32_t k = 0; k < nch; ++k) { nodes[n] = ch[k]; dist[n] = cd[k]; final_[n] = false; ++n; }
    }
    // nearest first
    for (uint32_t a = 1; a < n; ++a)
        for (uint32_t b = a; b > 0 && dist[b - 1] > dist[b]; --b)
        {
            const uint32_t tn = nodes[b]; nodes[b] = nodes[b - 1]; nodes[b - 1] = tn;
            const float td = dist[b]; dist[b] = dist[b - 1]; dist[b - 1] = td;
        }
    return n;
}
PTB_HD void tile_list_build(const RenderView& V, uint32_t tile, uint2* out)
{
    const float x0 = (float)((tile % V.tiles_x) * 8u), y0 = (float)((tile / V.tiles_x) * 4u);
    // pixel area of the tile with the jitter range (-0.5, 0.5) and a margin, as raygen's (u, v) (regen_slot)
    const float u0 = (x0 - 0.52f) / (float)V.W, u1 = (x0 + 7.52f) / (float)V.W, v0 = (y0 - 0.52f) / (float)V.H, v1 = (y0 + 3.52f) / (float)V.H;
    V3 e, c[4];
    camera_ray(V.cam, u0, v0, &e, &c[0]);
    camera_ray(V.cam, u1, v0, &e, &c[1]);
    camera_ray(V.cam, u1, v1, &e, &c[2]);
    camera_ray(V.cam, u0, v1, &e, &c[3]);
    const TileFrustum F = tile_frustum(e, c);
    // (1) the TLAS leaves the frustum touches, nearest first
    constexpr uint32_t INST_CAP = 12;
    uint32_t stack[48];
    int sp = 0;
    stack[sp++] = V.S.tlas[0].wide_root;
    uint2 best[INST_CAP];
    uint32_t cnt = 0;
    bool overflow = false;
    while (sp > 0 && !overflow)
    {
        const WideWords w = wide_words(V.S, stack[--sp]);
        const uint32_t imask = w.q0.w >> 24, lmask = w.q1.z & 0xffu;
        for (uint32_t s = 0; s < 8u; ++s)
        {
            if (!(((imask | lmask) >> s) & 1u)) continue;
            float dmin;
            if ((lmask >> s) & 1u)
            {
                const uint32_t hidx = w.q1.y + popc32(lmask & ((1u << s) - 1u));
                const float4* lp = V.S.leaf_headers + (size_t)hidx * LEAF_F4;
                const float4 h0 = PTB_LDG4F(lp), h1 = PTB_LDG4F(lp + 1);
                if (!frustum_touches(F, ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), &dmin)) continue;
                if (cnt == INST_CAP) { overflow = true; break; }
                uint32_t i = cnt++;
                for (; i > 0 && ex::from_bits(best[i - 1].y) > dmin; --i) best[i] = best[i - 1];
                best[i].x = hidx; best[i].y = ex::bits(dmin);
            }
            else
            {
                V3 lo, hi;
                wide_child_box(w, s, &lo, &hi);
                if (!frustum_touches(F, lo, hi, &dmin)) continue;
                if (sp == 48) { overflow = true; break; }
                stack[sp++] = w.q1.x + popc32(imask & ((1u << s) - 1u));
            }
        }
    }
    if (overflow) { out[0].x = TILE_LIST_NONE; out[0].y = 0u; return; }
    // (2) inside each instance: the frustum in BLAS space (Ray::transform's map applied to the apex and the corner directions)
    uint32_t n_out = 0;
    for (uint32_t i = 0; i < cnt; ++i)
    {
        const float4* lp = V.S.leaf_headers + (size_t)best[i].x * LEAF_F4;
        const uint32_t cw = ex::bits(PTB_LDG4F(lp).w);
        uint32_t nodes[TILE_NODES_MAX], k = 0;
        if (V.tile_nodes && !(cw & TLAS_DIRECT_FLAG))
        {
            const float4 i0 = PTB_LDG4F(lp + 2), i1 = PTB_LDG4F(lp + 3), i2 = PTB_LDG4F(lp + 4);
            const M3 A{ex::v3(i0.x, i0.y, i0.z), ex::v3(i0.w, i1.x, i1.y), ex::v3(i1.z, i1.w, i2.x)};
            const V3 t = ex::v3(i2.y, i2.z, i2.w);
            V3 cb[4];
            for (int q = 0; q < 4; ++q) cb[q] = mulm(A, c[q]);
            const TileFrustum FB = tile_frustum(mulm(A, e) + t, cb);
            float dist[TILE_NODES_MAX];
            k = tile_blas_frontier(V.S, FB, cw, nodes, dist);
            if (k == 0u) continue;                    // the frustum touches nothing inside: not a candidate
            if (k == 1u && nodes[0] == cw) k = 0u;    // still the root
        }
        if (n_out + 1u + k > TILE_LIST_CAP)
        {
            if (n_out + 1u > TILE_LIST_CAP) { out[0].x = TILE_LIST_NONE; out[0].y = 0u; return; }
            k = 0u;  // no room for its nodes: from the BLAS root
        }
        out[1u + n_out].x = best[i].x | (k << TILE_NODES_SHIFT);
        out[1u + n_out].y = best[i].y;
        ++n_out;
        for (uint32_t q = 0; q < k; ++q) { out[1u + n_out].x = nodes[q]; out[1u + n_out].y = 0u; ++n_out; }
    }
    out[0].x = n_out;
    out[0].y = 0u;
}

// ---- regen ---------------------------------------------------------------------------------------------------------------
// Finalises the slot's finished sample (if any), then starts the next one; the camera ray is returned in registers (and
// stored) so the caller can trace it right away.
PTB_HD uint32_t regen_slot(const RenderView& V, uint32_t slot, V3* o_out, V3* d_out)
{
    const uint32_t state = ex::bits(V.ray[slot].o.w);
    uint32_t started = ex::bits(V.ray[slot].d.w);
    if (state & ST_HAS_SAMPLE)
    {
        // integrator.rs:272-280 then accumulate.wgsl:23
        const float4 r4 = V.path[slot].radiance;
        V3 c = ex::v3(r4.x, r4.y, r4.z);
        if (finite3(c))
        {
            const float l2 = dot(c, c);
            if (l2 > 100.0f * 100.0f) c = c * ((1.0f / sqrtf(l2)) * 100.0f);  // clamp_length_max(100)
        }
        else c = ex::v3(0, 0, 0);
        float4 a = V.accum[slot];
        a.x += c.x; a.y += c.y; a.z += c.z; a.w += 1.0f;
        V.accum[slot] = a;
        if (V.accum_sq)
        {
            float4 q = V.accum_sq[slot];
            q.x += c.x * c.x; q.y += c.y * c.y; q.z += c.z * c.z; q.w += 1.0f;
            V.accum_sq[slot] = q;
        }
    }
    const uint32_t pixel = slot_pixel(V, slot % V.npix), sub = slot / V.npix;
    const uint32_t local = sub + started * V.K;  // index of the slot's next sample within this render call
    if (local >= V.sample_count)
    {
        V.ray[slot].o.w = ex::from_bits(0u);
        return 0u;
    }
    // main.rs:187-199
    const uint32_t smp = V.sample_begin + local;
    const uint32_t x = pixel % V.W, y = pixel / V.W;
    Rng rng = rng_stream(V.seed, pixel, smp);
    const uint32_t jseed = rng_u32(rng);
    float jx, jy;
    ss_sobol(smp == 0u ? 0u : 1u, jseed, &jx, &jy);  // index = c.w as u32: 0 on the first frame, 1 afterwards (main.rs:194,204)
    jx = ex::sub(jx, 0.5f);
    jy = ex::sub(jy, 0.5f);
    const float u = ex::div(ex::add((float)x, jx), (float)V.W), v = ex::div(ex::add((float)y, jy), (float)V.H);
    V3 o, d;
    camera_ray(V.cam, u, v, &o, &d);
    V.ray[slot].o = f4(o.x, o.y, o.z, ex::from_bits(ST_HAS_SAMPLE));  // bounce 0, last_delta false
    V.ray[slot].d = f4(d.x, d.y, d.z, ex::from_bits(started + 1u));
    V.path[slot].weight = f4(1.0f, 1.0f, 1.0f, 0.0f);
    V.path[slot].radiance = f4(0.0f, 0.0f, 0.0f, 0.0f);
    V.hit[slot].rng_lo = (uint32_t)rng.s;
    V.hit[slot].rng_hi = (uint32_t)(rng.s >> 32);
    if (V.volset) V.volset[slot] = 0ull;
    // integrator.rs:156-157 defaults; overwritten at b == 0 on a hit
    const V3 far = ex::ray_at(o, d, 1e5f);
    V.position[slot] = f4(far.x, far.y, far.z, 1e5f);
    V.id[slot] = (V.id[slot] << 16) | 255u;  // main.rs:206 with first_id defaulting to u8::MAX (integrator.rs:157)
    *o_out = o;
    *d_out = d;
    return WF_NEXT;
}
PTB_HD uint32_t regen_slot(const RenderView& V, uint32_t slot)
{
    V3 o, d;
    return regen_slot(V, slot, &o, &d);
}

// One shared, non-inlined copy of the 8-wide traversal for the per-slot bodies below (the single-launch tail kernels).  Those
// kernels hold the whole path loop; with the traversal inlined at its four call sites they were 190 KB of code, and their few
// resident warps — each role in a different copy — spent a third of their issue latency waiting for instructions
// (smsp stall no_instruction 1.8 per issued instruction, profiles/r1m_c3_tail_ncu.txt).  Closest-hit or any-hit is decided at
// run time (MODE 2); the arithmetic is the exact tier's, so results do not depend on which copy runs.
#if defined(__CUDA_ARCH__)
__device__ __noinline__
#else
inline
#endif
bool wide_traverse_shared(const SceneDev* S, int which, float ox, float oy, float oz, float dx, float dy, float dz, float t_max, int any, Hit* hit)
{
    uint2 stack[WIDE_STACK_CAP];
    WideTrav s;
    wide_begin(s, *S, which, ex::v3(ox, oy, oz), ex::v3(dx, dy, dz), t_max, any != 0);
    while (wide_step<2>(s, *S, stack)) {}
    if (s.hit_inst == TRAV_NONE) return false;
    if (!any) { hit->t = s.t_max; hit->u = wide_hit_u(s); hit->v = wide_hit_v(s); hit->slot_or_prim = s.hit_slot; hit->inst = s.hit_inst; }
    return true;
}
PTB_HD bool wide_traverse_call(const SceneDev& S, int which, V3 o, V3 d, float t_max, bool any, Hit* hit)
{
    return wide_traverse_shared(&S, which, o.x, o.y, o.z, d.x, d.y, d.z, t_max, any ? 1 : 0, hit);
}

// ---- extend --------------------------------------------------------------------------------------------------------------
// returns the shade-queue kind (material kind of the hit BLAS, or Q_MISS)
template <bool BIN>
PTB_HD uint32_t extend_ray_t(const RenderView& V, uint32_t slot, V3 o, V3 d)
{
    Hit h;
    h.inst = 0xffffffffu;
    const bool f = BIN ? bin_traverse<false>(V.S, 0, o, d, ex::inf(), &h) : wide_traverse_call(V.S, 0, o, d, ex::inf(), false, &h);
    if (!f)
    {
        V.hit[slot].info = 0xffffffffu;
        return Q_MISS;
    }
    const uint32_t prim = BIN ? h.slot_or_prim : V.S.tri_prim[h.slot_or_prim];
    V3 n;
    bool ff;
    finish_hit(V.S, d, h.inst, prim, h.u, h.v, &n, &ff);
    V.hit[slot].tn = f4(h.t, n.x, n.y, n.z);
    V.hit[slot].info = h.inst | (ff ? 0x80000000u : 0u);
    return V.S.mat[V.S.inst[h.inst].blas].kind;
}
PTB_HD uint32_t extend_ray(const RenderView& V, uint32_t slot, V3 o, V3 d) { return V.extend_kernel ? extend_ray_t<true>(V, slot, o, d) : extend_ray_t<false>(V, slot, o, d); }
PTB_HD uint32_t extend_slot(const RenderView& V, uint32_t slot)
{
    const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
    return extend_ray(V, slot, ex::v3(o4.x, o4.y, o4.z), ex::v3(d4.x, d4.y, d4.z));
}

// ---- shade ---------------------------------------------------------------------------------------------------------------
// KMASK: material kinds the scene can present (bit k = kind k), MEDIA: the scene has media.  The slot-ordered shade kernel is
// instantiated for a few scene classes (render.cu): with the kinds a scene does not hold compiled out — the body is 7.5 K instructions
// with all six kinds and media, and its warps stall on instruction fetch — the same source runs as a third to a half of that.
// 0x3f / true is the general body; the host never picks a mask that omits a kind of the scene.
template <uint32_t KMASK, bool MEDIA>
PTB_HD uint32_t shade_slot_t(const RenderView& V, uint32_t slot)
{
    const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
    const V3 o = ex::v3(o4.x, o4.y, o4.z), d = ex::v3(d4.x, d4.y, d4.z);
    uint32_t state = ex::bits(o4.w);
    const uint32_t b = state & ST_BOUNCE_MASK;
    const bool last_delta = (state & ST_LAST_DELTA) != 0u;
    const uint32_t info = V.hit[slot].info;
    const float4 w4 = V.path[slot].weight;
    V3 path_weight = ex::v3(w4.x, w4.y, w4.z);

    if (info == 0xffffffffu)
    {
        // miss: integrator.rs:254-269
        const float4 r4 = V.path[slot].radiance;
        V3 acc = ex::v3(r4.x, r4.y, r4.z);
        if (V.S.env)
        {
            const float u = fmaf(atan2f(d.x, d.z), INV_PI_F * 0.5f, 0.5f);
            const float v = fmaf(asinf(d.y), -INV_PI_F, 0.5f);
            acc = acc + env_bilinear(V.S.env, V.S.env_w, V.S.env_h, u, v) * path_weight;
        }
        else acc = acc + ex::v3(0.006f, 0.006f, 0.006f) * path_weight;
        V.path[slot].radiance = f4(acc.x, acc.y, acc.z, 0.0f);
        return WF_TERM;
    }

    const float4 h4 = V.hit[slot].tn;
    const float t = h4.x;
    const V3 normal = ex::v3(h4.y, h4.z, h4.w);
    const bool front = (info & 0x80000000u) != 0u;
    const uint32_t inst = info & 0x7fffffffu;
    const uint32_t blas = V.S.inst[inst].blas;
    Mat mat = load_mat(V.S.mat, blas);
    if (KMASK != 0x3fu)
    {
        // kinds outside the mask do not occur: fold them onto the lowest kind of the mask so that their code is provably dead
        if (!((KMASK >> mat.kind) & 1u)) mat.kind = (KMASK & 1u) ? 0u : 1u;
#if defined(__CUDA_ARCH__)
        __builtin_assume(mat.kind < 6u && ((KMASK >> mat.kind) & 1u));
#endif
    }
    Rng rng;
    rng.s = (unsigned long long)V.hit[slot].rng_lo | ((unsigned long long)V.hit[slot].rng_hi << 32);
    const V3 hit_p = ex::ray_at(o, d, t);  // Ray::at (fused, ray.rs:20)

    if (b == 0u)  // integrator.rs:181-185
    {
        V.position[slot] = f4(hit_p.x, hit_p.y, hit_p.z, t);
        V.id[slot] = (V.id[slot] & 0xffff0000u) | (blas & 0xffu);  // first_id = id.index() as u8 (integrator.rs:184); history shifted in regen
    }

    uint32_t flags = 0u;
    bool done = false, scattered = false;
    V3 new_d = d, new_o = hit_p;
    bool new_last_delta = last_delta;

    // media: integrator.rs:189-205 (volume set iterated in table order)
    unsigned long long vset = (MEDIA && V.volset) ? V.volset[slot] : 0ull;
    if (vset)
    {
        float best_t = 0.0f;
        V3 best_dir = ex::v3(0, 0, 0);
        for (uint32_t vi = 0; vi < V.S.n_volumes; ++vi)
        {
            if (!((vset >> vi) & 1ull) || !(V.S.vol[vi].flags & 2u)) continue;
            float ts;
            V3 ds;
            if (volume_scatter(V.S.vol[vi], rng, d, t, &ts, &ds))
                if (!scattered || ex::total_lt(ts, best_t)) { best_t = ts; best_dir = ds; scattered = true; }
        }
        V3 trans = ex::v3(1, 1, 1);
        for (uint32_t vi = 0; vi < V.S.n_volumes; ++vi)
            if (((vset >> vi) & 1ull) && (V.S.vol[vi].flags & 1u)) trans = trans * volume_transmission(V.S.vol[vi], scattered ? best_t : t);
        path_weight = path_weight * trans;
        if (scattered)
        {
            new_last_delta = true;
            new_o = ex::ray_at(o, d, best_t);
            new_d = best_dir;
        }
    }

    if (!scattered)
    {
        if (mat.kind == PTB_EMISSIVE)  // integrator.rs:207-214
        {
            if (!V.enable_nee || last_delta || b == 0u)
            {
                const float4 r4 = V.path[slot].radiance;
                V.path[slot].radiance = f4(fmaf(mat.colour.x, path_weight.x, r4.x), fmaf(mat.colour.y, path_weight.y, r4.y), fmaf(mat.colour.z, path_weight.z, r4.z), 0.0f);
            }
            done = true;
        }
        else
        {
            if (MEDIA && mat.volume >= 0 && V.volset)  // integrator.rs:217-227
            {
                if (front) vset |= (1ull << mat.volume);
                else vset &= ~(1ull << mat.volume);
                V.volset[slot] = vset;
            }
            const bool delta = is_delta(mat.kind);
            const V3 wi = -d;
            if (V.enable_nee && !delta)  // integrator.rs:231-234
            {
                // explicit light sample (integrator.rs:25-74): 1 + 2 draws, always
                const uint32_t li = light_pick(V.S.lights, V.S.n_lights, rng_f32(rng));
                const LightRow L = V.S.lights[li];
                float lu = rng_f32(rng), lv = rng_f32(rng);
                if (lu + lv > 1.0f) { lu = 1.0f - lu; lv = 1.0f - lv; }  // primitive.rs:77-91
                const float* pp = V.S.tri_positions + (size_t)L.tri * 9;
                const float* pn = V.S.tri_normals + (size_t)L.tri * 9;
                const float lw = 1.0f - lu - lv;
                const V3 point = (ex::v3(pp[0], pp[1], pp[2]) * lw + ex::v3(pp[3], pp[4], pp[5]) * lu) + ex::v3(pp[6], pp[7], pp[8]) * lv;
                const V3 lnormal = normalizef((ex::v3(pn[0], pn[1], pn[2]) * lw + ex::v3(pn[3], pn[4], pn[5]) * lu) + ex::v3(pn[6], pn[7], pn[8]) * lv);
                const V3 dl = point - hit_p;
                const float dist2 = dot(dl, dl), dist = sqrtf(dist2);
                const V3 ld = ex::normalize(dl);
                if (dot(ld, normal) > 0.0f)
                {
                    const BsdfPdf bp = bsdf_pdf(mat, wi, ld, normal, front);
                    const float4 n0 = V.S.planes_by_prim[(size_t)L.tri * 3];
                    const float area = 0.5f * sqrtf((n0.x * n0.x + n0.y * n0.y) + n0.z * n0.z);  // Triangle::area
                    const float sample_pdf = L.pdf / area;
                    const float cosine = fabsf(dot(ld, lnormal));
                    const float light_pdf = sample_pdf * (dist2 / cosine);
                    const float wgt = (light_pdf * light_pdf) / (light_pdf * light_pdf + bp.pdf * bp.pdf);  // mis_heuristic<2>
                    const Mat lmat = load_mat(V.S.mat, L.blas);
                    const V3 c = path_weight * ((lmat.colour * wgt * weakening(mat.kind, ld, normal)) * bp.bsdf / light_pdf);
                    V.shadow[slot].d = f4(ld.x, ld.y, ld.z, (1.0f - ex::EPS) * dist);
                    V.shadow[slot].c = f4(c.x, c.y, c.z, 0.0f);
                    flags |= WF_SHADOW;
                }
                // BSDF-sampled light (integrator.rs:77-130): the material's draws, always
                const V3 pd = scatter_direction(mat, rng, d, normal, front);
                if (dot(pd, normal) > 0.0f)
                {
                    const BsdfPdf bp = bsdf_pdf(mat, wi, pd, normal, front);
                    if (bp.pdf > 0.0f)  // MIN_PDF (checked upstream after the visibility tests; order is immaterial)
                    {
                        if (!V.extend_kernel && !tlas_may_hit(V.S, 1, hit_p, pd)) flags |= WF_PROBE_OUT;  // integrator.rs:100 would return None
                        else
                        {
                            const V3 f = path_weight * (weakening(mat.kind, pd, normal) * bp.bsdf / bp.pdf);
                            V.probe[slot].d = f4(pd.x, pd.y, pd.z, bp.pdf);
                            V.probe[slot].f = f4(f.x, f.y, f.z, 0.0f);
                            flags |= WF_PROBE;
                        }
                    }
                }
            }
            // continuation: integrator.rs:236-251
            new_d = scatter_direction(mat, rng, d, normal, front);
            const BsdfPdf mi = bsdf_pdf(mat, wi, new_d, normal, front);
            if (mi.pdf < 0.0f) done = true;
            else
            {
                path_weight = path_weight * (weakening(mat.kind, new_d, normal) * mi.bsdf / mi.pdf);
                new_last_delta = delta;
            }
        }
    }

    if (!done)
    {
        // next loop iteration: b+1 <= max_bounces, then Russian roulette (integrator.rs:163-177)
        const uint32_t nb = b + 1u;
        if (nb > V.max_bounces) done = true;
        else if (nb > 3u)
        {
            const float survive = fminf(max3(path_weight), 0.9999f);
            if (rng_f32(rng) > survive) done = true;
            else path_weight = path_weight / survive;
        }
        if (!done)
        {
            state = (state & ~(ST_BOUNCE_MASK | ST_LAST_DELTA)) | nb | (new_last_delta ? ST_LAST_DELTA : 0u);
            V.ray[slot].o = f4(new_o.x, new_o.y, new_o.z, ex::from_bits(state));
            V.ray[slot].d = f4(new_d.x, new_d.y, new_d.z, d4.w);
            V.path[slot].weight = f4(path_weight.x, path_weight.y, path_weight.z, 0.0f);
            V.hit[slot].rng_lo = (uint32_t)rng.s;
            V.hit[slot].rng_hi = (uint32_t)(rng.s >> 32);
            return flags | WF_NEXT;
        }
    }
    // the path ends here, but its last NEE rays still start at the hit point: keep it as the slot's ray origin (state bits unchanged)
    if (flags & (WF_SHADOW | WF_PROBE)) V.ray[slot].o = f4(hit_p.x, hit_p.y, hit_p.z, o4.w);
    return flags | WF_TERM;
}

PTB_HD uint32_t shade_slot(const RenderView& V, uint32_t slot) { return shade_slot_t<0x3fu, true>(V, slot); }

// ---- shadow --------------------------------------------------------------------------------------------------------------
// adds the explicit light sample's contribution to *r when the shadow ray is unoccluded (integrator.rs:56-70)
template <bool BIN>
PTB_HD void shadow_contrib_t(const RenderView& V, uint32_t slot, V3 o, float4* r)
{
    const float4 d4 = V.shadow[slot].d;
    Hit h;
    const V3 d = ex::v3(d4.x, d4.y, d4.z);
    const bool occluded = BIN ? bin_traverse<true>(V.S, 0, o, d, d4.w, &h) : wide_traverse_call(V.S, 0, o, d, d4.w, true, &h);
    if (occluded) return;
    const float4 c = V.shadow[slot].c;
    r->x += c.x; r->y += c.y; r->z += c.z;
}

// contribution of a light found by the BSDF-sampled probe and visible from the hit point (integrator.rs:105-123)
PTB_HD V3 probe_light_contrib(const RenderView& V, uint32_t slot, V3 d, float tl, float u, float v, uint32_t prim, uint32_t inst)
{
    V3 ln;
    bool ff;
    finish_hit(V.S, d, inst, prim, u, v, &ln, &ff);
    const InstanceRow& R = V.S.inst[inst];
    const Mat lmat = load_mat(V.S.mat, R.blas);
    // get_sample_pdf(light, material) / light.area()  (light_sampler.rs:39, integrator.rs:111)
    const float4 n0 = V.S.planes_by_prim[(size_t)(R.tri_base + prim) * 3];
    const float area = 0.5f * sqrtf((n0.x * n0.x + n0.y * n0.y) + n0.z * n0.z);
    const float sample_pdf = (area * sqrtf(dot(lmat.colour, lmat.colour)) / V.S.light_max) / area;
    const float cosine = fabsf(dot(d, ln));
    const float light_pdf = sample_pdf * (tl * tl / cosine);
    const float bp = V.probe[slot].d.w;
    const float wgt = (bp * bp) / (bp * bp + light_pdf * light_pdf);  // mis_heuristic<2>(bsdf_pdf, light_pdf)
    const float4 f = V.probe[slot].f;
    return ex::v3(lmat.colour.x * wgt * f.x, lmat.colour.y * wgt * f.y, lmat.colour.z * wgt * f.z);
}

// ---- probe ---------------------------------------------------------------------------------------------------------------
// returns 1 when the lights-TLAS probe hit a light and a world shadow ray was traced
template <bool BIN>
PTB_HD uint32_t probe_contrib_t(const RenderView& V, uint32_t slot, V3 o, float4* r)
{
    const float4 d4 = V.probe[slot].d;
    const V3 d = ex::v3(d4.x, d4.y, d4.z);
    if (!BIN && !tlas_may_hit(V.S, 1, o, d)) return 0u;  // cannot reach the bounds of the lights (the reference-order variant always traverses)
    Hit h;
    h.inst = 0xffffffffu;
    const bool lit = BIN ? bin_traverse<false>(V.S, 1, o, d, ex::inf(), &h) : wide_traverse_call(V.S, 1, o, d, ex::inf(), false, &h);  // integrator.rs:100
    if (!lit) return 0u;
    Hit dummy;
    const float tl = h.t;
    const bool occluded = BIN ? bin_traverse<true>(V.S, 0, o, d, tl * (1.0f - ex::EPS), &dummy) : wide_traverse_call(V.S, 0, o, d, tl * (1.0f - ex::EPS), true, &dummy);  // :103
    if (occluded) return 1u;
    const uint32_t prim = BIN ? h.slot_or_prim : V.S.tri_prim[h.slot_or_prim];
    const V3 c = probe_light_contrib(V, slot, d, tl, h.u, h.v, prim, h.inst);
    r->x += c.x; r->y += c.y; r->z += c.z;
    return 1u;
}

// ---- nee: both direct-light estimates of one bounce (estimate_direct, integrator.rs:138), one read-modify-write of the radiance --------
// flags: WF_SHADOW | WF_PROBE as returned by shade_slot; returns 1 when the probe traced its world shadow ray
template <bool BIN>
PTB_HD uint32_t nee_slot_t(const RenderView& V, uint32_t slot, uint32_t flags)
{
    const float4 o4 = V.ray[slot].o;
    const V3 o = ex::v3(o4.x, o4.y, o4.z);
    float4 r = V.path[slot].radiance;
    uint32_t traced = 0u;
    if (flags & WF_SHADOW) shadow_contrib_t<BIN>(V, slot, o, &r);
    if (flags & WF_PROBE) traced = probe_contrib_t<BIN>(V, slot, o, &r);
    V.path[slot].radiance = r;
    return traced;
}
PTB_HD uint32_t nee_slot(const RenderView& V, uint32_t slot, uint32_t flags) { return V.extend_kernel ? nee_slot_t<true>(V, slot, flags) : nee_slot_t<false>(V, slot, flags); }

}  // namespace ptb
