// team.cuh — warp-cooperative ("team") traversal of the 8-wide layout: EIGHT lanes on ONE ray, four rays per warp (device only).
//
// What the north star calls the warp-cooperative extend kernel, built to be measured against the one-ray-per-lane persistent kernel
// (persist.cuh) on the same fixed ray batches (ptb_trace_* with kernel = 2) and bit-compared with it and with the oracle:
//   N  lane j of a team tests quantised child box j of the node (6 byte loads + 6 PRMT + 6 FMA instead of the 244-instruction
//      unrolled test of eight boxes in one lane); the team's hit mask is one __ballot_sync;
//   L  every hit leaf of the node goes to its own lane: exact reference-leaf box, exact t_enter and the leaf's triangles
//      (exact Havel-Herout test, exact_math.h) — the leaves of a node side by side instead of one per step; the nearest hit is a
//      team-wide (t, lane) minimum over 3 shuffle rounds.  (t, u, v) of a triangle depends only on (ray, triangle, exact leaf box)
//      — traverse.h:9-14 — so results are those of the other kernels up to the resolution of exact ties;
//   P  pop / descend / enter an instance: the same decisions as wide_step_pop, executed redundantly by the eight lanes (the team's
//      state is replicated in their registers; the stack is one copy per team in shared memory).
// The stages sit at single code locations inside one loop, so the four teams of a warp re-converge at every stage, as the lanes
// of the one-ray-per-lane kernel do.  Measured result and what it is good for: DESIGN.md section 4 ("Team traversal").
#pragma once
#include "traverse.h"

namespace ptb {

constexpr int TEAM = 8;                      // lanes per ray
constexpr int TEAMS_PER_BLOCK = 128 / TEAM;  // blocks of 128 threads

struct TeamRay
{
    V3 ow, dw, invw;   // world ray
    uint32_t oiw;
    TravRay ray;       // current space
    float t_max;
    uint32_t node;     // wide node to test next, or TRAV_NONE
    uint2 G;           // inner children of the current node still to visit
    int sp, blas_sp;
    uint32_t cur_inst;
    float hit_ud, hit_vd, hit_det;
    uint32_t hit_slot, hit_inst;
};

__device__ __forceinline__ void team_begin(TeamRay& s, const SceneDev& S, int which, V3 o, V3 d, float t_max)
{
    s.ow = o; s.dw = d;
    set_ray(s.ray, o, d);
    s.invw = s.ray.inv;
    s.oiw = s.ray.oi;
    s.t_max = t_max;
    s.node = S.tlas[which].wide_root;
    s.G.x = 0; s.G.y = 0;
    s.sp = 0; s.blas_sp = 0;
    s.cur_inst = TRAV_NONE;
    s.hit_ud = s.hit_vd = 0.0f;
    s.hit_det = 1.0f;
    s.hit_slot = 0;
    s.hit_inst = TRAV_NONE;
}

// Lane `tl` tests quantised child box `tl` of the node: the arithmetic of wide_node_test (traverse.h) for one slot.
__device__ __forceinline__ bool team_child_test(const uint4& q0, const uint8_t* planes, unsigned tl, const TravRay& r, float t_max)
{
    const float px = ex::from_bits(q0.x), py = ex::from_bits(q0.y), pz = ex::from_bits(q0.z);
    const float sx0 = ex::from_bits((q0.w & 0xffu) << 23), sy0 = ex::from_bits(((q0.w >> 8) & 0xffu) << 23), sz0 = ex::from_bits(((q0.w >> 16) & 0xffu) << 23);
    const float sx = sx0 * r.idq.x, sy = sy0 * r.idq.y, sz = sz0 * r.idq.z;
    const float bx = (px - r.o.x) * r.idq.x, by = (py - r.o.y) * r.idq.y, bz = (pz - r.o.z) * r.idq.z;
    const float k = 9.5367432e-7f, kb = 7.8125e-3f;  // rounding slack, as wide_node_test
    const float ex_ = fmaf(fabsf(bx) + 255.0f * fabsf(sx), k, fabsf(sx) * kb), ey_ = fmaf(fabsf(by) + 255.0f * fabsf(sy), k, fabsf(sy) * kb),
                ez_ = fmaf(fabsf(bz) + 255.0f * fabsf(sz), k, fabsf(sz) * kb);
    const float bias = -32768.0f, inv256 = 0.00390625f;
    const float cnx = fmaf(bias, sx, bx - ex_), cfx = fmaf(bias, sx, bx + ex_), cny = fmaf(bias, sy, by - ey_), cfy = fmaf(bias, sy, by + ey_),
                cnz = fmaf(bias, sz, bz - ez_), cfz = fmaf(bias, sz, bz + ez_);
    const uint32_t kc = qbyte_const();
    // planes: qlo_x[8] qlo_y[8] qlo_z[8] qhi_x[8] qhi_y[8] qhi_z[8] (layout.h); one byte load per plane of this lane's slot
    const float lox = qbyte_biased(__ldg(planes + tl), 0, kc), loy = qbyte_biased(__ldg(planes + 8 + tl), 0, kc), loz = qbyte_biased(__ldg(planes + 16 + tl), 0, kc);
    const float hix = qbyte_biased(__ldg(planes + 24 + tl), 0, kc), hiy = qbyte_biased(__ldg(planes + 32 + tl), 0, kc), hiz = qbyte_biased(__ldg(planes + 40 + tl), 0, kc);
    const bool nx = r.idq.x < 0.0f, ny = r.idq.y < 0.0f, nz = r.idq.z < 0.0f;
    const float s8x = sx * inv256, s8y = sy * inv256, s8z = sz * inv256;
    const float tnx = fmaf(nx ? hix : lox, s8x, cnx), tfx = fmaf(nx ? lox : hix, s8x, cfx);
    const float tny = fmaf(ny ? hiy : loy, s8y, cny), tfy = fmaf(ny ? loy : hiy, s8y, cfy);
    const float tnz = fmaf(nz ? hiz : loz, s8z, cnz), tfz = fmaf(nz ? loz : hiz, s8z, cfz);
    const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, ex::EPS));
    const float tf = fminf(fminf(tfx, tfy), fminf(tfz, t_max));
    return tn <= tf;
}

// One reference leaf in this lane: exact box, exact t_enter, its triangles (stage L + T of traverse.h for a whole leaf).
// Closest hit: improves (bt, bud, bvd, bdet, bslot) and returns whether it did; any hit: returns at the first hit.
template <bool ANY>
__device__ __forceinline__ bool team_leaf(const SceneDev& S, uint32_t hidx, const TravRay& ray, float& bt, float& bud, float& bvd, float& bdet, uint32_t& bslot)
{
    const float4* lp = S.leaf_headers + (size_t)hidx * LEAF_F4;
    const float4 h0 = __ldg(lp), h1 = __ldg(lp + 1);
    float4 a = __ldg(lp + 2), b = __ldg(lp + 3), c = __ldg(lp + 4);
    const uint32_t cw = ex::bits(h0.w);
    float t_enter = 0.0f;
    if (ANY || !(cw & LEAF_ROOT_FLAG))  // blas.rs:217: the closest-hit query enters a root leaf untested, t_estimate = 0
        if (!ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), ray.o, ray.inv, bt, &t_enter)) return false;
    const uint32_t first = ex::bits(h1.w), cnt = cw & ~LEAF_ROOT_FLAG;
    bool got = false;
    for (uint32_t k = 0; k < cnt; ++k)
    {
        if (k)
        {
            const float4* pp = S.tri_planes + (size_t)(first + k) * 3;
            a = __ldg(pp); b = __ldg(pp + 1); c = __ldg(pp + 2);
        }
        const float n0[4] = {a.x, a.y, a.z, a.w}, n1[4] = {b.x, b.y, b.z, b.w}, n2[4] = {c.x, c.y, c.z, c.w};
        if (ANY)
        {
            if (ex::tri_hit_bool(n0, n1, n2, ray.o, ray.d, bt, t_enter)) return true;
        }
        else
        {
            float t, raw[3];
            if (ex::tri_hit_raw(n0, n1, n2, ray.o, ray.d, bt, t_enter, &t, raw))
            {
                bt = t; bud = raw[0]; bvd = raw[1]; bdet = raw[2]; bslot = first + k;
                got = true;
            }
        }
    }
    return got;
}

// All hit leaves of a BLAS node at once: lane tl takes slot tl when `leafhits` has its bit.  Returns true when an any-hit query is decided.
template <bool ANY>
__device__ __forceinline__ bool team_leaves(TeamRay& s, const SceneDev& S, uint32_t leafhits, uint32_t lmask, uint32_t leaf_base, unsigned tl, unsigned tmask, unsigned tbase)
{
    float bt = s.t_max, bud = 0.0f, bvd = 0.0f, bdet = 1.0f;
    uint32_t bslot = 0;
    bool got = false;
    if ((leafhits >> tl) & 1u) got = team_leaf<ANY>(S, leaf_base + popc32(lmask & ((1u << tl) - 1u)), s.ray, bt, bud, bvd, bdet, bslot);
    if (ANY)
    {
        if (__ballot_sync(tmask, got) & tmask)
        {
            s.hit_inst = s.cur_inst;
            return true;
        }
        return false;
    }
    // nearest hit of the team: (t, lane) minimum; among equal t the slot the one-lane kernel visits LAST wins (lowest priority p = tl ^ oi)
    float kt = got ? bt : ex::inf();
    unsigned kl = tl;
#pragma unroll
    for (int off = 4; off > 0; off >>= 1)
    {
        const float ot = __shfl_xor_sync(tmask, kt, off, TEAM);
        const unsigned ol = __shfl_xor_sync(tmask, kl, off, TEAM);
        const bool take = ot < kt || (ot == kt && ((ol ^ s.ray.oi) < (kl ^ s.ray.oi)));
        kt = take ? ot : kt;
        kl = take ? ol : kl;
    }
    if (kt < ex::inf())  // some lane found a hit nearer than t_max (tri_hit only accepts t <= current t_max)
    {
        const int src = (int)(tbase + kl);
        s.t_max = kt;
        s.hit_ud = __shfl_sync(tmask, bud, src);
        s.hit_vd = __shfl_sync(tmask, bvd, src);
        s.hit_det = __shfl_sync(tmask, bdet, src);
        s.hit_slot = __shfl_sync(tmask, bslot, src);
        s.hit_inst = s.cur_inst;
    }
    return false;
}

// One step of a team: P (pop / descend / instance entry) -> N (node test, one child per lane) -> L (hit leaves side by side).
// `stack` = this team's stack in shared memory.  Returns false when the traversal is complete.
template <bool ANY>
__device__ __forceinline__ bool team_step(TeamRay& s, const SceneDev& S, uint2* stack, unsigned tl, unsigned tmask, unsigned tbase)
{
    uint32_t direct_hits = 0u, direct_lmask = 0u, direct_base = 0u;
    __syncwarp(tmask);  // the lanes of a team went separate ways inside their leaves; the stack below is shared by the team
    // ---- P ---------------------------------------------------------------------------------------------------------------------
    if (s.node == TRAV_NONE)
    {
        if (s.G.y == 0u)
        {
            if (s.cur_inst != TRAV_NONE && s.sp == s.blas_sp)
            {
                s.cur_inst = TRAV_NONE;  // leave the instance
                s.ray.o = s.ow; s.ray.d = s.dw; s.ray.inv = s.invw;
                s.ray.idq = ex::v3(finite_recip_of(s.invw.x), finite_recip_of(s.invw.y), finite_recip_of(s.invw.z));
                s.ray.oi = s.oiw;
            }
            if (s.sp == 0) return false;
            const uint2 e = stack[--s.sp];
            if (e.y & 0x80000000u)
            {
                // TLAS leaf (tlas.rs:88-100 / :132-139): exact leaf box with the current t_max, then Ray::transform — all lanes alike
                const float4* lp = S.leaf_headers + (size_t)e.x * LEAF_F4;
                const float4 h0 = __ldg(lp), h1 = __ldg(lp + 1), i0 = __ldg(lp + 2), i1 = __ldg(lp + 3), i2 = __ldg(lp + 4);
                float t_enter;
                if (ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), s.ray.o, s.ray.inv, s.t_max, &t_enter))
                {
                    s.cur_inst = ex::bits(h1.w);
                    const uint32_t cw = ex::bits(h0.w);
                    const float inv12[12] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w, i2.x, i2.y, i2.z, i2.w};
                    transform_ray_reuse(s.ray, s.ow, s.dw, s.invw, inv12);
                    s.blas_sp = s.sp;
                    if (cw & TLAS_DIRECT_FLAG)
                    {
                        direct_lmask = (cw >> TLAS_DIRECT_BASE_BITS) & 0xffu;
                        direct_base = cw & ((1u << TLAS_DIRECT_BASE_BITS) - 1u);
                        direct_hits = direct_lmask;  // a BLAS of one or two leaves: straight to the exact leaf tests
                    }
                    else s.node = cw;
                }
            }
            else s.G = e;
        }
        if (s.node == TRAV_NONE && s.G.y)
        {
            const uint32_t ih = s.G.y & 0xffu, imask = (s.G.y >> 8) & 0xffu;
            const int p = hi_bit(ih);
            const uint32_t sl = (uint32_t)p ^ s.ray.oi;
            s.node = s.G.x + popc32(imask & ((1u << sl) - 1u));
            s.G.y &= ~(1u << p);
            if (s.G.y & 0xffu) stack[s.sp++] = s.G;
            s.G.y = 0u;
        }
    }
    // ---- N ---------------------------------------------------------------------------------------------------------------------
    uint32_t leafhits = direct_hits, lmask = direct_lmask, leaf_base = direct_base;
    if (s.node != TRAV_NONE)
    {
        const uint4* np = S.wide + (size_t)s.node * 5;
        const uint4 q0 = __ldg(np), q1 = __ldg(np + 1);
        s.node = TRAV_NONE;
        const bool hit = team_child_test(q0, (const uint8_t*)(np + 2), tl, s.ray, s.t_max);
        const uint32_t hits = (__ballot_sync(tmask, hit) >> tbase) & 0xffu;
        const uint32_t imask = q0.w >> 24;
        lmask = q1.z & 0xffu;
        const uint32_t ih = xor_permute16(hits & imask, s.ray.oi) & 0xffu;
        s.G.x = q1.x;
        s.G.y = ih ? (ih | (imask << 8)) : 0u;
        leaf_base = q1.y;
        leafhits = hits & lmask;
        if (s.cur_inst == TRAV_NONE)
        {
            // TLAS node: the inner group first, then the hit instances far to near (nearest on top); each hit lane writes its own entry
            if (s.G.y) { stack[s.sp++] = s.G; s.G.y = 0u; }
            const uint32_t lh = xor_permute16(leafhits, s.ray.oi) & 0xffu;  // bit p = priority (highest = nearest) of slot p ^ oi
            if ((leafhits >> tl) & 1u)
            {
                const uint32_t p = tl ^ s.ray.oi;
                uint2 e;
                e.x = leaf_base + popc32(lmask & ((1u << tl) - 1u));
                e.y = 0x80000000u;
                stack[s.sp + (int)popc32(lh & ((1u << p) - 1u))] = e;
            }
            s.sp += (int)popc32(lh);
            __syncwarp(tmask);  // the entries were written by different lanes of the team
            leafhits = 0u;
        }
    }
    // ---- L ---------------------------------------------------------------------------------------------------------------------
    if (leafhits)
        if (team_leaves<ANY>(s, S, leafhits, lmask, leaf_base, tl, tmask, tbase)) return false;
    return true;
}

}  // namespace ptb
