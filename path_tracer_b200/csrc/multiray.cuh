// multiray.cuh — K rays per lane with stage gating: the 8-wide traversal of traverse.h with the ray state in shared memory.
//
// Why.  In the one-ray-per-lane loop (persist.cuh) every iteration runs all stages once — P (pop / descend), I (instance entry:
// exact TLAS-leaf box + Ray::transform), N (one 8-wide node), L (one exact reference leaf), T (one triangle) — and each lane takes part
// in the stage its ray is at.  A node test hands 0.23 leaves to a lane on average, so L / T / I run at 4-5 of 32 lanes however the
// node stage is gated (profiles/r2m_c3_extend_sass_blocks.txt): a third of the warp instructions at an eighth of the width.  Making
// a lane WAIT for a fuller stage costs as much as it saves when the lane has nothing else to do (measured in round 1).  Here a lane
// owns K rays; a stage runs for the first of the lane's rays that wants it, and the thin stages (I, L) run only once enough lanes
// of the warp hold a ray waiting for them — meanwhile those lanes go on with their other rays.  SIMT cost model on C3's rays
// (tools/mrsim.py, calibrated against the capture above): K = 3-4 with quorums 16 / 12 / 12 needs 1.32-1.37 x fewer warp
// instructions per bounce ray, 1.2-1.3 x for primary and shadow rays, overheads of the shared-memory state included.
//
// State.  Registers hold only the stage at hand; the K ray records of a lane live in shared memory as [ray][word][thread]
// (conflict-free: a lane only ever touches its own column), the K stacks in local memory.  One 3-bit tag per ray in a lane
// register says which stage the ray wants next.  Arithmetic and visit order per ray are those of traverse.h (same helpers), so
// results are bit-identical to the one-ray-per-lane kernel up to exact ties.
#pragma once
#include "traverse.h"

namespace ptb {

// words of a ray record (25 with one job word: 100 bytes per ray; the world-space origin and direction are not kept — they are re-read
// from the job's own source when the ray leaves an instance, Job::world_ray)
enum : int
{
    MR_O = 0, MR_D = 3, MR_INV = 6,      // current space: world, or BLAS space after Ray::transform; exact reciprocal (ray.rs:1-8)
    MR_TMAX = 9,
    MR_NODE = 10,                        // tag N: wide node to fetch; tag I: leaf-header index of the TLAS leaf; tag L: leaf_base of the current node
    MR_GX = 11, MR_GY = 12,              // inner children of the current node still to visit
    MR_SP = 13,                          // sp | blas_sp << 8
    MR_INST = 14,                        // InstanceRow index while inside a BLAS
    MR_FL = 15,                          // lh | lmask << 8 | oi << 16 | oiw << 19 | any << 22 | in_blas << 23; bits 24.. are the job's (the traversal keeps them)
    MR_HUD = 16, MR_HVD = 17, MR_HDET = 18, MR_HSLOT = 19, MR_HINST = 20,
    MR_INVW = 21,                        // exact reciprocal of the world-space direction
    MR_JOB = 24                          // first word of the job's own per-ray data
};
enum : uint32_t { MT_IDLE = 0, MT_P = 1, MT_N = 2, MT_L = 3, MT_I = 4, MT_DONE = 5 };
constexpr uint32_t MR_FL_ANY = 1u << 22, MR_FL_BLAS = 1u << 23;

// one ray record: word w at q[w * STRIDE]
template <int STRIDE>
struct MrRay
{
    uint32_t* q;
    PTB_HD uint32_t ld(int w) const { return q[w * STRIDE]; }
    PTB_HD float ldf(int w) const { return ex::from_bits(q[w * STRIDE]); }
    PTB_HD V3 ld3(int w) const { return ex::v3(ldf(w), ldf(w + 1), ldf(w + 2)); }
    PTB_HD void st(int w, uint32_t v) const { q[w * STRIDE] = v; }
    PTB_HD void stf(int w, float v) const { q[w * STRIDE] = ex::bits(v); }
    PTB_HD void st3(int w, V3 v) const { stf(w, v.x); stf(w + 1, v.y); stf(w + 2, v.z); }
};

template <int K>
PTB_HD int mr_first(uint32_t tags, uint32_t x)
{
#pragma unroll
    for (int k = 0; k < K; ++k)
        if (((tags >> (3 * k)) & 7u) == x) return k;
    return -1;
}
template <int K>
PTB_HD bool mr_has(uint32_t tags, uint32_t x) { return mr_first<K>(tags, x) >= 0; }
template <int K>
PTB_HD bool mr_has2(uint32_t tags, uint32_t x, uint32_t y)
{
    bool r = false;
#pragma unroll
    for (int k = 0; k < K; ++k) r |= (((tags >> (3 * k)) & 7u) == x) || (((tags >> (3 * k)) & 7u) == y);
    return r;
}
template <int K>
PTB_HD int mr_count2(uint32_t tags, uint32_t x, uint32_t y)
{
    int n = 0;
#pragma unroll
    for (int k = 0; k < K; ++k) n += ((((tags >> (3 * k)) & 7u) == x) || (((tags >> (3 * k)) & 7u) == y)) ? 1 : 0;
    return n;
}
PTB_HD uint32_t mr_set(uint32_t tags, int k, uint32_t t) { return (tags & ~(7u << (3 * k))) | (t << (3 * k)); }

// wide_begin: returns the ray's first tag
// keep_hit: a follow-up any-hit traversal that leaves the previous closest hit's (ud, vd, det, slot) in place (an any-hit query only
// writes MR_HINST)
template <class R>
PTB_HD uint32_t mr_begin(const R& r, const SceneDev& S, int which, V3 o, V3 d, float t_max, bool any, bool keep_hit = false)
{
    TravRay tr;
    set_ray(tr, o, d);
    r.st3(MR_O, o); r.st3(MR_D, d); r.st3(MR_INV, tr.inv); r.st3(MR_INVW, tr.inv);
    r.stf(MR_TMAX, t_max);
    r.st(MR_NODE, S.tlas[which].wide_root);
    r.st(MR_GY, 0u);
    r.st(MR_SP, 0u);
    r.st(MR_INST, TRAV_NONE);
    r.st(MR_FL, (tr.oi << 16) | (tr.oi << 19) | (any ? MR_FL_ANY : 0u));
    if (!keep_hit)
    {
        r.stf(MR_HUD, 0.0f); r.stf(MR_HVD, 0.0f); r.stf(MR_HDET, 1.0f);
        r.st(MR_HSLOT, 0u);
    }
    r.st(MR_HINST, TRAV_NONE);
    return MT_N;
}
template <class R> PTB_HD bool mr_hit(const R& r) { return r.ld(MR_HINST) != TRAV_NONE; }
template <class R> PTB_HD float mr_hit_u(const R& r) { return ex::div(r.ldf(MR_HUD), r.ldf(MR_HDET)); }
template <class R> PTB_HD float mr_hit_v(const R& r) { return ex::div(r.ldf(MR_HVD), r.ldf(MR_HDET)); }

// ---- P: next node from the current group, or pop (wide_step_pop without the instance entry) ------------------------------------
// Job::world_ray(r, &o, &d): the ray's world-space origin and direction as they were given to mr_begin (re-read from the job's source)
template <class R, class Job>
PTB_HD uint32_t mr_stage_p(const R& r, uint2* stk, const Job& job)
{
    uint32_t gx = r.ld(MR_GX), gy = r.ld(MR_GY);
    const uint32_t spw = r.ld(MR_SP);
    uint32_t sp = spw & 0xffu;
    uint32_t fl = r.ld(MR_FL);
    if (gy == 0u)
    {
        if ((fl & MR_FL_BLAS) && sp == ((spw >> 8) & 0xffu))
        {
            // leave the instance: back to the world ray
            fl = (fl & ~(MR_FL_BLAS | (7u << 16))) | (((fl >> 19) & 7u) << 16);
            r.st(MR_FL, fl);
            r.st(MR_INST, TRAV_NONE);
            V3 o, d;
            job.world_ray(r, &o, &d);
            r.st3(MR_O, o); r.st3(MR_D, d); r.st3(MR_INV, r.ld3(MR_INVW));
        }
        if (sp == 0u) return MT_DONE;
        const uint2 e = stk[--sp];
        if (e.y & 0x80000000u)
        {
            r.st(MR_NODE, e.x);
            r.st(MR_SP, (spw & ~0xffu) | sp);
            return MT_I;
        }
        gx = e.x; gy = e.y;
    }
    // descend into the nearest remaining child of the group
    const uint32_t ih = gy & 0xffu, imask = (gy >> 8) & 0xffu;
    const int p = hi_bit(ih);
    const uint32_t sl = (uint32_t)p ^ ((fl >> 16) & 7u);
    r.st(MR_NODE, gx + popc32(imask & ((1u << sl) - 1u)));
    gy &= ~(1u << p);
    if (gy & 0xffu) { uint2 e; e.x = gx; e.y = gy; stk[sp++] = e; }
    r.st(MR_SP, (spw & ~0xffu) | sp);
    return MT_N;
}

// ---- I: TLAS leaf (tlas.rs:88-100 / :132-139): exact leaf box with the current t_max, then Ray::transform ------------------------
template <class R>
PTB_HD uint32_t mr_stage_i(const R& r, const SceneDev& S)
{
    const float4* lp = S.leaf_headers + (size_t)r.ld(MR_NODE) * LEAF_F4;
    const float4 h0 = PTB_LDG4F(lp), h1 = PTB_LDG4F(lp + 1), i0 = PTB_LDG4F(lp + 2), i1 = PTB_LDG4F(lp + 3), i2 = PTB_LDG4F(lp + 4);
    const V3 o = r.ld3(MR_O), inv = r.ld3(MR_INV);  // the world ray: instances are entered from the TLAS only
    float t_enter;
    PTB_STAT(3);
    if (!ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), o, inv, r.ldf(MR_TMAX), &t_enter)) return MT_P;
    PTB_STAT(4);
    const uint32_t cw = ex::bits(h0.w);
    const float inv12[12] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w, i2.x, i2.y, i2.z, i2.w};
    TravRay out;
    transform_ray_reuse(out, o, r.ld3(MR_D), inv, inv12);
    r.st3(MR_O, out.o); r.st3(MR_D, out.d); r.st3(MR_INV, out.inv);
    r.st(MR_INST, ex::bits(h1.w));
    const uint32_t spw = r.ld(MR_SP);
    r.st(MR_SP, (spw & 0xffu) | ((spw & 0xffu) << 8));  // blas_sp = sp
    const uint32_t fl = (r.ld(MR_FL) & ~((7u << 16) | 0xffffu)) | (out.oi << 16) | MR_FL_BLAS;  // G.y is 0: the entry was popped
    if (cw & TLAS_DIRECT_FLAG)
    {
        // a BLAS of one or two leaves: straight to the exact leaf tests
        const uint32_t lmask = (cw >> TLAS_DIRECT_BASE_BITS) & 0xffu;
        const uint32_t lh = xor_permute16(lmask, out.oi) & 0xffu;
        r.st(MR_NODE, cw & ((1u << TLAS_DIRECT_BASE_BITS) - 1u));
        r.st(MR_FL, fl | lh | (lmask << 8));
        return lh ? MT_L : MT_P;
    }
    r.st(MR_FL, fl);
    r.st(MR_NODE, cw);
    return MT_N;
}

// ---- N: fetch and test one wide node --------------------------------------------------------------------------------------------
template <class R>
PTB_HD uint32_t mr_stage_n(const R& r, const SceneDev& S, uint2* stk)
{
    const uint4* np = S.wide + (size_t)r.ld(MR_NODE) * 5;
    const uint4 q0 = PTB_LDG4U(np), q1 = PTB_LDG4U(np + 1), q2 = PTB_LDG4U(np + 2), q3 = PTB_LDG4U(np + 3), q4 = PTB_LDG4U(np + 4);
    uint32_t fl = r.ld(MR_FL);
    TravRay ray;
    ray.o = r.ld3(MR_O);
    const V3 inv = r.ld3(MR_INV);
    ray.idq = ex::v3(finite_recip_of(inv.x), finite_recip_of(inv.y), finite_recip_of(inv.z));
    ray.oi = (fl >> 16) & 7u;
    PTB_STAT((fl & MR_FL_BLAS) ? 0 : 5);
    const uint32_t hits = wide_node_test(q0, q2, q3, q4, ray, r.ldf(MR_TMAX));
    const uint32_t imask = q0.w >> 24, lmask = q1.z & 0xffu;
    const uint32_t perm = xor_permute16((hits & imask) | ((hits & lmask) << 8), ray.oi);
    const uint32_t ih = perm & 0xffu, lh = (perm >> 8) & 0xffu;
    const uint32_t gy = ih ? (ih | (imask << 8)) : 0u;
    if (fl & MR_FL_BLAS)
    {
        // triangle leaves: processed front to back before any inner child
        r.st(MR_GX, q1.x);
        r.st(MR_GY, gy);
        r.st(MR_FL, (fl & 0xffff0000u) | lh | (lmask << 8));
        r.st(MR_NODE, q1.y);
        return lh ? MT_L : MT_P;
    }
    // TLAS leaves: push the inner group first, then the instances far to near (nearest popped first)
    const uint32_t spw = r.ld(MR_SP);
    uint32_t sp = spw & 0xffu;
    if (gy) { uint2 e; e.x = q1.x; e.y = gy; stk[sp++] = e; }
    uint32_t m = lh;
    while (m)
    {
        const int p = lo_bit(m);
        m &= m - 1u;
        const uint32_t sl = (uint32_t)p ^ ray.oi;
        uint2 e;
        e.x = q1.y + popc32(lmask & ((1u << sl) - 1u));
        e.y = 0x80000000u;
        stk[sp++] = e;
    }
    r.st(MR_SP, (spw & ~0xffu) | sp);
    r.st(MR_GY, 0u);
    return MT_P;
}

// ---- L: one reference leaf of the current BLAS node: exact box, exact t_enter, its triangles (reference leaves hold one or two) ------
template <int MODE, class R>
PTB_HD uint32_t mr_stage_l(const R& r, const SceneDev& S)
{
    uint32_t fl = r.ld(MR_FL);
    const bool ANY = MODE == 2 ? ((fl & MR_FL_ANY) != 0u) : (MODE == 1);
    const V3 o = r.ld3(MR_O);
    float t_max = r.ldf(MR_TMAX);
    const uint32_t lh0 = fl & 0xffu, lmask = (fl >> 8) & 0xffu;
    const int p = hi_bit(lh0);
    fl &= ~(1u << p);
    r.st(MR_FL, fl);
    const uint32_t sl = (uint32_t)p ^ ((fl >> 16) & 7u);
    const uint32_t hidx = r.ld(MR_NODE) + popc32(lmask & ((1u << sl) - 1u));
    const float4* lp = S.leaf_headers + (size_t)hidx * LEAF_F4;
    const float4 h0 = PTB_LDG4F(lp), h1 = PTB_LDG4F(lp + 1);
    float4 a = PTB_LDG4F(lp + 2), b = PTB_LDG4F(lp + 3), c = PTB_LDG4F(lp + 4);  // planes of the leaf's first triangle, fetched with its header
    const uint32_t cw = ex::bits(h0.w);
    float t_enter = 0.0f;
    bool go = true;
    PTB_STAT(1);
    // blas.rs:137-138 / :264 — exact reference leaf box, exact entry distance (the BLAS root is entered untested with t_estimate = 0 by
    // the closest-hit query, blas.rs:217)
    if (ANY || !(cw & LEAF_ROOT_FLAG)) go = ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), o, r.ld3(MR_INV), t_max, &t_enter);
    if (go)
    {
        const V3 d = r.ld3(MR_D);
        uint32_t tri_i = ex::bits(h1.w);
        const uint32_t tri_end = tri_i + (cw & ~LEAF_ROOT_FLAG);
        for (;;)
        {
            PTB_STAT(2);
            const float n0[4] = {a.x, a.y, a.z, a.w}, n1[4] = {b.x, b.y, b.z, b.w}, n2[4] = {c.x, c.y, c.z, c.w};
            if (ANY)
            {
                if (ex::tri_hit_bool(n0, n1, n2, o, d, t_max, t_enter))
                {
                    r.st(MR_HINST, r.ld(MR_INST));
                    return MT_DONE;
                }
            }
            else
            {
                float t, raw[3];
                if (ex::tri_hit_raw(n0, n1, n2, o, d, t_max, t_enter, &t, raw))
                {
                    t_max = t;
                    r.stf(MR_TMAX, t);
                    r.stf(MR_HUD, raw[0]); r.stf(MR_HVD, raw[1]); r.stf(MR_HDET, raw[2]);
                    r.st(MR_HSLOT, tri_i);
                    r.st(MR_HINST, r.ld(MR_INST));
                }
            }
            if (++tri_i >= tri_end) break;
            const float4* pp = S.tri_planes + (size_t)tri_i * 3;
            a = PTB_LDG4F(pp); b = PTB_LDG4F(pp + 1); c = PTB_LDG4F(pp + 2);
        }
    }
    return (fl & 0xffu) ? MT_L : MT_P;
}

#if defined(__CUDACC__)
// ---- the persistent loop ----------------------------------------------------------------------------------------------------------
// Job interface (per lane; R = MrRay of one of the lane's rays, job words from MR_JOB on):
//   static constexpr int JOBW                         words of per-ray job data
//   uint32_t start(uint32_t index, const R& r)        fetch work item `index`; returns the ray's first tag (mr_begin) or MT_IDLE
//   uint32_t finish(const R& r, bool mine)            called by ALL lanes of the warp together; `mine` = r is a finished ray of this lane.
//                                                     Returns the tag of a follow-up traversal (mr_begin) or MT_IDLE.
//   void world_ray(const R& r, V3* o, V3* d)          the world-space ray given to mr_begin, re-read from the job's source
// free_quorum: finished / idle ray slots of the warp that trigger a finish + start pass; qn / ql / qi: lanes that must hold a ray
// waiting for the node / leaf / instance stage before the warp runs it (all stages run when none reaches its quorum).
template <int MODE, int K, class Job>
__device__ __forceinline__ void mr_persistent_trace(const SceneDev& S, Job& job, uint32_t* cursor, uint32_t total, int free_quorum, int qn, int ql, int qi)
{
    constexpr int NW = MR_JOB + Job::JOBW;
    constexpr int STRIDE = 128;
    using R = MrRay<STRIDE>;
    extern __shared__ uint32_t mr_smem[];
    uint32_t* const col = mr_smem + threadIdx.x;
    uint2 lstack[K][WIDE_STACK_CAP];
    auto ray = [&](int k) { return R{col + k * (NW * STRIDE)}; };
    uint32_t tags = 0u;
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    constexpr uint32_t BATCH = 64;
    const uint32_t n_warps = gridDim.x * blockDim.x >> 5;
    const uint32_t fine_below = n_warps * (2u * BATCH);
    const uint32_t fair = max(1u, (total + n_warps * K - 1u) / (n_warps * K));
    uint32_t wbase = 0, wend = 0, seen = 0;
    bool global_more = total > 0u;
    int n_free = 32 * K, n_done = 0;  // warp-uniform: ray slots that are idle or finished, and the finished ones among them
    for (;;)
    {
        // ---- finish + start --------------------------------------------------------------------------------------------------------
        bool more = wbase < wend || global_more;
        if (more ? (n_free >= free_quorum) : (n_done > 0 && (n_done >= 8 || n_free == 32 * K)))
        {
            if (n_done)
            {
                const int k = mr_first<K>(tags, MT_DONE);
                const uint32_t t = job.finish(ray(k < 0 ? 0 : k), k >= 0);
                if (k >= 0) tags = mr_set(tags, k, t);
                const int again = __popc(__ballot_sync(FULL, k >= 0 && t != MT_IDLE)), fin = __popc(__ballot_sync(FULL, k >= 0));
                n_done -= fin;
                n_free -= again;
            }
            for (int pass = 0; pass < K && (wbase < wend || global_more); ++pass)
            {
                const int k = mr_first<K>(tags, MT_IDLE);
                const unsigned idle = __ballot_sync(FULL, k >= 0);
                if (!idle) break;
                if (wbase >= wend)
                {
                    const uint32_t want = (total - seen > fine_below) ? BATCH : min((uint32_t)__popc(idle), fair);
                    uint32_t base = 0;
                    if (lane == 0u) base = atomicAdd(cursor, want);
                    base = __shfl_sync(FULL, base, 0);
                    if (base >= total) { global_more = false; wbase = wend = 0u; break; }
                    wbase = base; wend = min(base + want, total); seen = wend;
                    if (wend == total) global_more = false;
                }
                const uint32_t avail = wend - wbase, n = (uint32_t)__popc(idle);
                const uint32_t rank = (uint32_t)__popc(idle & ((1u << lane) - 1u));
                uint32_t t = MT_IDLE;
                if (k >= 0 && rank < avail) tags = mr_set(tags, k, t = job.start(wbase + rank, ray(k)));
                n_free -= __popc(__ballot_sync(FULL, t != MT_IDLE));
                wbase += min(n, avail);
            }
            more = wbase < wend || global_more;
        }
        if (n_free == 32 * K)
        {
            if (!more && n_done == 0) break;
            continue;
        }
        // ---- P ---------------------------------------------------------------------------------------------------------------------
        unsigned fin = 0u;
        {
            const int k = mr_first<K>(tags, MT_P);
            if (k >= 0)
            {
                const uint32_t t = mr_stage_p(ray(k), lstack[k], job);
                tags = mr_set(tags, k, t);
                fin = t == MT_DONE ? 1u : 0u;
            }
        }
        // ---- which stages run ------------------------------------------------------------------------------------------------------
        const int cn = __popc(__ballot_sync(FULL, mr_has<K>(tags, MT_N)));
        const int cl = __popc(__ballot_sync(FULL, mr_has<K>(tags, MT_L)));
        const int ci = __popc(__ballot_sync(FULL, mr_has<K>(tags, MT_I)));
        bool run_n = cn >= qn, run_l = cl >= ql, run_i = ci >= qi;
        if (!(run_n && cn) && !(run_l && cl) && !(run_i && ci)) run_n = run_l = run_i = true;
        if (run_i && ci)
        {
            const int k = mr_first<K>(tags, MT_I);
            if (k >= 0) tags = mr_set(tags, k, mr_stage_i(ray(k), S));
        }
        if (run_n && cn)
        {
            const int k = mr_first<K>(tags, MT_N);
            if (k >= 0) tags = mr_set(tags, k, mr_stage_n(ray(k), S, lstack[k]));
        }
        if (run_l)
        {
            const int k = mr_first<K>(tags, MT_L);
            if (k >= 0)
            {
                const uint32_t t = mr_stage_l<MODE>(ray(k), S);
                tags = mr_set(tags, k, t);
                fin += t == MT_DONE ? 1u : 0u;
            }
        }
        const int f = (int)__reduce_add_sync(FULL, fin);
        n_done += f;
        n_free += f;
    }
}
#endif

}  // namespace ptb
