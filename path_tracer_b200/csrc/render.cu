// render.cu — the frame loop of main.rs:181-207 as wavefront kernels over per-slot path state (wavefront.h holds the
// per-slot bodies) plus the camera / jitter / RNG probes of the C ABI.
//
// One wavefront iteration = three persistent kernels (grid = SMs x blocks_per_sm; every warp pulls chunks of 32 queue
// entries from a cursor):
//   k_extend  (a) continuing paths: closest hit on the world TLAS; (b) slots whose path ended in the previous iteration:
//             finalise + accumulate the finished sample (single writer per pixel, no atomics), raygen the next sample
//             and trace its primary ray in the same thread.  Pushes the slot into the shade queue of its material kind.
//   k_shade   per-material queues processed kind by kind (a warp never mixes kinds); emits shadow / probe entries and
//             the continue / terminated queues of the next iteration
//   k_nee     both direct-light estimates of a bounce in one thread: any-hit for the explicit light sample, lights-TLAS
//             closest hit + world any-hit for the BSDF-sampled one; one read-modify-write of the slot's radiance
// Queue counters live in four rotating sets (iteration it consumes set it%4, fills set (it+1)%4 and clears set (it+2)%4),
// so no separate bookkeeping launch is needed.  Compaction between stages is warp ballot / match + one atomicAdd per warp
// and queue.  When few paths remain (the tail of 1024-bounce paths) a single k_tail launch runs every remaining
// slot to completion, one thread per slot, instead of thousands of near-empty iterations.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "camera_host.h"
#include "context.h"
#include "multiray.cuh"
#include "persist.cuh"
#include "render_state.h"
#include "wavefront.h"

namespace ptb {

constexpr uint32_t WF_SETS = 4;  // rotating counter sets: iteration it consumes set it % 4, fills (it + 1) % 4, clears (it + 2) % 4; the fourth lets k_nee of iteration
                                // it (still reading set it % 4) run beside k_extend of iteration it + 1 (which clears set (it + 3) % 4)
struct WfSet  // 64 B
{
    uint32_t n_cont, n_term;  // consumed by k_extend (polled by the host: keep first)
    uint32_t n_nee, n_grp;    // n_grp: slot-ordered mode, live 32-slot groups listed for this iteration (polled with the two above)
    uint32_t n_mat[8];
    uint32_t cur_extend, cur_shade, cur_nee, cur_regen;  // cur_regen: cursor of the regeneration part when k_extend is launched in two parts
};
struct WfCounters
{
    WfSet set[WF_SETS];
    unsigned long long stat_extend, stat_shadow, stat_probe, stat_paths, stat_iters, stat_probe_shadow, stat_settled, stat_extend_path;
};

struct Queues
{
    uint32_t* cont[2];  // continuing paths, double-buffered by iteration parity
    uint32_t* term[2];  // slots whose path ended (to be finalised / regenerated)
    uint32_t* mat;      // 7 queues of npix entries each
    uint32_t* nee;      // slot | NEE_SHADOW_BIT | NEE_PROBE_BIT
    uint32_t* grp[2];   // slot-ordered mode: ascending list of the 32-slot groups that still hold a live slot, double-buffered like cont/term
    uint32_t* state;    // slot-ordered mode: per-slot WF_* flags of the last shade (WF_NEXT / WF_TERM / 0 = no samples left, | WF_SHADOW | WF_PROBE)
    uint32_t npix;       // queue capacity and stride of the material queues: the number of slots of this wavefront group
    uint32_t slot_base;  // the group's slots are [slot_base, slot_end); slot_base is a multiple of 32
    uint32_t slot_end;
    WfCounters* c;
};

constexpr uint32_t POLL_RING = 8192;
constexpr int MAX_GROUPS = 8;  // independent wavefronts per render (wavefront_groups), each on its own stream
constexpr uint32_t PATH_AUTO_PER_LANE = 8;  // automatic path_threshold: live slots per resident lane of k_path at which it takes over

struct RenderState
{
    uint32_t W = 0, H = 0, npix = 0;
    DBuf<float4> accum_own, accum_sq, position;   // per pixel
    DBuf<float4> s_accum, s_accum_sq, s_position;  // per slot (K slots per pixel)
    DBuf<uint32_t> s_id;
    uint32_t n_slots = 0;
    DBuf<RayRec> ray;
    DBuf<PathRec> path;
    DBuf<HitRec> hit;
    DBuf<ShadowRec> shadow;
    DBuf<ProbeRec> probe;
    DBuf<uint32_t> id, q_cont0, q_cont1, q_term0, q_term1, q_mat, q_nee, slot_state, q_grp0, q_grp1;
    DBuf<unsigned long long> volset;
    DBuf<uint2> tile_lists;  // per 8 x 4 pixel tile: the TLAS leaves its camera rays can reach (k_tile_lists)
    DBuf<WfCounters> counters;
    float4* accum = nullptr;  // bound or own
    uint32_t* h_flags = nullptr;  // pinned, MAX_GROUPS x POLL_RING x 4 (n_cont, n_term, n_nee, n_grp of every iteration of every group)
    std::vector<cudaStream_t> group_streams;  // streams of the wavefront groups 1.. (group 0 runs on the context stream)
    std::vector<cudaEvent_t> group_events;    // fork / join events
    std::vector<cudaStream_t> nee_streams;    // per group: the stream k_nee runs on while it overlaps the next iteration's k_extend
    std::vector<cudaEvent_t> nee_events;      // per group: [2g] shade done, [2g + 1] nee done
    uint32_t log_iters = 0;       // iterations recorded in h_flags by the last ptb_render
    std::vector<cudaEvent_t> poll_events;
    std::vector<cudaEvent_t> prof_events;
    // host-side statistics
    unsigned long long launches = 0;
    float t_total_ms = 0;
    float t_class_ms[8] = {0};
    uint32_t n_class[8] = {0};
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    std::vector<std::pair<int, float>> launch_log;  // (stage class, microseconds) of every launch of the last profiled ptb_render
    int tail_capacity = -1;  // slots the tail kernel (tail_kind: 1 = k_tail_roles, 0 = k_tail) keeps resident on this device
    int tail_kind = -1;
};

// ---- device helpers ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void warp_push(uint32_t* q, uint32_t* counter, bool pred, uint32_t value)
{
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0u) return;
    const unsigned lane = threadIdx.x & 31u;
    const int leader = __ffs((int)m) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (pred) q[base + __popc(m & ((1u << lane) - 1u))] = value;
}
// push into one of several queues selected per lane (key < 8; key == 0xff: no push): one atomicAdd per distinct key
__device__ __forceinline__ void warp_push_keyed(uint32_t* q, size_t stride, uint32_t* counters, uint32_t key, uint32_t value)
{
    const unsigned grp = __match_any_sync(0xffffffffu, key);
    const unsigned lane = threadIdx.x & 31u;
    const int leader = __ffs((int)grp) - 1;
    uint32_t base = 0;
    if ((int)lane == leader && key != 0xffu) base = atomicAdd(counters + key, (uint32_t)__popc(grp));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (key != 0xffu) q[(size_t)key * stride + base + __popc(grp & ((1u << lane) - 1u))] = value;
}
__device__ __forceinline__ bool fetch_chunk(uint32_t* cursor, uint32_t nchunks, uint32_t* chunk_out)
{
    uint32_t c = 0;
    if ((threadIdx.x & 31u) == 0u) c = atomicAdd(cursor, 1u);
    c = __shfl_sync(0xffffffffu, c, 0);
    *chunk_out = c;
    return c < nchunks;
}

__global__ void __launch_bounds__(128) k_init(RenderView V, Queues Q, const uint32_t* id_px)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= Q.npix) return;  // Q.npix = number of slots of this group
    const uint32_t i = Q.slot_base + j;
    V.ray[i].o.w = 0.0f;  // no sample in flight
    V.ray[i].d.w = 0.0f;  // samples started = 0
    V.accum[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (V.accum_sq) V.accum_sq[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    V.id[i] = V.K == 1u ? id_px[slot_pixel(V, i)] : 0u;  // one slot per pixel: the slot continues the pixel's id history (main.rs:206)
    Q.term[0][j] = i;
    Q.state[i] = WF_TERM;
    if ((j & 31u) == 0u) Q.grp[0][j >> 5] = i >> 5;
    if (j == 0)
    {
        for (uint32_t s = 0; s < WF_SETS; ++s) Q.c->set[s] = WfSet{};
        Q.c->set[0].n_term = Q.npix;
        Q.c->set[0].n_grp = (Q.npix + 31u) >> 5;
    }
}

// Fold the K per-slot partial results of every pixel into the per-pixel outputs, in slot order (deterministic):
// accum += sum_k accum_k (accumulate.wgsl:23); position / id of the LAST sample, id history shifted as main.rs:206.
__global__ void __launch_bounds__(256) k_combine(RenderView V, float4* accum_px, float4* accum_sq_px, float4* position_px, uint32_t* id_px, uint32_t sample_count)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= V.npix) return;
    const uint32_t j = pixel_slot(V, p);  // the pixel's slot index within a layer
    float4 a = accum_px[p];
    for (uint32_t k = 0; k < V.K; ++k)
    {
        const float4 b = V.accum[(size_t)k * V.npix + j];
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    accum_px[p] = a;
    if (accum_sq_px)
    {
        float4 q = accum_sq_px[p];
        for (uint32_t k = 0; k < V.K; ++k)
        {
            const float4 b = V.accum_sq[(size_t)k * V.npix + j];
            q.x += b.x; q.y += b.y; q.z += b.z; q.w += b.w;
        }
        accum_sq_px[p] = q;
    }
    const uint32_t k_last = (sample_count - 1u) % V.K;
    position_px[p] = V.position[(size_t)k_last * V.npix + j];
    if (V.K == 1u) id_px[p] = V.id[j];
    else
    {
        const uint32_t k_prev = (sample_count - 2u) % V.K;  // K <= sample_count, so sample_count >= 2 here
        id_px[p] = ((V.id[(size_t)k_prev * V.npix + j] & 0xffffu) << 16) | (V.id[(size_t)k_last * V.npix + j] & 0xffffu);
    }
}

// ---- candidate lists of the camera rays, one per 8 x 4 pixel tile ------------------------------------------------------------------
// Camera rays are half of the closest-hit rays of an instanced scene (C3: 132.7 M of 270 M per step), and every one of them walks the
// TLAS from its root (3.7 node tests of 7.8) although the 32 rays of a warp leave one point through one small patch of the image.  Per
// render call, one thread per tile walks the world TLAS once with the tile's frustum — the four rays through the corners of its pixel
// area, jitter included — and lists the TLAS leaves (instances) the frustum touches, nearest first, each with a lower bound of the
// distance at which a ray can enter it.  A regenerated slot's ray then starts from its tile's list (wide_begin_list, traverse.h)
// instead of the TLAS root.  The tests here only cull, with margins; every listed leaf is still entered through its exact box test.
__global__ void __launch_bounds__(128) k_tile_lists(RenderView V, uint2* lists, uint32_t n_tiles)
{
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile < n_tiles) tile_list_build(V, tile, lists + (size_t)tile * TILE_LIST_STRIDE);  // wavefront.h
}

// ---- extend: dynamic-fetch persistent kernel over the 8-wide layout --------------------------------------------------------
struct ExtendJob
{
    const RenderView& V;
    const Queues& Q;
    WfSet& S;
    uint32_t b, n_cont, n_term;
    int sort;  // 1: push the slot to the shade queue of its material kind, 0: to queue 0, -1: no push (k_shade_ordered walks the slots itself)
    uint32_t slot = 0;
    unsigned long long paths = 0, settled = 0;
    uint32_t idx0 = 0;  // first work item of this launch (two-part launches: the regeneration part starts at n_cont)

    __device__ __forceinline__ bool start(uint32_t idx, WideTrav& tr)
    {
        V3 o, d;
        idx += idx0;
        if (idx < n_cont)
        {
            slot = Q.cont[b][idx];
            const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
            o = ex::v3(o4.x, o4.y, o4.z);
            d = ex::v3(d4.x, d4.y, d4.z);
        }
        else
        {
            slot = Q.term[b][idx - n_cont];
            for (;;)
            {
                const uint32_t f = regen_slot(V, slot, &o, &d) & WF_NEXT;
                if (!f) { Q.state[slot] = 0u; return false; }  // out of samples
                ++paths;
                // A camera outside the scene bounds: a primary ray that cannot reach them is a miss without a traversal, and with the
                // constant background the whole path is that miss (integrator.rs:254-269 at b == 0: radiance = background * 1).  Settle
                // it here and start the slot's next sample at once; background pixels then never occupy the later kernels.
                // (Slot-ordered shading only: the queue shade kernel expects every traced slot in one of its queues.)
                if (sort >= 0 || V.S.env != nullptr) break;
                // ... and so is a camera ray whose tile's list of TLAS leaves is empty (k_tile_lists: no instance box touches the tile's frustum);
                // option tile_settle, off by default: such a ray then leaves the traced count, which the roofline figure is stated on
                if (tlas_may_hit(V.S, 0, o, d) && !(V.tile_settle && V.S.tile_lists && V.S.tile_lists[((slot % V.npix) >> 5) * TILE_LIST_STRIDE].x == 0u)) break;
                V.path[slot].radiance = f4(0.006f, 0.006f, 0.006f, 0.0f);
                ++settled;
            }
            Q.state[slot] = WF_NEXT;  // live: a ray is being traced for it this iteration
            if (V.S.tile_lists)
            {
                // a camera ray: from its tile's list of TLAS leaves instead of the TLAS root (k_tile_lists)
                const uint32_t first = ((slot % V.npix) >> 5) * TILE_LIST_STRIDE;
                const uint32_t n = V.S.tile_lists[first].x;
                if (n != TILE_LIST_NONE)
                {
                    wide_begin_list(tr, V.S, o, d, ex::inf(), first + 1u, first + 1u + n);
                    return true;
                }
            }
        }
        wide_begin(tr, V.S, 0, o, d, ex::inf(), false);
        return true;
    }
    __device__ __forceinline__ bool finish(WideTrav& tr, bool mine)
    {
        uint32_t kind = 0xffu;
        if (mine)
        {
            if (tr.hit_inst == TRAV_NONE)
            {
                V.hit[slot].info = 0xffffffffu;
                V.hit[slot].kind = kind = Q_MISS;
            }
            else
            {
                const uint32_t prim = V.S.tri_prim[tr.hit_slot];
                V3 n;
                bool ff;
                finish_hit(V.S, tr.dw, tr.hit_inst, prim, wide_hit_u(tr), wide_hit_v(tr), &n, &ff);
                V.hit[slot].tn = f4(tr.t_max, n.x, n.y, n.z);
                V.hit[slot].info = tr.hit_inst | (ff ? 0x80000000u : 0u);
                V.hit[slot].kind = kind = V.S.mat[V.S.inst[tr.hit_inst].blas].kind;
            }
            if (sort == 0) kind = 0u;
        }
        if (sort >= 0) warp_push_keyed(Q.mat, Q.npix, S.n_mat, kind, slot);
        return false;
    }
};

// part 0: the whole queue (continuing slots, then the slots to regenerate); part 1: the continuing slots only; part 2: the regeneration
// part only.  Two-part launches: (a) with the multiray kernels — bounce rays two per lane (k_extend_mr, part 1), primary rays one per lane;
// (b) thin iterations — part 1 of iteration it + 1 runs beside k_nee of iteration it (a continuing slot's ray does not depend on its
// pending direct-light rays; a slot to regenerate does: its sample is complete only when they have been added)
template <int MB, int SSTACK = 0>
__global__ void __launch_bounds__(128, MB) k_extend(RenderView V, Queues Q, uint32_t it, int sort, int refill_below, int node_quorum, int part = 0)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    const uint32_t n_cont = S.n_cont, n_term = S.n_term;
    if (part != 2 && blockIdx.x == 0 && threadIdx.x == 0)
    {
        Q.c->set[(it + 2u) % WF_SETS] = WfSet{};  // consumed in iteration it-1, refilled from iteration it+1 on
        if (n_cont + n_term) Q.c->stat_iters += 1;
    }
    ExtendJob job{V, Q, S, it & 1u, n_cont, n_term, sort};
    if (part == 2) job.idx0 = n_cont;
    if (V.S.tile_lists) persistent_trace<0, ExtendJob, false, SSTACK, true>(V.S, job, part == 2 ? &S.cur_regen : &S.cur_extend, part == 2 ? n_term : (part == 1 ? n_cont : n_cont + n_term), refill_below, node_quorum);
    else persistent_trace<0, ExtendJob, false, SSTACK>(V.S, job, part == 2 ? &S.cur_regen : &S.cur_extend, part == 2 ? n_term : (part == 1 ? n_cont : n_cont + n_term), refill_below, node_quorum);
    const unsigned long long p = warp_sum(job.paths), q = warp_sum(job.settled);
    if ((threadIdx.x & 31u) == 0u && p) atomicAdd(&Q.c->stat_paths, p);
    if ((threadIdx.x & 31u) == 0u && q) atomicAdd(&Q.c->stat_settled, q);  // primary rays settled in start() without a traversal: counted apart
}

// reference-order variant (binary trees, reference visit order): static chunks of 32, one ray per lane
__global__ void __launch_bounds__(128) k_extend_bin(RenderView V, Queues Q, uint32_t it)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    const uint32_t b = it & 1u;
    const uint32_t n_cont = S.n_cont, n_term = S.n_term;
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        Q.c->set[(it + 2u) % WF_SETS] = WfSet{};
        if (n_cont + n_term) Q.c->stat_iters += 1;
    }
    const uint32_t total = n_cont + n_term;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t c;
    while (fetch_chunk(&S.cur_extend, (total + 31u) >> 5, &c))
    {
        const uint32_t i = c * 32u + lane;
        uint32_t slot = 0;
        bool valid = i < total;
        V3 o = ex::v3(0, 0, 0), d = ex::v3(0, 0, 1);
        bool fresh = false;
        if (valid)
        {
            if (i < n_cont)
            {
                slot = Q.cont[b][i];
                const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
                o = ex::v3(o4.x, o4.y, o4.z);
                d = ex::v3(d4.x, d4.y, d4.z);
            }
            else
            {
                slot = Q.term[b][i - n_cont];
                valid = fresh = (regen_slot(V, slot, &o, &d) & WF_NEXT) != 0u;
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, fresh);
        if (lane == 0u && m) atomicAdd(&Q.c->stat_paths, (unsigned long long)__popc(m));
        const uint32_t kind = valid ? extend_ray_t<true>(V, slot, o, d) : 0xffu;
        warp_push_keyed(Q.mat, Q.npix, S.n_mat, kind, slot);
    }
}

__global__ void __launch_bounds__(128) k_shade(RenderView V, Queues Q, uint32_t it)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    WfSet& N = Q.c->set[(it + 1u) % WF_SETS];
    const uint32_t nb = (it + 1u) & 1u;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t cnt[7], chunks = 0, total = 0;
#pragma unroll
    for (int k = 0; k < 7; ++k) { cnt[k] = S.n_mat[k]; chunks += (cnt[k] + 31u) >> 5; total += cnt[k]; }
    if (blockIdx.x == 0 && threadIdx.x == 0) Q.c->stat_extend += total;
    uint32_t c;
    while (fetch_chunk(&S.cur_shade, chunks, &c))
    {
        uint32_t kind = 0, first = c;
#pragma unroll
        for (int k = 0; k < 7; ++k)
        {
            const uint32_t ck = (cnt[k] + 31u) >> 5;
            if (kind == (uint32_t)k && first >= ck) { first -= ck; kind = k + 1; }
        }
        const uint32_t i = first * 32u + lane;
        const bool valid = i < cnt[kind];
        const uint32_t slot = valid ? Q.mat[(size_t)kind * Q.npix + i] : 0u;
        const uint32_t flags = valid ? shade_slot(V, slot) : 0u;
        warp_push(Q.nee, &S.n_nee, (flags & (WF_SHADOW | WF_PROBE)) != 0u, slot | ((flags & WF_SHADOW) ? NEE_SHADOW_BIT : 0u) | ((flags & WF_PROBE) ? NEE_PROBE_BIT : 0u));
        warp_push(Q.cont[nb], &N.n_cont, (flags & WF_NEXT) != 0u, slot);
        warp_push(Q.term[nb], &N.n_term, (flags & WF_TERM) != 0u, slot);
        const unsigned mo = __ballot_sync(0xffffffffu, (flags & WF_PROBE_OUT) != 0u);
        if (lane == 0u && mo) atomicAdd(&Q.c->stat_probe, (unsigned long long)__popc(mo));
    }
}

// ---- nee: both direct-light estimates of a bounce as a chain of up to three traversals per lane -----------------------------------
//   stage 0  world any-hit towards the explicit light sample            (integrator.rs:56)
//   stage 1  lights-TLAS closest hit along the BSDF-sampled direction    (integrator.rs:100)
//   stage 2  world any-hit up to that light                              (integrator.rs:103)
struct NeeJob
{
    const RenderView& V;
    const Queues& Q;
    uint32_t slot = 0, flags = 0, stage = 0;
    V3 o;
    float4 r;
    // light hit of stage 1
    float lt = 0, lu = 0, lv = 0;
    uint32_t lslot = 0, linst = 0;
    unsigned long long n_shadow = 0, n_probe = 0, n_probe_shadow = 0;

    // false: the probe direction cannot reach the bounds of the lights, nothing to trace (integrator.rs:100 returns None)
    __device__ __forceinline__ bool begin_probe(WideTrav& tr)
    {
        const float4 d4 = V.probe[slot].d;
        const V3 d = ex::v3(d4.x, d4.y, d4.z);
        ++n_probe;
        if (!tlas_may_hit(V.S, 1, o, d)) return false;
        wide_begin(tr, V.S, 1, o, d, ex::inf(), false);
        stage = 1;
        return true;
    }
    __device__ __forceinline__ bool begin_slot(uint32_t slot_, uint32_t flags_, WideTrav& tr)
    {
        slot = slot_;
        flags = flags_;
        const float4 o4 = V.ray[slot].o;
        o = ex::v3(o4.x, o4.y, o4.z);
        r = V.path[slot].radiance;
        if (flags & WF_SHADOW)
        {
            const float4 d4 = V.shadow[slot].d;
            wide_begin(tr, V.S, 0, o, ex::v3(d4.x, d4.y, d4.z), d4.w, true);
            stage = 0;
            ++n_shadow;
        }
        else return begin_probe(tr);
        return true;
    }
    __device__ __forceinline__ bool start(uint32_t idx, WideTrav& tr)
    {
        const uint32_t e = Q.nee[idx];
        return begin_slot(e & NEE_SLOT_MASK, ((e & NEE_SHADOW_BIT) ? WF_SHADOW : 0u) | ((e & NEE_PROBE_BIT) ? WF_PROBE : 0u), tr);
    }
    __device__ __forceinline__ bool finish(WideTrav& tr, bool mine)
    {
        if (!mine) return false;
        if (stage == 0)
        {
            if (tr.hit_inst == TRAV_NONE)
            {
                const float4 c = V.shadow[slot].c;
                r.x += c.x; r.y += c.y; r.z += c.z;
            }
            if ((flags & WF_PROBE) && begin_probe(tr)) return true;
        }
        else if (stage == 1)
        {
            if (tr.hit_inst != TRAV_NONE)
            {
                lt = tr.t_max; lu = wide_hit_u(tr); lv = wide_hit_v(tr); lslot = tr.hit_slot; linst = tr.hit_inst;
                wide_begin(tr, V.S, 0, o, tr.dw, lt * (1.0f - ex::EPS), true);
                stage = 2;
                ++n_probe_shadow;
                return true;
            }
        }
        else if (tr.hit_inst == TRAV_NONE)
        {
            const V3 c = probe_light_contrib(V, slot, tr.dw, lt, lu, lv, V.S.tri_prim[lslot], linst);
            r.x += c.x; r.y += c.y; r.z += c.z;
        }
        V.path[slot].radiance = r;
        return false;
    }
};

template <int MB, int SSTACK = 0>
__global__ void __launch_bounds__(128, MB) k_nee(RenderView V, Queues Q, uint32_t it, int refill_below, int node_quorum)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    NeeJob job{V, Q};
    persistent_trace<2, NeeJob, false, SSTACK>(V.S, job, &S.cur_nee, S.n_nee, refill_below, node_quorum);
    const unsigned long long a = warp_sum(job.n_shadow), b = warp_sum(job.n_probe), c = warp_sum(job.n_probe_shadow);
    if ((threadIdx.x & 31u) == 0u)
    {
        if (a) atomicAdd(&Q.c->stat_shadow, a);
        if (b) atomicAdd(&Q.c->stat_probe, b);
        if (c) atomicAdd(&Q.c->stat_probe_shadow, c);
    }
}

// ---- the same two kernels with K rays per lane and gated stages (multiray.cuh; option `multiray` = K) ------------------------------
// Jobs as above with their per-ray data in the ray record's job words; per-ray arithmetic and its order are unchanged, so frames are
// bit-identical to the one-ray-per-lane kernels'.
struct MrExtendJob
{
    static constexpr int JOBW = 1;  // [0] slot
    const RenderView& V;
    const Queues& Q;
    WfSet& S;
    uint32_t b, n_cont, n_term;
    int sort;
    unsigned long long paths = 0, settled = 0;

    template <class R>
    __device__ __forceinline__ uint32_t start(uint32_t idx, const R& r)
    {
        V3 o, d;
        uint32_t slot;
        if (idx < n_cont)
        {
            slot = Q.cont[b][idx];
            const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
            o = ex::v3(o4.x, o4.y, o4.z);
            d = ex::v3(d4.x, d4.y, d4.z);
        }
        else
        {
            slot = Q.term[b][idx - n_cont];
            for (;;)
            {
                const uint32_t f = regen_slot(V, slot, &o, &d) & WF_NEXT;
                if (!f) { Q.state[slot] = 0u; return MT_IDLE; }  // out of samples
                ++paths;
                if (sort >= 0 || V.S.env != nullptr || tlas_may_hit(V.S, 0, o, d)) break;  // see ExtendJob::start
                V.path[slot].radiance = f4(0.006f, 0.006f, 0.006f, 0.0f);
                ++settled;
            }
            Q.state[slot] = WF_NEXT;
        }
        r.st(MR_JOB, slot);
        return mr_begin(r, V.S, 0, o, d, ex::inf(), false);
    }
    template <class R>
    __device__ __forceinline__ void world_ray(const R& r, V3* o, V3* d) const
    {
        const uint32_t slot = r.ld(MR_JOB);
        const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
        *o = ex::v3(o4.x, o4.y, o4.z);
        *d = ex::v3(d4.x, d4.y, d4.z);
    }
    template <class R>
    __device__ __forceinline__ uint32_t finish(const R& r, bool mine)
    {
        uint32_t kind = 0xffu, slot = 0u;
        if (mine)
        {
            slot = r.ld(MR_JOB);
            const uint32_t hit_inst = r.ld(MR_HINST);
            if (hit_inst == TRAV_NONE)
            {
                V.hit[slot].info = 0xffffffffu;
                V.hit[slot].kind = kind = Q_MISS;
            }
            else
            {
                const uint32_t prim = V.S.tri_prim[r.ld(MR_HSLOT)];
                V3 n;
                bool ff;
                finish_hit(V.S, r.ld3(MR_D), hit_inst, prim, mr_hit_u(r), mr_hit_v(r), &n, &ff);  // a finished closest-hit ray is back in world space
                V.hit[slot].tn = f4(r.ldf(MR_TMAX), n.x, n.y, n.z);
                V.hit[slot].info = hit_inst | (ff ? 0x80000000u : 0u);
                V.hit[slot].kind = kind = V.S.mat[V.S.inst[hit_inst].blas].kind;
            }
            if (sort == 0) kind = 0u;
        }
        if (sort >= 0) warp_push_keyed(Q.mat, Q.npix, S.n_mat, kind, slot);
        return MT_IDLE;
    }
};

template <int K>
__global__ void __launch_bounds__(128) k_extend_mr(RenderView V, Queues Q, uint32_t it, int sort, MrOpts mo, int part)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    const uint32_t n_cont = S.n_cont, n_term = S.n_term;
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        Q.c->set[(it + 2u) % WF_SETS] = WfSet{};
        if (n_cont + n_term) Q.c->stat_iters += 1;
    }
    MrExtendJob job{V, Q, S, it & 1u, n_cont, n_term, sort};
    mr_persistent_trace<0, K>(V.S, job, &S.cur_extend, part == 1 ? n_cont : n_cont + n_term, mo.free_quorum, mo.qn, mo.ql, mo.qi);
    const unsigned long long p = warp_sum(job.paths), q = warp_sum(job.settled);
    if ((threadIdx.x & 31u) == 0u && p) atomicAdd(&Q.c->stat_paths, p);
    if ((threadIdx.x & 31u) == 0u && q) atomicAdd(&Q.c->stat_settled, q);
}

struct MrNeeJob
{
    static constexpr int JOBW = 3;  // [0] nee queue entry (slot | flags), [1] t of the light hit (stage 1), [2] its instance; the stage (0 shadow ray, 1 light
                                    // probe, 2 probe visibility) rides in bits 24-25 of the record's flag word, which the traversal leaves alone
    static constexpr uint32_t STAGE_SHIFT = 24;
    template <class R>
    __device__ __forceinline__ uint32_t stage_of(const R& r) const { return (r.ld(MR_FL) >> STAGE_SHIFT) & 3u; }
    template <class R>
    __device__ __forceinline__ void set_stage(const R& r, uint32_t stage) const { r.st(MR_FL, r.ld(MR_FL) | (stage << STAGE_SHIFT)); }  // after mr_begin (which rewrites the word)
    const RenderView& V;
    const Queues& Q;
    unsigned long long n_shadow = 0, n_probe = 0, n_probe_shadow = 0;

    template <class R>
    __device__ __forceinline__ V3 origin(const R& r) const
    {
        const float4 o4 = V.ray[r.ld(MR_JOB) & NEE_SLOT_MASK].o;
        return ex::v3(o4.x, o4.y, o4.z);
    }
    template <class R>
    __device__ __forceinline__ void world_ray(const R& r, V3* o, V3* d) const
    {
        const uint32_t slot = r.ld(MR_JOB) & NEE_SLOT_MASK;
        *o = origin(r);
        const float4 d4 = stage_of(r) == 0u ? V.shadow[slot].d : V.probe[slot].d;
        *d = ex::v3(d4.x, d4.y, d4.z);
    }
    template <class R>
    __device__ __forceinline__ uint32_t begin_probe(const R& r, uint32_t slot, V3 o)
    {
        const float4 d4 = V.probe[slot].d;
        const V3 d = ex::v3(d4.x, d4.y, d4.z);
        ++n_probe;
        if (!tlas_may_hit(V.S, 1, o, d)) return MT_IDLE;
        const uint32_t t = mr_begin(r, V.S, 1, o, d, ex::inf(), false);
        set_stage(r, 1u);
        return t;
    }
    template <class R>
    __device__ __forceinline__ uint32_t start(uint32_t idx, const R& r)
    {
        const uint32_t e = Q.nee[idx], slot = e & NEE_SLOT_MASK;
        r.st(MR_JOB, e);
        const float4 o4 = V.ray[slot].o;
        const V3 o = ex::v3(o4.x, o4.y, o4.z);
        if (e & NEE_SHADOW_BIT)
        {
            const float4 d4 = V.shadow[slot].d;
            ++n_shadow;
            return mr_begin(r, V.S, 0, o, ex::v3(d4.x, d4.y, d4.z), d4.w, true);  // stage 0
        }
        return begin_probe(r, slot, o);
    }
    // the two estimates of a bounce are added to the radiance one after the other (nee_slot_t's order): two read-modify-writes here
    __device__ __forceinline__ void add_radiance(uint32_t slot, float x, float y, float z)
    {
        float4 rr = V.path[slot].radiance;
        rr.x += x; rr.y += y; rr.z += z;
        V.path[slot].radiance = rr;
    }
    template <class R>
    __device__ __forceinline__ uint32_t finish(const R& r, bool mine)
    {
        if (!mine) return MT_IDLE;
        const uint32_t e = r.ld(MR_JOB), slot = e & NEE_SLOT_MASK, stage = stage_of(r);
        const bool hit = r.ld(MR_HINST) != TRAV_NONE;
        if (stage == 0u)
        {
            if (!hit)
            {
                const float4 c = V.shadow[slot].c;
                add_radiance(slot, c.x, c.y, c.z);
            }
            if (e & NEE_PROBE_BIT) return begin_probe(r, slot, origin(r));
        }
        else if (stage == 1u)
        {
            if (hit)
            {
                const float lt = r.ldf(MR_TMAX);
                r.stf(MR_JOB + 1, lt);
                r.st(MR_JOB + 2, r.ld(MR_HINST));
                ++n_probe_shadow;
                const uint32_t t = mr_begin(r, V.S, 0, origin(r), r.ld3(MR_D), lt * (1.0f - ex::EPS), true, true);  // MR_D: the probe direction, back in world space
                set_stage(r, 2u);
                return t;
            }
        }
        else if (!hit)
        {
            const float4 d4 = V.probe[slot].d;
            const V3 c = probe_light_contrib(V, slot, ex::v3(d4.x, d4.y, d4.z), r.ldf(MR_JOB + 1), mr_hit_u(r), mr_hit_v(r), V.S.tri_prim[r.ld(MR_HSLOT)], r.ld(MR_JOB + 2));
            add_radiance(slot, c.x, c.y, c.z);
        }
        return MT_IDLE;
    }
};

template <int K>
__global__ void __launch_bounds__(128) k_nee_mr(RenderView V, Queues Q, uint32_t it, MrOpts mo)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    MrNeeJob job{V, Q};
    mr_persistent_trace<2, K>(V.S, job, &S.cur_nee, S.n_nee, mo.free_quorum, mo.qn, mo.ql, mo.qi);
    const unsigned long long a = warp_sum(job.n_shadow), b = warp_sum(job.n_probe), c = warp_sum(job.n_probe_shadow);
    if ((threadIdx.x & 31u) == 0u)
    {
        if (a) atomicAdd(&Q.c->stat_shadow, a);
        if (b) atomicAdd(&Q.c->stat_probe, b);
        if (c) atomicAdd(&Q.c->stat_probe_shadow, c);
    }
}

// ---- path: every remaining slot run to completion in ONE persistent launch -----------------------------------------------------
// Once the wavefront has thinned out (C3: from iteration ~17 on, 8 % of the frame's rays spread over 130 further iterations and
// a tail of 1024-bounce paths) every launch of the three-kernel iteration is floored by the latency of its slowest
// ray — 110-130 us per kernel however few rays it holds (profiles/r1l_c3_launch_log.txt) — and the frame pays that floor ~400
// times.  k_path replaces all of those launches and the tail kernel: a lane keeps its slot through extend -> shade -> shadow ray ->
// light probe -> probe visibility -> next bounce -> ... -> regeneration of the slot's next sample, as a chain of follow-up
// traversals in the persistent loop of persist.cuh (finish() of one traversal begins the next), and takes another slot from the
// cont / term queues of the switch-over iteration only when its slot has run out of samples.  The traversal is the one shared
// state machine (MODE 2: closest / any hit decided per lane), so lanes tracing primary, shadow and probe rays re-converge at every
// stage.  Shading runs inside finish(), gated by `finish_quorum` so that it executes with several lanes at a time.
// Per-slot arithmetic and its order (shade_slot, the NEE sum of nee_slot_t, regen_slot) are those of the wavefront kernels.
struct PathJob
{
    const RenderView& V;
    const Queues& Q;
    uint32_t b, n_cont;
    uint32_t slot = 0, flags = 0, stage = 0;  // stage: 0 extend, 1 shadow, 2 probe (lights TLAS), 3 probe visibility
    V3 o;
    float4 r;
    float lt = 0, lu = 0, lv = 0;
    uint32_t lslot = 0, linst = 0;
    unsigned long long n_ext = 0, n_shadow = 0, n_probe = 0, n_probe_shadow = 0, paths = 0, settled = 0;

    // regenerate the slot (finalise + accumulate the finished sample, raygen the next) and begin its primary ray; false: out of samples
    __device__ __forceinline__ bool begin_sample(WideTrav& tr)
    {
        V3 ro, rd;
        for (;;)
        {
            if (!(regen_slot(V, slot, &ro, &rd) & WF_NEXT)) { Q.state[slot] = 0u; return false; }
            ++paths;
            if (V.S.env != nullptr || tlas_may_hit(V.S, 0, ro, rd)) break;
            V.path[slot].radiance = f4(0.006f, 0.006f, 0.006f, 0.0f);  // background path settled without a traversal (as ExtendJob::start)
            ++settled;
        }
        wide_begin(tr, V.S, 0, ro, rd, ex::inf(), false);
        stage = 0;
        return true;
    }
    __device__ __forceinline__ bool begin_bounce(WideTrav& tr)
    {
        if (flags & WF_NEXT)
        {
            const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
            wide_begin(tr, V.S, 0, ex::v3(o4.x, o4.y, o4.z), ex::v3(d4.x, d4.y, d4.z), ex::inf(), false);
            stage = 0;
            return true;
        }
        return begin_sample(tr);
    }
    __device__ __forceinline__ bool begin_probe(WideTrav& tr)
    {
        const float4 d4 = V.probe[slot].d;
        const V3 d = ex::v3(d4.x, d4.y, d4.z);
        ++n_probe;
        if (!tlas_may_hit(V.S, 1, o, d)) return false;
        wide_begin(tr, V.S, 1, o, d, ex::inf(), false);
        stage = 2;
        return true;
    }
    __device__ __forceinline__ bool start(uint32_t idx, WideTrav& tr)
    {
        if (idx < n_cont)
        {
            slot = Q.cont[b][idx];
            flags = WF_NEXT;
            return begin_bounce(tr);
        }
        slot = Q.term[b][idx - n_cont];
        return begin_sample(tr);
    }
    __device__ __forceinline__ bool finish(WideTrav& tr, bool mine)
    {
        if (!mine) return false;
        if (stage == 0)
        {
            ++n_ext;
            if (tr.hit_inst == TRAV_NONE) V.hit[slot].info = 0xffffffffu;
            else
            {
                const uint32_t prim = V.S.tri_prim[tr.hit_slot];
                V3 n;
                bool ff;
                finish_hit(V.S, tr.dw, tr.hit_inst, prim, wide_hit_u(tr), wide_hit_v(tr), &n, &ff);
                V.hit[slot].tn = f4(tr.t_max, n.x, n.y, n.z);
                V.hit[slot].info = tr.hit_inst | (ff ? 0x80000000u : 0u);
            }
            flags = shade_slot(V, slot);
            n_probe += (flags & WF_PROBE_OUT) ? 1u : 0u;
            if (!(flags & (WF_SHADOW | WF_PROBE))) return begin_bounce(tr);
            const float4 o4 = V.ray[slot].o;
            o = ex::v3(o4.x, o4.y, o4.z);
            r = V.path[slot].radiance;
            if (flags & WF_SHADOW)
            {
                const float4 d4 = V.shadow[slot].d;
                wide_begin(tr, V.S, 0, o, ex::v3(d4.x, d4.y, d4.z), d4.w, true);
                stage = 1;
                ++n_shadow;
                return true;
            }
            if (begin_probe(tr)) return true;
        }
        else if (stage == 1)
        {
            if (tr.hit_inst == TRAV_NONE)
            {
                const float4 c = V.shadow[slot].c;
                r.x += c.x; r.y += c.y; r.z += c.z;
            }
            if ((flags & WF_PROBE) && begin_probe(tr)) return true;
        }
        else if (stage == 2)
        {
            if (tr.hit_inst != TRAV_NONE)
            {
                lt = tr.t_max; lu = wide_hit_u(tr); lv = wide_hit_v(tr); lslot = tr.hit_slot; linst = tr.hit_inst;
                wide_begin(tr, V.S, 0, o, tr.dw, lt * (1.0f - ex::EPS), true);
                stage = 3;
                ++n_probe_shadow;
                return true;
            }
        }
        else if (tr.hit_inst == TRAV_NONE)
        {
            const V3 c = probe_light_contrib(V, slot, tr.dw, lt, lu, lv, V.S.tri_prim[lslot], linst);
            r.x += c.x; r.y += c.y; r.z += c.z;
        }
        V.path[slot].radiance = r;  // both estimates of the bounce added in nee_slot_t's order
        return begin_bounce(tr);
    }
};

template <int MB>
__global__ void __launch_bounds__(128, MB) k_path(RenderView V, Queues Q, uint32_t it, int refill_below, int node_quorum, int finish_quorum)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    PathJob job{V, Q, it & 1u, S.n_cont};
    persistent_trace<2, PathJob, true>(V.S, job, &S.cur_extend, S.n_cont + S.n_term, refill_below, node_quorum, finish_quorum);
    const unsigned long long a = warp_sum(job.n_ext), b = warp_sum(job.n_shadow), c = warp_sum(job.n_probe), d = warp_sum(job.n_probe_shadow), e = warp_sum(job.paths),
                             f = warp_sum(job.settled);
    if ((threadIdx.x & 31u) == 0u)
    {
        if (a) { atomicAdd(&Q.c->stat_extend, a); atomicAdd(&Q.c->stat_extend_path, a); }
        if (b) atomicAdd(&Q.c->stat_shadow, b);
        if (c) atomicAdd(&Q.c->stat_probe, c);
        if (d) atomicAdd(&Q.c->stat_probe_shadow, d);
        if (e) atomicAdd(&Q.c->stat_paths, e);
        if (f) atomicAdd(&Q.c->stat_settled, f);
    }
}

// reference-order variant: static chunks, one entry per lane
__global__ void __launch_bounds__(128) k_nee_bin(RenderView V, Queues Q, uint32_t it)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    const uint32_t n = S.n_nee;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t c;
    while (fetch_chunk(&S.cur_nee, (n + 31u) >> 5, &c))
    {
        const uint32_t i = c * 32u + lane;
        const uint32_t e = i < n ? Q.nee[i] : 0u;
        const uint32_t flags = ((e & NEE_SHADOW_BIT) ? WF_SHADOW : 0u) | ((e & NEE_PROBE_BIT) ? WF_PROBE : 0u);
        const uint32_t traced = flags ? nee_slot_t<true>(V, e & NEE_SLOT_MASK, flags) : 0u;
        const unsigned ms = __ballot_sync(0xffffffffu, (flags & WF_SHADOW) != 0u), mp = __ballot_sync(0xffffffffu, (flags & WF_PROBE) != 0u), mt = __ballot_sync(0xffffffffu, traced != 0u);
        if (lane == 0u)
        {
            if (ms) atomicAdd(&Q.c->stat_shadow, (unsigned long long)__popc(ms));
            if (mp) atomicAdd(&Q.c->stat_probe, (unsigned long long)__popc(mp));
            if (mt) atomicAdd(&Q.c->stat_probe_shadow, (unsigned long long)__popc(mt));
        }
    }
}

// ---- slot-ordered shading ------------------------------------------------------------------------------------------------
// k_shade (above) takes its work from queues that k_extend fills in completion order: every queue push is an atomicAdd whose
// return value the warp waits for (four per 32 slots, on four hot addresses: 47 % of its stall samples, profiles/r1d_c2_*), and
// the queues it writes are in scrambled order, so the next kernels touch the 32-byte per-slot records as scattered sectors.
// k_shade_ordered walks the slots in index order instead (one warp per GROUP of 32 consecutive slots, only the groups that still
// hold a live slot) and writes all four outputs as ASCENDING lists with one reservation per block and span of 64 groups:
//   cont / term  the queues of the next k_extend,     nee  the queue of this iteration's k_nee,     grp  its own next work list.
// k_extend and k_nee are the queue kernels unchanged (k_extend just does not push, sort = -1); with ascending queues their
// record accesses are near-sequential.  Q.state holds one flag word per slot: WF_NEXT while a ray was traced for the slot in this
// iteration (set by the previous shade or by k_extend's regeneration), so that the walk knows which lanes of a group have a hit
// record.  The host may switch between the two shade kernels at any iteration: both leave valid cont / term / nee queues.
#ifndef PTB_SHADE_MINBLOCKS
#define PTB_SHADE_MINBLOCKS 4  // resident blocks per SM of k_shade_ordered = its register cap (4: 64 registers; 5 and 6 measured slower)
#endif
constexpr uint32_t SHADE_ROWS = 8, SHADE_GROUPS = SHADE_ROWS * 8;
template <bool SORT, uint32_t KMASK = 0x3fu, bool MEDIA = true>
__global__ void __launch_bounds__(256, PTB_SHADE_MINBLOCKS) k_shade_ordered(RenderView V, Queues Q, uint32_t it)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    WfSet& N = Q.c->set[(it + 1u) % WF_SETS];
    const uint32_t nb = (it + 1u) & 1u;
    const uint32_t* grp = Q.grp[it & 1u];
    uint32_t* out[4] = {Q.term[nb], Q.cont[nb], Q.nee, Q.grp[nb]};
    uint32_t* out_n[4] = {&N.n_term, &N.n_cont, &S.n_nee, &N.n_grp};
    const uint32_t n_grp = S.n_grp;
    // per list and (row, warp): entries of that warp-row; after the scan: exclusive prefix within the span
    __shared__ uint32_t cnt[4][SHADE_GROUPS];
    __shared__ uint32_t base[4], live_total, out_total;
    uint32_t n_out = 0;  // probes settled by the bounds pre-test in shade_slot (statistics)
    // span-local sort of the live slots by shade kind (7 kinds): counts per (kind, row, warp) -> offsets; the sorted slot numbers
    __shared__ uint32_t kcnt[7][SHADE_GROUPS], kbase[8];
    __shared__ uint32_t sorted[SHADE_GROUPS * 32u];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { live_total = 0u; out_total = 0u; }
    __syncthreads();
    for (uint32_t first = blockIdx.x * SHADE_GROUPS; first < n_grp; first += gridDim.x * SHADE_GROUPS)
    {
        if constexpr (SORT)
        {
            // (1) classify: one warp per listed group, lanes = its 32 slots
            uint32_t my_kind[SHADE_ROWS];
    #pragma unroll
            for (uint32_t row = 0; row < SHADE_ROWS; ++row)
            {
                const uint32_t gi = first + row * 8u + warp;
                const uint32_t slot = gi < n_grp ? grp[gi] * 32u + lane : 0xffffffffu;
                uint32_t kind = 7u;  // not live
                if (slot < Q.slot_end && (Q.state[slot] & WF_NEXT)) kind = min(V.hit[slot].kind, 6u);
                my_kind[row] = kind;
    #pragma unroll
                for (uint32_t k = 0; k < 7u; ++k)
                {
                    const unsigned m = __ballot_sync(0xffffffffu, kind == k);
                    if (lane == k) kcnt[k][row * 8u + warp] = (uint32_t)__popc(m);
                }
            }
            __syncthreads();
            // (2) offsets: warp k scans the 64 counts of kind k; then the seven totals are chained
            if (warp < 7u)
            {
                uint32_t* c = kcnt[warp];
                const uint32_t a = c[2u * lane], b = c[2u * lane + 1u];
                uint32_t incl = a + b;
    #pragma unroll
                for (int o = 1; o < 32; o <<= 1)
                {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                    if ((int)lane >= o) incl += t;
                }
                const uint32_t excl = incl - (a + b);
                c[2u * lane] = excl;
                c[2u * lane + 1u] = excl + a;
                if (lane == 31u) kbase[warp + 1u] = incl;
            }
            __syncthreads();
            if (threadIdx.x == 0)
            {
                kbase[0] = 0u;
                for (int k = 1; k <= 7; ++k) kbase[k] += kbase[k - 1];
            }
            __syncthreads();
            // (3) scatter the slot numbers in kind order (ascending slot order within a kind)
    #pragma unroll
            for (uint32_t row = 0; row < SHADE_ROWS; ++row)
            {
                const uint32_t kind = my_kind[row];
                const unsigned m = __match_any_sync(0xffffffffu, kind);
                if (kind < 7u)
                {
                    const uint32_t gi = first + row * 8u + warp;
                    sorted[kbase[kind] + kcnt[kind][row * 8u + warp] + (uint32_t)__popc(m & ((1u << lane) - 1u))] = grp[gi] * 32u + lane;
                }
            }
            __syncthreads();
            // (4) shade: consecutive threads take consecutive sorted entries, so a warp shades one kind (but for the few warps on a boundary)
            const uint32_t span_live = kbase[7];
    #pragma unroll 1
            for (uint32_t i = threadIdx.x; i < span_live; i += 256u)
            {
                const uint32_t slot = sorted[i];
                const uint32_t flags = shade_slot_t<KMASK, MEDIA>(V, slot);
                Q.state[slot] = flags;
                n_out += (flags & WF_PROBE_OUT) ? 1u : 0u;
            }
            if (threadIdx.x == 0) live_total += span_live;
            __syncthreads();
        }
        else
        {
            // all kinds alike (one surface material in the scene): shade in place, one warp per listed group
            uint32_t n_live = 0;
#pragma unroll 1
            for (uint32_t row = 0; row < SHADE_ROWS; ++row)
            {
                const uint32_t gi = first + row * 8u + warp;
                const uint32_t slot = gi < n_grp ? grp[gi] * 32u + lane : 0xffffffffu;
                if (slot < Q.slot_end && (Q.state[slot] & WF_NEXT))
                {
                    const uint32_t flags = shade_slot_t<KMASK, MEDIA>(V, slot);
                    Q.state[slot] = flags;
                    n_out += (flags & WF_PROBE_OUT) ? 1u : 0u;
                    ++n_live;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) n_live += __shfl_xor_sync(0xffffffffu, n_live, o);
            if (lane == 0u && n_live) atomicAdd(&live_total, n_live);
        }
        // (5) back in slot order: the four ascending output lists
#pragma unroll 1
        for (uint32_t row = 0; row < SHADE_ROWS; ++row)
        {
            const uint32_t gi = first + row * 8u + warp;
            const uint32_t slot = gi < n_grp ? grp[gi] * 32u + lane : 0xffffffffu;
            // every non-zero flag word of a listed group now is this pass's: the pass ran on the WF_NEXT slots and the others were 0
            const uint32_t flags = slot < Q.slot_end ? Q.state[slot] : 0u;
            const unsigned mt = __ballot_sync(0xffffffffu, (flags & WF_TERM) != 0u), mc = __ballot_sync(0xffffffffu, (flags & WF_NEXT) != 0u);
            const unsigned mn = __ballot_sync(0xffffffffu, (flags & (WF_SHADOW | WF_PROBE)) != 0u);
            if (lane == 0u)
            {
                const uint32_t e = row * 8u + warp;
                cnt[0][e] = (uint32_t)__popc(mt); cnt[1][e] = (uint32_t)__popc(mc); cnt[2][e] = (uint32_t)__popc(mn); cnt[3][e] = (mt | mc) ? 1u : 0u;
            }
        }
        __syncthreads();
        if (warp < 4u)  // warp w scans list w (64 counts, two per lane) and makes the span's one reservation on that list's counter
        {
            uint32_t* c = cnt[warp];
            const uint32_t a = c[2u * lane], b = c[2u * lane + 1u];
            uint32_t incl = a + b;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if ((int)lane >= o) incl += t;
            }
            const uint32_t excl = incl - (a + b);
            c[2u * lane] = excl;
            c[2u * lane + 1u] = excl + a;
            if (lane == 31u) base[warp] = incl ? atomicAdd(out_n[warp], incl) : 0u;
        }
        __syncthreads();
#pragma unroll 1
        for (uint32_t row = 0; row < SHADE_ROWS; ++row)
        {
            const uint32_t gi = first + row * 8u + warp, e = row * 8u + warp;
            const uint32_t g = gi < n_grp ? grp[gi] : 0xffffffffu;
            const uint32_t slot = gi < n_grp ? g * 32u + lane : 0xffffffffu;
            const uint32_t flags = slot < Q.slot_end ? Q.state[slot] : 0u;  // written by this thread in the first pass
            const unsigned lt = (1u << lane) - 1u;
            const unsigned mt = __ballot_sync(0xffffffffu, (flags & WF_TERM) != 0u), mc = __ballot_sync(0xffffffffu, (flags & WF_NEXT) != 0u);
            const unsigned mn = __ballot_sync(0xffffffffu, (flags & (WF_SHADOW | WF_PROBE)) != 0u);
            if (flags & WF_TERM) out[0][base[0] + cnt[0][e] + (uint32_t)__popc(mt & lt)] = slot;
            if (flags & WF_NEXT) out[1][base[1] + cnt[1][e] + (uint32_t)__popc(mc & lt)] = slot;
            if (flags & (WF_SHADOW | WF_PROBE))
                out[2][base[2] + cnt[2][e] + (uint32_t)__popc(mn & lt)] = slot | ((flags & WF_SHADOW) ? NEE_SHADOW_BIT : 0u) | ((flags & WF_PROBE) ? NEE_PROBE_BIT : 0u);
            if (lane == 0u && (mt | mc)) out[3][base[3] + cnt[3][e]] = g;
        }
        __syncthreads();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n_out += __shfl_xor_sync(0xffffffffu, n_out, o);
    if (lane == 0u && n_out) atomicAdd(&out_total, n_out);
    __syncthreads();
    if (threadIdx.x == 0 && live_total) atomicAdd(&Q.c->stat_extend, (unsigned long long)live_total);
    if (threadIdx.x == 0 && out_total) atomicAdd(&Q.c->stat_probe, (unsigned long long)out_total);
}

// Scene classes of k_shade_ordered (shade_slot_t): 0 the general body; 1 Lambertian + Emissive, 2 + GGX metal, 3 all six kinds — 1 to 3
// without media.
constexpr uint32_t SHADE_KMASK_1 = (1u << PTB_LAMBERTIAN) | (1u << PTB_EMISSIVE), SHADE_KMASK_2 = SHADE_KMASK_1 | (1u << PTB_GGX_METAL);
static void launch_shade_ordered(bool sort, int cls, int grid, cudaStream_t st, const RenderView& V, const Queues& Q, uint32_t it)
{
    switch (cls * 2 + (sort ? 1 : 0))
    {
    case 2: k_shade_ordered<false, SHADE_KMASK_1, false><<<grid, 256, 0, st>>>(V, Q, it); break;
    case 3: k_shade_ordered<true, SHADE_KMASK_1, false><<<grid, 256, 0, st>>>(V, Q, it); break;
    case 4: k_shade_ordered<false, SHADE_KMASK_2, false><<<grid, 256, 0, st>>>(V, Q, it); break;
    case 5: k_shade_ordered<true, SHADE_KMASK_2, false><<<grid, 256, 0, st>>>(V, Q, it); break;
    case 6: k_shade_ordered<false, 0x3fu, false><<<grid, 256, 0, st>>>(V, Q, it); break;
    case 7: k_shade_ordered<true, 0x3fu, false><<<grid, 256, 0, st>>>(V, Q, it); break;
    case 1: k_shade_ordered<true><<<grid, 256, 0, st>>>(V, Q, it); break;
    default: k_shade_ordered<false><<<grid, 256, 0, st>>>(V, Q, it); break;
    }
}

// The tail: one thread per remaining slot runs regen -> extend -> shade -> shadow -> probe until the slot has no samples left.
template <bool BIN>
__global__ void __launch_bounds__(64) k_tail(RenderView V, Queues Q, uint32_t it)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    const uint32_t b = it & 1u;
    const uint32_t n_cont = S.n_cont, n_term = S.n_term;
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long n_ext = 0, n_sh = 0, n_pr = 0, n_ps = 0, n_paths = 0;
    if (idx < n_cont + n_term)
    {
        const uint32_t slot = idx < n_cont ? Q.cont[b][idx] : Q.term[b][idx - n_cont];
        bool need_regen = idx >= n_cont;
        for (;;)
        {
            V3 o, d;
            if (need_regen)
            {
                if (!(regen_slot(V, slot, &o, &d) & WF_NEXT)) break;
                ++n_paths;
            }
            else
            {
                const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
                o = ex::v3(o4.x, o4.y, o4.z);
                d = ex::v3(d4.x, d4.y, d4.z);
            }
            extend_ray_t<BIN>(V, slot, o, d);
            ++n_ext;
            const uint32_t flags = shade_slot(V, slot);
            n_pr += (flags & WF_PROBE_OUT) ? 1u : 0u;
            if (flags & (WF_SHADOW | WF_PROBE))
            {
                n_ps += nee_slot_t<BIN>(V, slot, flags);
                n_sh += (flags & WF_SHADOW) ? 1u : 0u;
                n_pr += (flags & WF_PROBE) ? 1u : 0u;
            }
            need_regen = (flags & WF_TERM) != 0u;
        }
    }
    if (n_ext | n_paths)
    {
        atomicAdd(&Q.c->stat_extend, n_ext);
        atomicAdd(&Q.c->stat_extend_path, n_ext);
        atomicAdd(&Q.c->stat_shadow, n_sh);
        atomicAdd(&Q.c->stat_probe, n_pr);
        atomicAdd(&Q.c->stat_probe_shadow, n_ps);
        atomicAdd(&Q.c->stat_paths, n_paths);
    }
}

// The tail with the three traversals of a bounce side by side.  What is left in the tail are a few thousand slots, some of them
// on paths of up to 1024 bounces: the launch is as long as the longest chain of bounces, every bounce a chain of dependent loads
// (L2 latency, not throughput).  Shadow ray, light probe and the NEXT bounce's closest-hit ray all depend only on the shade step
// before them, so a block of three warps keeps 32 slots and runs them concurrently: warp 0 shades and extends, warp 1 traces the
// shadow rays, warp 2 the probes (and their visibility rays).  Warp 0 adds the two contributions to the path's radiance in the
// order of nee_slot_t, before the next shade or regeneration reads it: results are bit-identical to k_tail.
template <bool BIN>
__global__ void __launch_bounds__(96) k_tail_roles(RenderView V, Queues Q, uint32_t it, uint32_t spb)
{
    WfSet& S = Q.c->set[it % WF_SETS];
    const uint32_t b = it & 1u;
    const uint32_t n_cont = S.n_cont, n_term = S.n_term;
    const uint32_t lane = threadIdx.x & 31u, role = threadIdx.x >> 5;
    // spb slots per block (lanes >= spb idle): a bounce of the block lasts as long as the slowest of its slots, so fewer slots
    // per block shorten the chain when the device has warps to spare
    const uint32_t idx = blockIdx.x * spb + lane;
    __shared__ uint32_t s_flags[32];
    __shared__ float4 s_r[2][32];
    unsigned long long n_ext = 0, n_sh = 0, n_pr = 0, n_ps = 0, n_paths = 0;
    bool alive = lane < spb && idx < n_cont + n_term;
    const uint32_t slot = !alive ? 0u : idx < n_cont ? Q.cont[b][idx] : Q.term[b][idx - n_cont];
    bool need_regen = idx >= n_cont;
    bool have_hit = false;  // warp 0: the slot's current ray was traced during the previous bounce's NEE
    for (;;)
    {
        uint32_t flags = 0u;
        if (role == 0u && alive)
        {
            if (!have_hit)
            {
                V3 o, d;
                if (need_regen)
                {
                    if (regen_slot(V, slot, &o, &d) & WF_NEXT) ++n_paths;
                    else alive = false;
                }
                else
                {
                    const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
                    o = ex::v3(o4.x, o4.y, o4.z);
                    d = ex::v3(d4.x, d4.y, d4.z);
                }
                if (alive)
                {
                    extend_ray_t<BIN>(V, slot, o, d);
                    ++n_ext;
                }
            }
            if (alive)
            {
                flags = shade_slot(V, slot);
                n_pr += (flags & WF_PROBE_OUT) ? 1u : 0u;
            }
        }
        if (role == 0u) s_flags[lane] = flags;
        if (!__syncthreads_or(role == 0u && alive)) break;  // also publishes the records shade_slot wrote
        flags = s_flags[lane];
        if (role == 0u)
        {
            have_hit = false;
            if (flags & WF_NEXT)
            {
                const float4 o4 = V.ray[slot].o, d4 = V.ray[slot].d;
                extend_ray_t<BIN>(V, slot, ex::v3(o4.x, o4.y, o4.z), ex::v3(d4.x, d4.y, d4.z));
                ++n_ext;
                have_hit = true;
            }
        }
        else if (flags & (role == 1u ? WF_SHADOW : WF_PROBE))
        {
            const float4 o4 = V.ray[slot].o;
            const V3 o = ex::v3(o4.x, o4.y, o4.z);
            float4 r = f4(0.0f, 0.0f, 0.0f, 0.0f);
            if (role == 1u)
            {
                shadow_contrib_t<BIN>(V, slot, o, &r);
                ++n_sh;
            }
            else
            {
                n_ps += probe_contrib_t<BIN>(V, slot, o, &r);
                ++n_pr;
            }
            s_r[role - 1u][lane] = r;
        }
        __syncthreads();
        if (role == 0u)
        {
            if (flags & (WF_SHADOW | WF_PROBE))
            {
                float4 r = V.path[slot].radiance;
                if (flags & WF_SHADOW) { const float4 c = s_r[0][lane]; r.x += c.x; r.y += c.y; r.z += c.z; }
                if (flags & WF_PROBE) { const float4 c = s_r[1][lane]; r.x += c.x; r.y += c.y; r.z += c.z; }
                V.path[slot].radiance = r;
            }
            need_regen = (flags & WF_TERM) != 0u;
        }
    }
    if (n_ext | n_paths | n_sh | n_pr)
    {
        if (n_ext) { atomicAdd(&Q.c->stat_extend, n_ext); atomicAdd(&Q.c->stat_extend_path, n_ext); }
        if (n_sh) atomicAdd(&Q.c->stat_shadow, n_sh);
        if (n_pr) atomicAdd(&Q.c->stat_probe, n_pr);
        if (n_ps) atomicAdd(&Q.c->stat_probe_shadow, n_ps);
        if (n_paths) atomicAdd(&Q.c->stat_paths, n_paths);
    }
}

// ---- probes (device code of raygen, run on small batches) -------------------------------------------------------------------
__global__ void k_camera_rays(CameraDev cam, unsigned long long n, const float* st, float* o3, float* d3)
{
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 o, d;
    camera_ray(cam, st[i * 2], st[i * 2 + 1], &o, &d);
    o3[i * 3] = o.x; o3[i * 3 + 1] = o.y; o3[i * 3 + 2] = o.z;
    d3[i * 3] = d.x; d3[i * 3 + 1] = d.y; d3[i * 3 + 2] = d.z;
}
__global__ void k_ss_sobol(unsigned long long n, const uint32_t* index, const uint32_t* seed, float* xy)
{
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ss_sobol(index[i], seed[i], &xy[i * 2], &xy[i * 2 + 1]);
}
__global__ void k_rng_stream(unsigned long long seed, uint32_t pixel, uint32_t sample, uint32_t n, uint32_t* out)
{
    if (blockIdx.x || threadIdx.x) return;
    Rng r = rng_stream(seed, pixel, sample);
    for (uint32_t i = 0; i < n; ++i) out[i] = rng_u32(r);
}

bool render_state_frame(ptb_context* ctx, float4** accum, uint32_t* W, uint32_t* H)
{
    if (!ctx->rs || !ctx->rs->accum) return false;
    *accum = ctx->rs->accum;
    *W = ctx->rs->W;
    *H = ctx->rs->H;
    return true;
}

bool render_state_aux(ptb_context* ctx, float4** position, uint32_t** id)
{
    if (!ctx->rs || !ctx->rs->position.p || !ctx->rs->id.p) return false;
    *position = ctx->rs->position.p;
    *id = ctx->rs->id.p;
    return true;
}

void render_state_destroy(ptb_context* ctx)
{
    RenderState* rs = ctx->rs;
    if (!rs) return;
    for (auto e : rs->poll_events) cudaEventDestroy(e);
    for (auto e : rs->prof_events) cudaEventDestroy(e);
    for (auto e : rs->group_events) cudaEventDestroy(e);
    for (auto s : rs->group_streams) cudaStreamDestroy(s);
    for (auto s : rs->nee_streams) cudaStreamDestroy(s);
    for (auto e : rs->nee_events) cudaEventDestroy(e);
    if (rs->ev_begin) cudaEventDestroy(rs->ev_begin);
    if (rs->ev_end) cudaEventDestroy(rs->ev_end);
    if (rs->h_flags) cudaFreeHost(rs->h_flags);
    rs->accum_own.release(); rs->accum_sq.release(); rs->position.release();
    rs->s_accum.release(); rs->s_accum_sq.release(); rs->s_position.release(); rs->s_id.release();
    rs->ray.release(); rs->path.release(); rs->hit.release(); rs->shadow.release(); rs->probe.release();
    DBuf<uint32_t>* u32s[] = {&rs->id, &rs->q_cont0, &rs->q_cont1, &rs->q_term0, &rs->q_term1, &rs->q_mat, &rs->q_nee, &rs->slot_state, &rs->q_grp0, &rs->q_grp1};
    for (auto* b : u32s) b->release();
    rs->volset.release();
    rs->counters.release();
    delete rs;
    ctx->rs = nullptr;
}

}  // namespace ptb

using namespace ptb;

#define REQUIRE_DEVICE(ctx)                                                                                                   \
    if (!(ctx)) return PTB_ERR_INVALID;                                                                                       \
    if ((ctx)->host_only) return (ctx)->fail(PTB_ERR_NO_DEVICE, "host-only context: no CUDA device bound (no CPU fallback)")

extern "C" {

int ptb_set_camera(ptb_context* ctx, const float origin[3], const float target[3], float fov_deg, float aspect_ratio)
{
    if (!ctx) return PTB_ERR_INVALID;
    if (!origin || !target || !(fov_deg > 0.0f) || !(aspect_ratio > 0.0f)) return ctx->fail(PTB_ERR_INVALID, "bad camera");
    camera_setup(origin, target, fov_deg, aspect_ratio, ctx->cam_m, ctx->cam_eye);
    ctx->cam_set = true;
    return PTB_OK;
}

static CameraDev cam_dev(const ptb_context* ctx)
{
    CameraDev c;
    std::memcpy(c.m, ctx->cam_m, 64);
    std::memcpy(c.eye, ctx->cam_eye, 12);
    return c;
}

int ptb_camera_rays(ptb_context* ctx, uint64_t n, const float* st2, float* o3, float* d3)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->cam_set) return ctx->fail(PTB_ERR_INVALID, "camera not set");
    if (n == 0) return PTB_OK;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    DBuf<float> st, o, d;
    PTB_CUDA(ctx, st.resize(n * 2));
    PTB_CUDA(ctx, o.resize(n * 3));
    PTB_CUDA(ctx, d.resize(n * 3));
    PTB_CUDA(ctx, cudaMemcpyAsync(st.p, st2, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    k_camera_rays<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(cam_dev(ctx), n, st.p, o.p, d.p);
    PTB_CUDA(ctx, cudaGetLastError());
    PTB_CUDA(ctx, cudaMemcpyAsync(o3, o.p, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(d3, d.p, n * 12, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    st.release(); o.release(); d.release();
    return PTB_OK;
}

int ptb_ss_sobol(ptb_context* ctx, uint64_t n, const uint32_t* index, const uint32_t* seed, float* xy2)
{
    REQUIRE_DEVICE(ctx);
    if (n == 0) return PTB_OK;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    DBuf<uint32_t> di, ds;
    DBuf<float> dxy;
    PTB_CUDA(ctx, di.resize(n));
    PTB_CUDA(ctx, ds.resize(n));
    PTB_CUDA(ctx, dxy.resize(n * 2));
    PTB_CUDA(ctx, cudaMemcpyAsync(di.p, index, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(ds.p, seed, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    k_ss_sobol<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(n, di.p, ds.p, dxy.p);
    PTB_CUDA(ctx, cudaGetLastError());
    PTB_CUDA(ctx, cudaMemcpyAsync(xy2, dxy.p, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    di.release(); ds.release(); dxy.release();
    return PTB_OK;
}

int ptb_rng_stream(ptb_context* ctx, uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t n, uint32_t* out)
{
    REQUIRE_DEVICE(ctx);
    if (n == 0) return PTB_OK;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    DBuf<uint32_t> d;
    PTB_CUDA(ctx, d.resize(n));
    k_rng_stream<<<1, 1, 0, ctx->stream>>>(seed, pixel, sample, n, d.p);
    PTB_CUDA(ctx, cudaGetLastError());
    PTB_CUDA(ctx, cudaMemcpyAsync(out, d.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    d.release();
    return PTB_OK;
}

int ptb_enable_second_moment(ptb_context* ctx, int enable)
{
    if (!ctx) return PTB_ERR_INVALID;
    ctx->second_moment = enable;
    return PTB_OK;
}

int ptb_frame_begin(ptb_context* ctx, uint32_t width, uint32_t height)
{
    REQUIRE_DEVICE(ctx);
    if (width == 0 || height == 0 || (uint64_t)width * height > (uint64_t)NEE_SLOT_MASK / 16u) return ctx->fail(PTB_ERR_INVALID, "bad frame size");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool fresh = ctx->rs == nullptr;
    if (fresh) ctx->rs = new RenderState();
    RenderState& rs = *ctx->rs;
    const uint32_t npix = width * height;
    const bool bound = rs.accum && rs.accum != rs.accum_own.p;
    if (bound && npix != rs.npix) return ctx->fail(PTB_ERR_INVALID, "bound accumulation buffer has a different size; rebind it");
    rs.W = width; rs.H = height; rs.npix = npix;
    PTB_CUDA(ctx, rs.accum_own.resize(npix));
    PTB_CUDA(ctx, rs.position.resize(npix));
    PTB_CUDA(ctx, rs.id.resize(npix));
    if (ctx->second_moment) PTB_CUDA(ctx, rs.accum_sq.resize(npix));
    PTB_CUDA(ctx, rs.counters.resize(MAX_GROUPS));
    if (!rs.h_flags) PTB_CUDA(ctx, cudaHostAlloc((void**)&rs.h_flags, (size_t)MAX_GROUPS * POLL_RING * 4 * sizeof(uint32_t), cudaHostAllocDefault));
    if (!rs.ev_begin) { PTB_CUDA(ctx, cudaEventCreate(&rs.ev_begin)); PTB_CUDA(ctx, cudaEventCreate(&rs.ev_end)); }
    if (!bound) rs.accum = rs.accum_own.p;
    PTB_CUDA(ctx, cudaMemsetAsync(rs.accum, 0, (size_t)npix * 16, ctx->stream));
    if (ctx->second_moment) PTB_CUDA(ctx, cudaMemsetAsync(rs.accum_sq.p, 0, (size_t)npix * 16, ctx->stream));
    PTB_CUDA(ctx, cudaMemsetAsync(rs.position.p, 0, (size_t)npix * 16, ctx->stream));
    PTB_CUDA(ctx, cudaMemsetAsync(rs.id.p, 0, (size_t)npix * 4, ctx->stream));
    if (fresh) PTB_CUDA(ctx, cudaMemsetAsync(rs.counters.p, 0, sizeof(WfCounters) * MAX_GROUPS, ctx->stream));  // statistics persist until ptb_reset_counters
    return PTB_OK;
}

int ptb_frame_next(ptb_context* ctx)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs || !ctx->rs->accum) return ctx->fail(PTB_ERR_INVALID, "call ptb_frame_begin first");
    RenderState& rs = *ctx->rs;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    // the reference overwrites `data` and `position` every frame and shifts `id` (main.rs:204-206): only the id history survives
    PTB_CUDA(ctx, cudaMemsetAsync(rs.accum, 0, (size_t)rs.npix * 16, ctx->stream));
    if (ctx->second_moment) PTB_CUDA(ctx, cudaMemsetAsync(rs.accum_sq.p, 0, (size_t)rs.npix * 16, ctx->stream));
    return PTB_OK;
}

int ptb_camera_view_projection(ptb_context* ctx, float m16[16])
{
    if (!ctx) return PTB_ERR_INVALID;
    if (!ctx->cam_set || !m16) return ctx->fail(PTB_ERR_INVALID, "camera not set");
    mat4_inverse(ctx->cam_m, m16);  // (cam.matrix * cam.inv_projection).inverse(), main.rs:209-216
    return PTB_OK;
}

int ptb_bind_accum(ptb_context* ctx, void* dptr, size_t bytes)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "call ptb_frame_begin first");
    RenderState& rs = *ctx->rs;
    if (!dptr) { rs.accum = rs.accum_own.p; return PTB_OK; }
    if (bytes != (size_t)rs.npix * 16) return ctx->fail(PTB_ERR_INVALID, "accumulation buffer must be W*H*16 bytes");
    rs.accum = (float4*)dptr;
    return PTB_OK;
}

int ptb_accum_device_ptr(ptb_context* ctx, void** dptr, size_t* bytes)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "call ptb_frame_begin first");
    if (dptr) *dptr = ctx->rs->accum;
    if (bytes) *bytes = (size_t)ctx->rs->npix * 16;
    return PTB_OK;
}

int ptb_render(ptb_context* ctx, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->scene.built || !ctx->device_ready) return ctx->fail(PTB_ERR_NOT_BUILT, "scene not built");
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "call ptb_frame_begin first");
    if (!ctx->cam_set) return ctx->fail(PTB_ERR_INVALID, "camera not set");
    if (max_bounces > ST_BOUNCE_MASK - 1) return ctx->fail(PTB_ERR_UNSUPPORTED, "max_bounces too large");
    if (sample_count == 0) return PTB_OK;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    RenderState& rs = *ctx->rs;
    // K slots per pixel (never more than samples, so every slot renders at least one)
    const uint32_t K = std::max(1u, std::min<uint32_t>(std::min<uint32_t>((uint32_t)std::max(1, ctx->opt_slots_per_pixel), 16u), sample_count));
    const size_t n_slots = (size_t)rs.npix * K;
    if (n_slots > (size_t)NEE_SLOT_MASK) return ctx->fail(PTB_ERR_UNSUPPORTED, "too many path slots");
    PTB_CUDA(ctx, rs.s_accum.resize(n_slots));
    PTB_CUDA(ctx, rs.s_position.resize(n_slots));
    PTB_CUDA(ctx, rs.s_id.resize(n_slots));
    if (ctx->second_moment) PTB_CUDA(ctx, rs.s_accum_sq.resize(n_slots));
    PTB_CUDA(ctx, rs.ray.resize(n_slots));
    PTB_CUDA(ctx, rs.path.resize(n_slots));
    PTB_CUDA(ctx, rs.hit.resize(n_slots));
    PTB_CUDA(ctx, rs.shadow.resize(n_slots));
    PTB_CUDA(ctx, rs.probe.resize(n_slots));
    {
        DBuf<uint32_t>* u32s[] = {&rs.q_cont0, &rs.q_cont1, &rs.q_term0, &rs.q_term1, &rs.q_nee, &rs.slot_state};
        for (auto* b : u32s) PTB_CUDA(ctx, b->resize(n_slots));
    }
    PTB_CUDA(ctx, rs.q_mat.resize(n_slots * 7));
    PTB_CUDA(ctx, rs.q_grp0.resize((n_slots + 31) / 32));
    PTB_CUDA(ctx, rs.q_grp1.resize((n_slots + 31) / 32));
    PTB_CUDA(ctx, rs.volset.resize(n_slots));
    rs.n_slots = (uint32_t)n_slots;
    RenderView V{};
    V.S = ctx->S;
    V.cam = cam_dev(ctx);
    V.W = rs.W; V.H = rs.H;
    V.npix = rs.npix;
    V.tiles_x = (ctx->opt_tile_pixels && rs.W % 8u == 0u && rs.H % 4u == 0u) ? rs.W / 8u : 0u;
    // camera rays from per-tile lists of TLAS leaves (k_tile_lists): scenes of many instances (automatic: >= 64), wide kernels, tiled slots
    const uint32_t n_tiles = V.tiles_x * (rs.H / 4u);
    const bool use_tile_lists = V.tiles_x > 0u && !ctx->opt_extend_kernel && (ctx->opt_tile_lists < 0 ? ctx->scene.inst_rows.size() >= 64u : ctx->opt_tile_lists != 0);
    if (use_tile_lists)
    {
        PTB_CUDA(ctx, rs.tile_lists.resize((size_t)n_tiles * TILE_LIST_STRIDE));
        V.S.tile_lists = rs.tile_lists.p;
        V.tile_settle = ctx->opt_tile_settle ? 1u : 0u;
        V.tile_nodes = ctx->opt_tile_nodes ? 1u : 0u;
    }
    V.K = K;
    V.sample_begin = sample_begin; V.sample_count = sample_count;
    V.seed = seed;
    V.max_bounces = max_bounces;
    V.enable_nee = enable_nee ? 1u : 0u;
    V.extend_kernel = ctx->opt_extend_kernel ? 1u : 0u;
    V.accum = rs.s_accum.p;
    V.accum_sq = ctx->second_moment ? rs.s_accum_sq.p : nullptr;
    V.position = rs.s_position.p; V.id = rs.s_id.p;
    V.ray = rs.ray.p; V.path = rs.path.p; V.hit = rs.hit.p; V.shadow = rs.shadow.p; V.probe = rs.probe.p;
    V.volset = ctx->scene.volumes.empty() ? nullptr : rs.volset.p;
    const int threads = 128;
    const int grid = ctx->sm_count * std::max(1, ctx->opt_blocks_per_sm);
    cudaStream_t st = ctx->stream;
    const bool prof = ctx->profiling != 0;
    const bool sort = ctx->opt_sort_materials != 0, bin = V.extend_kernel != 0u;

    // Wavefront groups: the slots are split into G contiguous ranges (boundaries at multiples of 32), each an independent
    // wavefront with its own queues and counters on its own stream.  Slots never interact, so results do not depend on G.  While
    // most slots are live the groups simply share the device; once few paths remain — every launch then lasts as long as the
    // slowest ray in it, a chain of dependent loads, whatever the number of rays — the G chains overlap instead of following one
    // another.  Per-launch profiling (CUDA events between consecutive launches of one stream) needs a single group.
    // automatic (-1): two rays per lane when the scene's trees are large enough for the stages of a warp to diverge and for a ray to live
    // long enough to repay its record set-up (measured: C3, 519 wide nodes, + 6.4 % per frame, C4, 1.03 M nodes, + 5.5 %; C1, 332 nodes, - 3 %,
    // C2, 8 nodes, - 3.5 % — the threshold sits between those, the option overrides it)
    const int mr_want = ctx->opt_multiray < 0 ? (ctx->scene.wide.size() >= 384 ? 2 : 0) : ctx->opt_multiray;
    const int mr_k = (!bin && threads == 128 && (mr_want == 2 || mr_want == 3)) ? mr_want : 0;
    // wavefront groups and the refill threshold of the one-ray kernels: by default (< 0) 1 group / 4 lanes with the multiray kernels, 2 / 8
    // without (sweeps: profiles/r3g_*, r3h_*: C3 256.5 -> 252.6 ms, C4 172.4 -> 170.3; C2 keeps 2 / 8: 162.3 against 167.5 with one group)
    const int groups_want = ctx->opt_wavefront_groups < 0 ? (mr_k ? 1 : 2) : ctx->opt_wavefront_groups;
    const int refill_want = ctx->opt_refill_below < 0 ? (mr_k ? 4 : 8) : ctx->opt_refill_below;
    const int G = prof ? 1 : std::max(1, std::min<int>(std::min(MAX_GROUPS, groups_want), (int)(n_slots / 4096u) + 1));
    while ((int)rs.group_streams.size() < G - 1)
    {
        cudaStream_t s2;
        PTB_CUDA(ctx, cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
        rs.group_streams.push_back(s2);
    }
    while ((int)rs.group_events.size() < MAX_GROUPS + 1)
    {
        cudaEvent_t e;
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        rs.group_events.push_back(e);
    }
    // k_nee beside the next iteration's k_extend (thin iterations; option nee_overlap): needs a stream and two events per group
    const bool overlap_on = !prof && !bin && ctx->opt_nee_overlap != 0 && enable_nee;
    while (overlap_on && (int)rs.nee_streams.size() < G)
    {
        cudaStream_t s2;
        PTB_CUDA(ctx, cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
        rs.nee_streams.push_back(s2);
        for (int k = 0; k < 2; ++k)
        {
            cudaEvent_t e;
            PTB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            rs.nee_events.push_back(e);
        }
    }
    struct GroupRun
    {
        Queues Q;
        cudaStream_t st;
        bool nee_pending = false;  // a k_nee of this group runs on its own stream: whatever needs its results waits for nee_events[2g + 1]
        uint32_t it = 0;
        bool dense = false, done = false;
        uint32_t* h_flags = nullptr;
        uint32_t tail_threshold = 0;
        uint32_t bound = 0;  // upper bound of the group's live slots (the last polled count; live counts never grow)
    };
    GroupRun gr[MAX_GROUPS];
    for (int g = 0; g < G; ++g)
    {
        const uint32_t b0 = (uint32_t)((n_slots * (size_t)g / (size_t)G) & ~(size_t)31), b1 = g + 1 == G ? (uint32_t)n_slots : (uint32_t)((n_slots * (size_t)(g + 1) / (size_t)G) & ~(size_t)31);
        Queues& Q = gr[g].Q;
        Q.cont[0] = rs.q_cont0.p + b0; Q.cont[1] = rs.q_cont1.p + b0; Q.term[0] = rs.q_term0.p + b0; Q.term[1] = rs.q_term1.p + b0;
        Q.mat = rs.q_mat.p + (size_t)b0 * 7; Q.nee = rs.q_nee.p + b0; Q.state = rs.slot_state.p; Q.grp[0] = rs.q_grp0.p + (b0 >> 5); Q.grp[1] = rs.q_grp1.p + (b0 >> 5);
        Q.npix = b1 - b0;  // queue capacity and stride: one entry per slot of the group
        Q.slot_base = b0; Q.slot_end = b1;
        Q.c = rs.counters.p + g;
        gr[g].st = g == 0 ? st : rs.group_streams[g - 1];
        gr[g].h_flags = rs.h_flags + (size_t)g * POLL_RING * 4;
    }

    size_t prof_used = 0;
    std::vector<int> prof_class;
    auto mark = [&](int cls) -> cudaError_t {
        if (!prof) return cudaSuccess;
        if (prof_used == rs.prof_events.size())
        {
            cudaEvent_t e;
            cudaError_t r = cudaEventCreate(&e);
            if (r != cudaSuccess) return r;
            rs.prof_events.push_back(e);
        }
        prof_class.push_back(cls);
        return cudaEventRecord(rs.prof_events[prof_used++], st);
    };
    const uint32_t NEV = 64;
    while (rs.poll_events.size() < NEV * MAX_GROUPS)
    {
        cudaEvent_t e;
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        rs.poll_events.push_back(e);
    }
    if (rs.tail_capacity < 0 || rs.tail_kind != (ctx->opt_tail_roles ? 1 : 0))
    {
        rs.tail_kind = ctx->opt_tail_roles ? 1 : 0;
        int per_sm = 0;
        if (ctx->opt_tail_roles) PTB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_tail_roles<false>, 96, 0));
        else PTB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_tail<false>, 64, 0));
        rs.tail_capacity = per_sm * (ctx->opt_tail_roles ? 32 : 64) * ctx->sm_count;
    }
    // switch to the tail kernel once the remaining slots (of all groups together) are a few waves of its resident capacity (0 disables,
    // <0 = automatic: 2.5 x capacity, the measured optimum on C3 — below it the thin wavefront iterations cost more per bounce than
    // the tail kernel, above it the tail kernel's one-thread-per-ray throughput does)
    const uint32_t tail_threshold = (ctx->opt_tail_threshold < 0 ? (uint32_t)rs.tail_capacity * 5u / 2u : (uint32_t)ctx->opt_tail_threshold) / (uint32_t)G;
    // The persistent path kernel takes over earlier and replaces thin iterations and tail alike (wide kernels only): once the live
    // slots are down to a few rays per resident lane, a three-kernel iteration costs its latency floor, not its work (automatic:
    // PATH_AUTO_PER_LANE slots per lane of the k_path grid, see DESIGN.md for the sweep)
    const uint32_t path_lanes = (uint32_t)(ctx->sm_count * ctx->opt_path_blocks * threads);
    const uint32_t path_threshold = bin ? 0u : (ctx->opt_path_threshold < 0 ? path_lanes * PATH_AUTO_PER_LANE : (uint32_t)ctx->opt_path_threshold) / (uint32_t)G;

    PTB_CUDA(ctx, cudaEventRecord(rs.ev_begin, st));
    PTB_CUDA(ctx, mark(-1));
    if (use_tile_lists)
    {
        k_tile_lists<<<(n_tiles + 127u) / 128u, 128, 0, st>>>(V, rs.tile_lists.p, n_tiles);
        rs.launches += 1;
        if (std::getenv("PTB_TILE_STATS"))
        {
            // development aid: histogram of the list lengths (stderr)
            std::vector<uint2> h((size_t)n_tiles * TILE_LIST_STRIDE);
            PTB_CUDA(ctx, cudaMemcpyAsync(h.data(), rs.tile_lists.p, h.size() * sizeof(uint2), cudaMemcpyDeviceToHost, st));
            PTB_CUDA(ctx, cudaStreamSynchronize(st));
            uint32_t hist[TILE_LIST_STRIDE + 1] = {0};
            unsigned long long sum = 0;
            for (uint32_t t = 0; t < n_tiles; ++t)
            {
                const uint32_t n = h[(size_t)t * TILE_LIST_STRIDE].x;
                hist[n == TILE_LIST_NONE ? TILE_LIST_STRIDE : n] += 1;
                if (n != TILE_LIST_NONE) sum += n;
            }
            std::fprintf(stderr, "tile lists: %u tiles, mean length %.2f, overflow %u;", n_tiles, (double)sum / n_tiles, hist[TILE_LIST_STRIDE]);
            for (uint32_t k = 0; k < TILE_LIST_STRIDE; ++k) std::fprintf(stderr, " %u:%u", k, hist[k]);
            std::fprintf(stderr, "\n");
        }
    }
    for (int g = 0; g < G; ++g) k_init<<<(gr[g].Q.npix + threads - 1) / threads, threads, 0, st>>>(V, gr[g].Q, rs.id.p);
    PTB_CUDA(ctx, mark(1));
    rs.launches += G;
    if (G > 1)
    {
        PTB_CUDA(ctx, cudaEventRecord(rs.group_events[MAX_GROUPS], st));  // fork
        for (int g = 1; g < G; ++g) PTB_CUDA(ctx, cudaStreamWaitEvent(gr[g].st, rs.group_events[MAX_GROUPS], 0));
    }
    const uint32_t LAG = 8;
    rs.log_iters = 0;
    // k_shade_ordered while the listed groups are at least dense_threshold % full of live slots (never for the reference-order variant)
    const uint64_t dense_pct = (uint64_t)std::max(0, std::min(100, ctx->opt_dense_threshold));
    const int shade_grid = ctx->sm_count * 8;
    // span-local sort by shade kind: worth its extra passes when warps would otherwise mix surface materials (auto: >= 2 non-emissive kinds)
    bool shade_sort = ctx->opt_shade_sort > 0;
    if (ctx->opt_shade_sort < 0)
    {
        uint32_t kinds = 0;
        for (const auto& b : ctx->scene.blas_desc) kinds |= 1u << std::min<uint32_t>(b.material.kind, 7u);
        shade_sort = __builtin_popcount(kinds & ~2u) >= 2;  // kind 1 = Emissive
    }
    for (int g = 0; g < G; ++g) gr[g].dense = !bin && ctx->opt_dense != 0;
    // scene class of the slot-ordered shade kernel: the instantiation without the material kinds (and media) the scene does not hold
    int shade_class = 0;
    if (ctx->opt_shade_class != 0)
    {
        uint32_t kinds = 0;
        for (const auto& b : ctx->scene.blas_desc) kinds |= 1u << std::min<uint32_t>(b.material.kind, 7u);
        const bool media = !ctx->scene.volumes.empty();
        if (!media && !(kinds & ~SHADE_KMASK_1)) shade_class = 1;
        else if (!media && !(kinds & ~SHADE_KMASK_2)) shade_class = 2;
        else if (!media && !(kinds & ~0x3fu)) shade_class = 3;
    }

    // multiray kernels (option multiray = 2 / 3 rays per lane, multiray.cuh): dynamic shared memory for the ray records, as many resident
    // blocks per SM as fit (or mr_blocks)
    MrOpts mro = ctx->opt_mr;
    if (mro.free_quorum <= 0) mro.free_quorum = 24;
    const size_t mr_es = (size_t)mr_k * (MR_JOB + MrExtendJob::JOBW) * 128 * sizeof(uint32_t), mr_ns = (size_t)mr_k * (MR_JOB + MrNeeJob::JOBW) * 128 * sizeof(uint32_t);
    int mr_eb = 0, mr_nb = 0;
    // thin iterations (fewer live slots than mr_min_per_lane per resident lane of the one-ray kernels) go to the one-ray-per-lane kernels:
    // a lane that holds one ray only waits at every gate
    const uint32_t mr_min = (uint32_t)std::max(0, ctx->opt_mr_min_per_lane) * (uint32_t)(ctx->sm_count * 7 * 128);
    // k_nee beside the next iteration's k_extend once a group's live slots are below nee_overlap per lane of the one-ray grid (0 = never)
    const uint32_t ov_below = ctx->opt_nee_overlap < 0 ? 0xffffffffu : (uint32_t)ctx->opt_nee_overlap * (uint32_t)(ctx->sm_count * 7 * 128);
    if (mr_k)
    {
        if (mr_k == 3)
        {
            PTB_CUDA(ctx, cudaFuncSetAttribute(k_extend_mr<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mr_es));
            PTB_CUDA(ctx, cudaFuncSetAttribute(k_nee_mr<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mr_ns));
            PTB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&mr_eb, k_extend_mr<3>, 128, mr_es));
            PTB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&mr_nb, k_nee_mr<3>, 128, mr_ns));
        }
        else
        {
            PTB_CUDA(ctx, cudaFuncSetAttribute(k_extend_mr<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mr_es));
            PTB_CUDA(ctx, cudaFuncSetAttribute(k_nee_mr<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mr_ns));
            PTB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&mr_eb, k_extend_mr<2>, 128, mr_es));
            PTB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&mr_nb, k_nee_mr<2>, 128, mr_ns));
        }
        if (mro.blocks_per_sm > 0) { mr_eb = std::min(mr_eb, mro.blocks_per_sm); mr_nb = std::min(mr_nb, mro.blocks_per_sm); }
        if (mr_eb < 1 || mr_nb < 1) return ctx->fail(PTB_ERR_CUDA, "multiray kernels do not fit on this device");
    }
    // Grid of a persistent traversal / queue kernel compiled for `blocks` resident blocks per SM, given an upper bound of the work items.
    // shared_device = 1 (two contexts rendering alternate frames on one device, ptb_render_async): the dense launches leave one block
    // slot per SM free and the thin ones take only the blocks their work fills at thin_rays_per_lane items per lane, so the other
    // context's kernels become resident at once instead of at this context's next launch boundary.
    auto sized_grid = [&](int blocks, uint32_t bound) -> int {
        int per_sm = blocks;
        if (ctx->opt_shared_device && per_sm > 1) per_sm -= 1;
        int g = ctx->sm_count * per_sm;
        if (ctx->opt_thin_rays_per_lane > 0)
        {
            const uint64_t per_block = (uint64_t)threads * (uint64_t)ctx->opt_thin_rays_per_lane;
            g = (int)std::min<uint64_t>((uint64_t)g, std::max<uint64_t>(1u, ((uint64_t)bound + per_block - 1u) / per_block));
        }
        return g;
    };
    for (int g = 0; g < G; ++g) gr[g].bound = gr[g].Q.npix;
    auto launch_path = [&](int g, uint32_t it) -> int {
        const int pg = ctx->sm_count * ctx->opt_path_blocks, fq = ctx->opt_finish_quorum, rb = refill_want, nq = ctx->opt_node_quorum;
        switch (ctx->opt_path_blocks)
        {
        case 4: k_path<4><<<pg, threads, 0, gr[g].st>>>(V, gr[g].Q, it, rb, nq, fq); break;
        case 6: k_path<6><<<pg, threads, 0, gr[g].st>>>(V, gr[g].Q, it, rb, nq, fq); break;
        default: k_path<5><<<ctx->sm_count * 5, threads, 0, gr[g].st>>>(V, gr[g].Q, it, rb, nq, fq); break;
        }
        PTB_CUDA(ctx, mark(5));
        PTB_CUDA(ctx, cudaGetLastError());
        rs.launches += 1;
        gr[g].done = true;
        return PTB_OK;
    };
    // one wavefront iteration of group g (three launches on its stream) and its poll; sets done when the group has finished
    auto step_group = [&](int g) -> int {
        GroupRun& R = gr[g];
        const Queues& Q = R.Q;
        cudaStream_t gs = R.st;
        const uint32_t it = R.it;
        const bool dense = R.dense;
        const int rb = refill_want, rbn = ctx->opt_nee_refill_below, nq = ctx->opt_node_quorum;
        // thin iterations: k_nee of iteration it on its own stream, beside part 1 (continuing slots) of iteration it + 1's k_extend
        const bool ov = overlap_on && R.bound < ov_below && !(mr_k && R.bound >= mr_min);
        cudaStream_t ns = overlap_on ? rs.nee_streams[g] : gs;
        auto join_nee = [&]() -> cudaError_t {
            if (!R.nee_pending) return cudaSuccess;
            R.nee_pending = false;
            return cudaStreamWaitEvent(gs, rs.nee_events[2 * g + 1], 0);
        };
        if (!ov) PTB_CUDA(ctx, join_nee());
        if (bin) k_extend_bin<<<grid, threads, 0, gs>>>(V, Q, it);
        else if (ov)
        {
            const int so = dense ? -1 : (sort ? 1 : 0);
            const int eg = sized_grid(7, R.bound);
            k_extend<7><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq, 1);
            PTB_CUDA(ctx, join_nee());  // a slot to regenerate is complete only with its last direct-light rays
            k_extend<7><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq, 2);
            rs.launches += 1;
        }
        else if (mr_k && R.bound >= mr_min)
        {
            const int so = dense ? -1 : (sort ? 1 : 0);
            const int eg = sized_grid(mr_eb, R.bound);
            const int part = ctx->opt_mr_split ? 1 : 0;
            if (mr_k == 3) k_extend_mr<3><<<eg, 128, mr_es, gs>>>(V, Q, it, so, mro, part);
            else k_extend_mr<2><<<eg, 128, mr_es, gs>>>(V, Q, it, so, mro, part);
            if (part)
            {
                switch (ctx->opt_regen_blocks)
                {
                case 5: k_extend<5><<<sized_grid(5, R.bound), threads, 0, gs>>>(V, Q, it, so, rb, nq, 2); break;
                case 6: k_extend<6><<<sized_grid(6, R.bound), threads, 0, gs>>>(V, Q, it, so, rb, nq, 2); break;
                case 8: k_extend<8><<<sized_grid(8, R.bound), threads, 0, gs>>>(V, Q, it, so, rb, nq, 2); break;
                default: k_extend<7><<<sized_grid(7, R.bound), threads, 0, gs>>>(V, Q, it, so, rb, nq, 2); break;
                }
                rs.launches += 1;
            }
        }
        else
        {
            const int so = dense ? -1 : (sort ? 1 : 0);
            const int eb = (ctx->opt_extend_blocks == 5 || ctx->opt_extend_blocks == 7 || ctx->opt_extend_blocks == 8) ? ctx->opt_extend_blocks : 6;
            const int eg = sized_grid(eb, R.bound);
            switch (eb)
            {
            case 5: k_extend<5><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq); break;
            case 7:
                if (ctx->opt_stack_shared == 8) k_extend<7, 8><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq);
                else if (ctx->opt_stack_shared == 12) k_extend<7, 12><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq);
                else if (ctx->opt_stack_shared == 16) k_extend<7, 16><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq);
                else k_extend<7><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq);
                break;
            case 8: k_extend<8><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq); break;
            default: k_extend<6><<<eg, threads, 0, gs>>>(V, Q, it, so, rb, nq); break;
            }
        }
        PTB_CUDA(ctx, mark(2));
        if (dense) launch_shade_ordered(shade_sort, shade_class, shade_grid, gs, V, Q, it);
        else k_shade<<<sized_grid(std::max(1, ctx->opt_blocks_per_sm), R.bound), threads, 0, gs>>>(V, Q, it);
        PTB_CUDA(ctx, mark(3));
        rs.launches += 2;
        if (V.enable_nee)
        {
            if (bin) k_nee_bin<<<grid, threads, 0, gs>>>(V, Q, it);
            else if (ov)
            {
                PTB_CUDA(ctx, cudaEventRecord(rs.nee_events[2 * g], gs));
                PTB_CUDA(ctx, cudaStreamWaitEvent(ns, rs.nee_events[2 * g], 0));
                k_nee<8><<<sized_grid(8, R.bound), threads, 0, ns>>>(V, Q, it, rbn, nq);
                PTB_CUDA(ctx, cudaEventRecord(rs.nee_events[2 * g + 1], ns));
                R.nee_pending = true;
            }
            else if (mr_k && R.bound >= mr_min)
            {
                const int ng = sized_grid(mr_nb, R.bound);
                if (mr_k == 3) k_nee_mr<3><<<ng, 128, mr_ns, gs>>>(V, Q, it, mro);
                else k_nee_mr<2><<<ng, 128, mr_ns, gs>>>(V, Q, it, mro);
            }
            else
            {
                const int nb = (ctx->opt_nee_blocks == 4 || ctx->opt_nee_blocks == 6 || ctx->opt_nee_blocks == 7 || ctx->opt_nee_blocks == 8) ? ctx->opt_nee_blocks : 5;
                const int ng = sized_grid(nb, R.bound);
                switch (nb)
                {
                case 4: k_nee<4><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq); break;
                case 7: k_nee<7><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq); break;
                case 8:
                    if (ctx->opt_stack_shared == 8) k_nee<8, 8><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq);
                    else if (ctx->opt_stack_shared == 12) k_nee<8, 12><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq);
                    else if (ctx->opt_stack_shared == 16) k_nee<8, 16><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq);
                    else k_nee<8><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq);
                    break;
                case 6: k_nee<6><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq); break;
                default: k_nee<5><<<ng, threads, 0, gs>>>(V, Q, it, rbn, nq); break;
                }
            }
            PTB_CUDA(ctx, mark(4));
            rs.launches += 1;
        }
        PTB_CUDA(ctx, cudaGetLastError());
        // active-slot count of the next iteration, read LAG iterations behind the launch front
        PTB_CUDA(ctx, cudaMemcpyAsync(&R.h_flags[(it % POLL_RING) * 4], &Q.c->set[(it + 1u) % WF_SETS].n_cont, 16, cudaMemcpyDeviceToHost, gs));
        PTB_CUDA(ctx, cudaEventRecord(rs.poll_events[(size_t)g * NEV + it % NEV], gs));
        if (g == 0) rs.log_iters = it + 1;
        R.it = it + 1;
        if (it >= LAG)
        {
            PTB_CUDA(ctx, cudaEventSynchronize(rs.poll_events[(size_t)g * NEV + (it - LAG) % NEV]));
            const uint32_t* f = &R.h_flags[((it - LAG) % POLL_RING) * 4];
            const uint32_t active = f[0] + f[1];
            R.bound = active;
            if (active == 0u) { R.done = true; return PTB_OK; }  // a pending k_nee is joined with the group (below)
            const bool to_path = active <= path_threshold;
            const bool to_tail = !to_path && active <= tail_threshold;
            // live slots scattered thinly over the listed groups: the queue shade kernel from the next iteration on (the queues are valid as they are)
            if (dense && (uint64_t)active * 100u < dense_pct * 32u * (uint64_t)f[3]) R.dense = false;
            if (to_path || to_tail) PTB_CUDA(ctx, join_nee());
            if (to_path)
            {
                const int rc = launch_path(g, it + 1u);
                if (rc != PTB_OK) return rc;
            }
            if (to_tail)
            {
                // active counts never grow: `active` bounds the queue at it+1
                const uint32_t spb = (uint32_t)std::max(1, std::min(32, ctx->opt_tail_slots_per_block));
                if (ctx->opt_tail_roles && bin) k_tail_roles<true><<<(active + spb - 1u) / spb, 96, 0, gs>>>(V, Q, it + 1u, spb);
                else if (ctx->opt_tail_roles) k_tail_roles<false><<<(active + spb - 1u) / spb, 96, 0, gs>>>(V, Q, it + 1u, spb);
                else if (bin) k_tail<true><<<(active + 63u) / 64u, 64, 0, gs>>>(V, Q, it + 1u);
                else k_tail<false><<<(active + 63u) / 64u, 64, 0, gs>>>(V, Q, it + 1u);
                PTB_CUDA(ctx, mark(5));
                PTB_CUDA(ctx, cudaGetLastError());
                rs.launches += 1;
                R.done = true;
            }
        }
        return PTB_OK;
    };
    // a group that starts at or below the path threshold never enters the wavefront loop: k_init's queues are k_path's input
    // error exit from the launch loop: kernels of the forked group streams may still be running — drain them before handing the
    // context back, or a later call on ctx->stream would race with them
    auto abandon = [&](int rc) -> int {
        for (auto ns : rs.nee_streams) cudaStreamSynchronize(ns);
        for (int g = 1; g < G; ++g) cudaStreamSynchronize(gr[g].st);
        cudaStreamSynchronize(st);
        return rc;
    };
    for (int g = 0; g < G; ++g)
        if (gr[g].Q.npix <= path_threshold)
        {
            const int rc = launch_path(g, 0u);
            if (rc != PTB_OK) return abandon(rc);
        }
    for (int running = G; running > 0;)
    {
        running = 0;
        for (int g = 0; g < G; ++g)
        {
            if (gr[g].done) continue;
            const int rc = step_group(g);
            if (rc != PTB_OK) return abandon(rc);
            if (!gr[g].done) ++running;
        }
    }
    for (int g = 0; g < G; ++g)
        if (gr[g].nee_pending) PTB_CUDA(ctx, cudaStreamWaitEvent(gr[g].st, rs.nee_events[2 * g + 1], 0));
    for (int g = 1; g < G; ++g)  // join
    {
        PTB_CUDA(ctx, cudaEventRecord(rs.group_events[g], gr[g].st));
        PTB_CUDA(ctx, cudaStreamWaitEvent(st, rs.group_events[g], 0));
    }
    k_combine<<<(rs.npix + 255) / 256, 256, 0, st>>>(V, rs.accum, ctx->second_moment ? rs.accum_sq.p : nullptr, rs.position.p, rs.id.p, sample_count);
    PTB_CUDA(ctx, mark(6));
    PTB_CUDA(ctx, cudaGetLastError());
    rs.launches += 1;
    PTB_CUDA(ctx, cudaEventRecord(rs.ev_end, st));
    PTB_CUDA(ctx, cudaEventSynchronize(rs.ev_end));
    float ms = 0;
    PTB_CUDA(ctx, cudaEventElapsedTime(&ms, rs.ev_begin, rs.ev_end));
    rs.t_total_ms += ms;
    rs.t_class_ms[0] += ms;
    rs.n_class[0] += 1;
    if (prof)
    {
        // development aid: PTB_LAUNCH_LOG=<file> lists every launch of the render (stage class, microseconds)
        const char* log_path = std::getenv("PTB_LAUNCH_LOG");
        std::FILE* launch_log = log_path ? std::fopen(log_path, "w") : nullptr;
        rs.launch_log.clear();
        for (size_t i = 1; i < prof_used; ++i)
        {
            float d = 0;
            PTB_CUDA(ctx, cudaEventElapsedTime(&d, rs.prof_events[i - 1], rs.prof_events[i]));
            const int cls = prof_class[i];
            if (cls >= 1 && cls < 8) { rs.t_class_ms[cls] += d; rs.n_class[cls] += 1; }
            rs.launch_log.emplace_back(cls, d * 1000.0f);
            if (launch_log) std::fprintf(launch_log, "%d %.3f\n", cls, d * 1000.0f);
        }
        if (launch_log) std::fclose(launch_log);
    }
    return PTB_OK;
}

// Frame-after-frame hosts (the reference's loop, main.rs:179-219): the wavefront of one frame thins out into ~10^2 latency-bound
// iterations and a tail during which most of the device idles (C3: 30 % of the frame for 8 % of its rays).  A second context on
// the same device fills that time with the dense phase of the next frame: ptb_render_async returns at once, the render loop runs
// on a worker thread of the context, ptb_wait joins it.  Between the two calls the context must not be used otherwise.
int ptb_render_async(ptb_context* ctx, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee)
{
    REQUIRE_DEVICE(ctx);
    if (ctx->async_thread) return ctx->fail(PTB_ERR_INVALID, "ptb_render_async: a render is already in flight on this context (call ptb_wait)");
    ctx->async_rc = PTB_OK;
    ctx->async_thread = new (std::nothrow) std::thread([=] { ctx->async_rc = ptb_render(ctx, sample_begin, sample_count, seed, max_bounces, enable_nee); });
    if (!ctx->async_thread) return ctx->fail(PTB_ERR_OOM, "ptb_render_async: no worker thread");
    return PTB_OK;
}
int ptb_wait(ptb_context* ctx)
{
    if (!ctx) return PTB_ERR_INVALID;
    return ctx->wait_async();  // the worker's status; its message is ptb_last_error
}

int ptb_get_launch_log(ptb_context* ctx, uint32_t capacity, int32_t* stage, float* microseconds, uint32_t* count)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs || !count) return ctx->fail(PTB_ERR_INVALID, "no frame");
    const uint32_t n = std::min<uint32_t>((uint32_t)ctx->rs->launch_log.size(), capacity);
    for (uint32_t i = 0; i < n; ++i)
    {
        if (stage) stage[i] = ctx->rs->launch_log[i].first;
        if (microseconds) microseconds[i] = ctx->rs->launch_log[i].second;
    }
    *count = n;
    return PTB_OK;
}

int ptb_get_iteration_log(ptb_context* ctx, uint32_t capacity, uint32_t* active, uint32_t* count)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs || !count) return ctx->fail(PTB_ERR_INVALID, "no frame");
    const uint32_t n = std::min<uint32_t>(std::min<uint32_t>(ctx->rs->log_iters, POLL_RING), capacity);
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (uint32_t i = 0; i < n && active; ++i) active[i] = ctx->rs->h_flags[i * 4] + ctx->rs->h_flags[i * 4 + 1];
    *count = n;
    return PTB_OK;
}

int ptb_read_accum(ptb_context* ctx, float* rgba)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs || !rgba) return ctx->fail(PTB_ERR_INVALID, "no frame");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    PTB_CUDA(ctx, cudaMemcpyAsync(rgba, ctx->rs->accum, (size_t)ctx->rs->npix * 16, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}
int ptb_read_second_moment(ptb_context* ctx, float* rgba)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs || !rgba || !ctx->second_moment || !ctx->rs->accum_sq.p) return ctx->fail(PTB_ERR_INVALID, "second moment not enabled");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    PTB_CUDA(ctx, cudaMemcpyAsync(rgba, ctx->rs->accum_sq.p, (size_t)ctx->rs->npix * 16, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}
int ptb_read_aux(ptb_context* ctx, float* position4, uint32_t* id)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "no frame");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (position4) PTB_CUDA(ctx, cudaMemcpyAsync(position4, ctx->rs->position.p, (size_t)ctx->rs->npix * 16, cudaMemcpyDeviceToHost, ctx->stream));
    if (id) PTB_CUDA(ctx, cudaMemcpyAsync(id, ctx->rs->id.p, (size_t)ctx->rs->npix * 4, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_get_counters(ptb_context* ctx, uint64_t out[8])
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "no frame");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    WfCounters cg[MAX_GROUPS], c{};
    PTB_CUDA(ctx, cudaMemcpyAsync(cg, ctx->rs->counters.p, sizeof(cg), cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int g = 0; g < MAX_GROUPS; ++g)  // one counter block per wavefront group; iterations: the longest group
    {
        c.stat_extend += cg[g].stat_extend; c.stat_shadow += cg[g].stat_shadow; c.stat_probe += cg[g].stat_probe; c.stat_paths += cg[g].stat_paths;
        c.stat_probe_shadow += cg[g].stat_probe_shadow; c.stat_iters = std::max(c.stat_iters, cg[g].stat_iters); c.stat_settled += cg[g].stat_settled; c.stat_extend_path += cg[g].stat_extend_path;
    }
    out[0] = c.stat_extend;
    out[1] = c.stat_shadow + c.stat_probe_shadow;
    out[2] = c.stat_probe;
    out[3] = c.stat_paths;
    out[4] = c.stat_iters;
    out[5] = ctx->rs->launches;
    out[6] = c.stat_extend_path;  // part of [0] traced by the single-launch kernels (k_path, tail) rather than by k_extend launches
    out[7] = c.stat_settled;  // primary rays of a camera outside the scene bounds that the bounds pre-test settled without a traversal (not in [0])
    return PTB_OK;
}
int ptb_reset_counters(ptb_context* ctx)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "no frame");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    PTB_CUDA(ctx, cudaMemsetAsync(ctx->rs->counters.p, 0, sizeof(WfCounters) * MAX_GROUPS, ctx->stream));
    ctx->rs->launches = 0;
    ctx->rs->t_total_ms = 0;
    for (int k = 0; k < 8; ++k) { ctx->rs->t_class_ms[k] = 0; ctx->rs->n_class[k] = 0; }
    return PTB_OK;
}
int ptb_set_profiling(ptb_context* ctx, int enable)
{
    if (!ctx) return PTB_ERR_INVALID;
    ctx->profiling = enable;
    return PTB_OK;
}
int ptb_get_timing(ptb_context* ctx, float times_ms[8], uint32_t launches[8])
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "no frame");
    for (int k = 0; k < 8; ++k)
    {
        if (times_ms) times_ms[k] = ctx->rs->t_class_ms[k];
        if (launches) launches[k] = ctx->rs->n_class[k];
    }
    return PTB_OK;
}

}  // extern "C"
