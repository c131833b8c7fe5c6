// persist.cuh — persistent-thread ray loop with dynamic fetch (device only).
//
// Every warp keeps 32 rays in flight: a lane whose traversal has finished hands its result to the job (`finish`, executed
// convergently so the job can use warp-aggregated queue pushes), and as soon as fewer than `refill_below` lanes are still
// traversing the idle lanes fetch new work items with one atomicAdd for the whole warp.  Between refills all busy lanes run
// wide_step() (traverse.h), whose stages sit at single code locations, so the lanes of a warp re-converge at every step
// whatever mix of TLAS nodes, BLAS nodes, leaves and triangles they are at.
//
// Job interface (per lane):
//   bool start(uint32_t index, WideTrav& tr)   fetch work item `index`, call wide_begin; false = the item needs no traversal
//   bool finish(WideTrav& tr, bool mine)       called by ALL lanes of the warp together; `mine` = this lane's traversal just
//                                              ended.  May begin a follow-up traversal for the lane (return true).
// `finish_quorum` (jobs whose finish() is expensive — k_path shades in it): the stepping loop also ends, and finish() runs, as
// soon as that many lanes are waiting with a finished traversal — or a quarter of the warp's live lanes, whichever is smaller, so
// that a warp down to a few long paths does not make a finished lane wait for the others.  Compiled in only with GATED.
#pragma once
#include "traverse.h"

namespace ptb {

template <int MODE, class Job, bool GATED = false, int SSTACK = 0, bool LIST = false>
__device__ __forceinline__ void persistent_trace(const SceneDev& S, Job& job, uint32_t* cursor, uint32_t total, int refill_below, int node_quorum, int finish_quorum = 33)
{
    // traversal stack: local memory, or (SSTACK > 0, blocks of 128 threads) its first SSTACK entries in shared memory
    uint2 lstack[WIDE_STACK_CAP - SSTACK];
    __shared__ uint2 s_stack[SSTACK > 0 ? SSTACK : 1][SSTACK > 0 ? 128 : 1];
    HybridStack<SSTACK, 128> hstack{&s_stack[0][SSTACK > 0 ? threadIdx.x : 0], lstack};
    auto stack = [&]() {
        if constexpr (SSTACK > 0) return hstack;
        else return (uint2*)lstack;
    }();
    WideTrav tr;
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    int state = 0;  // 0 idle, 1 traversing, 2 traversal finished (result not yet consumed)
    // Work distribution: the warp takes BATCH consecutive items from the global cursor at a time (one atomicAdd per BATCH
    // items, not per refill) and deals them out to its idle lanes from the warp-uniform range [wbase, wend).
    // Near the end of the queue the batch shrinks to what the idle lanes need right now, so that no warp is left holding a
    // long private backlog while the others have run dry.
    constexpr uint32_t BATCH = 64;
    const uint32_t n_warps = gridDim.x * blockDim.x >> 5;
    const uint32_t fine_below = n_warps * (2u * BATCH);
    // a queue shorter than one full set of lanes is spread evenly over the warps of the grid: a warp runs as long as the slowest
    // of its rays, and a launch as long as its slowest warp
    const uint32_t fair = max(1u, (total + n_warps - 1u) / n_warps);
    uint32_t wbase = 0, wend = 0, seen = 0;
    bool global_more = total > 0u;
    for (;;)
    {
        if (__any_sync(FULL, state == 2))
        {
            const bool again = job.finish(tr, state == 2);
            if (state == 2) state = again ? 1 : 0;
        }
        // deal work to the idle lanes; an item may need no traversal (start() false), so repeat while lanes stay idle and items remain
        for (int pass = 0; pass < 4 && (wbase < wend || global_more); ++pass)
        {
            const unsigned idle = __ballot_sync(FULL, state == 0);
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
            if (state == 0 && rank < avail && job.start(wbase + rank, tr)) state = 1;
            wbase += min(n, avail);
        }
        const bool more = wbase < wend || global_more;
        unsigned busy = __ballot_sync(FULL, state == 1);
        if (!busy)
        {
            if (!more && !__any_sync(FULL, state == 2)) break;
            continue;
        }
        const int keep = more ? refill_below : 1;
        unsigned done = 0u;
        do
        {
            if (state == 1 && !wide_step_pop<LIST>(tr, S, stack)) state = 2;
            // the node stage runs when at least node_quorum lanes want it, or when no lane has leaf/triangle work left
            const unsigned wn = __ballot_sync(FULL, state == 1 && wide_wants_node(tr));
            const unsigned wl = __ballot_sync(FULL, state == 1 && wide_has_leaf_work(tr));
            const bool allow = __popc(wn) >= node_quorum || wl == 0u;
            if (state == 1 && !wide_step_work<MODE>(tr, S, stack, allow)) state = 2;
            busy = __ballot_sync(FULL, state == 1);
            if constexpr (GATED) done = __ballot_sync(FULL, state == 2);
        } while (__popc(busy) >= keep && (!GATED || __popc(done) < min(finish_quorum, (__popc(busy | done) + 3) >> 2)));
    }
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace ptb
