// kernels_trace.cu — fixed-batch TLAS::intersect / TLAS::any_intersect kernels (ptb_trace_*).
// 8-wide layout: the dynamic-fetch persistent loop of persist.cuh (a warp keeps 32 rays in flight, lanes whose ray has
// finished write their result and take the next ray from a global cursor, 64 rays per atomic).  Reference-order binary
// layout (kernel = 1, the checker for traversal order): one ray per thread, grid-stride.  The per-ray work is traverse.h.
#include "context.h"
#include "persist.cuh"
#include "team.cuh"
#include "multiray.cuh"

namespace ptb {

__device__ __forceinline__ void write_result(const SceneDev& S, int which, const TraceArgs& a, unsigned long long i, bool any, bool f, const Hit& h, uint32_t prim_raw_is_slot, V3 d)
{
    a.hit[i] = f ? 1 : 0;
    if (any) return;
    uint32_t prim = 0xffffffffu, blas = 0xffffffffu, inst = 0xffffffffu;
    V3 nn = ex::v3(0.0f, 0.0f, 0.0f);
    bool ff = false;
    if (f)
    {
        prim = prim_raw_is_slot ? S.tri_prim[h.slot_or_prim] : h.slot_or_prim;
        blas = S.inst[h.inst].blas;
        inst = h.inst - S.tlas[which].inst_base;
        finish_hit(S, d, h.inst, prim, h.u, h.v, &nn, &ff);
    }
    if (a.t) a.t[i] = f ? h.t : ex::inf();
    if (a.u) a.u[i] = f ? h.u : 0.0f;
    if (a.v) a.v[i] = f ? h.v : 0.0f;
    if (a.prim) a.prim[i] = prim;
    if (a.blas) a.blas[i] = blas;
    if (a.inst) a.inst[i] = inst;
    if (a.front) a.front[i] = ff ? 1 : 0;
    if (a.normal) { a.normal[i * 3] = nn.x; a.normal[i * 3 + 1] = nn.y; a.normal[i * 3 + 2] = nn.z; }
}

// 8-wide layout: dynamic-fetch persistent kernel (persist.cuh)
template <bool ANY>
struct TraceJob
{
    const SceneDev& S;
    const TraceArgs& a;
    int which;
    unsigned long long base;
    unsigned long long i = 0;
    __device__ __forceinline__ bool start(uint32_t idx, WideTrav& tr)
    {
        i = base + idx;
        wide_begin(tr, S, which, ex::v3(a.o[i * 3], a.o[i * 3 + 1], a.o[i * 3 + 2]), ex::v3(a.d[i * 3], a.d[i * 3 + 1], a.d[i * 3 + 2]), a.tmax[i], ANY);
        return true;
    }
    __device__ __forceinline__ bool finish(WideTrav& tr, bool mine)
    {
        if (!mine) return false;
        Hit h;
        h.t = tr.t_max; h.u = wide_hit_u(tr); h.v = wide_hit_v(tr); h.slot_or_prim = tr.hit_slot; h.inst = tr.hit_inst;
        write_result(S, which, a, i, ANY, tr.hit_inst != TRAV_NONE, h, 1u, tr.dw);
        return false;
    }
};

template <bool ANY>
__global__ void __launch_bounds__(128) trace_wide_kernel(SceneDev S, int which, unsigned long long base, uint32_t n, TraceArgs a, int refill_below, int node_quorum)
{
    TraceJob<ANY> job{S, a, which, base};
    persistent_trace<ANY ? 1 : 0>(S, job, (uint32_t*)a.counter, n, refill_below, node_quorum);
}

// reference-order binary kernel: every warp pulls batches of 32 rays from a global counter, one ray per lane
template <bool ANY>
__global__ void __launch_bounds__(128) trace_bin_kernel(SceneDev S, int which, unsigned long long n, TraceArgs a)
{
    const unsigned lane = threadIdx.x & 31u;
    for (;;)
    {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(a.counter, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const unsigned long long i = base + lane;
        if (i >= n) continue;
        const V3 o = ex::v3(a.o[i * 3], a.o[i * 3 + 1], a.o[i * 3 + 2]);
        const V3 d = ex::v3(a.d[i * 3], a.d[i * 3 + 1], a.d[i * 3 + 2]);
        Hit h;
        h.inst = 0xffffffffu;
        const bool f = bin_traverse<ANY>(S, which, o, d, a.tmax[i], &h);
        write_result(S, which, a, i, ANY, f, h, 0u, d);
    }
}

// team kernel (kernel = 2): eight lanes per ray, four rays per warp (team.cuh); a warp draws rays from a global cursor and deals them
// to its teams as they finish, like persist.cuh does per lane
template <bool ANY>
__global__ void __launch_bounds__(128) trace_team_kernel(SceneDev S, int which, unsigned long long base, uint32_t n, TraceArgs a)
{
    __shared__ uint2 s_stack[TEAMS_PER_BLOCK][WIDE_STACK_CAP];
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u, tl = lane & 7u, tbase = lane & 24u, tmask = 0xffu << tbase;
    uint2* stack = s_stack[threadIdx.x >> 3];
    uint32_t* cursor = (uint32_t*)a.counter;
    TeamRay tr;
    unsigned long long i = 0;
    bool active = false;
    uint32_t wbase = 0, wend = 0;
    bool more = n > 0u;
    for (;;)
    {
        // deal rays to the idle teams of the warp
        const unsigned idle = __ballot_sync(FULL, !active) & 0x01010101u;  // one bit per team (its lane 0)
        if (idle && (wbase < wend || more))
        {
            if (wbase >= wend)
            {
                // 16 rays per atomic while the batch is much larger than the grid, else just what the idle teams need now: a warp must
                // not sit on a private backlog while other warps have run dry (small batches are latency tests)
                const uint32_t want = n > (gridDim.x * blockDim.x >> 5) * 64u ? 16u : (uint32_t)__popc(idle);
                uint32_t b = 0;
                if (lane == 0u) b = atomicAdd(cursor, want);
                b = __shfl_sync(FULL, b, 0);
                if (b >= n) more = false;
                else { wbase = b; wend = min(b + want, n); if (wend == n) more = false; }
            }
            const uint32_t avail = wend - wbase, rank = (uint32_t)__popc(idle & ((1u << tbase) - 1u));
            if (!active && rank < avail)
            {
                i = base + wbase + rank;
                team_begin(tr, S, which, ex::v3(a.o[i * 3], a.o[i * 3 + 1], a.o[i * 3 + 2]), ex::v3(a.d[i * 3], a.d[i * 3 + 1], a.d[i * 3 + 2]), a.tmax[i]);
                active = true;
            }
            wbase += min((uint32_t)__popc(idle), avail);
        }
        if (!__any_sync(FULL, active))
        {
            if (wbase >= wend && !more) break;
            continue;
        }
        if (active && !team_step<ANY>(tr, S, stack, tl, tmask, tbase))
        {
            if (tl == 0u)
            {
                Hit h;
                h.t = tr.t_max; h.u = ex::div(tr.hit_ud, tr.hit_det); h.v = ex::div(tr.hit_vd, tr.hit_det); h.slot_or_prim = tr.hit_slot; h.inst = tr.hit_inst;
                write_result(S, which, a, i, ANY, tr.hit_inst != TRAV_NONE, h, 1u, tr.dw);
            }
            active = false;
        }
    }
}

// multiray kernels (kernel = 3 / 4 / 5: two / three / four rays per lane; 6: one ray per lane, i.e. the cost of the shared-memory ray
// records alone): multiray.cuh
template <bool ANY>
struct MrTraceJob
{
    static constexpr int JOBW = 1;
    const SceneDev& S;
    const TraceArgs& a;
    int which;
    unsigned long long base;
    template <class R>
    __device__ __forceinline__ uint32_t start(uint32_t idx, const R& r)
    {
        const unsigned long long i = base + idx;
        r.st(MR_JOB, idx);
        return mr_begin(r, S, which, ex::v3(a.o[i * 3], a.o[i * 3 + 1], a.o[i * 3 + 2]), ex::v3(a.d[i * 3], a.d[i * 3 + 1], a.d[i * 3 + 2]), a.tmax[i], ANY);
    }
    template <class R>
    __device__ __forceinline__ void world_ray(const R& r, V3* o, V3* d) const
    {
        const unsigned long long i = base + r.ld(MR_JOB);
        *o = ex::v3(a.o[i * 3], a.o[i * 3 + 1], a.o[i * 3 + 2]);
        *d = ex::v3(a.d[i * 3], a.d[i * 3 + 1], a.d[i * 3 + 2]);
    }
    template <class R>
    __device__ __forceinline__ uint32_t finish(const R& r, bool mine)
    {
        if (!mine) return MT_IDLE;
        Hit h;
        h.t = r.ldf(MR_TMAX); h.u = mr_hit_u(r); h.v = mr_hit_v(r); h.slot_or_prim = r.ld(MR_HSLOT); h.inst = r.ld(MR_HINST);
        V3 o, d;
        world_ray(r, &o, &d);
        write_result(S, which, a, base + r.ld(MR_JOB), ANY, h.inst != TRAV_NONE, h, 1u, d);
        return MT_IDLE;
    }
};

template <bool ANY, int K>
__global__ void __launch_bounds__(128) trace_mr_kernel(SceneDev S, int which, unsigned long long base, uint32_t n, TraceArgs a, MrOpts mo)
{
    MrTraceJob<ANY> job{S, a, which, base};
    mr_persistent_trace<ANY ? 1 : 0, K>(S, job, (uint32_t*)a.counter, n, mo.free_quorum, mo.qn, mo.ql, mo.qi);
}

template <bool ANY, int K>
static cudaError_t launch_mr(const SceneDev& S, int which, uint64_t n, const TraceArgs& a, int sm_count, cudaStream_t stream, const MrOpts& mo)
{
    const size_t smem = (size_t)K * (MR_JOB + MrTraceJob<ANY>::JOBW) * 128 * sizeof(uint32_t);
    auto kern = trace_mr_kernel<ANY, K>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem);
    if (e != cudaSuccess) return e;
    if (mo.blocks_per_sm > 0) per_sm = std::min(per_sm, mo.blocks_per_sm);
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    for (uint64_t base = 0; base < n; base += 0x40000000ull)
    {
        e = cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
        const uint32_t cnt = (uint32_t)std::min<uint64_t>(n - base, 0x40000000ull);
        kern<<<sm_count * per_sm, 128, smem, stream>>>(S, which, base, cnt, a, mo);
    }
    return cudaGetLastError();
}

cudaError_t launch_trace(const SceneDev& S, int which, int kernel, int any_hit, uint64_t n, const TraceArgs& a, int sm_count, cudaStream_t stream, int refill_below, int node_quorum, const MrOpts& mo)
{
    if (kernel >= 3)
    {
        MrOpts m = mo;
        if (m.free_quorum <= 0) m.free_quorum = 24;
        switch (kernel)
        {
        case 3: return any_hit ? launch_mr<true, 2>(S, which, n, a, sm_count, stream, m) : launch_mr<false, 2>(S, which, n, a, sm_count, stream, m);
        case 4: return any_hit ? launch_mr<true, 3>(S, which, n, a, sm_count, stream, m) : launch_mr<false, 3>(S, which, n, a, sm_count, stream, m);
        case 5: return any_hit ? launch_mr<true, 4>(S, which, n, a, sm_count, stream, m) : launch_mr<false, 4>(S, which, n, a, sm_count, stream, m);
        default: return any_hit ? launch_mr<true, 1>(S, which, n, a, sm_count, stream, m) : launch_mr<false, 1>(S, which, n, a, sm_count, stream, m);
        }
    }
    const int blocks = sm_count * 8;
    if (kernel == 0)
    {
        // the persistent loop indexes work items with 32 bits: split larger batches
        for (uint64_t base = 0; base < n; base += 0x40000000ull)
        {
            cudaError_t e = cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), stream);
            if (e != cudaSuccess) return e;
            const uint32_t cnt = (uint32_t)std::min<uint64_t>(n - base, 0x40000000ull);
            if (any_hit) trace_wide_kernel<true><<<blocks, 128, 0, stream>>>(S, which, base, cnt, a, refill_below, node_quorum);
            else trace_wide_kernel<false><<<blocks, 128, 0, stream>>>(S, which, base, cnt, a, refill_below, node_quorum);
        }
    }
    else if (kernel == 2)
    {
        for (uint64_t base = 0; base < n; base += 0x40000000ull)
        {
            cudaError_t e = cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), stream);
            if (e != cudaSuccess) return e;
            const uint32_t cnt = (uint32_t)std::min<uint64_t>(n - base, 0x40000000ull);
            if (any_hit) trace_team_kernel<true><<<blocks, 128, 0, stream>>>(S, which, base, cnt, a);
            else trace_team_kernel<false><<<blocks, 128, 0, stream>>>(S, which, base, cnt, a);
        }
    }
    else
    {
        cudaError_t e = cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), stream);
        if (e != cudaSuccess) return e;
        if (any_hit) trace_bin_kernel<true><<<blocks, 128, 0, stream>>>(S, which, n, a);
        else trace_bin_kernel<false><<<blocks, 128, 0, stream>>>(S, which, n, a);
    }
    return cudaGetLastError();
}

}  // namespace ptb
