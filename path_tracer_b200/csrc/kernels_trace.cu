// kernels_trace.cu — fixed-batch TLAS::intersect / TLAS::any_intersect kernels (ptb_trace_*).
// Persistent threads: the grid is a multiple of the SM count and every warp pulls batches of 32 rays from a global
// counter until the batch is exhausted (one atomic per warp-batch).  The per-ray work is traverse.h.
#include "context.h"
#include "traverse.h"

namespace ptb {

template <bool WIDE, bool ANY>
__global__ void __launch_bounds__(128) trace_kernel(SceneDev S, int which, unsigned long long n, TraceArgs a)
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
        const bool f = WIDE ? wide_traverse<ANY>(S, which, o, d, a.tmax[i], &h) : bin_traverse<ANY>(S, which, o, d, a.tmax[i], &h);
        a.hit[i] = f ? 1 : 0;
        if (ANY) continue;
        uint32_t prim = 0xffffffffu, blas = 0xffffffffu, inst = 0xffffffffu;
        V3 nn = ex::v3(0.0f, 0.0f, 0.0f);
        bool ff = false;
        if (f)
        {
            prim = WIDE ? S.tri_prim[h.slot_or_prim] : h.slot_or_prim;
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
}

cudaError_t launch_trace(const SceneDev& S, int which, int kernel, int any_hit, uint64_t n, const TraceArgs& a, int sm_count, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(a.counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    const int blocks = sm_count * 8;
    if (kernel == 0)
    {
        if (any_hit) trace_kernel<true, true><<<blocks, 128, 0, stream>>>(S, which, n, a);
        else trace_kernel<true, false><<<blocks, 128, 0, stream>>>(S, which, n, a);
    }
    else
    {
        if (any_hit) trace_kernel<false, true><<<blocks, 128, 0, stream>>>(S, which, n, a);
        else trace_kernel<false, false><<<blocks, 128, 0, stream>>>(S, which, n, a);
    }
    return cudaGetLastError();
}

}  // namespace ptb
