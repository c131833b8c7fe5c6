// render.cu — the frame loop of main.rs:181-207 as wavefront kernels over SoA path state (wavefront.h holds the per-slot
// bodies) plus the camera / jitter / RNG probes of the C ABI.
//
// Kernels (all persistent: grid = SMs x blocks_per_sm, every warp pulls chunks of 32 queue entries from a cursor):
//   k_regen   finalise + accumulate finished samples (single writer per pixel, no atomics) and raygen the next sample
//   k_extend  closest hit on the world TLAS, then pushes the slot into the shade queue of its material kind
//   k_shade   per-material queues processed kind by kind (a warp never mixes kinds); emits shadow / probe / next / term
//   k_shadow  any-hit for explicit light samples
//   k_probe   lights-TLAS closest hit + world any-hit for BSDF-sampled direct light
//   k_advance 1 thread: swaps queue counts, folds ray statistics, resets cursors
// Compaction between stages is a warp ballot + one atomicAdd per warp and queue.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "camera_host.h"
#include "context.h"
#include "wavefront.h"

namespace ptb {

struct WfCounters
{
    uint32_t n_ext, n_next, n_shadow, n_probe, n_term, pad0;
    uint32_t n_mat[8];
    uint32_t cur_regen, cur_extend, cur_shade, cur_shadow, cur_probe, pad1;
    unsigned long long stat_extend, stat_shadow, stat_probe, stat_paths, stat_iters, stat_primary, stat_probe_shadow;
};

struct Queues
{
    uint32_t* ext;   // current extend queue
    uint32_t* next;  // extend queue being filled for the next iteration
    uint32_t* mat;   // 7 queues of npix entries each
    uint32_t* shadow;
    uint32_t* probe;
    uint32_t* term;
    uint32_t npix;
    WfCounters* c;
};

struct RenderState
{
    uint32_t W = 0, H = 0, npix = 0;
    DBuf<float4> accum_own, accum_sq, position, ray_o, ray_d, weight, radiance, hit_tn, nee_o, sh_d, sh_c, pr_d, pr_f;
    DBuf<uint32_t> id, hit_info, q_a, q_b, q_mat, q_shadow, q_probe, q_term;
    DBuf<uint2> rng;
    DBuf<unsigned long long> volset;
    DBuf<WfCounters> counters;
    float4* accum = nullptr;  // bound or own
    uint32_t* h_flags = nullptr;  // pinned
    std::vector<cudaEvent_t> poll_events;
    std::vector<cudaEvent_t> prof_events;
    // host-side statistics
    unsigned long long launches = 0;
    float t_total_ms = 0;
    float t_class_ms[8] = {0};
    uint32_t n_class[8] = {0};
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
};

// ---- device helpers ------------------------------------------------------------------------------------------------------
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
__device__ __forceinline__ bool fetch_chunk(uint32_t* cursor, uint32_t n, uint32_t* base_out)
{
    uint32_t base = 0;
    if ((threadIdx.x & 31u) == 0u) base = atomicAdd(cursor, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    *base_out = base;
    return base < n;
}

__global__ void __launch_bounds__(128) k_init(RenderView V, Queues Q)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Q.npix) return;
    V.ray_o[i].w = 0.0f;  // no sample in flight
    V.ray_d[i].w = 0.0f;  // samples started = 0
    Q.term[i] = i;
    if (i == 0)
    {
        Q.c->n_term = Q.npix;
        Q.c->n_ext = Q.c->n_next = Q.c->n_shadow = Q.c->n_probe = 0;
        for (int k = 0; k < 8; ++k) Q.c->n_mat[k] = 0;
        Q.c->cur_regen = Q.c->cur_extend = Q.c->cur_shade = Q.c->cur_shadow = Q.c->cur_probe = 0;
    }
}

__global__ void __launch_bounds__(128) k_regen(RenderView V, Queues Q)
{
    const uint32_t n = Q.c->n_term;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base;
    while (fetch_chunk(&Q.c->cur_regen, n, &base))
    {
        const uint32_t i = base + lane;
        const bool valid = i < n;
        const uint32_t slot = valid ? Q.term[i] : 0u;
        const uint32_t flags = valid ? regen_slot(V, slot) : 0u;
        warp_push(Q.next, &Q.c->n_next, (flags & WF_NEXT) != 0u, slot);
        const unsigned m = __ballot_sync(0xffffffffu, (flags & WF_NEXT) != 0u);
        if (lane == 0u && m) atomicAdd(&Q.c->stat_paths, (unsigned long long)__popc(m));
    }
}

template <bool SORT>
__global__ void __launch_bounds__(128) k_extend(RenderView V, Queues Q)
{
    const uint32_t n = Q.c->n_ext;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base;
    while (fetch_chunk(&Q.c->cur_extend, n, &base))
    {
        const uint32_t i = base + lane;
        const bool valid = i < n;
        const uint32_t slot = valid ? Q.ext[i] : 0u;
        uint32_t kind = valid ? extend_slot(V, slot) : 0xffu;
        if (!SORT && valid) kind = 0u;
#pragma unroll
        for (uint32_t k = 0; k < (SORT ? 7u : 1u); ++k) warp_push(Q.mat + (size_t)k * Q.npix, &Q.c->n_mat[k], kind == k, slot);
    }
}

__global__ void __launch_bounds__(128) k_shade(RenderView V, Queues Q)
{
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t cnt[7], chunks = 0;
#pragma unroll
    for (int k = 0; k < 7; ++k) { cnt[k] = Q.c->n_mat[k]; chunks += (cnt[k] + 31u) >> 5; }
    for (;;)
    {
        uint32_t c = 0;
        if (lane == 0u) c = atomicAdd(&Q.c->cur_shade, 1u);
        c = __shfl_sync(0xffffffffu, c, 0);
        if (c >= chunks) break;
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
        warp_push(Q.shadow, &Q.c->n_shadow, (flags & WF_SHADOW) != 0u, slot);
        warp_push(Q.probe, &Q.c->n_probe, (flags & WF_PROBE) != 0u, slot);
        warp_push(Q.next, &Q.c->n_next, (flags & WF_NEXT) != 0u, slot);
        warp_push(Q.term, &Q.c->n_term, (flags & WF_TERM) != 0u, slot);
    }
}

__global__ void __launch_bounds__(128) k_shadow(RenderView V, Queues Q)
{
    const uint32_t n = Q.c->n_shadow;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base;
    while (fetch_chunk(&Q.c->cur_shadow, n, &base))
    {
        const uint32_t i = base + lane;
        if (i < n) shadow_slot(V, Q.shadow[i]);
    }
}

__global__ void __launch_bounds__(128) k_probe(RenderView V, Queues Q)
{
    const uint32_t n = Q.c->n_probe;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base;
    while (fetch_chunk(&Q.c->cur_probe, n, &base))
    {
        const uint32_t i = base + lane;
        const uint32_t traced = (i < n) ? probe_slot(V, Q.probe[i]) : 0u;
        const unsigned m = __ballot_sync(0xffffffffu, traced != 0u);
        if (lane == 0u && m) atomicAdd(&Q.c->stat_probe_shadow, (unsigned long long)__popc(m));
    }
}

// end of an iteration (1 thread): fold statistics, make q_next current, reset counts and cursors
__global__ void k_advance(Queues Q)
{
    WfCounters& c = *Q.c;
    c.stat_extend += c.n_ext;
    c.stat_shadow += c.n_shadow;
    c.stat_probe += c.n_probe;
    c.stat_iters += 1;
    c.n_ext = c.n_next;
    c.n_next = c.n_term = c.n_shadow = c.n_probe = 0;
    for (int k = 0; k < 8; ++k) c.n_mat[k] = 0;
    c.cur_regen = c.cur_extend = c.cur_shade = c.cur_shadow = c.cur_probe = 0;
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

void render_state_destroy(ptb_context* ctx)
{
    RenderState* rs = ctx->rs;
    if (!rs) return;
    for (auto e : rs->poll_events) cudaEventDestroy(e);
    for (auto e : rs->prof_events) cudaEventDestroy(e);
    if (rs->ev_begin) cudaEventDestroy(rs->ev_begin);
    if (rs->ev_end) cudaEventDestroy(rs->ev_end);
    if (rs->h_flags) cudaFreeHost(rs->h_flags);
    DBuf<float4>* f4s[] = {&rs->accum_own, &rs->accum_sq, &rs->position, &rs->ray_o, &rs->ray_d, &rs->weight, &rs->radiance, &rs->hit_tn, &rs->nee_o, &rs->sh_d, &rs->sh_c, &rs->pr_d, &rs->pr_f};
    for (auto* b : f4s) b->release();
    DBuf<uint32_t>* u32s[] = {&rs->id, &rs->hit_info, &rs->q_a, &rs->q_b, &rs->q_mat, &rs->q_shadow, &rs->q_probe, &rs->q_term};
    for (auto* b : u32s) b->release();
    rs->rng.release();
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
    if (width == 0 || height == 0 || (uint64_t)width * height >= (1ull << 31)) return ctx->fail(PTB_ERR_INVALID, "bad frame size");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    const bool fresh = ctx->rs == nullptr;
    if (fresh) ctx->rs = new RenderState();
    RenderState& rs = *ctx->rs;
    const uint32_t npix = width * height;
    const bool bound = rs.accum && rs.accum != rs.accum_own.p;
    if (bound && npix != rs.npix) return ctx->fail(PTB_ERR_INVALID, "bound accumulation buffer has a different size; rebind it");
    rs.W = width; rs.H = height; rs.npix = npix;
    DBuf<float4>* f4s[] = {&rs.accum_own, &rs.position, &rs.ray_o, &rs.ray_d, &rs.weight, &rs.radiance, &rs.hit_tn, &rs.nee_o, &rs.sh_d, &rs.sh_c, &rs.pr_d, &rs.pr_f};
    for (auto* b : f4s) PTB_CUDA(ctx, b->resize(npix));
    if (ctx->second_moment) PTB_CUDA(ctx, rs.accum_sq.resize(npix));
    DBuf<uint32_t>* u32s[] = {&rs.id, &rs.hit_info, &rs.q_a, &rs.q_b, &rs.q_shadow, &rs.q_probe, &rs.q_term};
    for (auto* b : u32s) PTB_CUDA(ctx, b->resize(npix));
    PTB_CUDA(ctx, rs.q_mat.resize((size_t)npix * 7));
    PTB_CUDA(ctx, rs.rng.resize(npix));
    PTB_CUDA(ctx, rs.volset.resize(npix));
    PTB_CUDA(ctx, rs.counters.resize(1));
    if (!rs.h_flags) PTB_CUDA(ctx, cudaHostAlloc((void**)&rs.h_flags, 256 * sizeof(uint32_t), cudaHostAllocDefault));
    if (!rs.ev_begin) { PTB_CUDA(ctx, cudaEventCreate(&rs.ev_begin)); PTB_CUDA(ctx, cudaEventCreate(&rs.ev_end)); }
    if (!bound) rs.accum = rs.accum_own.p;
    PTB_CUDA(ctx, cudaMemsetAsync(rs.accum, 0, (size_t)npix * 16, ctx->stream));
    if (ctx->second_moment) PTB_CUDA(ctx, cudaMemsetAsync(rs.accum_sq.p, 0, (size_t)npix * 16, ctx->stream));
    PTB_CUDA(ctx, cudaMemsetAsync(rs.position.p, 0, (size_t)npix * 16, ctx->stream));
    PTB_CUDA(ctx, cudaMemsetAsync(rs.id.p, 0, (size_t)npix * 4, ctx->stream));
    if (fresh) PTB_CUDA(ctx, cudaMemsetAsync(rs.counters.p, 0, sizeof(WfCounters), ctx->stream));  // statistics persist until ptb_reset_counters
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
    RenderView V{};
    V.S = ctx->S;
    V.cam = cam_dev(ctx);
    V.W = rs.W; V.H = rs.H;
    V.sample_begin = sample_begin; V.sample_count = sample_count;
    V.seed = seed;
    V.max_bounces = max_bounces;
    V.enable_nee = enable_nee ? 1u : 0u;
    V.extend_kernel = ctx->opt_extend_kernel ? 1u : 0u;
    V.accum = rs.accum;
    V.accum_sq = ctx->second_moment ? rs.accum_sq.p : nullptr;
    V.position = rs.position.p; V.id = rs.id.p;
    V.ray_o = rs.ray_o.p; V.ray_d = rs.ray_d.p; V.weight = rs.weight.p; V.radiance = rs.radiance.p; V.rng = rs.rng.p;
    V.volset = ctx->scene.volumes.empty() ? nullptr : rs.volset.p;
    V.hit_tn = rs.hit_tn.p; V.hit_info = rs.hit_info.p; V.nee_o = rs.nee_o.p; V.sh_d = rs.sh_d.p; V.sh_c = rs.sh_c.p; V.pr_d = rs.pr_d.p; V.pr_f = rs.pr_f.p;
    Queues Q{};
    Q.ext = rs.q_a.p; Q.next = rs.q_b.p; Q.mat = rs.q_mat.p; Q.shadow = rs.q_shadow.p; Q.probe = rs.q_probe.p; Q.term = rs.q_term.p;
    Q.npix = rs.npix;
    Q.c = rs.counters.p;

    const int threads = 128;
    const int grid = ctx->sm_count * std::max(1, ctx->opt_blocks_per_sm);
    cudaStream_t st = ctx->stream;
    const bool prof = ctx->profiling != 0;
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
    while (rs.poll_events.size() < 64)
    {
        cudaEvent_t e;
        PTB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        rs.poll_events.push_back(e);
    }

    PTB_CUDA(ctx, cudaEventRecord(rs.ev_begin, st));
    PTB_CUDA(ctx, mark(-1));
    k_init<<<(rs.npix + threads - 1) / threads, threads, 0, st>>>(V, Q);
    // first regen fills q_next from q_term = all slots
    Q.next = rs.q_a.p;  // the very first extend queue
    k_regen<<<grid, threads, 0, st>>>(V, Q);
    PTB_CUDA(ctx, mark(1));
    k_advance<<<1, 1, 0, st>>>(Q);
    PTB_CUDA(ctx, mark(7));
    rs.launches += 3;
    uint32_t* cur = rs.q_a.p;
    uint32_t* other = rs.q_b.p;
    const bool sort = ctx->opt_sort_materials != 0;
    const uint32_t LAG = 8;
    for (uint32_t it = 0;; ++it)
    {
        Q.ext = cur;
        Q.next = other;
        if (sort) k_extend<true><<<grid, threads, 0, st>>>(V, Q);
        else k_extend<false><<<grid, threads, 0, st>>>(V, Q);
        PTB_CUDA(ctx, mark(2));
        k_shade<<<grid, threads, 0, st>>>(V, Q);
        PTB_CUDA(ctx, mark(3));
        if (V.enable_nee)
        {
            k_shadow<<<grid, threads, 0, st>>>(V, Q);
            PTB_CUDA(ctx, mark(4));
            k_probe<<<grid, threads, 0, st>>>(V, Q);
            PTB_CUDA(ctx, mark(5));
            rs.launches += 2;
        }
        k_regen<<<grid, threads, 0, st>>>(V, Q);
        PTB_CUDA(ctx, mark(1));
        k_advance<<<1, 1, 0, st>>>(Q);
        PTB_CUDA(ctx, mark(7));
        rs.launches += 4;
        PTB_CUDA(ctx, cudaGetLastError());
        std::swap(cur, other);
        // termination poll, LAG iterations behind the launch front
        PTB_CUDA(ctx, cudaMemcpyAsync(&rs.h_flags[it % 64], &rs.counters.p->n_ext, 4, cudaMemcpyDeviceToHost, st));
        PTB_CUDA(ctx, cudaEventRecord(rs.poll_events[it % 64], st));
        if (it >= LAG)
        {
            PTB_CUDA(ctx, cudaEventSynchronize(rs.poll_events[(it - LAG) % 64]));
            if (rs.h_flags[(it - LAG) % 64] == 0u) break;
        }
    }
    PTB_CUDA(ctx, cudaEventRecord(rs.ev_end, st));
    PTB_CUDA(ctx, cudaEventSynchronize(rs.ev_end));
    float ms = 0;
    PTB_CUDA(ctx, cudaEventElapsedTime(&ms, rs.ev_begin, rs.ev_end));
    rs.t_total_ms += ms;
    rs.t_class_ms[0] += ms;
    rs.n_class[0] += 1;
    if (prof)
    {
        for (size_t i = 1; i < prof_used; ++i)
        {
            float d = 0;
            PTB_CUDA(ctx, cudaEventElapsedTime(&d, rs.prof_events[i - 1], rs.prof_events[i]));
            const int cls = prof_class[i];
            if (cls >= 1 && cls < 8) { rs.t_class_ms[cls] += d; rs.n_class[cls] += 1; }
        }
    }
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
    WfCounters c;
    PTB_CUDA(ctx, cudaMemcpyAsync(&c, ctx->rs->counters.p, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out[0] = c.stat_extend;
    out[1] = c.stat_shadow + c.stat_probe_shadow;
    out[2] = c.stat_probe;
    out[3] = c.stat_paths;
    out[4] = c.stat_iters;
    out[5] = ctx->rs->launches;
    out[6] = c.stat_paths;  // every path starts with exactly one primary ray
    out[7] = 0;
    return PTB_OK;
}
int ptb_reset_counters(ptb_context* ctx)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->rs) return ctx->fail(PTB_ERR_INVALID, "no frame");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    PTB_CUDA(ctx, cudaMemsetAsync(ctx->rs->counters.p, 0, sizeof(WfCounters), ctx->stream));
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
