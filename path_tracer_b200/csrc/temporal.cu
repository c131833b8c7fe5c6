// temporal.cu — State::update (state.rs:505-579) without the graphics API: the accumulation "texture" lives in a device
// buffer, the three compute passes are CUDA kernels (bodies: temporal.cuh).
//   inv_projection == last_inv_projection (state.rs:561)  ->  k_accumulate            (accumulate.wgsl)
//   otherwise                                             ->  k_resolve               (velocity.wgsl + compute.wgsl, one kernel:
//                                                             the velocity a texel needs is a pure function of the position
//                                                             texel it is read at, so the velocity pass is evaluated in place
//                                                             and the rg32float texture is only written, never re-read)
//   then output -> accumulation (state.rs:577): the two history buffers swap roles instead of being copied.
// Both kernels are streaming image kernels bound by HBM: per pixel k_resolve reads 16 B input (+ halo from L1/L2), 16 B
// position, 8 B ids, <= 16 history texels (L1/L2 hits after the first) and writes 16 + 8 B — about 72 algorithmic bytes.
#include "context.h"
#include "render_state.h"
#include "temporal.cuh"

namespace ptb {

struct TemporalState
{
    DBuf<float4> history[2];  // accumulation texture (cur) and output texture (1 - cur)
    DBuf<float2> velocity;
    DBuf<float4> in_data, in_position;  // staging for the host-buffer entry point
    DBuf<uint32_t> in_id;
    uint32_t W = 0, H = 0;
    int cur = 0;
    uint64_t updates = 0, resolves = 0;
    // with ptb_set_profiling: device time of the update kernels (CUDA events on the context's stream)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float ms_total = 0.0f;
    uint32_t timed = 0;
};

struct Mat4Arg { float m[16]; };

__global__ void __launch_bounds__(256) k_accumulate(const float4* __restrict__ input, const float4* __restrict__ history, float4* __restrict__ out, uint32_t npix, int input_is_sum)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    const float4 d = input[i], h = history[i];
    // accumulate.wgsl:20-23: accumulation + vec4(current_diffuse, 1.0); a resident frame of n samples adds (sum, n)
    out[i] = make_float4(ex::add(h.x, d.x), ex::add(h.y, d.y), ex::add(h.z, d.z), ex::add(h.w, input_is_sum ? d.w : 1.0f));
}

// 32 x 8 pixel blocks: a warp covers 32 consecutive texels of a row (512 B of each rgba32float image per load)
__global__ void __launch_bounds__(256, 4) k_resolve(const float4* __restrict__ input, const float4* __restrict__ history, const float4* __restrict__ position,
                                                 const uint32_t* __restrict__ id, Mat4Arg last, int W, int H, int input_is_sum, float4* __restrict__ out, float2* __restrict__ velocity)
{
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    if (x >= W || y >= H) return;
    const tp::Image in{reinterpret_cast<const tp::F4*>(input), W, H, input_is_sum}, hist{reinterpret_cast<const tp::F4*>(history), W, H, 0};
    const tp::Image pos{reinterpret_cast<const tp::F4*>(position), W, H, 0};
    const tp::F2 v = tp::velocity_at(pos, last.m, tp::F2{(float)W, (float)H}, x, y);
    const tp::F4 r = tp::resolve_at(in, hist, pos, id, last.m, x, y);
    const size_t i = (size_t)y * W + x;
    velocity[i] = make_float2(v.x, v.y);
    out[i] = make_float4(r.x, r.y, r.z, r.w);
}

void temporal_state_destroy(ptb_context* ctx)
{
    TemporalState* ts = ctx->ts;
    if (!ts) return;
    ts->history[0].release(); ts->history[1].release(); ts->velocity.release();
    ts->in_data.release(); ts->in_position.release(); ts->in_id.release();
    if (ts->ev0) cudaEventDestroy(ts->ev0);
    if (ts->ev1) cudaEventDestroy(ts->ev1);
    delete ts;
    ctx->ts = nullptr;
}

bool temporal_state_image(ptb_context* ctx, const float4** history, uint32_t* W, uint32_t* H)
{
    if (!ctx->ts || !ctx->ts->history[0].p) return false;
    *history = ctx->ts->history[ctx->ts->cur].p;
    *W = ctx->ts->W;
    *H = ctx->ts->H;
    return true;
}

static int temporal_prepare(ptb_context* ctx, uint32_t W, uint32_t H)
{
    if (!ctx->ts) ctx->ts = new TemporalState();
    TemporalState& ts = *ctx->ts;
    if (ts.W == W && ts.H == H && ts.history[0].p) return PTB_OK;
    const size_t npix = (size_t)W * H;
    ts.history[0].release(); ts.history[1].release(); ts.velocity.release();
    PTB_CUDA(ctx, ts.history[0].resize(npix));
    PTB_CUDA(ctx, ts.history[1].resize(npix));
    PTB_CUDA(ctx, ts.velocity.resize(npix));
    // a new State starts from zero-initialised textures (state.rs:112-166)
    PTB_CUDA(ctx, cudaMemsetAsync(ts.history[0].p, 0, npix * 16, ctx->stream));
    PTB_CUDA(ctx, cudaMemsetAsync(ts.history[1].p, 0, npix * 16, ctx->stream));
    PTB_CUDA(ctx, cudaMemsetAsync(ts.velocity.p, 0, npix * 8, ctx->stream));
    ts.W = W; ts.H = H; ts.cur = 0;
    return PTB_OK;
}

// the two dispatch alternatives of state.rs:561-575 on device-resident inputs
static int temporal_update_device(ptb_context* ctx, const float4* data, const float4* position, const uint32_t* id, const float* inv_projection, const float* last_inv_projection,
                                  int input_is_sum)
{
    TemporalState& ts = *ctx->ts;
    const uint32_t npix = ts.W * ts.H;
    bool same = true;  // glam Mat4 == : all 16 lanes compare equal (NaN != NaN, -0 == +0)
    for (int k = 0; k < 16; ++k) same = same && (inv_projection[k] == last_inv_projection[k]);
    const float4* hist = ts.history[ts.cur].p;
    float4* out = ts.history[1 - ts.cur].p;
    const bool timed = ctx->profiling != 0;
    if (timed)
    {
        if (!ts.ev0) { PTB_CUDA(ctx, cudaEventCreate(&ts.ev0)); PTB_CUDA(ctx, cudaEventCreate(&ts.ev1)); }
        PTB_CUDA(ctx, cudaEventRecord(ts.ev0, ctx->stream));
    }
    if (same) k_accumulate<<<(npix + 255) / 256, 256, 0, ctx->stream>>>(data, hist, out, npix, input_is_sum);
    else
    {
        Mat4Arg last;
        for (int k = 0; k < 16; ++k) last.m[k] = last_inv_projection[k];
        const dim3 grid((ts.W + 31) / 32, (ts.H + 7) / 8), block(32, 8);
        k_resolve<<<grid, block, 0, ctx->stream>>>(data, hist, position, id, last, (int)ts.W, (int)ts.H, input_is_sum, out, ts.velocity.p);
        ++ts.resolves;
    }
    PTB_CUDA(ctx, cudaGetLastError());
    if (timed)
    {
        float ms = 0.0f;
        PTB_CUDA(ctx, cudaEventRecord(ts.ev1, ctx->stream));
        PTB_CUDA(ctx, cudaEventSynchronize(ts.ev1));
        PTB_CUDA(ctx, cudaEventElapsedTime(&ms, ts.ev0, ts.ev1));
        ts.ms_total += ms;
        ++ts.timed;
    }
    ts.cur = 1 - ts.cur;  // copy_texture_to_texture(output -> accumulation), state.rs:577
    ++ts.updates;
    return PTB_OK;
}

}  // namespace ptb

using namespace ptb;

#define REQUIRE_DEVICE(ctx)                                                                                                   \
    if (!(ctx)) return PTB_ERR_INVALID;                                                                                       \
    if ((ctx)->host_only) return (ctx)->fail(PTB_ERR_NO_DEVICE, "host-only context: no CUDA device bound (no CPU fallback)")

extern "C" {

int ptb_state_reset(ptb_context* ctx, uint32_t width, uint32_t height)
{
    REQUIRE_DEVICE(ctx);
    if (width == 0 || height == 0 || (uint64_t)width * height > 0x3fffffffull) return ctx->fail(PTB_ERR_INVALID, "bad image size");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ctx->ts) { ctx->ts->W = 0; ctx->ts->H = 0; }
    return temporal_prepare(ctx, width, height);
}

int ptb_state_update(ptb_context* ctx, uint32_t width, uint32_t height, const float* data, const float* position, const uint32_t* id, const float inv_projection[16],
                     const float last_inv_projection[16])
{
    REQUIRE_DEVICE(ctx);
    if (!data || !position || !id || !inv_projection || !last_inv_projection) return ctx->fail(PTB_ERR_INVALID, "null argument");
    if (width == 0 || height == 0 || (uint64_t)width * height > 0x3fffffffull) return ctx->fail(PTB_ERR_INVALID, "bad image size");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (int e = temporal_prepare(ctx, width, height)) return e;
    TemporalState& ts = *ctx->ts;
    const size_t npix = (size_t)width * height;
    PTB_CUDA(ctx, ts.in_data.resize(npix));
    PTB_CUDA(ctx, ts.in_position.resize(npix));
    PTB_CUDA(ctx, ts.in_id.resize(npix));
    // queue.write_texture x3 (state.rs:519-556)
    PTB_CUDA(ctx, cudaMemcpyAsync(ts.in_data.p, data, npix * 16, cudaMemcpyHostToDevice, ctx->stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(ts.in_position.p, position, npix * 16, cudaMemcpyHostToDevice, ctx->stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(ts.in_id.p, id, npix * 4, cudaMemcpyHostToDevice, ctx->stream));
    // the host arrays are borrowed for the duration of the call only (ptb.h): with pinned memory the copies above are truly
    // asynchronous, and the reference's caller overwrites these arrays for the next frame (main.rs:204-206) — wait for them here
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return temporal_update_device(ctx, ts.in_data.p, ts.in_position.p, ts.in_id.p, inv_projection, last_inv_projection, 0);
}

int ptb_state_update_resident(ptb_context* ctx, const float inv_projection[16], const float last_inv_projection[16])
{
    REQUIRE_DEVICE(ctx);
    if (!inv_projection || !last_inv_projection) return ctx->fail(PTB_ERR_INVALID, "null argument");
    float4 *accum = nullptr, *position = nullptr;
    uint32_t* id = nullptr;
    uint32_t W = 0, H = 0;
    if (!render_state_frame(ctx, &accum, &W, &H) || !render_state_aux(ctx, &position, &id)) return ctx->fail(PTB_ERR_INVALID, "no frame");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (int e = temporal_prepare(ctx, W, H)) return e;
    return temporal_update_device(ctx, accum, position, id, inv_projection, last_inv_projection, 1);
}

int ptb_state_read(ptb_context* ctx, float* rgba, float* velocity2)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->ts || !ctx->ts->history[0].p) return ctx->fail(PTB_ERR_INVALID, "no temporal state");
    TemporalState& ts = *ctx->ts;
    const size_t npix = (size_t)ts.W * ts.H;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (rgba) PTB_CUDA(ctx, cudaMemcpyAsync(rgba, ts.history[ts.cur].p, npix * 16, cudaMemcpyDeviceToHost, ctx->stream));
    if (velocity2) PTB_CUDA(ctx, cudaMemcpyAsync(velocity2, ts.velocity.p, npix * 8, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_state_timing(ptb_context* ctx, float* ms_total, uint32_t* launches, int reset)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->ts) return ctx->fail(PTB_ERR_INVALID, "no temporal state");
    if (ms_total) *ms_total = ctx->ts->ms_total;
    if (launches) *launches = ctx->ts->timed;
    if (reset) { ctx->ts->ms_total = 0.0f; ctx->ts->timed = 0; }
    return PTB_OK;
}

int ptb_state_device_ptr(ptb_context* ctx, void** dptr, size_t* bytes)
{
    REQUIRE_DEVICE(ctx);
    if (!ctx->ts || !ctx->ts->history[0].p) return ctx->fail(PTB_ERR_INVALID, "no temporal state");
    if (dptr) *dptr = ctx->ts->history[ctx->ts->cur].p;
    if (bytes) *bytes = (size_t)ctx->ts->W * ctx->ts->H * 16;
    return PTB_OK;
}

}  // extern "C"
