// output.cu — the step directly after the render path: linear accumulation buffer -> display image.
// Headless replacement for the present pass (shader.wgsl:58-64: rgb / w, then tonemap) and for the reference's unused CPU
// writer ImageHelper::write_image (image_helper.rs:37-58): per channel
//     out = gt_tonemap(x, p=1, a=1, m=0.22, l=0.4, c=1.33, b=0).powf(1/2.2) * 255  as u8        (image_helper.rs:44-46)
// with gt_tonemap = Uchimura's "Gran Turismo" curve exactly as written in image_helper/tonemapping.rs:68-96 (including
// its degenerate shoulder weight gt_lerp(x, m+l0, m+l0), :84, which is 0 up to m+l0 and 1 above).  `as u8` saturates
// (NaN -> 0).  Image rows: the accumulation buffer has row 0 at the bottom (camera.rs:96); flip_y writes the top row first
// as image files want it.
#include "context.h"
#include "render_state.h"

namespace ptb {

__device__ __forceinline__ float gt_smoothstep(float x, float e0, float e1)  // tonemapping.rs:39-56
{
    if (x < e0) return 0.0f;
    if (x > e1) return 1.0f;
    const float a = (x - e0) / (e1 - e0);
    const float b = 3.0f - 2.0f * a;
    return a * a * b;
}
__device__ __forceinline__ float gt_lerp(float x, float e0, float e1)  // tonemapping.rs:21-35
{
    if (x < e0) return 0.0f;
    if (x > e1) return 1.0f;
    return (x - e0) / (e1 - e0);
}
__device__ __forceinline__ float gt_tonemap(float x, float p, float a, float m, float l, float c, float b)  // tonemapping.rs:68-96
{
    if (x < 0.0f) return b;
    const float l0 = (p - m) * l / a;
    const float w0 = 1.0f - gt_smoothstep(x, 0.0f, m);
    const float w2 = gt_lerp(x, m + l0, m + l0);
    const float w1 = 1.0f - w0 - w2;
    const float toe = m * powf(x / m, c) + b;                                     // gt_toe :5
    const float lin = m + a * (x - m);                                            // gt_linear :2
    const float s0 = m + l0, s1 = m + a * l0, c2 = a * p / (p - s1);
    const float sho = p - (p - s1) * expf(-c2 * (x - s0) / p);                    // gt_shoulder :8-16
    return toe * w0 + lin * w1 + sho * w2;
}
__device__ __forceinline__ unsigned char to_u8(float v)  // Rust `as u8`: saturating, NaN -> 0
{
    if (!(v > 0.0f)) return 0;
    return v >= 255.0f ? 255 : (unsigned char)v;
}

__global__ void __launch_bounds__(256) k_tonemap_rgb8(const float4* accum, uint32_t W, uint32_t H, int flip_y, unsigned char* rgb)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    const float4 a = accum[i];
    const float iw = a.w;  // shader.wgsl:63: rgb / w
    const float ch[3] = {a.x / iw, a.y / iw, a.z / iw};
    const uint32_t x = i % W, y = i / W;
    const uint32_t row = flip_y ? (H - 1u - y) : y;
    unsigned char* o = rgb + ((size_t)row * W + x) * 3;
#pragma unroll
    for (int k = 0; k < 3; ++k) o[k] = to_u8(powf(gt_tonemap(ch[k], 1.0f, 1.0f, 0.22f, 0.4f, 1.33f, 0.0f), 1.0f / 2.2f) * 255.0f);
}

}  // namespace ptb

using namespace ptb;

static int read_rgb8(ptb_context* ctx, const float4* src, uint32_t W, uint32_t H, uint8_t* rgb, int flip_y)
{
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    DBuf<unsigned char> d;
    PTB_CUDA(ctx, d.resize((size_t)W * H * 3));
    k_tonemap_rgb8<<<(W * H + 255) / 256, 256, 0, ctx->stream>>>(src, W, H, flip_y, d.p);
    PTB_CUDA(ctx, cudaGetLastError());
    PTB_CUDA(ctx, cudaMemcpyAsync(rgb, d.p, (size_t)W * H * 3, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    d.release();
    return PTB_OK;
}

extern "C" int ptb_read_rgb8(ptb_context* ctx, uint8_t* rgb, int flip_y)
{
    if (!ctx) return PTB_ERR_INVALID;
    if (ctx->host_only) return ctx->fail(PTB_ERR_NO_DEVICE, "host-only context: no CUDA device bound (no CPU fallback)");
    float4* accum = nullptr;
    uint32_t W = 0, H = 0;
    if (!render_state_frame(ctx, &accum, &W, &H) || !rgb) return ctx->fail(PTB_ERR_INVALID, "no frame");
    return read_rgb8(ctx, accum, W, H, rgb, flip_y);
}

extern "C" int ptb_state_read_rgb8(ptb_context* ctx, uint8_t* rgb, int flip_y)
{
    if (!ctx) return PTB_ERR_INVALID;
    if (ctx->host_only) return ctx->fail(PTB_ERR_NO_DEVICE, "host-only context: no CUDA device bound (no CPU fallback)");
    const float4* hist = nullptr;
    uint32_t W = 0, H = 0;
    if (!temporal_state_image(ctx, &hist, &W, &H) || !rgb) return ctx->fail(PTB_ERR_INVALID, "no temporal state");
    return read_rgb8(ctx, hist, W, H, rgb, flip_y);
}
