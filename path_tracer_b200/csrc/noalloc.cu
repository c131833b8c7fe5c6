// noalloc.cu — ptb_trace_closest_static: ptb_trace_closest (TLAS::intersect on a fixed ray batch, tlas.rs:66-110) for callers that
// must not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// ptb_trace_closest stages the batch in context-owned device buffers that grow on demand (cudaMalloc) and keeps the work cursor of
// the persistent kernel in another allocation.  Here all of that scratch — the cursor, the staged rays and the staged results — lives
// in ONE file-scope device array of fixed size, so the call allocates nothing: no cudaMalloc / cudaFree, no allocating library call.
// Consequences, all documented in include/ptb.h:
//   * at most PTB_STATIC_TRACE_MAX_RAYS rays per call (larger batches are rejected, PTB_ERR_INVALID);
//   * every GPU that loads libptb.so holds the array (PTB_STATIC_SCRATCH_BYTES, ~4.3 MB) until the process exits;
//   * every call on a device shares it and it keeps the previous call's contents: the cursor is cleared by every call (the only part
//     that must start at zero; everything else is overwritten before it is read), and two calls that overlap on one GPU — from two
//     contexts, threads or processes' streams — race.  Callers serialise them.
// In host code the array's NAME is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress, once per
// device with that device current, and only that pointer is handed to the kernel launch.
#include <mutex>

#include "context.h"

using namespace ptb;

namespace {

constexpr uint64_t MAX_RAYS = PTB_STATIC_TRACE_MAX_RAYS;
constexpr size_t align256(size_t x) { return (x + 255u) & ~(size_t)255u; }
// fixed layout: [cursor | o | d | tmax | t | u | v | prim | blas | inst | normal | hit | front], every part 256-byte aligned
constexpr size_t OFF_CURSOR = 0;
constexpr size_t OFF_O = 256;
constexpr size_t OFF_D = OFF_O + align256(MAX_RAYS * 12);
constexpr size_t OFF_TMAX = OFF_D + align256(MAX_RAYS * 12);
constexpr size_t OFF_T = OFF_TMAX + align256(MAX_RAYS * 4);
constexpr size_t OFF_U = OFF_T + align256(MAX_RAYS * 4);
constexpr size_t OFF_V = OFF_U + align256(MAX_RAYS * 4);
constexpr size_t OFF_PRIM = OFF_V + align256(MAX_RAYS * 4);
constexpr size_t OFF_BLAS = OFF_PRIM + align256(MAX_RAYS * 4);
constexpr size_t OFF_INST = OFF_BLAS + align256(MAX_RAYS * 4);
constexpr size_t OFF_NORMAL = OFF_INST + align256(MAX_RAYS * 4);
constexpr size_t OFF_HIT = OFF_NORMAL + align256(MAX_RAYS * 12);
constexpr size_t OFF_FRONT = OFF_HIT + align256(MAX_RAYS);
constexpr size_t SCRATCH_BYTES = OFF_FRONT + align256(MAX_RAYS);
static_assert(SCRATCH_BYTES == PTB_STATIC_SCRATCH_BYTES, "ptb.h states the size of the static scratch array");

static __device__ __align__(256) unsigned char g_static_trace_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[MAX_DEVICES] = {nullptr};  // device address of the array on each device, resolved once

// device address of g_static_trace_scratch on `device` (made current by the caller)
cudaError_t scratch_address(int device, unsigned char** out)
{
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[device])
    {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_trace_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = (unsigned char*)p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" {

int ptb_trace_closest_static(ptb_context* ctx, int which, uint64_t n, const float* o3, const float* d3, const float* tmax, uint8_t* hit, float* t, float* u, float* v,
                             uint32_t* prim, uint32_t* blas, uint32_t* inst, uint8_t* front, float* normal3)
{
    if (!ctx) return PTB_ERR_INVALID;
    if (ctx->host_only) return ctx->fail(PTB_ERR_NO_DEVICE, "host-only context: no CUDA device bound (no CPU fallback)");
    if (!ctx->scene.built || !ctx->device_ready) return ctx->fail(PTB_ERR_NOT_BUILT, "scene not built: call ptb_scene_upload and ptb_build first");
    if (!hit) return ctx->fail(PTB_ERR_INVALID, "hit output is required");
    if (which < 0 || which > 1) return ctx->fail(PTB_ERR_INVALID, "bad which");
    if (n > MAX_RAYS)
        return ctx->fail(PTB_ERR_INVALID, "ptb_trace_closest_static: " + std::to_string(n) + " rays exceed the static scratch capacity of " + std::to_string(MAX_RAYS) +
                                              " (PTB_STATIC_TRACE_MAX_RAYS); split the batch or use ptb_trace_closest");
    if (n == 0) return PTB_OK;
    if (!o3 || !d3 || !tmax) return ctx->fail(PTB_ERR_INVALID, "null ray arrays");
    if (ctx->device < 0 || ctx->device >= MAX_DEVICES) return ctx->fail(PTB_ERR_INVALID, "device ordinal beyond the static table");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    PTB_CUDA(ctx, scratch_address(ctx->device, &base));
    TraceArgs a;
    a.counter = (unsigned long long*)(base + OFF_CURSOR);
    a.o = (const float*)(base + OFF_O); a.d = (const float*)(base + OFF_D); a.tmax = (const float*)(base + OFF_TMAX);
    a.t = (float*)(base + OFF_T); a.u = (float*)(base + OFF_U); a.v = (float*)(base + OFF_V);
    a.prim = (uint32_t*)(base + OFF_PRIM); a.blas = (uint32_t*)(base + OFF_BLAS); a.inst = (uint32_t*)(base + OFF_INST);
    a.normal = (float*)(base + OFF_NORMAL); a.hit = base + OFF_HIT; a.front = base + OFF_FRONT;
    cudaStream_t st = ctx->stream;
    PTB_CUDA(ctx, cudaMemcpyAsync(base + OFF_O, o3, n * 12, cudaMemcpyHostToDevice, st));
    PTB_CUDA(ctx, cudaMemcpyAsync(base + OFF_D, d3, n * 12, cudaMemcpyHostToDevice, st));
    PTB_CUDA(ctx, cudaMemcpyAsync(base + OFF_TMAX, tmax, n * 4, cudaMemcpyHostToDevice, st));
    // the 8-wide persistent kernel; launch_trace clears the cursor (the one part of the scratch that must start at zero) before the launch
    PTB_CUDA(ctx, launch_trace(ctx->S, which, 0, 0, n, a, ctx->sm_count, st, ctx->opt_refill_below < 0 ? 8 : ctx->opt_refill_below, ctx->opt_node_quorum));
#define DL(dst, off, bytes) \
    if (dst) PTB_CUDA(ctx, cudaMemcpyAsync(dst, base + off, bytes, cudaMemcpyDeviceToHost, st))
    DL(hit, OFF_HIT, n);
    DL(t, OFF_T, n * 4);
    DL(u, OFF_U, n * 4);
    DL(v, OFF_V, n * 4);
    DL(prim, OFF_PRIM, n * 4);
    DL(blas, OFF_BLAS, n * 4);
    DL(inst, OFF_INST, n * 4);
    DL(front, OFF_FRONT, n);
    DL(normal3, OFF_NORMAL, n * 12);
#undef DL
    PTB_CUDA(ctx, cudaStreamSynchronize(st));
    return PTB_OK;
}

}  // extern "C"
