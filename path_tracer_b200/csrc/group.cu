// group.cu — multi-GPU rendering behind the C ABI (ptb_group_* of include/ptb.h).
//
// The path shards by global sample index (main.rs:181-207 is a pure map over (pixel, sample), accumulate.wgsl:23 a commutative
// float sum): every member renders a contiguous block of the batch's samples for all pixels into its own accumulation buffer,
// then ONE ncclReduce(sum, float32) per member brings the W x H x 4 buffers onto the root over NVLink / NVSwitch.  There is no
// other data-path exchange, so the collective is NCCL's own reduce kernel on each member's render stream (stream order makes it
// follow that member's k_combine without a host synchronisation).
//
// Two topologies share the code: ONE process driving n devices (ptb_group_create: ncclCommInitAll, one host thread per member for
// the blocking calls) and one process per GPU (ptb_group_join: ncclCommInitRank with an id the host passes around).
#include <nccl.h>

#include <cstring>
#include <functional>
#include <thread>

#include "context.h"

namespace ptb {
bool render_state_frame(ptb_context* ctx, float4** accum, uint32_t* W, uint32_t* H);  // render.cu
}
using namespace ptb;

struct ptb_group
{
    int n_ranks = 0;
    std::vector<ptb_context*> ctx;   // local members
    std::vector<int> rank;           // global member number of each local member
    std::vector<ncclComm_t> comm;    // one per local member
    bool owns_contexts = false;
    std::string err;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;  // around the first local member's reduce
    bool reduce_timed = false;

    int fail(int code, const std::string& m)
    {
        err = m;
        return code;
    }
};

static thread_local std::string g_group_err;

#define GRP_NCCL(g, call)                                                                              \
    do                                                                                                 \
    {                                                                                                  \
        ncclResult_t _r = (call);                                                                      \
        if (_r != ncclSuccess) return (g)->fail(PTB_ERR_NCCL, std::string(#call) + ": " + ncclGetErrorString(_r)); \
    } while (0)

// Runs f(local index) for every local member, in parallel when there are several (the per-context calls block); returns the
// first failure and copies that member's message.
static int for_members(ptb_group* g, const std::function<int(int)>& f)
{
    const int n = (int)g->ctx.size();
    std::vector<int> rc(n, PTB_OK);
    if (n == 1) rc[0] = f(0);
    else
    {
        std::vector<std::thread> th;
        th.reserve(n);
        for (int i = 0; i < n; ++i) th.emplace_back([&, i] { rc[i] = f(i); });
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < n; ++i)
        if (rc[i] != PTB_OK) return g->fail(rc[i], "member " + std::to_string(g->rank[i]) + ": " + g->ctx[i]->err);
    return PTB_OK;
}

extern "C" {

int ptb_group_sample_range(uint32_t begin, uint32_t count, int rank, int n_ranks, uint32_t* first, uint32_t* mine)
{
    if (n_ranks <= 0 || rank < 0 || rank >= n_ranks || !first || !mine) return PTB_ERR_INVALID;
    const uint32_t n = (uint32_t)n_ranks, r = (uint32_t)rank;
    const uint32_t base = count / n, extra = count % n;
    *mine = base + (r < extra ? 1u : 0u);
    *first = begin + r * base + (r < extra ? r : extra);
    return PTB_OK;
}

int ptb_group_unique_id(uint8_t id[128])
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    if (!id) return PTB_ERR_INVALID;
    ncclUniqueId u;
    const ncclResult_t r = ncclGetUniqueId(&u);
    if (r != ncclSuccess)
    {
        g_group_err = std::string("ncclGetUniqueId: ") + ncclGetErrorString(r);
        return PTB_ERR_NCCL;
    }
    std::memcpy(id, &u, 128);
    return PTB_OK;
}

int ptb_group_create(int n_devices, const int* device_ordinals, ptb_group** out)
{
    if (!out) return PTB_ERR_INVALID;
    *out = nullptr;
    if (n_devices <= 0 || !device_ordinals)
    {
        g_group_err = "ptb_group_create: need at least one device ordinal";
        return PTB_ERR_INVALID;
    }
    for (int i = 0; i < n_devices; ++i)
        for (int j = 0; j < i; ++j)
            if (device_ordinals[i] == device_ordinals[j])
            {
                g_group_err = "ptb_group_create: a device may appear once";
                return PTB_ERR_INVALID;
            }
    ptb_group* g = new (std::nothrow) ptb_group();
    if (!g) return PTB_ERR_OOM;
    g->n_ranks = n_devices;
    g->owns_contexts = true;
    for (int i = 0; i < n_devices; ++i)
    {
        ptb_context* c = nullptr;
        const int rc = ptb_create(device_ordinals[i], &c);
        if (rc != PTB_OK)
        {
            g_group_err = std::string("ptb_create: ") + ptb_last_error(nullptr);
            ptb_group_destroy(g);
            return rc;
        }
        g->ctx.push_back(c);
        g->rank.push_back(i);
    }
    g->comm.assign(n_devices, nullptr);
    const ncclResult_t r = ncclCommInitAll(g->comm.data(), n_devices, device_ordinals);
    if (r != ncclSuccess)
    {
        g_group_err = std::string("ncclCommInitAll: ") + ncclGetErrorString(r);
        g->comm.clear();
        ptb_group_destroy(g);
        return PTB_ERR_NCCL;
    }
    *out = g;
    return PTB_OK;
}

int ptb_group_join(ptb_context* ctx, int n_ranks, int rank, const uint8_t id[128], ptb_group** out)
{
    if (!out) return PTB_ERR_INVALID;
    *out = nullptr;
    if (!ctx || !id || n_ranks <= 0 || rank < 0 || rank >= n_ranks)
    {
        g_group_err = "ptb_group_join: bad arguments";
        return PTB_ERR_INVALID;
    }
    if (ctx->host_only)
    {
        g_group_err = "ptb_group_join: host-only context (no CPU fallback)";
        return PTB_ERR_NO_DEVICE;
    }
    ptb_group* g = new (std::nothrow) ptb_group();
    if (!g) return PTB_ERR_OOM;
    g->n_ranks = n_ranks;
    g->ctx.push_back(ctx);
    g->rank.push_back(rank);
    g->comm.assign(1, nullptr);
    ncclUniqueId u;
    std::memcpy(&u, id, 128);
    cudaSetDevice(ctx->device);
    const ncclResult_t r = ncclCommInitRank(&g->comm[0], n_ranks, u, rank);
    if (r != ncclSuccess)
    {
        g_group_err = std::string("ncclCommInitRank: ") + ncclGetErrorString(r);
        g->comm.clear();
        ptb_group_destroy(g);
        return PTB_ERR_NCCL;
    }
    *out = g;
    return PTB_OK;
}

void ptb_group_destroy(ptb_group* g)
{
    if (!g) return;
    for (size_t i = 0; i < g->ctx.size(); ++i)
    {
        cudaSetDevice(g->ctx[i]->device);
        cudaStreamSynchronize(g->ctx[i]->stream);
        if (i < g->comm.size() && g->comm[i]) ncclCommDestroy(g->comm[i]);
    }
    if (g->ev0) cudaEventDestroy(g->ev0);
    if (g->ev1) cudaEventDestroy(g->ev1);
    if (g->owns_contexts)
        for (ptb_context* c : g->ctx) ptb_destroy(c);
    delete g;
}

const char* ptb_group_last_error(const ptb_group* g) { return g ? g->err.c_str() : g_group_err.c_str(); }

int ptb_group_size(const ptb_group* g, int* n_local, int* n_ranks)
{
    if (!g) return PTB_ERR_INVALID;
    if (n_local) *n_local = (int)g->ctx.size();
    if (n_ranks) *n_ranks = g->n_ranks;
    return PTB_OK;
}

ptb_context* ptb_group_context(ptb_group* g, int local_index)
{
    if (!g || local_index < 0 || local_index >= (int)g->ctx.size()) return nullptr;
    return g->ctx[local_index];
}

int ptb_group_scene_upload(ptb_group* g, const float* verts, uint64_t ntri, const ptb_blas_desc* blas, uint32_t nblas, const ptb_instance* inst, uint32_t ninst)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) { return ptb_scene_upload(g->ctx[i], verts, ntri, blas, nblas, inst, ninst); });
}
int ptb_group_build(ptb_group* g)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) { return ptb_build(g->ctx[i]); });
}
int ptb_group_set_camera(ptb_group* g, const float origin[3], const float target[3], float fov_deg, float aspect_ratio)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) { return ptb_set_camera(g->ctx[i], origin, target, fov_deg, aspect_ratio); });
}
int ptb_group_set_option(ptb_group* g, const char* name, int value)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) { return ptb_set_option(g->ctx[i], name, value); });
}
int ptb_group_frame_begin(ptb_group* g, uint32_t width, uint32_t height)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) { return ptb_frame_begin(g->ctx[i], width, height); });
}

int ptb_group_render(ptb_group* g, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) {
        uint32_t first = 0, mine = 0;
        ptb_group_sample_range(sample_begin, sample_count, g->rank[i], g->n_ranks, &first, &mine);
        return mine ? ptb_render(g->ctx[i], first, mine, seed, max_bounces, enable_nee) : (int)PTB_OK;
    });
}

int ptb_group_reduce(ptb_group* g, int root)
{
    if (!g) return PTB_ERR_INVALID;
    if (root < 0 || root >= g->n_ranks) return g->fail(PTB_ERR_INVALID, "ptb_group_reduce: root out of range");
    const int n = (int)g->ctx.size();
    std::vector<float4*> acc(n, nullptr);
    uint32_t W0 = 0, H0 = 0;
    for (int i = 0; i < n; ++i)
    {
        uint32_t W = 0, H = 0;
        if (!render_state_frame(g->ctx[i], &acc[i], &W, &H)) return g->fail(PTB_ERR_INVALID, "ptb_group_reduce: call ptb_group_frame_begin first");
        if (i == 0) { W0 = W; H0 = H; }
        else if (W != W0 || H != H0) return g->fail(PTB_ERR_INVALID, "ptb_group_reduce: members have different frame sizes");
    }
    const size_t count = (size_t)W0 * H0 * 4;
    if (!g->ev0)
    {
        cudaSetDevice(g->ctx[0]->device);
        if (cudaEventCreate(&g->ev0) != cudaSuccess || cudaEventCreate(&g->ev1) != cudaSuccess) return g->fail(PTB_ERR_CUDA, "ptb_group_reduce: event creation");
    }
    cudaSetDevice(g->ctx[0]->device);
    cudaEventRecord(g->ev0, g->ctx[0]->stream);
    GRP_NCCL(g, ncclGroupStart());
    for (int i = 0; i < n; ++i)
    {
        const ncclResult_t r = ncclReduce(acc[i], acc[i], count, ncclFloat, ncclSum, root, g->comm[i], g->ctx[i]->stream);
        if (r != ncclSuccess)
        {
            ncclGroupEnd();
            return g->fail(PTB_ERR_NCCL, std::string("ncclReduce: ") + ncclGetErrorString(r));
        }
    }
    GRP_NCCL(g, ncclGroupEnd());
    cudaSetDevice(g->ctx[0]->device);
    cudaEventRecord(g->ev1, g->ctx[0]->stream);
    g->reduce_timed = true;
    return PTB_OK;
}

int ptb_group_reduce_ms(ptb_group* g, float* ms)
{
    if (!g || !ms) return PTB_ERR_INVALID;
    if (!g->reduce_timed) return g->fail(PTB_ERR_INVALID, "ptb_group_reduce_ms: no reduce yet");
    cudaSetDevice(g->ctx[0]->device);
    if (cudaEventSynchronize(g->ev1) != cudaSuccess || cudaEventElapsedTime(ms, g->ev0, g->ev1) != cudaSuccess) return g->fail(PTB_ERR_CUDA, "ptb_group_reduce_ms: event timing");
    return PTB_OK;
}

static int local_of(ptb_group* g, int member)
{
    for (size_t i = 0; i < g->rank.size(); ++i)
        if (g->rank[i] == member) return (int)i;
    return -1;
}

int ptb_group_read_accum(ptb_group* g, int root, float* rgba)
{
    if (!g) return PTB_ERR_INVALID;
    const int i = local_of(g, root);
    if (i < 0) return g->fail(PTB_ERR_INVALID, "ptb_group_read_accum: that member is not local to this process");
    const int rc = ptb_read_accum(g->ctx[i], rgba);
    return rc == PTB_OK ? PTB_OK : g->fail(rc, g->ctx[i]->err);
}
int ptb_group_read_aux(ptb_group* g, int root, float* position4, uint32_t* id)
{
    if (!g) return PTB_ERR_INVALID;
    const int i = local_of(g, root);
    if (i < 0) return g->fail(PTB_ERR_INVALID, "ptb_group_read_aux: that member is not local to this process");
    const int rc = ptb_read_aux(g->ctx[i], position4, id);
    return rc == PTB_OK ? PTB_OK : g->fail(rc, g->ctx[i]->err);
}

int ptb_group_get_counters(ptb_group* g, uint64_t out[8])
{
    if (!g || !out) return PTB_ERR_INVALID;
    for (int k = 0; k < 8; ++k) out[k] = 0;
    for (ptb_context* c : g->ctx)
    {
        uint64_t v[8];
        const int rc = ptb_get_counters(c, v);
        if (rc != PTB_OK) return g->fail(rc, c->err);
        for (int k = 0; k < 8; ++k) out[k] = k == 4 ? std::max(out[k], v[k]) : out[k] + v[k];
    }
    return PTB_OK;
}
int ptb_group_reset_counters(ptb_group* g)
{
    if (!g) return PTB_ERR_INVALID;
    return for_members(g, [&](int i) { return ptb_reset_counters(g->ctx[i]); });
}

}  // extern "C"
