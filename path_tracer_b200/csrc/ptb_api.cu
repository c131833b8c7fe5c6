// ptb_api.cu — the extern "C" boundary of include/ptb.h: context, Scene::new, build download, fixed ray batches.
// The frame loop lives in render.cu.  No compute entry point has a CPU path: without a device they return
// PTB_ERR_NO_DEVICE.  A host-only context (device_ordinal < 0) exists solely so that the *host* part of the product
// — Scene::new, which is host code upstream too — can be exercised and its build output downloaded.
#include <cstring>

#include "context.h"

using namespace ptb;

static thread_local std::string g_create_err;

#define REQUIRE_CTX(ctx) \
    if (!(ctx)) return PTB_ERR_INVALID
#define REQUIRE_DEVICE(ctx)                                                                                                   \
    REQUIRE_CTX(ctx);                                                                                                        \
    if ((ctx)->host_only) return (ctx)->fail(PTB_ERR_NO_DEVICE, "host-only context: no CUDA device bound (no CPU fallback)")
#define REQUIRE_BUILT(ctx) \
    if (!(ctx)->scene.built) return (ctx)->fail(PTB_ERR_NOT_BUILT, "scene not built: call ptb_scene_upload and ptb_build first")

extern "C" {

int ptb_abi_version(void) { return PTB_ABI_VERSION; }

int ptb_create(int device_ordinal, ptb_context** out)
{
    if (!out) return PTB_ERR_INVALID;
    *out = nullptr;
    ptb_context* ctx = new (std::nothrow) ptb_context();
    if (!ctx) return PTB_ERR_OOM;
    if (device_ordinal < 0)
    {
        ctx->host_only = true;
        *out = ctx;
        return PTB_OK;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || device_ordinal >= count)
    {
        g_create_err = e != cudaSuccess ? std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e) : "device ordinal out of range";
        delete ctx;
        return PTB_ERR_NO_DEVICE;
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device_ordinal);
    if (prop.major != 10)
    {
        g_create_err = "libptb is built for sm_100a only; device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor);
        delete ctx;
        return PTB_ERR_NO_DEVICE;
    }
    ctx->device = device_ordinal;
    ctx->sm_count = prop.multiProcessorCount;
    e = cudaSetDevice(device_ordinal);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess)
    {
        g_create_err = std::string("stream creation: ") + cudaGetErrorString(e);
        delete ctx;
        return PTB_ERR_CUDA;
    }
    *out = ctx;
    return PTB_OK;
}

static void free_device(ptb_context* c)
{
    c->d_wide.release(); c->d_leaf_headers.release(); c->d_tri_planes.release(); c->d_planes_by_prim.release(); c->d_tri_prim.release();
    c->d_leaf_prims.release(); c->d_bin.release(); c->d_tri_normals.release(); c->d_tri_positions.release(); c->d_env.release(); c->d_inst.release();
    c->d_mat.release(); c->d_vol.release(); c->d_lights.release();
    c->t_o.release(); c->t_d.release(); c->t_tmax.release(); c->t_t.release(); c->t_u.release(); c->t_v.release(); c->t_normal.release();
    c->t_prim.release(); c->t_blas.release(); c->t_inst.release(); c->t_hit.release(); c->t_front.release(); c->t_counter.release();
}

void ptb_destroy(ptb_context* ctx)
{
    if (!ctx) return;
    ctx->wait_async();
    if (!ctx->host_only)
    {
        cudaSetDevice(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        render_state_destroy(ctx);
        temporal_state_destroy(ctx);
        free_device(ctx);
        cudaStreamDestroy(ctx->stream);
    }
    delete ctx;
}

const char* ptb_last_error(const ptb_context* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int ptb_synchronize(ptb_context* ctx)
{
    REQUIRE_DEVICE(ctx);
    const int rc = ctx->wait_async();
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return rc;
}

int ptb_set_build_threads(ptb_context* ctx, int threads)
{
    REQUIRE_CTX(ctx);
    ctx->build_threads = threads;
    return PTB_OK;
}

int ptb_scene_upload(ptb_context* ctx, const float* verts, uint64_t ntri, const ptb_blas_desc* blas, uint32_t nblas, const ptb_instance* inst, uint32_t ninst)
{
    REQUIRE_CTX(ctx);
    ctx->device_ready = false;
    std::string e;
    int rc = ctx->scene.upload(verts, ntri, blas, nblas, inst, ninst, e);
    return rc == PTB_OK ? PTB_OK : ctx->fail(rc, e);
}

int ptb_scene_set_env(ptb_context* ctx, uint32_t w, uint32_t h, const float* rgb)
{
    REQUIRE_CTX(ctx);
    if ((w == 0) != (h == 0) || (w && !rgb)) return ctx->fail(PTB_ERR_INVALID, "bad environment map");
    ctx->scene.env_w = w;
    ctx->scene.env_h = h;
    ctx->scene.env.assign(rgb, rgb + (size_t)w * h * 3);
    ctx->device_ready = false;
    return PTB_OK;
}

}  // extern "C"

template <class T, class U>
static int to_device(ptb_context* ctx, DBuf<T>& d, const std::vector<U>& h, size_t count_T)
{
    PTB_CUDA(ctx, d.resize(std::max<size_t>(count_T, 1)));
    if (count_T) PTB_CUDA(ctx, cudaMemcpyAsync(d.p, h.data(), count_T * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    ctx->device_bvh_bytes += count_T * sizeof(T);
    return PTB_OK;
}

static int upload_device(ptb_context* ctx)
{
    HostScene& sc = ctx->scene;
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->device_bvh_bytes = 0;
    int rc;
#define UP(dst, src, cnt) \
    if ((rc = to_device(ctx, dst, src, cnt)) != PTB_OK) return rc
    UP(ctx->d_wide, sc.wide, sc.wide.size() * 5);
    UP(ctx->d_leaf_headers, sc.leaf_headers, sc.leaf_headers.size() * ptb::LEAF_F4);
    UP(ctx->d_tri_planes, sc.tri_planes, sc.tri_planes.size() / 4);
    UP(ctx->d_tri_prim, sc.tri_prim, sc.tri_prim.size());
    UP(ctx->d_bin, sc.bin, sc.bin.size());
    UP(ctx->d_leaf_prims, sc.leaf_prims, sc.leaf_prims.size());
    UP(ctx->d_planes_by_prim, sc.planes_by_prim, sc.planes_by_prim.size() / 4);
    UP(ctx->d_tri_normals, sc.tri_normals, sc.tri_normals.size());
    UP(ctx->d_tri_positions, sc.tri_positions, sc.tri_positions.size());
    UP(ctx->d_inst, sc.inst_rows, sc.inst_rows.size());
    UP(ctx->d_mat, sc.materials, sc.materials.size());
    UP(ctx->d_vol, sc.volumes, sc.volumes.size());
    UP(ctx->d_lights, sc.light_rows, sc.light_rows.size());
    UP(ctx->d_env, sc.env, sc.env.size());
#undef UP
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    SceneDev S = host_view(sc);  // scalars; pointers replaced below
    S.wide = ctx->d_wide.p;
    S.leaf_headers = ctx->d_leaf_headers.p;
    S.tri_planes = ctx->d_tri_planes.p;
    S.tri_prim = ctx->d_tri_prim.p;
    S.bin = ctx->d_bin.p;
    S.leaf_prims = ctx->d_leaf_prims.p;
    S.planes_by_prim = ctx->d_planes_by_prim.p;
    S.tri_normals = ctx->d_tri_normals.p;
    S.tri_positions = ctx->d_tri_positions.p;
    S.inst = ctx->d_inst.p;
    S.mat = ctx->d_mat.p;
    S.vol = ctx->d_vol.p;
    S.lights = ctx->d_lights.p;
    S.env = sc.env.empty() ? nullptr : ctx->d_env.p;
    ctx->S = S;
    ctx->device_ready = true;
    return PTB_OK;
}

extern "C" {

int ptb_build(ptb_context* ctx)
{
    REQUIRE_CTX(ctx);
    std::string e;
    int rc = ctx->scene.build(ctx->build_threads, e);
    if (rc != PTB_OK) return ctx->fail(rc, e);
    if (!ctx->host_only) return upload_device(ctx);
    return PTB_OK;
}

// ---- build download -------------------------------------------------------------------------------------------------
int ptb_bvh_blas_counts(ptb_context* ctx, uint32_t blas, uint32_t* node_count, uint32_t* leaf_prim_count, uint32_t* root)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    if (blas >= ctx->scene.blas.size()) return ctx->fail(PTB_ERR_INVALID, "blas index out of range");
    const HostBlas& B = ctx->scene.blas[blas];
    if (node_count) *node_count = (uint32_t)B.nodes.size();
    if (leaf_prim_count) *leaf_prim_count = (uint32_t)B.leaf_prims.size();
    if (root) *root = B.root;
    return PTB_OK;
}
int ptb_bvh_blas_nodes(ptb_context* ctx, uint32_t blas, float* boxes6, uint32_t* meta3, uint32_t* leaf_prims)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    if (blas >= ctx->scene.blas.size()) return ctx->fail(PTB_ERR_INVALID, "blas index out of range");
    const HostBlas& B = ctx->scene.blas[blas];
    for (size_t i = 0; i < B.nodes.size(); ++i)
    {
        const HostNode& n = B.nodes[i];
        if (boxes6) { std::memcpy(boxes6 + i * 6, n.lo, 12); std::memcpy(boxes6 + i * 6 + 3, n.hi, 12); }
        if (meta3) { meta3[i * 3] = n.kind; meta3[i * 3 + 1] = n.a; meta3[i * 3 + 2] = n.b; }
    }
    if (leaf_prims) std::memcpy(leaf_prims, B.leaf_prims.data(), B.leaf_prims.size() * 4);
    return PTB_OK;
}
int ptb_bvh_blas_triangles(ptb_context* ctx, uint32_t blas, float* out30)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    if (blas >= ctx->scene.blas.size() || !out30) return ctx->fail(PTB_ERR_INVALID, "bad arguments");
    const HostBlas& B = ctx->scene.blas[blas];
    for (uint32_t i = 0; i < B.tri_count; ++i)
    {
        const HostTri& t = ctx->scene.tris[B.tri_base + i];
        float* o = out30 + (size_t)i * 30;
        std::memcpy(o, t.p, 36);
        std::memcpy(o + 9, t.n, 36);
        std::memcpy(o + 18, t.n0, 16);
        std::memcpy(o + 22, t.n1, 16);
        std::memcpy(o + 26, t.n2, 16);
    }
    return PTB_OK;
}
int ptb_bvh_tlas_counts(ptb_context* ctx, int which, uint32_t* node_count, uint32_t* root, uint32_t* nblas)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    if (which < 0 || which > 1) return ctx->fail(PTB_ERR_INVALID, "which must be 0 (world) or 1 (lights)");
    const HostTlas& T = ctx->scene.tlas[which];
    if (node_count) *node_count = (uint32_t)T.nodes.size();
    if (root) *root = T.root;
    if (nblas) *nblas = (uint32_t)T.blas_ids.size();
    return PTB_OK;
}
int ptb_bvh_tlas_nodes(ptb_context* ctx, int which, float* boxes6, uint32_t* meta3, float* matrices24, uint32_t* blas_ids)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    if (which < 0 || which > 1) return ctx->fail(PTB_ERR_INVALID, "which must be 0 (world) or 1 (lights)");
    const HostTlas& T = ctx->scene.tlas[which];
    for (size_t i = 0; i < T.nodes.size(); ++i)
    {
        const HostTlasNode& n = T.nodes[i];
        if (boxes6) { std::memcpy(boxes6 + i * 6, n.lo, 12); std::memcpy(boxes6 + i * 6 + 3, n.hi, 12); }
        if (meta3) { meta3[i * 3] = n.kind; meta3[i * 3 + 1] = n.a; meta3[i * 3 + 2] = n.b; }
        if (matrices24)
        {
            if (n.kind == 1) { std::memcpy(matrices24 + i * 24, n.matrix, 48); std::memcpy(matrices24 + i * 24 + 12, n.inv, 48); }
            else std::memset(matrices24 + i * 24, 0, 96);
        }
    }
    if (blas_ids) std::memcpy(blas_ids, T.blas_ids.data(), T.blas_ids.size() * 4);
    return PTB_OK;
}
int ptb_light_count(ptb_context* ctx, uint32_t* count)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    if (count) *count = (uint32_t)ctx->scene.lights.size();
    return PTB_OK;
}
int ptb_light_sampler(ptb_context* ctx, uint32_t* items2, float* pdf, float* cdf, float* max)
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    const auto& L = ctx->scene.lights;
    for (size_t i = 0; i < L.size(); ++i)
    {
        if (items2) { items2[i * 2] = L[i].blas_local; items2[i * 2 + 1] = L[i].prim; }
        if (pdf) pdf[i] = L[i].pdf;
        if (cdf) cdf[i] = L[i].cdf;
    }
    if (max) *max = ctx->scene.light_max;
    return PTB_OK;
}
int ptb_bvh_stats(ptb_context* ctx, uint64_t out[8])
{
    REQUIRE_CTX(ctx);
    REQUIRE_BUILT(ctx);
    const HostScene& sc = ctx->scene;
    uint64_t bd = 0;
    for (auto& B : sc.blas) bd = std::max<uint64_t>(bd, B.wide_depth);
    out[0] = sc.wide.size();
    out[1] = sc.leaf_headers.size();
    out[2] = sc.tri_prim.size();
    out[3] = sc.wide.size() * sizeof(WideNode) + sc.leaf_headers.size() * sizeof(LeafHeader) + sc.tri_prim.size() * 52 + sc.inst_rows.size() * sizeof(InstanceRow);
    out[4] = bd + sc.tlas[0].wide_depth;
    out[5] = sc.bin.size();
    out[6] = sc.inst_rows.size();
    out[7] = sc.wide_stack_bound;
    return PTB_OK;
}

// ---- fixed ray batches ----------------------------------------------------------------------------------------------------
int ptb_trace_upload(ptb_context* ctx, uint64_t n, const float* o3, const float* d3, const float* tmax)
{
    REQUIRE_DEVICE(ctx);
    if (!o3 || !d3 || !tmax || n == 0) return ctx->fail(PTB_ERR_INVALID, "empty ray batch");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    PTB_CUDA(ctx, ctx->t_o.resize(n * 3));
    PTB_CUDA(ctx, ctx->t_d.resize(n * 3));
    PTB_CUDA(ctx, ctx->t_tmax.resize(n));
    PTB_CUDA(ctx, ctx->t_t.resize(n));
    PTB_CUDA(ctx, ctx->t_u.resize(n));
    PTB_CUDA(ctx, ctx->t_v.resize(n));
    PTB_CUDA(ctx, ctx->t_normal.resize(n * 3));
    PTB_CUDA(ctx, ctx->t_prim.resize(n));
    PTB_CUDA(ctx, ctx->t_blas.resize(n));
    PTB_CUDA(ctx, ctx->t_inst.resize(n));
    PTB_CUDA(ctx, ctx->t_hit.resize(n));
    PTB_CUDA(ctx, ctx->t_front.resize(n));
    PTB_CUDA(ctx, ctx->t_counter.resize(1));
    PTB_CUDA(ctx, cudaMemcpyAsync(ctx->t_o.p, o3, n * 12, cudaMemcpyHostToDevice, ctx->stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(ctx->t_d.p, d3, n * 12, cudaMemcpyHostToDevice, ctx->stream));
    PTB_CUDA(ctx, cudaMemcpyAsync(ctx->t_tmax.p, tmax, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    ctx->t_n = n;
    return PTB_OK;
}

static TraceArgs trace_args(ptb_context* ctx)
{
    TraceArgs a;
    a.o = ctx->t_o.p; a.d = ctx->t_d.p; a.tmax = ctx->t_tmax.p; a.hit = ctx->t_hit.p; a.t = ctx->t_t.p; a.u = ctx->t_u.p; a.v = ctx->t_v.p;
    a.prim = ctx->t_prim.p; a.blas = ctx->t_blas.p; a.inst = ctx->t_inst.p; a.front = ctx->t_front.p; a.normal = ctx->t_normal.p; a.counter = ctx->t_counter.p;
    return a;
}

int ptb_trace_resident(ptb_context* ctx, int which, int kernel, int any_hit, int repeats, float* mean_ms)
{
    REQUIRE_DEVICE(ctx);
    REQUIRE_BUILT(ctx);
    if (ctx->t_n == 0) return ctx->fail(PTB_ERR_INVALID, "no resident ray batch: call ptb_trace_upload");
    if (which < 0 || which > 1 || kernel < 0 || kernel > 6 || repeats < 1) return ctx->fail(PTB_ERR_INVALID, "bad arguments");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    PTB_CUDA(ctx, cudaEventCreate(&e0));
    PTB_CUDA(ctx, cudaEventCreate(&e1));
    TraceArgs a = trace_args(ctx);
    PTB_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    for (int r = 0; r < repeats; ++r) PTB_CUDA(ctx, launch_trace(ctx->S, which, kernel, any_hit, ctx->t_n, a, ctx->sm_count, ctx->stream, any_hit ? ctx->opt_nee_refill_below : (ctx->opt_refill_below < 0 ? 8 : ctx->opt_refill_below), ctx->opt_node_quorum, ctx->opt_mr));
    PTB_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
    PTB_CUDA(ctx, cudaEventSynchronize(e1));
    float ms = 0;
    PTB_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (mean_ms) *mean_ms = ms / repeats;
    return PTB_OK;
}

int ptb_trace_download(ptb_context* ctx, uint8_t* hit, float* t, uint32_t* prim, uint32_t* inst)
{
    REQUIRE_DEVICE(ctx);
    const uint64_t n = ctx->t_n;
    if (n == 0) return ctx->fail(PTB_ERR_INVALID, "no resident ray batch");
    PTB_CUDA(ctx, cudaSetDevice(ctx->device));
    if (hit) PTB_CUDA(ctx, cudaMemcpyAsync(hit, ctx->t_hit.p, n, cudaMemcpyDeviceToHost, ctx->stream));
    if (t) PTB_CUDA(ctx, cudaMemcpyAsync(t, ctx->t_t.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (prim) PTB_CUDA(ctx, cudaMemcpyAsync(prim, ctx->t_prim.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (inst) PTB_CUDA(ctx, cudaMemcpyAsync(inst, ctx->t_inst.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_trace_closest(ptb_context* ctx, int which, int kernel, uint64_t n, const float* o3, const float* d3, const float* tmax, uint8_t* hit, float* t, float* u, float* v,
                      uint32_t* prim, uint32_t* blas, uint32_t* inst, uint8_t* front, float* normal3)
{
    REQUIRE_DEVICE(ctx);
    REQUIRE_BUILT(ctx);
    if (!hit) return ctx->fail(PTB_ERR_INVALID, "hit output is required");
    if (which < 0 || which > 1 || kernel < 0 || kernel > 6) return ctx->fail(PTB_ERR_INVALID, "bad which/kernel");
    if (n == 0) return PTB_OK;  // empty batch
    int rc = ptb_trace_upload(ctx, n, o3, d3, tmax);
    if (rc != PTB_OK) return rc;
    PTB_CUDA(ctx, launch_trace(ctx->S, which, kernel, 0, n, trace_args(ctx), ctx->sm_count, ctx->stream, ctx->opt_refill_below < 0 ? 8 : ctx->opt_refill_below, ctx->opt_node_quorum, ctx->opt_mr));
#define DL(dst, src, bytes) \
    if (dst) PTB_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream))
    DL(hit, ctx->t_hit.p, n);
    DL(t, ctx->t_t.p, n * 4);
    DL(u, ctx->t_u.p, n * 4);
    DL(v, ctx->t_v.p, n * 4);
    DL(prim, ctx->t_prim.p, n * 4);
    DL(blas, ctx->t_blas.p, n * 4);
    DL(inst, ctx->t_inst.p, n * 4);
    DL(front, ctx->t_front.p, n);
    DL(normal3, ctx->t_normal.p, n * 12);
#undef DL
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_trace_any(ptb_context* ctx, int which, int kernel, uint64_t n, const float* o3, const float* d3, const float* tmax, uint8_t* hit)
{
    REQUIRE_DEVICE(ctx);
    REQUIRE_BUILT(ctx);
    if (!hit) return ctx->fail(PTB_ERR_INVALID, "hit output is required");
    if (which < 0 || which > 1 || kernel < 0 || kernel > 6) return ctx->fail(PTB_ERR_INVALID, "bad which/kernel");
    if (n == 0) return PTB_OK;
    int rc = ptb_trace_upload(ctx, n, o3, d3, tmax);
    if (rc != PTB_OK) return rc;
    PTB_CUDA(ctx, launch_trace(ctx->S, which, kernel, 1, n, trace_args(ctx), ctx->sm_count, ctx->stream, ctx->opt_nee_refill_below, ctx->opt_node_quorum, ctx->opt_mr));
    PTB_CUDA(ctx, cudaMemcpyAsync(hit, ctx->t_hit.p, n, cudaMemcpyDeviceToHost, ctx->stream));
    PTB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}

int ptb_set_option(ptb_context* ctx, const char* name, int value)
{
    REQUIRE_CTX(ctx);
    if (!name) return ctx->fail(PTB_ERR_INVALID, "null option name");
    std::string n(name);
    if (n == "sort_materials") ctx->opt_sort_materials = value;
    else if (n == "extend_kernel") ctx->opt_extend_kernel = value;
    else if (n == "block_threads") ctx->opt_block_threads = value;
    else if (n == "blocks_per_sm") ctx->opt_blocks_per_sm = value;
    else if (n == "node_quorum") ctx->opt_node_quorum = value;
    else if (n == "shade_class") ctx->opt_shade_class = value;
    else if (n == "multiray") ctx->opt_multiray = (value == 2 || value == 3 || value < 0) ? value : 0;
    else if (n == "tile_nodes") ctx->opt_tile_nodes = value ? 1 : 0;
    else if (n == "tile_settle") ctx->opt_tile_settle = value ? 1 : 0;
    else if (n == "tile_lists") ctx->opt_tile_lists = value;
    else if (n == "tile_pixels") ctx->opt_tile_pixels = value ? 1 : 0;
    else if (n == "nee_overlap") ctx->opt_nee_overlap = value;
    else if (n == "regen_blocks") ctx->opt_regen_blocks = value;
    else if (n == "mr_split") ctx->opt_mr_split = value ? 1 : 0;
    else if (n == "mr_min_per_lane") ctx->opt_mr_min_per_lane = value;
    else if (n == "mr_free_quorum") ctx->opt_mr.free_quorum = value;
    else if (n == "mr_qn") ctx->opt_mr.qn = value;
    else if (n == "mr_ql") ctx->opt_mr.ql = value;
    else if (n == "mr_qi") ctx->opt_mr.qi = value;
    else if (n == "mr_blocks") ctx->opt_mr.blocks_per_sm = value;
    else if (n == "extend_blocks") ctx->opt_extend_blocks = value;
    else if (n == "nee_blocks") ctx->opt_nee_blocks = value;
    else if (n == "slots_per_pixel") ctx->opt_slots_per_pixel = value;
    else if (n == "tail_threshold") ctx->opt_tail_threshold = value;
    else if (n == "tail_roles") ctx->opt_tail_roles = value;
    else if (n == "path_threshold") ctx->opt_path_threshold = value;
    else if (n == "shared_device") ctx->opt_shared_device = value ? 1 : 0;
    else if (n == "stack_shared") ctx->opt_stack_shared = (value == 8 || value == 12 || value == 16) ? value : 0;
    else if (n == "thin_rays_per_lane") ctx->opt_thin_rays_per_lane = std::max(0, value);
    else if (n == "path_blocks") ctx->opt_path_blocks = std::max(4, std::min(6, value));
    else if (n == "finish_quorum") ctx->opt_finish_quorum = std::max(1, std::min(32, value));
    else if (n == "tail_slots_per_block") ctx->opt_tail_slots_per_block = value;
    else if (n == "wide_collapse") ctx->scene.wide_collapse = value;  // takes effect at the next ptb_build
    else if (n == "dense") ctx->opt_dense = value;
    else if (n == "wavefront_groups") ctx->opt_wavefront_groups = value;
    else if (n == "shade_sort") ctx->opt_shade_sort = value;
    else if (n == "dense_threshold") ctx->opt_dense_threshold = value;
    else if (n == "nee_refill_below") ctx->opt_nee_refill_below = value < 1 ? 1 : (value > 32 ? 32 : value);
    else if (n == "refill_below") ctx->opt_refill_below = value < 0 ? -1 : (value < 1 ? 1 : (value > 32 ? 32 : value));
    else return ctx->fail(PTB_ERR_INVALID, "unknown option " + n);
    return PTB_OK;
}

}  // extern "C"
