// context.h — internal state behind the opaque ptb_context of include/ptb.h.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <thread>
#include <vector>

#include "host_scene.h"

namespace ptb {

template <class T>
struct DBuf  // owning device array
{
    T* p = nullptr;
    size_t n = 0;
    cudaError_t resize(size_t count)
    {
        if (count <= n && p) return cudaSuccess;
        release();
        if (count == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
        if (e == cudaSuccess) n = count;
        else p = nullptr;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    size_t bytes() const { return n * sizeof(T); }
};

// SoA wavefront state of the renderer (render.cu); one slot per pixel (path regeneration keeps a slot on its pixel, so
// per-pixel accumulation needs no atomics).
struct RenderState;
// accumulation texture + velocity of the temporal pass (temporal.cu)
struct TemporalState;

}  // namespace ptb

// options of the multiray kernels (multiray.cuh): finished / idle ray slots of a warp that trigger a finish + start pass, lanes that must
// hold a ray waiting for the node / leaf-triangle / instance stage before the warp runs it, resident blocks per SM (0 = what fits)
struct MrOpts
{
    int free_quorum = 24, qn = 12, ql = 8, qi = 8, blocks_per_sm = 0;
};
struct ptb_context
{
    int device = -1;
    bool host_only = false;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    int build_threads = 0;

    ptb::HostScene scene;
    // device copies of the layout
    ptb::DBuf<uint4> d_wide;
    ptb::DBuf<float4> d_leaf_headers, d_tri_planes, d_planes_by_prim;
    ptb::DBuf<uint32_t> d_tri_prim, d_leaf_prims;
    ptb::DBuf<ptb::BinNode> d_bin;
    ptb::DBuf<float> d_tri_normals, d_tri_positions, d_env;
    ptb::DBuf<ptb::InstanceRow> d_inst;
    ptb::DBuf<ptb::MaterialRow> d_mat;
    ptb::DBuf<ptb::VolumeRow> d_vol;
    ptb::DBuf<ptb::LightRow> d_lights;
    ptb::SceneDev S{};
    bool device_ready = false;
    uint64_t device_bvh_bytes = 0;

    // resident ray batch (ptb_trace_upload / ptb_trace_resident)
    ptb::DBuf<float> t_o, t_d, t_tmax, t_t, t_u, t_v, t_normal;
    ptb::DBuf<uint32_t> t_prim, t_blas, t_inst;
    ptb::DBuf<uint8_t> t_hit, t_front;
    ptb::DBuf<unsigned long long> t_counter;
    uint64_t t_n = 0;

    // camera (Camera::new): ray matrix = matrix * inv_projection (column-major 4x4) and eye
    float cam_m[16] = {0};
    float cam_eye[3] = {0};
    bool cam_set = false;

    // options
    int opt_sort_materials = 1;
    int opt_extend_kernel = 0;  // 0 wide, 1 reference-order
    int opt_block_threads = 128;
    int opt_blocks_per_sm = 8;
    int opt_refill_below = -1, opt_nee_refill_below = 8;   // persistent kernels refill idle lanes when fewer than this many lanes of a warp are traversing
    int opt_shade_class = -1;     // k_shade_ordered instantiated for the scene's material kinds (-1, default) or the general body (0)
    int opt_multiray = -1;        // rays per lane of k_extend / k_nee: 0 = one (persist.cuh), 2 / 3 = multiray.cuh, -1 = by scene size (render.cu)
    int opt_tile_nodes = 1;       // tile lists also name the one to three BLAS nodes a tile's frustum stays within: camera rays start their BLAS walk there
    int opt_tile_settle = 0;      // camera rays of tiles with an empty list are settled without a traversal (C4 -1.6 % per frame; off: they stay traced rays)
    int opt_tile_lists = -1;      // camera rays start from their tile's list of TLAS leaves (k_tile_lists): 1, 0, < 0 = scenes of >= 64 instances
    int opt_tile_pixels = 1;      // slots laid over the frame in 8 x 4 pixel tiles (a warp's 32 rays cover a compact patch) instead of row-major lines
    int opt_nee_overlap = 0;      // live slots per lane of the one-ray grid below which k_nee runs beside the next iteration's k_extend (0 never, < 0 always)
    int opt_regen_blocks = 7;     // resident blocks per SM (= register cap) of the regeneration part of a two-part k_extend launch (5..8)
    int opt_mr_split = 1;         // with the multiray kernels: continuing slots (bounce rays) two per lane, regenerated slots (primary rays) one per lane, two launches
    int opt_mr_min_per_lane = 16; // multiray kernels only while the live slots are at least this many per lane of the one-ray grid
    MrOpts opt_mr;                // multiray kernels (mr_free_quorum, mr_qn, mr_ql, mr_qi, mr_blocks)
    int opt_node_quorum = 8;      // lanes of a warp that must want the node stage before it runs (1 = never postponed)
    int opt_extend_blocks = 7, opt_nee_blocks = 8;  // resident blocks per SM (launch bounds -> register cap) of k_extend / k_nee
    int opt_slots_per_pixel = 16;  // path slots per pixel (1..16): more slots = more paths in flight, fewer and larger wavefront iterations
    int opt_wavefront_groups = -1; // independent wavefronts per render, each over 1/G of the slots on its own stream (1..8; < 0: 1 with the multiray kernels, else 2)
    int opt_dense = 1;            // slot-ordered wavefront kernels while most slots are live (0 = queue kernels throughout)
    int opt_dense_threshold = 60; // live slots per listed 32-slot group, in percent, below which ptb_render continues with the queue kernels
    int opt_shade_sort = -1;      // k_shade_ordered sorts each span's slots by shade kind: 1 yes, 0 no, <0 automatic (scenes with >= 2 surface material kinds)
    int opt_tail_roles = 1;       // tail kernel with shadow / probe / next-bounce traversals on three warps side by side (0: one thread per slot)
    int opt_tail_slots_per_block = 32;  // k_tail_roles: slots per block of three warps (1..32)
    int opt_tail_threshold = -1;  // active slots at which ptb_render switches to the tail kernel (<0 automatic, 0 never)
    int opt_path_threshold = 0;   // active slots at which ptb_render hands ALL remaining slots to the persistent path kernel k_path (0 = never, the default: measured slower than thin iterations + tail kernel; <0 automatic)
    int opt_path_blocks = 5;      // resident blocks per SM (register cap) of k_path (4..6)
    int opt_shared_device = 0;    // 1: another context renders on this device at the same time (ptb_render_async pairs): leave it room, see sized_grid in render.cu
    int opt_thin_rays_per_lane = 0;  // > 0: persistent kernels take only ceil(live / (128 * this)) blocks once that is fewer than their full grid
    int opt_stack_shared = 0;     // first N entries (0, 8, 12, 16) of each thread's traversal stack in shared memory (k_extend<7> / k_nee<8>), the rest in local memory
    int opt_finish_quorum = 8;    // k_path: lanes with a finished traversal that make the warp leave the stepping loop and shade
    int profiling = 0;
    int second_moment = 0;

    ptb::RenderState* rs = nullptr;
    ptb::TemporalState* ts = nullptr;

    // ptb_render_async: the blocking render loop on a worker thread of this context (joined by ptb_wait / ptb_synchronize / ptb_destroy)
    std::thread* async_thread = nullptr;
    int async_rc = 0;
    int wait_async()
    {
        if (!async_thread) return 0;
        async_thread->join();
        delete async_thread;
        async_thread = nullptr;
        return async_rc;
    }

    int fail(int code, const std::string& m)
    {
        err = m;
        return code;
    }
    int cuda_fail(cudaError_t e, const char* where)
    {
        err = std::string(where) + ": " + cudaGetErrorString(e);
        return e == cudaErrorMemoryAllocation ? PTB_ERR_OOM : PTB_ERR_CUDA;
    }
};

#define PTB_CUDA(ctx, call)                                         \
    do                                                              \
    {                                                               \
        cudaError_t _e = (call);                                    \
        if (_e != cudaSuccess) return (ctx)->cuda_fail(_e, #call);  \
    } while (0)

namespace ptb {
// render.cu
void render_state_destroy(ptb_context* ctx);
void temporal_state_destroy(ptb_context* ctx);
// trace kernels (kernels_trace.cu)
struct TraceArgs
{
    const float* o;
    const float* d;
    const float* tmax;
    uint8_t* hit;
    float* t;
    float* u;
    float* v;
    uint32_t* prim;
    uint32_t* blas;
    uint32_t* inst;
    uint8_t* front;
    float* normal;
    unsigned long long* counter;
};
cudaError_t launch_trace(const SceneDev& S, int which, int kernel, int any_hit, uint64_t n, const TraceArgs& a, int sm_count, cudaStream_t stream, int refill_below, int node_quorum,
                         const MrOpts& mo = MrOpts());
}  // namespace ptb
