// host_scene.h — host side of Scene::new (scene.rs:21-35): triangle setup (primitive.rs:31-54), the bit-exact BLAS
// builder (blas_bvh.rs:62-136), the TLAS builder (tlas_bvh.rs:85-138), LightSampler::new (light_sampler.rs:41-61) and
// the derivation of the 8-wide device layout (layout.h).  Product code: shares nothing with oracle/.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/ptb.h"
#include "layout.h"

namespace ptb {

struct HostTri  // Triangle, primitive.rs:14-27
{
    float p[9], n[9], n0[4], n1[4], n2[4];
};
struct HostNode  // canonical (Appendix A) node: box + kind,a,b
{
    float lo[3], hi[3];
    uint32_t kind, a, b;
};
struct HostBlas
{
    uint32_t tri_base = 0, tri_count = 0;
    std::vector<HostNode> nodes;       // arena (post-order) order
    std::vector<uint32_t> leaf_prims;
    uint32_t root = 0;
    uint32_t bin_depth = 0;
    // device placement
    uint32_t bin_base = 0, leafprim_base = 0, wide_root = 0, wide_depth = 0;
};
struct HostTlasNode
{
    float lo[3], hi[3];
    uint32_t kind, a, b;
    float matrix[12], inv[12];
};
struct HostTlas
{
    std::vector<uint32_t> blas_ids;  // TLAS-local -> scene blas
    std::vector<HostTlasNode> nodes;
    uint32_t root = 0;
    uint32_t bin_depth = 0, wide_root = 0, wide_depth = 0, bin_base = 0, inst_base = 0;
};
struct HostLight
{
    uint32_t blas_local, prim;
    float pdf, cdf;
};

struct HostScene
{
    // inputs (copied)
    std::vector<float> verts;  // ntri*18
    std::vector<ptb_blas_desc> blas_desc;
    std::vector<ptb_instance> instances;
    std::vector<float> env;
    uint32_t env_w = 0, env_h = 0;
    int wide_collapse = 1;  // 8-wide collapse of the reference trees: 1 = minimum summed node area (dynamic programme), 0 = greedy largest-area-first
    // build products
    std::vector<HostTri> tris;  // global triangle order
    std::vector<HostBlas> blas;
    HostTlas tlas[2];
    std::vector<HostLight> lights;
    float light_max = 0;
    std::vector<MaterialRow> materials;
    std::vector<VolumeRow> volumes;
    // device layout (host copies)
    std::vector<WideNode> wide;
    std::vector<LeafHeader> leaf_headers;
    std::vector<float> tri_planes;    // 12 floats per slot
    std::vector<uint32_t> tri_prim;
    std::vector<BinNode> bin;
    std::vector<uint32_t> leaf_prims;
    std::vector<InstanceRow> inst_rows;  // world instances then lights instances
    std::vector<LightRow> light_rows;
    std::vector<float> planes_by_prim;  // 12 floats per global triangle (n0,n1,n2)
    std::vector<float> tri_normals;     // 9 floats per global triangle
    std::vector<float> tri_positions;   // 9 floats per global triangle
    uint32_t wide_stack_bound = 0, bin_stack_bound = 0;
    bool built = false;

    // returns ptb_status; message in err
    int upload(const float* v, uint64_t ntri, const ptb_blas_desc* b, uint32_t nblas, const ptb_instance* in, uint32_t ninst, std::string& err);
    int build(int threads, std::string& err);
};

// Host view of the device layout (pointers into the vectors above) — used by the upload and by tests/hostemu.
SceneDev host_view(const HostScene& sc);

constexpr uint32_t WIDE_STACK_CAP = 64;  // entries (uint2) of the per-thread traversal stack of the wide kernels
constexpr uint32_t BIN_STACK_CAP = 160;  // entries (id, t_enter) of the reference-order kernels

}  // namespace ptb
