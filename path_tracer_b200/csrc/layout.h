// layout.h — device data layout of the acceleration structures and scene tables (host builds, device reads).
//
// Two representations of the same reference trees live in HBM:
//  (1) the reference's own binary BVHs in arena order (BinNode) — walked by the reference-order kernel, and what
//      ptb_bvh_* downloads for the bit-exact build comparison;
//  (2) an 8-wide compressed layout (WideNode, 80 B, 5 x 128-bit loads) derived from (1) by collapsing, used by the
//      persistent extend/shadow kernels.  Child boxes are 8-bit quantised (conservative); every *reference leaf* keeps its
//      exact f32 AABB in a 32-byte LeafHeader so the entry distance t_enter — which the reference feeds into the
//      triangle test (primitive.rs:147-155) — is recomputed bit-exactly.
#pragma once
#include <cstdint>
#include <vector_types.h>  // uint4 / float4 (CUDA toolkit header, host-safe)

namespace ptb {

// ---- (1) reference-order binary nodes, 32 B -----------------------------------------------------------------
enum : uint32_t { BIN_BRANCH = 0, BIN_LEAF_SINGLE = 1, BIN_LEAF = 2 };
struct BinNode
{
    float lo[3];
    uint32_t a;  // bits 31..30 kind; branch: left | leaf single: prim | leaf: first index into leaf_prims   (all BLAS/TLAS-local)
    float hi[3];
    uint32_t b;  // branch: right | leaf: count | TLAS leaf: instance ordinal
};
static_assert(sizeof(BinNode) == 32, "BinNode must be 32 bytes");
constexpr uint32_t BIN_KIND_SHIFT = 30;
constexpr uint32_t BIN_INDEX_MASK = (1u << 30) - 1;

// ---- (2) 8-wide compressed node, 80 B ------------------------------------------------------------------------
// word layout (uint4 x 5):
//   q0 = { px, py, pz, ex | ey<<8 | ez<<16 | imask<<24 }
//   q1 = { child_base, leaf_base, lmask | 0, 0 }            (leaf_base in LeafHeader units)
//   q2 = { qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7] }
//   q3 = { qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7] }
//   q4 = { qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7] }
// Slot s holds: an inner child (imask bit s; node index child_base + popc(imask & ((1<<s)-1))), a leaf (lmask bit s;
// header index leaf_base + popc(lmask & ((1<<s)-1))), or nothing (qlo = 255 > qhi = 0, never hit).
// Decoded plane = p + q * 2^(e-127); slots are assigned so that (s ^ ray_octant) orders children front to back.
struct WideNode
{
    float p[3];
    uint8_t e[3];
    uint8_t imask;
    uint32_t child_base;
    uint32_t leaf_base;
    uint8_t lmask;
    uint8_t pad[7];
    uint8_t qlo[3][8];
    uint8_t qhi[3][8];
};
static_assert(sizeof(WideNode) == 80, "WideNode must be 80 bytes");

// ---- leaf header, 32 B + 48 B payload (5 x 128-bit loads) ----------------------------------------------------------------------
// BLAS level: exact AABB of one reference leaf (LeafSingle: the primitive's own box, blas_bvh.rs:69-74; Leaf: the union,
// :78), count = number of primitives (bit 31 set: the leaf is the BLAS root -> closest-hit uses t_estimate = 0 and no
// box test, blas.rs:217), first = first triangle slot (planes at tri_planes[3*slot..], prim id at tri_prim[slot]).
// TLAS level: exact AABB of the reference TLAS leaf (AABB::transform of the BLAS root box, tlas_bvh.rs:94), count = wide root
// of the BLAS (or the direct-leaf encoding, TLAS_DIRECT_FLAG below),
// first = instance ordinal (row of the instance table).
// The record carries what the traversal needs next in the same 80 bytes (5 x 128-bit loads issued together, no dependent
// second fetch): BLAS level — the planes n0|d0, n1|d1, n2|d2 of the leaf's FIRST triangle (a copy of tri_planes[3*first..];
// the other triangles of a multi-triangle leaf are read from tri_planes); TLAS level — the instance's inv_matrix.
struct LeafHeader
{
    float lo[3];
    uint32_t count;
    float hi[3];
    uint32_t first;
    float payload[12];
};
static_assert(sizeof(LeafHeader) == 80, "LeafHeader must be 80 bytes");
constexpr uint32_t LEAF_F4 = 5;  // float4 per LeafHeader
constexpr uint32_t LEAF_ROOT_FLAG = 0x80000000u;
// TLAS-level header: count = wide root of the BLAS — or, for a BLAS whose root node holds nothing but at most
// TLAS_DIRECT_MAX_LEAVES leaves (a quad, a single triangle), this flag | lmask << 23 | leaf_base: the traversal then goes
// straight to the exact leaf tests and never fetches or tests that node (its boxes could only repeat what the instance box said)
constexpr uint32_t TLAS_DIRECT_FLAG = 0x80000000u;
constexpr uint32_t TLAS_DIRECT_MAX_LEAVES = 2;
constexpr uint32_t TLAS_DIRECT_BASE_BITS = 23;

// ---- instance table row, 128 B ------------------------------------------------------------------------------------
struct InstanceRow
{
    float inv[12];        // inv_matrix, glam column-major (x_axis,y_axis,z_axis,translation)   tlas_bvh.rs:100
    float fwd[12];        // matrix
    uint32_t blas;        // scene blas index
    uint32_t wide_root;   // WideNode index of the BLAS root
    uint32_t bin_base;    // first BinNode of the BLAS in the concatenated array
    uint32_t bin_root;    // BLAS-local root id
    uint32_t tri_base;    // first global triangle of the BLAS (prim id + tri_base = global triangle)
    uint32_t leafprim_base;
    uint32_t pad[2];
};
static_assert(sizeof(InstanceRow) == 128, "InstanceRow must be 128 bytes");

// ---- materials / media / lights --------------------------------------------------------------------------------------
struct MaterialRow  // 32 B
{
    uint32_t kind;
    float colour[3];
    float a;        // GGX alpha = clamp(roughness^2, 1e-4, 0.9999)
    float ior;
    int32_t volume; // row of the volume table or -1
    uint32_t pad;
};
struct VolumeRow  // 32 B
{
    float absorption[3];  // absorption * k
    float c, g;
    uint32_t flags;       // bit0 has_absorption, bit1 has_scatter
    uint32_t pad[2];
};
struct LightRow  // 16 B
{
    uint32_t tri;   // global triangle index
    uint32_t blas;  // scene blas index (material lookup)
    float pdf;
    float cdf;
};

// One TLAS (world or lights) on the device
struct TlasDev
{
    uint32_t wide_root;   // WideNode index
    uint32_t bin_base;    // first TLAS BinNode in the concatenated bin array
    uint32_t bin_root;    // TLAS-local
    uint32_t inst_base;   // first InstanceRow of this TLAS
    uint32_t inst_count;
    // world bounds of everything under this TLAS, widened by a rounding margin (host_wide.cpp): a ray that misses them
    // under an approximate slab test cannot hit any of its triangles (tlas_may_hit, traverse.h)
    float box_lo[3], box_hi[3];
    uint32_t pad;
};

// Everything the kernels need, passed by value
struct SceneDev
{
    const uint4* wide;           // WideNode[]
    const float4* leaf_headers;  // LeafHeader[] as LEAF_F4 x float4
    const float4* tri_planes;    // 3 float4 per triangle slot (n0,n1,n2), leaf order
    const uint32_t* tri_prim;    // triangle slot -> BLAS-local prim id
    const BinNode* bin;          // all BLAS binary nodes, then TLAS nodes
    const uint32_t* leaf_prims;  // concatenated Leaf.primitive_ids of all BLASes
    const float4* planes_by_prim;// 3 float4 per global triangle (n0,n1,n2), prim order — reference-order kernel & lights
    const float* tri_normals;    // 9 floats per global triangle
    const float* tri_positions;  // 9 floats per global triangle
    const InstanceRow* inst;
    const MaterialRow* mat;      // per scene blas
    const VolumeRow* vol;
    const LightRow* lights;
    const float* env;            // w*h*3 or null
    TlasDev tlas[2];
    uint32_t n_lights;
    uint32_t n_volumes;
    uint32_t env_w, env_h;
    float light_max;             // LightSampler::max
    // per 8 x 4 pixel tile of the frame (render.cu: k_tile_lists): the TLAS leaves a camera ray of the tile can reach, nearest first —
    // TILE_LIST_STRIDE uint2 per tile: [0] = (count | TILE_LIST_NONE, 0), then (leaf-header index, lower bound of the entry distance)
    const uint2* tile_lists;
};
constexpr uint32_t TILE_LIST_STRIDE = 16, TILE_LIST_NONE = 0xffffffffu;

}  // namespace ptb
