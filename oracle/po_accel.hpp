// ORACLE — TEST INFRASTRUCTURE ONLY (see po_math.hpp header).  PARITY UNPINNED.
// po_accel.hpp — acceleration structures of the reference, restated:
//   AABB            src/tlas/tlas_bvh/blas/blas_bvh/boundingbox.rs
//   Triangle        src/tlas/tlas_bvh/blas/primitive.rs
//   BLAS build      src/tlas/tlas_bvh/blas/blas_bvh.rs:62-136, blas.rs:174-201
//   TLAS build      src/tlas/tlas_bvh.rs:56-138, tlas.rs:24-53
//   traversal       blas.rs:133-162,214-294, tlas.rs:66-144
#pragma once
#include <algorithm>
#include <vector>

#include "po_math.hpp"

namespace po {

// ---- boundingbox.rs ---------------------------------------------------------------------------
struct AABB
{
    V3 mn, mx;
};
inline AABB aabb_identity() { return {splat(INF), splat(-INF)}; }                                  // :59-65
inline AABB surrounding_box(const AABB& a, const AABB& b) { return {vmin(a.mn, b.mn), vmax(a.mx, b.mx)}; }  // :134-139
inline AABB aabb_transform(const AABB& b, const Aff3& m)                                           // :51-57 (two corners only)
{
    V3 a = transform_point(m, b.mn);
    V3 c = transform_point(m, b.mx);
    return {vmin(a, c), vmax(a, c)};
}
inline uint8_t longest_axis(const AABB& b)                                                         // :71-88
{
    V3 len = b.mx - b.mn;
    float m = max_element(len);
    if (len.x == m) return 0;
    if (len.y == m) return 1;
    return 2;
}
inline float surface_area(const AABB& b)                                                           // :90-95
{
    V3 v = b.mx - b.mn;
    return 2.0f * dot3(v, V3{v.z, v.x, v.y});
}
// :115-131.  Returns true and the entry distance (>= EPSILON, independent of t_max) on a hit.
inline bool aabb_intersect_t(const AABB& b, const Ray& r, float t_max, float* t_enter)
{
    V3 t0 = (b.mn - r.o) * r.inv;
    V3 t1 = (b.mx - r.o) * r.inv;
    V3 tmin_v = splat(EPSILON), tmax_v = splat(t_max);
    V3 t_smaller = vmin(vmax(t0, tmin_v), vmax(t1, tmin_v));
    V3 t_bigger = vmax(vmin(t0, tmax_v), vmin(t1, tmax_v));
    float t = max_element(t_smaller);
    *t_enter = t;
    return t <= min_element(t_bigger);
}
inline bool aabb_intersect(const AABB& b, const Ray& r, float t_max)                               // :97-113
{
    float t;
    return aabb_intersect_t(b, r, t_max, &t);
}

// ---- primitive.rs -----------------------------------------------------------------------------
struct Triangle
{
    V3 p[3];       // positions A,B,C   (:17)
    V3 n[3];       // vertex normals    (:19)
    V4 n0, n1, n2; // Havel-Herout planes (:21-26)
};
inline Triangle triangle_new(const V3 pos[3], const V3 nrm[3])                                      // :31-54
{
    Triangle t;
    V3 ab = pos[1] - pos[0];
    V3 ac = pos[2] - pos[0];
    V3 n0 = cross(ab, ac);
    float d0 = dot3(n0, pos[0]);
    float scale = length_squared(n0);
    V3 n1 = cross(ac, n0) / scale;
    float d1 = -dot3(n1, pos[0]);
    V3 n2 = cross(n0, ab) / scale;
    float d2 = -dot3(n2, pos[0]);
    for (int i = 0; i < 3; ++i) { t.p[i] = pos[i]; t.n[i] = nrm[i]; }
    t.n0 = {n0.x, n0.y, n0.z, d0};
    t.n1 = {n1.x, n1.y, n1.z, d1};
    t.n2 = {n2.x, n2.y, n2.z, d2};
    return t;
}
inline V3 tri_normal(const Triangle& t, float u, float v)                                           // :57-63
{
    float w = 1.0f - u - v;
    return normalize(mul(M3{t.n[0], t.n[1], t.n[2]}, V3{w, u, v}));
}
inline V3 tri_position(const Triangle& t, float u, float v)                                         // :66-70
{
    float w = 1.0f - u - v;
    return mul(M3{t.p[0], t.p[1], t.p[2]}, V3{w, u, v});
}
inline float tri_area(const Triangle& t) { return 0.5f * length(V3{t.n0.x, t.n0.y, t.n0.z}); }      // :94
inline AABB tri_bbox(const Triangle& t)                                                             // :97-103
{
    return {vmin(vmin(t.p[0], t.p[1]), t.p[2]), vmax(vmax(t.p[0], t.p[1]), t.p[2])};
}
// :117-144.  out = (t*det, u*det, v*det, det)
inline bool tri_intersect_naive(const Triangle& t, V3 o, V3 d, float t_min, float t_max, V4* out)
{
    float det = dot3(d, V3{t.n0.x, t.n0.y, t.n0.z});
    float td = -dot4(V4{o.x, o.y, o.z, -1.0f}, t.n0);
    if (!same_signum(td - det * t_min, det * t_max - td)) return false;
    V3 p3 = det * o + td * d;
    V4 p = {p3.x, p3.y, p3.z, det};
    float ud = dot4(p, t.n1);
    if (!same_signum(ud, det - ud)) return false;
    float vd = dot4(p, t.n2);
    if (!same_signum(vd, det - ud - vd)) return false;
    *out = {td, ud, vd, det};
    return true;
}
struct HitInfo                                                                                      // model.rs:5-11
{
    V3 normal;
    float u, v;
    float t;
    bool front_facing;
};
inline bool tri_intersect(const Triangle& tri, const Ray& r, float t_max, float t_est, HitInfo* hi)  // :147-178
{
    V3 moved = ray_at(r, t_est);
    V4 tuvd;
    if (!tri_intersect_naive(tri, moved, r.d, EPSILON - t_est, t_max - t_est, &tuvd)) return false;
    V3 q = V3{tuvd.x, tuvd.y, tuvd.z} / tuvd.w;
    V3 n = tri_normal(tri, q.y, q.z);
    bool ff = dot3(r.d, n) < 0.0f;
    hi->normal = ff ? n : -n;
    hi->u = q.y;
    hi->v = q.z;
    hi->t = q.x + t_est;
    hi->front_facing = ff;
    return true;
}
inline bool tri_intersect_bool(const Triangle& tri, const Ray& r, float t_max, float t_est)          // :181-189
{
    V3 moved = ray_at(r, t_est);
    V4 tuvd;
    return tri_intersect_naive(tri, moved, r.d, EPSILON - t_est, t_max - t_est, &tuvd);
}

// ---- blas_bvh.rs ------------------------------------------------------------------------------
constexpr size_t DESIRED_BINS = 64;      // :13
constexpr float TRAVERSAL_COST = 1.0f;   // :15
constexpr float INTERSECTION_COST = 2.0f; // :16

struct PrimInfo                          // :18-22
{
    AABB box;
    uint32_t prim;
};
enum : uint32_t { NODE_BRANCH = 0, NODE_LEAF_SINGLE = 1, NODE_LEAF = 2 };
// Canonical serialisation: a/b = left/right | prim,0 | first index into leaf_prims,count
struct BNode
{
    AABB box;
    uint32_t kind, a, b;
};
struct BlasTree
{
    std::vector<BNode> nodes;          // arena (allocation = post-order) order
    std::vector<uint32_t> leaf_prims;  // concatenated Leaf.primitive_ids in sorted-span order
    uint32_t root = 0;
};

inline AABB prim_fold_box(const PrimInfo* p, size_t n)                                              // :28-33 (left fold)
{
    AABB a = aabb_identity();
    for (size_t i = 0; i < n; ++i) a = surrounding_box(a, p[i].box);
    return a;
}

// :62-136.  Deliberately the naive O(k*n) candidate evaluation of the reference (both side boxes
// re-folded per candidate) — the product builder uses scans, so the two are independent derivations.
inline uint32_t generate_blas(BlasTree& tree, PrimInfo* info, size_t n, uint8_t last_axis)
{
    if (n == 1)
    {
        tree.nodes.push_back({info[0].box, NODE_LEAF_SINGLE, info[0].prim, 0});
        return (uint32_t)tree.nodes.size() - 1;
    }
    AABB box = prim_fold_box(info, n);
    float bb_sa = surface_area(box);
    uint8_t axis = longest_axis(box);
    if (axis != last_axis)
    {
        // glidesort::sort_by is a stable sort; comparator = total_cmp on minimum[axis] (boundingbox.rs:67)
        std::stable_sort(info, info + n, [axis](const PrimInfo& a, const PrimInfo& b) { return total_lt(a.box.mn[axis], b.box.mn[axis]); });
    }
    size_t bin_size = std::max<size_t>(n / DESIRED_BINS, 1);
    size_t num_bins = (n / bin_size) - 1;
    size_t best_split = 0;
    float best_sah = 0.0f;
    for (size_t i = 0; i < num_bins; ++i)
    {
        size_t j = (i + 1) * bin_size;
        AABB l = prim_fold_box(info, j);
        AABB r = prim_fold_box(info + j, n - j);
        float sah = TRAVERSAL_COST + (((float)j) * surface_area(l) + ((float)(n - j)) * surface_area(r)) * INTERSECTION_COST / bb_sa;
        if (i == 0 || total_lt(sah, best_sah))  // min_by(total_cmp): first minimum wins
        {
            best_sah = sah;
            best_split = j;
        }
    }
    float no_split = INTERSECTION_COST * (float)n;
    if (no_split < best_sah)
    {
        uint32_t first = (uint32_t)tree.leaf_prims.size();
        for (size_t i = 0; i < n; ++i) tree.leaf_prims.push_back(info[i].prim);
        tree.nodes.push_back({box, NODE_LEAF, first, (uint32_t)n});
        return (uint32_t)tree.nodes.size() - 1;
    }
    uint32_t left = generate_blas(tree, info, best_split, axis);
    uint32_t right = generate_blas(tree, info + best_split, n - best_split, axis);
    tree.nodes.push_back({box, NODE_BRANCH, left, right});
    return (uint32_t)tree.nodes.size() - 1;
}

struct Blas                                                                                         // blas.rs:164-170
{
    std::vector<Triangle> prims;
    BlasTree tree;
    uint32_t material;  // index into the scene's material table (one Material per BLAS, blas.rs:167)
};
inline void blas_build(Blas& b)                                                                     // blas.rs:174-201
{
    std::vector<PrimInfo> info(b.prims.size());
    for (size_t i = 0; i < b.prims.size(); ++i) info[i] = {tri_bbox(b.prims[i]), (uint32_t)i};
    b.tree = BlasTree{};
    b.tree.root = generate_blas(b.tree, info.data(), info.size(), 4);
}

// ---- tlas_bvh.rs ------------------------------------------------------------------------------
enum : uint32_t { TNODE_BRANCH = 0, TNODE_LEAF = 1 };
struct TNode                                                                                        // :32-52
{
    AABB box;
    uint32_t kind, a, b;  // branch: left,right ; leaf: blas index (within this TLAS), instance ordinal
    Aff3 matrix, inv_matrix;
};
struct Tlas
{
    std::vector<uint32_t> blas_ids;  // TLAS-local blas index -> scene blas index
    std::vector<TNode> nodes;        // arena order: leaves first, then branches in creation order
    uint32_t root = 0;
};
inline size_t find_best_match(const std::vector<TNode>& arena, const std::vector<uint32_t>& nodes, size_t a_index)  // :56-83
{
    const AABB& a = arena[nodes[a_index]].box;
    float best_sa = INF;
    size_t best = (size_t)-1;
    for (size_t i = 0; i < nodes.size(); ++i)
    {
        if (i == a_index) continue;
        float sa = surface_area(surrounding_box(a, arena[nodes[i]].box));
        if (sa < best_sa) { best = i; best_sa = sa; }
    }
    return best;
}
template <class T>
inline T swap_remove(std::vector<T>& v, size_t i)
{
    T x = v[i];
    v[i] = v.back();
    v.pop_back();
    return x;
}
struct InstanceIn
{
    uint32_t blas;  // TLAS-local blas index
    Aff3 matrix;
};
// :85-138.  `instances` must be in (BLAS order x matrix order); returns false when empty (the reference
// panics there: find_best_match indexes nodes[0], tlas_bvh.rs:58).
inline bool generate_tlas(Tlas& t, const std::vector<AABB>& blas_root_boxes, const std::vector<InstanceIn>& instances)
{
    if (instances.empty()) return false;
    std::vector<uint32_t> nodes;
    for (size_t k = 0; k < instances.size(); ++k)
    {
        const InstanceIn& in = instances[k];
        TNode n{};
        n.box = aabb_transform(blas_root_boxes[in.blas], in.matrix);
        n.kind = TNODE_LEAF;
        n.a = in.blas;
        n.b = (uint32_t)k;
        n.matrix = in.matrix;
        n.inv_matrix = inverse(in.matrix);
        t.nodes.push_back(n);
        nodes.push_back((uint32_t)t.nodes.size() - 1);
    }
    size_t a = 0;
    size_t b = find_best_match(t.nodes, nodes, a);
    while (nodes.size() > 1)
    {
        size_t c = find_best_match(t.nodes, nodes, b);
        if (a == c)
        {
            uint32_t n1, n2;
            if (a > b) { n1 = swap_remove(nodes, a); n2 = swap_remove(nodes, b); }
            else { n1 = swap_remove(nodes, b); n2 = swap_remove(nodes, a); }
            a = nodes.size();
            TNode n{};
            n.box = surrounding_box(t.nodes[n1].box, t.nodes[n2].box);
            n.kind = TNODE_BRANCH;
            n.a = n1;
            n.b = n2;
            t.nodes.push_back(n);
            nodes.push_back((uint32_t)t.nodes.size() - 1);
            b = find_best_match(t.nodes, nodes, a);
        }
        else
        {
            a = b;
            b = c;
        }
    }
    t.root = nodes[0];
    return true;
}

// ---- traversal --------------------------------------------------------------------------------
// Work counters for SURVEY §8(d): B = slab tests, T = intersect_naive calls, I = TLAS leaves entered.
struct Counters
{
    uint64_t boxes = 0, tris = 0, insts = 0;
};
struct StackItem
{
    uint32_t id;
    float t;
};
// Reusable traversal stacks (the reference allocates them from a per-path bump arena, integrator.rs:151).
struct Scratch
{
    std::vector<StackItem> tlas_st, blas_st;
    std::vector<uint32_t> tlas_any, blas_any;
};
inline Scratch& scratch()
{
    static thread_local Scratch s;
    return s;
}
// blas.rs:133-162 generic over node arenas
template <class N>
inline void push_to_stack(const Ray& r, float t_max, std::vector<StackItem>& st, const std::vector<N>& arena, uint32_t left, uint32_t right, Counters* c)
{
    float tl, tr;
    bool hl = aabb_intersect_t(arena[left].box, r, t_max, &tl);
    bool hr = aabb_intersect_t(arena[right].box, r, t_max, &tr);
    if (c) c->boxes += 2;
    if (hl && hr)
    {
        if (tl < tr) { st.push_back({right, tr}); st.push_back({left, tl}); }
        else { st.push_back({left, tl}); st.push_back({right, tr}); }
    }
    else if (hl) st.push_back({left, tl});
    else if (hr) st.push_back({right, tr});
}

inline bool blas_intersect(const Blas& b, const Ray& r, float t_max, HitInfo* out, uint32_t* prim, Counters* c)  // blas.rs:214-256
{
    std::vector<StackItem>& st = scratch().blas_st;
    st.clear();
    st.push_back({b.tree.root, 0.0f});
    bool found = false;
    while (!st.empty())
    {
        StackItem cur = st.back();
        st.pop_back();
        if (cur.t > t_max) continue;
        const BNode& n = b.tree.nodes[cur.id];
        if (n.kind == NODE_BRANCH) push_to_stack(r, t_max, st, b.tree.nodes, n.a, n.b, c);
        else if (n.kind == NODE_LEAF_SINGLE)
        {
            if (c) c->tris += 1;
            HitInfo hi;
            if (tri_intersect(b.prims[n.a], r, t_max, cur.t, &hi)) { t_max = hi.t; *out = hi; *prim = n.a; found = true; }
        }
        else
        {
            for (uint32_t k = 0; k < n.b; ++k)
            {
                uint32_t id = b.tree.leaf_prims[n.a + k];
                if (c) c->tris += 1;
                HitInfo hi;
                if (tri_intersect(b.prims[id], r, t_max, cur.t, &hi)) { t_max = hi.t; *out = hi; *prim = id; found = true; }
            }
        }
    }
    return found;
}
inline bool blas_any_intersect(const Blas& b, const Ray& r, float t_max, Counters* c)               // blas.rs:257-294
{
    std::vector<uint32_t>& st = scratch().blas_any;
    st.clear();
    st.push_back(b.tree.root);
    while (!st.empty())
    {
        uint32_t cur = st.back();
        st.pop_back();
        const BNode& n = b.tree.nodes[cur];
        float t_enter;
        if (c) c->boxes += 1;
        if (!aabb_intersect_t(n.box, r, t_max, &t_enter)) continue;
        if (n.kind == NODE_BRANCH) { st.push_back(n.a); st.push_back(n.b); }
        else if (n.kind == NODE_LEAF_SINGLE)
        {
            if (c) c->tris += 1;
            if (tri_intersect_bool(b.prims[n.a], r, t_max, t_enter)) return true;
        }
        else
        {
            for (uint32_t k = 0; k < n.b; ++k)
            {
                if (c) c->tris += 1;
                if (tri_intersect_bool(b.prims[b.tree.leaf_prims[n.a + k]], r, t_max, t_enter)) return true;
            }
        }
    }
    return false;
}

struct TlasHit
{
    HitInfo hi;
    uint32_t blas;      // TLAS-local blas index
    uint32_t instance;  // instance ordinal (leaf index)
    uint32_t prim;
};
inline bool tlas_intersect(const Tlas& t, const std::vector<const Blas*>& blases, const Ray& r, float t_max, TlasHit* out, Counters* c)  // tlas.rs:66-110
{
    if (c) c->boxes += 1;
    if (!aabb_intersect(t.nodes[t.root].box, r, t_max)) return false;
    std::vector<StackItem>& st = scratch().tlas_st;
    st.clear();
    st.push_back({t.root, 0.0f});
    bool found = false;
    const Aff3* hit_matrix = nullptr;
    while (!st.empty())
    {
        StackItem cur = st.back();
        st.pop_back();
        if (cur.t > t_max) continue;
        const TNode& n = t.nodes[cur.id];
        if (n.kind == TNODE_BRANCH) push_to_stack(r, t_max, st, t.nodes, n.a, n.b, c);
        else
        {
            if (c) c->insts += 1;
            Ray lr = ray_transform(r, n.inv_matrix);
            HitInfo hi;
            uint32_t prim;
            if (blas_intersect(*blases[n.a], lr, t_max, &hi, &prim, c))
            {
                t_max = hi.t;
                out->hi = hi;
                out->blas = n.a;
                out->instance = n.b;
                out->prim = prim;
                hit_matrix = &n.matrix;
                found = true;
            }
        }
    }
    if (found) out->hi.normal = transform_vector(*hit_matrix, out->hi.normal);  // tlas.rs:103-105
    return found;
}
inline bool tlas_any_intersect(const Tlas& t, const std::vector<const Blas*>& blases, const Ray& r, float t_max, Counters* c)  // tlas.rs:111-144
{
    std::vector<uint32_t>& st = scratch().tlas_any;
    st.clear();
    st.push_back(t.root);
    while (!st.empty())
    {
        uint32_t cur = st.back();
        st.pop_back();
        const TNode& n = t.nodes[cur];
        if (c) c->boxes += 1;
        if (!aabb_intersect(n.box, r, t_max)) continue;
        if (n.kind == TNODE_BRANCH) { st.push_back(n.a); st.push_back(n.b); }
        else
        {
            if (c) c->insts += 1;
            Ray lr = ray_transform(r, n.inv_matrix);
            if (blas_any_intersect(*blases[n.a], lr, t_max, c)) return true;
        }
    }
    return false;
}

}  // namespace po
