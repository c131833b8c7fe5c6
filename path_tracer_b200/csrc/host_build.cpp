// host_build.cpp — Scene::new on the host: triangle setup, bit-exact BLAS/TLAS build, light sampler.
// Build with -ffp-contract=off (see exact_math.h).  The 8-wide collapse lives in host_wide.cpp.
//
// BLAS builder (blas_bvh.rs:62-136).  Same decisions, different evaluation strategy than the reference's O(k*n) loop:
//   * node box      : left fold over the span in its pre-sort order (blas_bvh.rs:78)
//   * sort          : stable, key = total order of minimum[axis] (boundingbox.rs:67), only when the axis changes (:86-91)
//   * split boxes   : one forward scan and one backward scan give every candidate's left/right box.  min/max are
//                     order-independent in value; the scans are written so that even the sign of a zero matches the
//                     reference's left folds (forward: acc = (acc<x)?acc:x, backward: acc = (x<acc)?x:acc).
//   * cost          : 1 + ((j*SA(L) + (n-j)*SA(R)) * 2 / SA(box)), first minimum under total_cmp (:101-110)
//   * ids           : sub-trees are built in parallel into a pool, then numbered in post-order (left, right, parent;
//                     :127-133) so node ids equal the reference's arena indices.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>

#include "exact_math.h"
#include "host_scene.h"

namespace ptb {

using ex::V3;

static inline V3 ld3(const float* f) { return ex::v3(f[0], f[1], f[2]); }

// ---- Triangle::new (primitive.rs:31-54) --------------------------------------------------------------------
static void make_triangle(const float* v18, HostTri& t)
{
    V3 p[3], n[3];
    for (int k = 0; k < 3; ++k) { p[k] = ld3(v18 + k * 6); n[k] = ld3(v18 + k * 6 + 3); }
    V3 ab = ex::sub(p[1], p[0]);
    V3 ac = ex::sub(p[2], p[0]);
    V3 n0 = ex::cross(ab, ac);
    float d0 = ex::dot3(n0, p[0]);
    float scale = ex::dot3(n0, n0);
    V3 n1 = ex::div(ex::cross(ac, n0), scale);
    float d1 = -ex::dot3(n1, p[0]);
    V3 n2 = ex::div(ex::cross(n0, ab), scale);
    float d2 = -ex::dot3(n2, p[0]);
    for (int k = 0; k < 3; ++k)
    {
        t.p[k * 3] = p[k].x; t.p[k * 3 + 1] = p[k].y; t.p[k * 3 + 2] = p[k].z;
        t.n[k * 3] = n[k].x; t.n[k * 3 + 1] = n[k].y; t.n[k * 3 + 2] = n[k].z;
    }
    t.n0[0] = n0.x; t.n0[1] = n0.y; t.n0[2] = n0.z; t.n0[3] = d0;
    t.n1[0] = n1.x; t.n1[1] = n1.y; t.n1[2] = n1.z; t.n1[3] = d1;
    t.n2[0] = n2.x; t.n2[1] = n2.y; t.n2[2] = n2.z; t.n2[3] = d2;
}

// ---- a tiny task pool (sub-tree parallelism; the reference builds sequentially, blas_bvh.rs:127-128) -----------
class TaskPool
{
  public:
    explicit TaskPool(int n) : pending_(0), stop_(false)
    {
        for (int i = 0; i < n; ++i) th_.emplace_back([this] { run(); });
    }
    ~TaskPool()
    {
        { std::lock_guard<std::mutex> l(m_); stop_ = true; }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    void submit(std::function<void()> f)
    {
        { std::lock_guard<std::mutex> l(m_); q_.push_back(std::move(f)); ++pending_; }
        cv_.notify_one();
    }
    void wait()
    {
        std::unique_lock<std::mutex> l(m_);
        done_.wait(l, [this] { return pending_ == 0; });
    }

  private:
    void run()
    {
        for (;;)
        {
            std::function<void()> f;
            {
                std::unique_lock<std::mutex> l(m_);
                cv_.wait(l, [this] { return stop_ || !q_.empty(); });
                if (stop_ && q_.empty()) return;
                f = std::move(q_.back());
                q_.pop_back();
            }
            f();
            {
                std::lock_guard<std::mutex> l(m_);
                if (--pending_ == 0) done_.notify_all();
            }
        }
    }
    std::vector<std::thread> th_;
    std::deque<std::function<void()>> q_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    size_t pending_;
    bool stop_;
};

// ---- BLAS builder ----------------------------------------------------------------------------------------------
struct Item  // PrimitiveInfo (blas_bvh.rs:18-22)
{
    float lo[3], hi[3];
    uint32_t prim;
};
struct PoolNode
{
    float lo[3], hi[3];
    uint32_t kind;
    uint32_t l, r;          // pool indices (branch)
    uint32_t first, count;  // item span (leaves)
};
struct BlasBuilder
{
    std::vector<Item> items;
    std::vector<PoolNode> pool;
    std::atomic<uint32_t> next{0};
    TaskPool* tp = nullptr;
    size_t spawn_threshold = 1u << 14;

    uint32_t alloc() { return next.fetch_add(1); }

    static inline void fold_fwd(float* lo, float* hi, const Item& x)  // surrounding_box(acc, x), boundingbox.rs:134-139
    {
        for (int k = 0; k < 3; ++k) { lo[k] = ex::mn(lo[k], x.lo[k]); hi[k] = ex::mx(hi[k], x.hi[k]); }
    }
    static inline void fold_bwd(float* lo, float* hi, const Item& x)  // same value; zero-sign of a later-in-order tie kept
    {
        for (int k = 0; k < 3; ++k) { lo[k] = (x.lo[k] < lo[k]) ? x.lo[k] : lo[k]; hi[k] = (x.hi[k] > hi[k]) ? x.hi[k] : hi[k]; }
    }
    static inline float sa(const float* lo, const float* hi) { return ex::surface_area(ld3(lo), ld3(hi)); }

    // Candidate evaluation kept out of the recursive frame (it owns ~6 KB of scratch).
    static __attribute__((noinline)) void find_split(const Item* it, size_t n, float bb_sa, size_t* out_j, float* out_best)
    {
        const float INF = ex::inf();
        const size_t bin = std::max<size_t>(n / 64, 1);  // DESIRED_BINS, blas_bvh.rs:13,93
        const size_t num_bins = n / bin - 1;             // <= 126
        float Llo[128][3], Lhi[128][3], Rlo[128][3], Rhi[128][3];
        {
            float a[3] = {INF, INF, INF}, b[3] = {-INF, -INF, -INF};
            size_t k = 0, stop = num_bins * bin;
            for (size_t i = 0; i < stop; ++i)
            {
                fold_fwd(a, b, it[i]);
                if (i + 1 == (k + 1) * bin) { std::memcpy(Llo[k], a, 12); std::memcpy(Lhi[k], b, 12); ++k; }
            }
        }
        {
            float a[3] = {INF, INF, INF}, b[3] = {-INF, -INF, -INF};
            size_t k = num_bins;
            for (size_t i = n; i-- > bin;)
            {
                fold_bwd(a, b, it[i]);
                if (i == k * bin) { --k; std::memcpy(Rlo[k], a, 12); std::memcpy(Rhi[k], b, 12); if (k == 0) break; }
            }
        }
        size_t best_j = 0;
        float best = 0.0f;
        for (size_t k = 0; k < num_bins; ++k)
        {
            size_t j = (k + 1) * bin;
            float num = ex::add(ex::mul((float)j, sa(Llo[k], Lhi[k])), ex::mul((float)(n - j), sa(Rlo[k], Rhi[k])));
            float sah = ex::add(1.0f, ex::div(ex::mul(num, 2.0f), bb_sa));
            if (k == 0 || ex::total_lt(sah, best)) { best = sah; best_j = j; }
        }
        *out_j = best_j;
        *out_best = best;
    }

    void build(uint32_t node, size_t first, size_t n, uint8_t last_axis)
    {
        PoolNode& pn = pool[node];
        Item* it = items.data() + first;
        if (n == 1)  // blas_bvh.rs:67-75
        {
            std::memcpy(pn.lo, it[0].lo, 12);
            std::memcpy(pn.hi, it[0].hi, 12);
            pn.kind = BIN_LEAF_SINGLE;
            pn.first = (uint32_t)first;
            pn.count = 1;
            return;
        }
        const float INF = ex::inf();
        float lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
        for (size_t i = 0; i < n; ++i) fold_fwd(lo, hi, it[i]);
        std::memcpy(pn.lo, lo, 12);
        std::memcpy(pn.hi, hi, 12);
        const float bb_sa = sa(lo, hi);
        // longest_axis (boundingbox.rs:71-88): x wins ties, then y
        V3 len = ex::sub(ld3(hi), ld3(lo));
        float ml = ex::mx(ex::mx(len.x, len.y), len.z);
        uint8_t axis = (len.x == ml) ? 0 : ((len.y == ml) ? 1 : 2);
        if (axis != last_axis)
            std::stable_sort(it, it + n, [axis](const Item& a, const Item& b) { return ex::total_key(a.lo[axis]) < ex::total_key(b.lo[axis]); });
        size_t best_j = 0;
        float best = 0.0f;
        find_split(it, n, bb_sa, &best_j, &best);
        const float no_split = ex::mul(2.0f, (float)n);
        if (no_split < best)  // blas_bvh.rs:114-122
        {
            pn.kind = BIN_LEAF;
            pn.first = (uint32_t)first;
            pn.count = (uint32_t)n;
            return;
        }
        uint32_t l = alloc(), r = alloc();
        pool[node].kind = BIN_BRANCH;
        pool[node].l = l;
        pool[node].r = r;
        const size_t nl = best_j, nr = n - best_j;
        if (tp && n >= spawn_threshold)
        {
            tp->submit([this, l, first, nl, axis] { build(l, first, nl, axis); });
            build(r, first + nl, nr, axis);
        }
        else
        {
            build(l, first, nl, axis);
            build(r, first + nl, nr, axis);
        }
    }
};

static void build_blas(HostScene& sc, HostBlas& B, TaskPool* tp)
{
    BlasBuilder bb;
    bb.tp = tp;
    const uint32_t n = B.tri_count;
    bb.items.resize(n);
    for (uint32_t i = 0; i < n; ++i)  // Triangle::create_bounding_box, primitive.rs:97-103
    {
        const HostTri& t = sc.tris[B.tri_base + i];
        Item& it = bb.items[i];
        for (int k = 0; k < 3; ++k)
        {
            it.lo[k] = ex::mn(ex::mn(t.p[k], t.p[3 + k]), t.p[6 + k]);
            it.hi[k] = ex::mx(ex::mx(t.p[k], t.p[3 + k]), t.p[6 + k]);
        }
        it.prim = i;
    }
    bb.pool.resize(2 * (size_t)n);
    uint32_t root = bb.alloc();
    bb.build(root, 0, n, 4);  // initial last_split_axis = 4, blas.rs:191
    if (tp) tp->wait();
    // post-order numbering (left subtree, right subtree, parent) == reference arena order
    const uint32_t used = bb.next.load();
    B.nodes.clear();
    B.nodes.reserve(used);
    B.leaf_prims.clear();
    std::vector<uint32_t> new_id(used, 0xffffffffu);
    struct Frame { uint32_t node; uint8_t state; uint32_t depth; };
    std::vector<Frame> st;
    st.push_back({root, 0, 1});
    uint32_t max_depth = 1;
    while (!st.empty())
    {
        Frame& f = st.back();
        const PoolNode& pn = bb.pool[f.node];
        max_depth = std::max(max_depth, f.depth);
        if (pn.kind == BIN_BRANCH && f.state == 0) { f.state = 1; st.push_back({pn.l, 0, f.depth + 1}); continue; }
        if (pn.kind == BIN_BRANCH && f.state == 1) { f.state = 2; st.push_back({pn.r, 0, f.depth + 1}); continue; }
        HostNode hn;
        std::memcpy(hn.lo, pn.lo, 12);
        std::memcpy(hn.hi, pn.hi, 12);
        hn.kind = pn.kind;
        if (pn.kind == BIN_BRANCH) { hn.a = new_id[pn.l]; hn.b = new_id[pn.r]; }
        else if (pn.kind == BIN_LEAF_SINGLE) { hn.a = bb.items[pn.first].prim; hn.b = 0; }
        else
        {
            hn.a = (uint32_t)B.leaf_prims.size();
            hn.b = pn.count;
            for (uint32_t k = 0; k < pn.count; ++k) B.leaf_prims.push_back(bb.items[pn.first + k].prim);
        }
        new_id[f.node] = (uint32_t)B.nodes.size();
        B.nodes.push_back(hn);
        st.pop_back();
    }
    B.root = (uint32_t)B.nodes.size() - 1;
    B.bin_depth = max_depth;
}

// ---- glam Affine3A helpers (column-major 12 floats) -------------------------------------------------------------
struct Aff
{
    V3 c0, c1, c2, t;
};
static Aff aff_load(const float* f) { return {ld3(f), ld3(f + 3), ld3(f + 6), ld3(f + 9)}; }
static void aff_store(const Aff& a, float* f)
{
    const V3* c[4] = {&a.c0, &a.c1, &a.c2, &a.t};
    for (int k = 0; k < 4; ++k) { f[k * 3] = c[k]->x; f[k * 3 + 1] = c[k]->y; f[k * 3 + 2] = c[k]->z; }
}
// Affine3A::inverse = Mat3A::inverse (cross products * 1/det, transposed), translation = -(M^-1 * t)
static Aff aff_inverse(const Aff& a)
{
    V3 t0 = ex::cross(a.c1, a.c2), t1 = ex::cross(a.c2, a.c0), t2 = ex::cross(a.c0, a.c1);
    float det = ex::dot3(a.c2, t2);
    float id = ex::div(1.0f, det);
    V3 r0 = ex::mul(t0, id), r1 = ex::mul(t1, id), r2 = ex::mul(t2, id);
    Aff o;
    o.c0 = ex::v3(r0.x, r1.x, r2.x);
    o.c1 = ex::v3(r0.y, r1.y, r2.y);
    o.c2 = ex::v3(r0.z, r1.z, r2.z);
    o.t = ex::neg(ex::m3_mul(o.c0, o.c1, o.c2, a.t));
    return o;
}
static V3 aff_point(const Aff& a, V3 p) { return ex::add(ex::m3_mul(a.c0, a.c1, a.c2, p), a.t); }
// Model::new (model.rs:36-51): to_scale_rotation_translation().0 == Vec3::ONE exactly
static bool unit_scale(const Aff& a)
{
    float det = ex::dot3(a.c2, ex::cross(a.c0, a.c1));
    float sgn = ex::is_nan(det) ? det : std::copysign(1.0f, det);
    return ex::mul(ex::length(a.c0), sgn) == 1.0f && ex::length(a.c1) == 1.0f && ex::length(a.c2) == 1.0f;
}

// ---- TLAS builder (tlas_bvh.rs:56-138) ------------------------------------------------------------------------------
static size_t find_best_match(const std::vector<HostTlasNode>& arena, const std::vector<uint32_t>& nodes, size_t a_index)
{
    const HostTlasNode& a = arena[nodes[a_index]];
    float best_sa = ex::inf();
    size_t best = (size_t)-1;
    for (size_t i = 0; i < nodes.size(); ++i)
    {
        if (i == a_index) continue;
        const HostTlasNode& b = arena[nodes[i]];
        V3 lo = ex::vmin(ld3(a.lo), ld3(b.lo)), hi = ex::vmax(ld3(a.hi), ld3(b.hi));
        float sa = ex::surface_area(lo, hi);
        if (sa < best_sa) { best = i; best_sa = sa; }
    }
    return best;
}
static uint32_t swap_remove(std::vector<uint32_t>& v, size_t i)
{
    uint32_t x = v[i];
    v[i] = v.back();
    v.pop_back();
    return x;
}
// leaves: (TLAS-local blas index, matrix) in Model x matrices order
static void build_tlas(HostTlas& T, const std::vector<HostBlas>& blas, const std::vector<std::pair<uint32_t, Aff>>& leaves)
{
    std::vector<uint32_t> nodes;
    T.nodes.clear();
    for (size_t k = 0; k < leaves.size(); ++k)
    {
        const HostBlas& B = blas[T.blas_ids[leaves[k].first]];
        const HostNode& rootn = B.nodes[B.root];
        const Aff& m = leaves[k].second;
        V3 a = aff_point(m, ld3(rootn.lo)), c = aff_point(m, ld3(rootn.hi));  // AABB::transform, boundingbox.rs:51-57
        V3 lo = ex::vmin(a, c), hi = ex::vmax(a, c);
        HostTlasNode n{};
        n.lo[0] = lo.x; n.lo[1] = lo.y; n.lo[2] = lo.z; n.hi[0] = hi.x; n.hi[1] = hi.y; n.hi[2] = hi.z;
        n.kind = 1;
        n.a = leaves[k].first;
        n.b = (uint32_t)k;
        aff_store(m, n.matrix);
        aff_store(aff_inverse(m), n.inv);
        T.nodes.push_back(n);
        nodes.push_back((uint32_t)T.nodes.size() - 1);
    }
    size_t a = 0;
    size_t b = nodes.size() > 1 ? find_best_match(T.nodes, nodes, a) : 0;
    while (nodes.size() > 1)
    {
        size_t c = find_best_match(T.nodes, nodes, b);
        if (a == c)
        {
            uint32_t n1, n2;
            if (a > b) { n1 = swap_remove(nodes, a); n2 = swap_remove(nodes, b); }
            else { n1 = swap_remove(nodes, b); n2 = swap_remove(nodes, a); }
            a = nodes.size();
            HostTlasNode n{};
            V3 lo = ex::vmin(ld3(T.nodes[n1].lo), ld3(T.nodes[n2].lo)), hi = ex::vmax(ld3(T.nodes[n1].hi), ld3(T.nodes[n2].hi));
            n.lo[0] = lo.x; n.lo[1] = lo.y; n.lo[2] = lo.z; n.hi[0] = hi.x; n.hi[1] = hi.y; n.hi[2] = hi.z;
            n.kind = 0;
            n.a = n1;
            n.b = n2;
            T.nodes.push_back(n);
            nodes.push_back((uint32_t)T.nodes.size() - 1);
            if (nodes.size() > 1) b = find_best_match(T.nodes, nodes, a);
        }
        else { a = b; b = c; }
    }
    T.root = nodes[0];
    // depth (for device stack bounds)
    std::vector<std::pair<uint32_t, uint32_t>> st{{T.root, 1}};
    T.bin_depth = 1;
    while (!st.empty())
    {
        auto [id, d] = st.back();
        st.pop_back();
        T.bin_depth = std::max(T.bin_depth, d);
        if (T.nodes[id].kind == 0) { st.push_back({T.nodes[id].a, d + 1}); st.push_back({T.nodes[id].b, d + 1}); }
    }
}

// ---- upload: validation of what the reference asserts / panics on ------------------------------------------------
int HostScene::upload(const float* v, uint64_t ntri, const ptb_blas_desc* b, uint32_t nblas, const ptb_instance* in, uint32_t ninst, std::string& err)
{
    built = false;
    if (!v || !b || !in || ntri == 0 || nblas == 0 || ninst == 0) { err = "empty scene"; return PTB_ERR_INVALID; }
    if (ntri >= (1ull << 30)) { err = "more than 2^30 triangles"; return PTB_ERR_UNSUPPORTED; }
    for (uint64_t i = 0; i < ntri * 18; ++i)
        if (!std::isfinite(v[i])) { err = "non-finite vertex data"; return PTB_ERR_INVALID; }
    bool any_emissive = false;
    for (uint32_t i = 0; i < nblas; ++i)
    {
        if (b[i].tri_count == 0 || (uint64_t)b[i].first_tri + b[i].tri_count > ntri) { err = "BLAS " + std::to_string(i) + ": empty or out-of-range triangle span"; return PTB_ERR_INVALID; }
        if (b[i].material.kind > PTB_DIELECTRIC) { err = "unknown material kind"; return PTB_ERR_INVALID; }
        any_emissive |= b[i].material.kind == PTB_EMISSIVE;
    }
    if (!any_emissive) { err = "scene has no emissive model: Scene::new builds the lights TLAS unconditionally and panics (scene.rs:26, tlas_bvh.rs:58)"; return PTB_ERR_INVALID; }
    uint32_t prev = 0;
    for (uint32_t i = 0; i < ninst; ++i)
    {
        if (in[i].blas >= nblas) { err = "instance references a missing BLAS"; return PTB_ERR_INVALID; }
        if (in[i].blas < prev) { err = "instances must be grouped in BLAS (Model) order"; return PTB_ERR_INVALID; }
        prev = in[i].blas;
        if (!unit_scale(aff_load(in[i].affine))) { err = "Model matrix can only contain translation and rotation (model.rs:43)"; return PTB_ERR_INVALID; }
    }
    verts.assign(v, v + ntri * 18);
    blas_desc.assign(b, b + nblas);
    instances.assign(in, in + ninst);
    return PTB_OK;
}

int collapse_scene(HostScene& sc, std::string& err);  // host_wide.cpp

int HostScene::build(int threads, std::string& err)
{
    if (blas_desc.empty()) { err = "no scene uploaded"; return PTB_ERR_NOT_BUILT; }
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    const uint32_t nblas = (uint32_t)blas_desc.size();
    // triangles in global order = concatenation of the BLAS spans in Model order
    blas.assign(nblas, HostBlas{});
    uint64_t total = 0;
    for (uint32_t i = 0; i < nblas; ++i) { blas[i].tri_base = (uint32_t)total; blas[i].tri_count = blas_desc[i].tri_count; total += blas_desc[i].tri_count; }
    if (total >= (1ull << 30)) { err = "more than 2^30 instanced-unique triangles"; return PTB_ERR_UNSUPPORTED; }
    tris.resize(total);
    {
        std::vector<std::thread> th;
        std::atomic<uint64_t> next{0};
        auto work = [&] {
            for (;;)
            {
                uint64_t c = next.fetch_add(1 << 14);
                if (c >= total) break;
                uint64_t e = std::min<uint64_t>(total, c + (1 << 14));
                for (uint64_t g = c; g < e; ++g)
                {
                    // locate the BLAS of global triangle g
                    uint32_t bi = (uint32_t)(std::upper_bound(blas.begin(), blas.end(), g, [](uint64_t x, const HostBlas& B) { return x < B.tri_base; }) - blas.begin()) - 1;
                    make_triangle(verts.data() + ((uint64_t)blas_desc[bi].first_tri + (g - blas[bi].tri_base)) * 18, tris[g]);
                }
            }
        };
        for (int t = 0; t < threads; ++t) th.emplace_back(work);
        for (auto& t : th) t.join();
    }
    {
        TaskPool tp(std::max(1, threads - 1));
        for (uint32_t i = 0; i < nblas; ++i) build_blas(*this, blas[i], threads > 1 ? &tp : nullptr);  // tlas.rs:31-33 (Model order)
    }
    // materials / media
    materials.clear();
    volumes.clear();
    for (uint32_t i = 0; i < nblas; ++i)
    {
        const ptb_material& m = blas_desc[i].material;
        MaterialRow r{};
        r.kind = m.kind;
        std::memcpy(r.colour, m.colour, 12);
        r.ior = m.ior;
        r.volume = -1;
        if (m.kind == PTB_GGX_METAL || m.kind == PTB_GGX_DIELECTRIC)
        {
            float r2 = ex::mul(m.roughness, m.roughness);  // roughness.powi(2).clamp(0.0001, 0.9999), material.rs:294,309
            r.a = std::fmin(std::fmax(r2, 0.0001f), 0.9999f);
        }
        if (m.has_volume && (m.kind == PTB_GGX_DIELECTRIC || m.kind == PTB_DIELECTRIC))
        {
            if (volumes.size() >= 64) { err = "more than 64 participating media"; return PTB_ERR_UNSUPPORTED; }
            VolumeRow vr{};  // Volume::new, volume.rs:136-142
            for (int k = 0; k < 3; ++k) vr.absorption[k] = ex::mul(m.vol_absorption[k], m.vol_k);
            vr.c = m.vol_c;
            vr.g = std::fmin(std::fmax(m.vol_g, -0.999f), 0.999f);
            vr.flags = (m.vol_k == 0.0f ? 0u : 1u) | (m.vol_c == 0.0f ? 0u : 2u);
            r.volume = (int32_t)volumes.size();
            volumes.push_back(vr);
        }
        materials.push_back(r);
    }
    // world TLAS over all models, lights TLAS over the emissive subset (scene.rs:23-26)
    std::vector<int> light_local(nblas, -1);
    tlas[0] = HostTlas{};
    tlas[1] = HostTlas{};
    for (uint32_t i = 0; i < nblas; ++i)
    {
        tlas[0].blas_ids.push_back(i);
        if (materials[i].kind == PTB_EMISSIVE) { light_local[i] = (int)tlas[1].blas_ids.size(); tlas[1].blas_ids.push_back(i); }
    }
    std::vector<std::pair<uint32_t, Aff>> wl, ll;
    for (const ptb_instance& in : instances)
    {
        Aff m = aff_load(in.affine);
        wl.push_back({in.blas, m});
        if (light_local[in.blas] >= 0) ll.push_back({(uint32_t)light_local[in.blas], m});
    }
    if (ll.empty()) { err = "no instance of an emissive model (lights TLAS would be empty, tlas_bvh.rs:58)"; return PTB_ERR_INVALID; }
    build_tlas(tlas[0], blas, wl);
    build_tlas(tlas[1], blas, ll);
    // LightSampler::new (tlas.rs:55-64, blas.rs:203-212, light_sampler.rs:41-61): sequential sum and scan
    lights.clear();
    std::vector<float> w;
    for (uint32_t lb = 0; lb < tlas[1].blas_ids.size(); ++lb)
    {
        const uint32_t bi = tlas[1].blas_ids[lb];
        const float le = ex::length(ld3(materials[bi].colour));
        for (uint32_t p = 0; p < blas[bi].tri_count; ++p)
        {
            const HostTri& t = tris[blas[bi].tri_base + p];
            float area = ex::mul(0.5f, ex::length(ld3(t.n0)));  // Triangle::area, primitive.rs:94
            lights.push_back({lb, p, 0.0f, 0.0f});
            w.push_back(ex::mul(area, le));
        }
    }
    float mx = 0.0f;
    for (float x : w) mx = ex::add(mx, x);
    light_max = mx;
    float state = 0.0f;
    for (size_t i = 0; i < lights.size(); ++i)
    {
        lights[i].pdf = ex::div(w[i], mx);
        state = ex::add(state, lights[i].pdf);
        lights[i].cdf = state;
    }
    int rc = collapse_scene(*this, err);
    if (rc != PTB_OK) return rc;
    built = true;
    return PTB_OK;
}

}  // namespace ptb
