// host_wide.cpp — derive the device layout (layout.h) from the reference-order trees built in host_build.cpp:
// concatenated BinNode arrays for the reference-order kernel, and the 8-wide compressed nodes + exact leaf headers +
// leaf-ordered triangle planes for the persistent kernels.
//
// Collapse (wide_collapse = 0): start from a binary node's two children and repeatedly open the branch child with the largest surface area
// until 8 children or only leaves remain (reference leaves are never split or merged: each keeps its exact AABB, H2).
// Slots are assigned greedily so that slot index bits (x=4,y=2,z=1) follow the child's centroid offset; traversal
// visits hit slots in descending (slot ^ (7 ^ octant)) order, which is front to back.  Child boxes are quantised to
// 8 bits per plane on a per-node power-of-two grid (floor for lo, ceil for hi, verified in double).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>

#include "exact_math.h"
#include "host_scene.h"

namespace ptb {

namespace {

struct TreeView  // uniform view over a BLAS or TLAS binary tree
{
    size_t count;
    const float* (*lo)(const void*, uint32_t);
    const float* (*hi)(const void*, uint32_t);
    bool (*is_branch)(const void*, uint32_t);
    uint32_t (*left)(const void*, uint32_t);
    uint32_t (*right)(const void*, uint32_t);
    const void* self;
};

double box_sa(const float* lo, const float* hi)
{
    double x = (double)hi[0] - lo[0], y = (double)hi[1] - lo[1], z = (double)hi[2] - lo[2];
    return 2.0 * (x * y + y * z + z * x);
}

void quantise_axis(const float p, const double extent, int nchild, const float* clo, const float* chi, uint8_t* e_out, uint8_t* qlo, uint8_t* qhi)
{
    int e;
    if (!(extent > 0.0)) e = -126;
    else
    {
        int ex2;
        std::frexp(extent / 255.0, &ex2);  // extent/255 = m * 2^ex2, m in [0.5,1)  -> 2^ex2 >= extent/255
        e = std::max(ex2, -126);
    }
    for (;; ++e)
    {
        const double scale = std::ldexp(1.0, e);
        bool ok = true;
        for (int c = 0; c < nchild && ok; ++c)
        {
            double l = std::floor(((double)clo[c] - (double)p) / scale);
            double h = std::ceil(((double)chi[c] - (double)p) / scale);
            if (l < 0) l = 0;
            while (l > 0 && (double)p + l * scale > (double)clo[c]) l -= 1;
            while ((double)p + h * scale < (double)chi[c]) h += 1;
            if (h > 255 || l > 255) { ok = false; break; }
            qlo[c] = (uint8_t)l;
            qhi[c] = (uint8_t)h;
        }
        if (ok) break;
    }
    *e_out = (uint8_t)(e + 127);
}

struct Collapser
{
    HostScene& sc;
    const TreeView& tv;
    bool is_tlas;
    // BLAS context
    const HostBlas* blas = nullptr;
    uint32_t inst_base = 0;  // TLAS: first InstanceRow
    uint32_t max_depth = 0;

    // Optimal collapse (after Ylitie, Karras, Laine 2017, section 3.1), specialised to fixed leaves: every reference leaf stays a leaf
    // (its exact box is part of the parity contract), so the only freedom is WHICH binary branch nodes become 8-wide nodes, and the
    // SAH cost to minimise is the summed surface area of those.  plan[n].cost[i-1] = least cost of representing the subtree of
    // binary node n by at most i child slots of its parent; split[j-2] = how many of j slots go to the left child when n is dissolved.
    struct Plan
    {
        float cost[7];     // [i-1]: least cost with at most i slots
        uint8_t split[7];  // [j-2]: slots given to the left child when the node is dissolved into j slots
        uint8_t used[7];   // [i-1]: slots actually taken at that optimum (1 = the node stays an 8-wide node of its own)
    };
    std::vector<Plan> plan;

    void make_plan()
    {
        plan.assign(tv.count, Plan{});
        for (uint32_t n = 0; n < tv.count; ++n)  // both trees store children before their parent
        {
            Plan& P = plan[n];
            if (!tv.is_branch(tv.self, n)) { for (int i = 0; i < 7; ++i) { P.cost[i] = 0.0f; P.split[i] = 0; P.used[i] = 1; } continue; }
            const Plan& L = plan[tv.left(tv.self, n)];
            const Plan& R = plan[tv.right(tv.self, n)];
            float dist[9];  // dist[j] = n dissolved into j >= 2 slots
            for (int j = 2; j <= 8; ++j)
            {
                float best = 3.0e38f;
                int bk = j / 2;
                for (int t = 0; t < j; ++t)  // candidates from the balanced split outwards: ties (degenerate zero-area trees) stay balanced
                {
                    const int k = j / 2 + ((t & 1) ? (t + 1) / 2 : -(t / 2));
                    if (k < 1 || k > j - 1 || k > 7 || j - k > 7) continue;
                    const float c = L.cost[k - 1] + R.cost[j - k - 1];
                    if (c < best) { best = c; bk = k; }
                }
                dist[j] = best;
                if (j <= 8) P.split[j - 2] = (uint8_t)bk;
            }
            P.cost[0] = (float)box_sa(tv.lo(tv.self, n), tv.hi(tv.self, n)) + dist[8];  // n becomes a wide node of its own
            P.used[0] = 1;
            float best = P.cost[0];
            uint8_t used = 1;
            for (int i = 2; i <= 7; ++i)
            {
                if (dist[i] <= best) { best = dist[i]; used = (uint8_t)i; }  // on a tie rather dissolve: fewer, fuller nodes
                P.cost[i - 1] = best;
                P.used[i - 1] = used;
            }
        }
    }
    // children of the wide node rooted at binary node `bin` (budget slots): follow the plan
    void gather(uint32_t bin, int budget, uint32_t* child, int& n)
    {
        const int k = plan[bin].split[budget - 2];
        place(tv.left(tv.self, bin), k, child, n);
        place(tv.right(tv.self, bin), budget - k, child, n);
    }
    void place(uint32_t bin, int budget, uint32_t* child, int& n)
    {
        if (!tv.is_branch(tv.self, bin)) { child[n++] = bin; return; }
        const int j = plan[bin].used[budget - 1];
        if (j == 1) { child[n++] = bin; return; }  // stays a node of its own
        gather(bin, j, child, n);                   // dissolved into j slots
    }

    void emit_leaf(uint32_t bin_id, bool root_leaf)
    {
        LeafHeader h{};
        std::memcpy(h.lo, tv.lo(tv.self, bin_id), 12);
        std::memcpy(h.hi, tv.hi(tv.self, bin_id), 12);
        if (is_tlas)
        {
            const HostTlasNode& n = ((const HostTlas*)tv.self)->nodes[bin_id];
            // count = wide root of the instance's BLAS, or the direct-leaf encoding when that root holds only one or two leaves
            const HostBlas& B = sc.blas[((const HostTlas*)tv.self)->blas_ids[n.a]];
            const WideNode& rn = sc.wide[B.wide_root];
            if (rn.imask == 0 && (uint32_t)__builtin_popcount(rn.lmask) <= TLAS_DIRECT_MAX_LEAVES && rn.leaf_base < (1u << TLAS_DIRECT_BASE_BITS))
                h.count = TLAS_DIRECT_FLAG | ((uint32_t)rn.lmask << TLAS_DIRECT_BASE_BITS) | rn.leaf_base;
            else h.count = B.wide_root;
            h.first = inst_base + n.b;
            std::memcpy(h.payload, n.inv, sizeof(h.payload));
        }
        else
        {
            const HostNode& n = blas->nodes[bin_id];
            const uint32_t cnt = n.kind == BIN_LEAF_SINGLE ? 1 : n.b;
            h.count = cnt | (root_leaf ? LEAF_ROOT_FLAG : 0u);
            h.first = (uint32_t)sc.tri_prim.size();
            for (uint32_t k = 0; k < cnt; ++k)
            {
                const uint32_t prim = n.kind == BIN_LEAF_SINGLE ? n.a : blas->leaf_prims[n.a + k];
                const HostTri& t = sc.tris[blas->tri_base + prim];
                sc.tri_planes.insert(sc.tri_planes.end(), t.n0, t.n0 + 4);
                sc.tri_planes.insert(sc.tri_planes.end(), t.n1, t.n1 + 4);
                sc.tri_planes.insert(sc.tri_planes.end(), t.n2, t.n2 + 4);
                sc.tri_prim.push_back(prim);
                if (k == 0) { std::memcpy(h.payload, t.n0, 16); std::memcpy(h.payload + 4, t.n1, 16); std::memcpy(h.payload + 8, t.n2, 16); }
            }
        }
        sc.leaf_headers.push_back(h);
    }

    // returns index of the wide root
    uint32_t run(uint32_t bin_root)
    {
        const uint32_t wroot = (uint32_t)sc.wide.size();
        sc.wide.push_back(WideNode{});
        struct Work { uint32_t bin, wide, depth; bool root_is_leaf; };
        std::deque<Work> q;
        q.push_back({bin_root, wroot, 1, !tv.is_branch(tv.self, bin_root)});
        while (!q.empty())
        {
            Work w = q.front();
            q.pop_front();
            max_depth = std::max(max_depth, w.depth);
            uint32_t child[8];
            int n = 0;
            if (w.root_is_leaf) child[n++] = w.bin;
            else
            {
                if (plan.empty())
                {
                    // greedy: open the branch child with the largest surface area until the node is full
                    child[n++] = tv.left(tv.self, w.bin);
                    child[n++] = tv.right(tv.self, w.bin);
                    while (n < 8)
                    {
                        int best = -1;
                        double best_sa = -1.0;
                        for (int c = 0; c < n; ++c)
                            if (tv.is_branch(tv.self, child[c]))
                            {
                                double s = box_sa(tv.lo(tv.self, child[c]), tv.hi(tv.self, child[c]));
                                if (s > best_sa) { best_sa = s; best = c; }
                            }
                        if (best < 0) break;
                        uint32_t b = child[best];
                        child[best] = tv.left(tv.self, b);
                        child[n++] = tv.right(tv.self, b);
                    }
                }
                else gather(w.bin, 8, child, n);
            }
            // node box = box of the binary node (exact union of everything below)
            const float* nlo = tv.lo(tv.self, w.bin);
            const float* nhi = tv.hi(tv.self, w.bin);
            // greedy slot assignment
            int slot_of[8];
            bool slot_used[8] = {false, false, false, false, false, false, false, false};
            bool child_done[8] = {false, false, false, false, false, false, false, false};
            double off[8][3];
            for (int c = 0; c < n; ++c)
                for (int k = 0; k < 3; ++k)
                    off[c][k] = 0.5 * ((double)tv.lo(tv.self, child[c])[k] + (double)tv.hi(tv.self, child[c])[k]) - 0.5 * ((double)nlo[k] + (double)nhi[k]);
            for (int round = 0; round < n; ++round)
            {
                double best = -1e300;
                int bc = -1, bs = -1;
                for (int c = 0; c < n; ++c)
                {
                    if (child_done[c]) continue;
                    for (int s = 0; s < 8; ++s)
                    {
                        if (slot_used[s]) continue;
                        double cost = ((s & 4) ? off[c][0] : -off[c][0]) + ((s & 2) ? off[c][1] : -off[c][1]) + ((s & 1) ? off[c][2] : -off[c][2]);
                        if (cost > best) { best = cost; bc = c; bs = s; }
                    }
                }
                slot_of[bc] = bs;
                slot_used[bs] = true;
                child_done[bc] = true;
            }
            uint32_t in_slot[8];
            for (int s = 0; s < 8; ++s) in_slot[s] = 0xffffffffu;
            for (int c = 0; c < n; ++c) in_slot[slot_of[c]] = child[c];
            WideNode node{};
            std::memcpy(node.p, nlo, 12);
            node.child_base = (uint32_t)sc.wide.size();
            node.leaf_base = (uint32_t)sc.leaf_headers.size();
            float clo[3][8], chi[3][8];
            int order[8], m = 0;
            for (int s = 0; s < 8; ++s)
            {
                node.qlo[0][s] = node.qlo[1][s] = node.qlo[2][s] = 255;
                node.qhi[0][s] = node.qhi[1][s] = node.qhi[2][s] = 0;
                if (in_slot[s] == 0xffffffffu) continue;
                order[m] = s;
                for (int k = 0; k < 3; ++k) { clo[k][m] = tv.lo(tv.self, in_slot[s])[k]; chi[k][m] = tv.hi(tv.self, in_slot[s])[k]; }
                ++m;
            }
            for (int k = 0; k < 3; ++k)
            {
                uint8_t ql[8], qh[8];
                quantise_axis(nlo[k], (double)nhi[k] - (double)nlo[k], m, clo[k], chi[k], &node.e[k], ql, qh);
                for (int i = 0; i < m; ++i) { node.qlo[k][order[i]] = ql[i]; node.qhi[k][order[i]] = qh[i]; }
            }
            for (int s = 0; s < 8; ++s)
            {
                if (in_slot[s] == 0xffffffffu) continue;
                if (!w.root_is_leaf && tv.is_branch(tv.self, in_slot[s]))
                {
                    node.imask |= (uint8_t)(1u << s);
                    q.push_back({in_slot[s], (uint32_t)sc.wide.size(), w.depth + 1, false});
                    sc.wide.push_back(WideNode{});
                }
                else
                {
                    node.lmask |= (uint8_t)(1u << s);
                    emit_leaf(in_slot[s], w.root_is_leaf);
                }
            }
            sc.wide[w.wide] = node;
        }
        return wroot;
    }
};

const float* blas_lo(const void* s, uint32_t i) { return ((const HostBlas*)s)->nodes[i].lo; }
const float* blas_hi(const void* s, uint32_t i) { return ((const HostBlas*)s)->nodes[i].hi; }
bool blas_branch(const void* s, uint32_t i) { return ((const HostBlas*)s)->nodes[i].kind == BIN_BRANCH; }
uint32_t blas_left(const void* s, uint32_t i) { return ((const HostBlas*)s)->nodes[i].a; }
uint32_t blas_right(const void* s, uint32_t i) { return ((const HostBlas*)s)->nodes[i].b; }
const float* tlas_lo(const void* s, uint32_t i) { return ((const HostTlas*)s)->nodes[i].lo; }
const float* tlas_hi(const void* s, uint32_t i) { return ((const HostTlas*)s)->nodes[i].hi; }
bool tlas_branch(const void* s, uint32_t i) { return ((const HostTlas*)s)->nodes[i].kind == 0; }
uint32_t tlas_left(const void* s, uint32_t i) { return ((const HostTlas*)s)->nodes[i].a; }
uint32_t tlas_right(const void* s, uint32_t i) { return ((const HostTlas*)s)->nodes[i].b; }

}  // namespace

int collapse_scene(HostScene& sc, std::string& err)
{
    sc.wide.clear();
    sc.leaf_headers.clear();
    sc.tri_planes.clear();
    sc.tri_prim.clear();
    sc.bin.clear();
    sc.leaf_prims.clear();
    sc.inst_rows.clear();
    sc.light_rows.clear();
    uint32_t max_blas_wide = 0, max_blas_bin = 0;
    for (HostBlas& B : sc.blas)
    {
        B.bin_base = (uint32_t)sc.bin.size();
        B.leafprim_base = (uint32_t)sc.leaf_prims.size();
        for (const HostNode& n : B.nodes)
        {
            BinNode bn;
            std::memcpy(bn.lo, n.lo, 12);
            std::memcpy(bn.hi, n.hi, 12);
            bn.a = n.a | (n.kind << BIN_KIND_SHIFT);
            bn.b = n.b;
            sc.bin.push_back(bn);
        }
        sc.leaf_prims.insert(sc.leaf_prims.end(), B.leaf_prims.begin(), B.leaf_prims.end());
        TreeView tv{B.nodes.size(), blas_lo, blas_hi, blas_branch, blas_left, blas_right, &B};
        Collapser c{sc, tv, false};
        c.blas = &B;
        if (sc.wide_collapse) c.make_plan();
        B.wide_root = c.run(B.root);
        B.wide_depth = c.max_depth;
        max_blas_wide = std::max(max_blas_wide, B.wide_depth);
        max_blas_bin = std::max(max_blas_bin, B.bin_depth);
    }
    uint32_t max_tlas_wide = 0, max_tlas_bin = 0;
    for (int w = 0; w < 2; ++w)
    {
        HostTlas& T = sc.tlas[w];
        T.inst_base = (uint32_t)sc.inst_rows.size();
        T.bin_base = (uint32_t)sc.bin.size();
        for (const HostTlasNode& n : T.nodes)
        {
            BinNode bn;
            std::memcpy(bn.lo, n.lo, 12);
            std::memcpy(bn.hi, n.hi, 12);
            if (n.kind == 1)
            {
                const HostBlas& B = sc.blas[T.blas_ids[n.a]];
                InstanceRow r{};
                std::memcpy(r.inv, n.inv, 48);
                std::memcpy(r.fwd, n.matrix, 48);
                r.blas = T.blas_ids[n.a];
                r.wide_root = B.wide_root;
                r.bin_base = B.bin_base;
                r.bin_root = B.root;
                r.tri_base = B.tri_base;
                r.leafprim_base = B.leafprim_base;
                sc.inst_rows.push_back(r);  // leaves are allocated first, in ordinal order (tlas_bvh.rs:90-104)
                bn.a = n.a | (1u << BIN_KIND_SHIFT);
                bn.b = T.inst_base + n.b;
            }
            else { bn.a = n.a; bn.b = n.b; }
            sc.bin.push_back(bn);
        }
        TreeView tv{T.nodes.size(), tlas_lo, tlas_hi, tlas_branch, tlas_left, tlas_right, &T};
        Collapser c{sc, tv, true};
        c.inst_base = T.inst_base;
        if (sc.wide_collapse) c.make_plan();
        T.wide_root = c.run(T.root);
        T.wide_depth = c.max_depth;
        max_tlas_wide = std::max(max_tlas_wide, T.wide_depth);
        max_tlas_bin = std::max(max_tlas_bin, T.bin_depth);
    }
    for (const HostLight& l : sc.lights)
    {
        const uint32_t bi = sc.tlas[1].blas_ids[l.blas_local];
        sc.light_rows.push_back({sc.blas[bi].tri_base + l.prim, bi, l.pdf, l.cdf});
    }
    sc.planes_by_prim.resize(sc.tris.size() * 12);
    sc.tri_normals.resize(sc.tris.size() * 9);
    sc.tri_positions.resize(sc.tris.size() * 9);
    for (size_t i = 0; i < sc.tris.size(); ++i)
    {
        const HostTri& t = sc.tris[i];
        std::memcpy(&sc.planes_by_prim[i * 12], t.n0, 16);
        std::memcpy(&sc.planes_by_prim[i * 12 + 4], t.n1, 16);
        std::memcpy(&sc.planes_by_prim[i * 12 + 8], t.n2, 16);
        std::memcpy(&sc.tri_normals[i * 9], t.n, 36);
        std::memcpy(&sc.tri_positions[i * 9], t.p, 36);
    }
    // device stack bounds (see traverse.h): wide = one group entry per level + up to 8 instance entries at the TLAS node in flight
    sc.wide_stack_bound = max_tlas_wide + 8 + max_blas_wide + 2;
    sc.bin_stack_bound = max_tlas_bin + max_blas_bin + 4;
    if (sc.wide_stack_bound > WIDE_STACK_CAP || sc.bin_stack_bound > BIN_STACK_CAP)
    {
        err = "BVH too deep for the device traversal stacks (wide bound " + std::to_string(sc.wide_stack_bound) + "/" + std::to_string(WIDE_STACK_CAP) + ", binary bound " +
              std::to_string(sc.bin_stack_bound) + "/" + std::to_string(BIN_STACK_CAP) + ")";
        return PTB_ERR_UNSUPPORTED;
    }
    return PTB_OK;
}

SceneDev host_view(const HostScene& sc)
{
    SceneDev S{};
    S.wide = (const uint4*)sc.wide.data();
    S.leaf_headers = (const float4*)sc.leaf_headers.data();
    S.tri_planes = (const float4*)sc.tri_planes.data();
    S.tri_prim = sc.tri_prim.data();
    S.bin = sc.bin.data();
    S.leaf_prims = sc.leaf_prims.data();
    S.planes_by_prim = (const float4*)sc.planes_by_prim.data();
    S.tri_normals = sc.tri_normals.data();
    S.tri_positions = sc.tri_positions.data();
    S.inst = sc.inst_rows.data();
    S.mat = sc.materials.data();
    S.vol = sc.volumes.data();
    S.lights = sc.light_rows.data();
    S.env = sc.env.empty() ? nullptr : sc.env.data();
    for (int w = 0; w < 2; ++w)
    {
        S.tlas[w].wide_root = sc.tlas[w].wide_root;
        S.tlas[w].bin_base = sc.tlas[w].bin_base;
        S.tlas[w].bin_root = sc.tlas[w].root;
        S.tlas[w].inst_base = sc.tlas[w].inst_base;
        S.tlas[w].inst_count = (uint32_t)(w == 0 ? (sc.tlas[1].inst_base - sc.tlas[0].inst_base) : (sc.inst_rows.size() - sc.tlas[1].inst_base));
        const HostTlasNode& rn = sc.tlas[w].nodes[sc.tlas[w].root];
        for (int k = 0; k < 3; ++k)
        {
            const float m = 1e-4f * ((rn.hi[k] - rn.lo[k]) + std::fabs(rn.lo[k]) + std::fabs(rn.hi[k]) + 1.0f);
            S.tlas[w].box_lo[k] = rn.lo[k] - m;
            S.tlas[w].box_hi[k] = rn.hi[k] + m;
        }
    }
    S.n_lights = (uint32_t)sc.light_rows.size();
    S.n_volumes = (uint32_t)sc.volumes.size();
    S.env_w = sc.env_w;
    S.env_h = sc.env_h;
    S.light_max = sc.light_max;
    return S;
}

}  // namespace ptb
