// ORACLE — TEST INFRASTRUCTURE ONLY (see po_math.hpp header).  PARITY UNPINNED.
// po_capi.cpp — flat C entry points (ctypes) over the restatement: scene build, canonical BVH dump
// (SURVEY Appendix A), fixed-batch traversal with work counters, brute force, Sobol/camera probes and
// the multi-threaded frame driver restating src/main.rs:181-207.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <string>
#include <thread>

#include "po_obj.hpp"
#include "po_shading.hpp"
#include "po_temporal.hpp"

using namespace po;

extern "C" {

// Same POD layout as include/ptb.h (ptb_material / ptb_blas_desc / ptb_instance) so the harness builds one set of arrays.
struct po_material
{
    uint32_t kind;
    float colour[3];
    float roughness;  // GGX: linear roughness (constructor applies powi(2).clamp, material.rs:294)
    float ior;
    uint32_t has_volume;
    float vol_absorption[3];
    float vol_k, vol_c, vol_g;
};
struct po_blas_desc
{
    uint32_t first_tri, tri_count;
    po_material material;
};
struct po_instance
{
    uint32_t blas;
    float affine[12];  // glam Affine3A column-major: x_axis, y_axis, z_axis, translation
};

struct po_scene
{
    Scene sc;
    std::vector<uint32_t> world_inst_blas;  // instance ordinal -> scene blas
    std::string error;
};

static thread_local std::string g_err;
const char* po_last_error() { return g_err.c_str(); }

// model.rs:36-51 — matrices must have scale exactly (1,1,1): glam to_scale_rotation_translation
static bool unit_scale(const Aff3& a)
{
    float det = dot3(a.m.c2, cross(a.m.c0, a.m.c1));
    float sx = length(a.m.c0) * signum(det), sy = length(a.m.c1), sz = length(a.m.c2);
    return sx == 1.0f && sy == 1.0f && sz == 1.0f;
}

po_scene* po_scene_create(const float* verts, uint64_t ntri, const po_blas_desc* blas, uint32_t nblas, const po_instance* inst, uint32_t ninst)
{
    auto* s = new po_scene();
    Scene& sc = s->sc;
    auto fail = [&](const std::string& m) -> po_scene* { g_err = m; delete s; return nullptr; };
    if (nblas == 0 || ninst == 0) return fail("empty scene");
    for (uint32_t b = 0; b < nblas; ++b)
    {
        const po_blas_desc& d = blas[b];
        if (d.tri_count == 0 || (uint64_t)d.first_tri + d.tri_count > ntri) return fail("bad BLAS triangle range");
        auto bl = std::make_unique<Blas>();
        bl->prims.reserve(d.tri_count);
        for (uint32_t t = 0; t < d.tri_count; ++t)
        {
            const float* f = verts + ((uint64_t)d.first_tri + t) * 18;
            V3 p[3], n[3];
            for (int k = 0; k < 3; ++k) { p[k] = {f[k * 6 + 0], f[k * 6 + 1], f[k * 6 + 2]}; n[k] = {f[k * 6 + 3], f[k * 6 + 4], f[k * 6 + 5]}; }
            bl->prims.push_back(triangle_new(p, n));
        }
        bl->material = b;
        blas_build(*bl);
        Material m;
        m.kind = d.material.kind;
        m.colour = {d.material.colour[0], d.material.colour[1], d.material.colour[2]};
        m.ior = d.material.ior;
        if (m.kind == MAT_GGX_METAL || m.kind == MAT_GGX_DIELECTRIC)
        {
            float r2 = d.material.roughness * d.material.roughness;
            m.a = std::fmin(std::fmax(r2, 0.0001f), 0.9999f);
        }
        if (d.material.has_volume && (m.kind == MAT_GGX_DIELECTRIC || m.kind == MAT_DIELECTRIC))
        {
            if (sc.volumes.size() >= 64) return fail("more than 64 volumes");
            m.volume = (int)sc.volumes.size();
            sc.volumes.push_back(volume_new({d.material.vol_absorption[0], d.material.vol_absorption[1], d.material.vol_absorption[2]}, d.material.vol_k, d.material.vol_c, d.material.vol_g));
        }
        sc.materials.push_back(m);
        sc.blases.push_back(std::move(bl));
    }
    // instances must arrive in (BLAS order x matrix order), tlas_bvh.rs:90-104
    std::vector<AABB> roots;
    std::vector<InstanceIn> win, lin;
    std::vector<int> light_local(nblas, -1);
    std::vector<AABB> lroots;
    for (uint32_t b = 0; b < nblas; ++b)
    {
        const Blas& bl = *sc.blases[b];
        roots.push_back(bl.tree.nodes[bl.tree.root].box);
        sc.world_blases.push_back(&bl);
        sc.world.blas_ids.push_back(b);
        if (mat_is_emissive(sc.materials[b]))
        {
            light_local[b] = (int)sc.light_blases.size();
            sc.light_blases.push_back(&bl);
            sc.lights.blas_ids.push_back(b);
            lroots.push_back(roots.back());
        }
    }
    uint32_t prev = 0;
    for (uint32_t i = 0; i < ninst; ++i)
    {
        if (inst[i].blas >= nblas) return fail("instance references missing BLAS");
        if (inst[i].blas < prev) return fail("instances must be grouped in BLAS order");
        prev = inst[i].blas;
        Aff3 m = aff_from_cols12(inst[i].affine);
        if (!unit_scale(m)) return fail("Model matrix can only contain translation and rotation");
        win.push_back({inst[i].blas, m});
        s->world_inst_blas.push_back(inst[i].blas);
        if (light_local[inst[i].blas] >= 0) lin.push_back({(uint32_t)light_local[inst[i].blas], m});
    }
    generate_tlas(sc.world, roots, win);
    // scene.rs:26 builds the lights TLAS unconditionally; with no emissive model the reference panics.
    if (!generate_tlas(sc.lights, lroots, lin)) return fail("scene has no emissive model (reference panics in Scene::new)");
    light_sampler_new(sc);
    return s;
}
void po_scene_destroy(po_scene* s) { delete s; }

int po_scene_set_env(po_scene* s, uint32_t w, uint32_t h, const float* rgb)
{
    s->sc.env.w = w;
    s->sc.env.h = h;
    s->sc.env.data.resize((size_t)w * h);
    for (size_t i = 0; i < (size_t)w * h; ++i) s->sc.env.data[i] = {rgb[i * 3], rgb[i * 3 + 1], rgb[i * 3 + 2]};
    return 0;
}

// ---- canonical dump -----------------------------------------------------------
// triangle derived data: 15 float quads per triangle? -> we emit 30 floats: p[9], n[9], n0[4], n1[4], n2[4]
void po_blas_triangles(const po_scene* s, uint32_t blas, float* out30)
{
    const Blas& b = *s->sc.blases[blas];
    for (size_t i = 0; i < b.prims.size(); ++i)
    {
        const Triangle& t = b.prims[i];
        float* o = out30 + i * 30;
        for (int k = 0; k < 3; ++k) { o[k * 3] = t.p[k].x; o[k * 3 + 1] = t.p[k].y; o[k * 3 + 2] = t.p[k].z; o[9 + k * 3] = t.n[k].x; o[9 + k * 3 + 1] = t.n[k].y; o[9 + k * 3 + 2] = t.n[k].z; }
        const V4* q[3] = {&t.n0, &t.n1, &t.n2};
        for (int k = 0; k < 3; ++k) { o[18 + k * 4] = q[k]->x; o[18 + k * 4 + 1] = q[k]->y; o[18 + k * 4 + 2] = q[k]->z; o[18 + k * 4 + 3] = q[k]->w; }
    }
}
void po_blas_counts(const po_scene* s, uint32_t blas, uint32_t* node_count, uint32_t* leaf_prim_count, uint32_t* root)
{
    const BlasTree& t = s->sc.blases[blas]->tree;
    *node_count = (uint32_t)t.nodes.size();
    *leaf_prim_count = (uint32_t)t.leaf_prims.size();
    *root = t.root;
}
// boxes: 6 floats per node (min xyz, max xyz); meta: 3 u32 per node (kind, a, b)
void po_blas_nodes(const po_scene* s, uint32_t blas, float* boxes, uint32_t* meta, uint32_t* leaf_prims)
{
    const BlasTree& t = s->sc.blases[blas]->tree;
    for (size_t i = 0; i < t.nodes.size(); ++i)
    {
        const BNode& n = t.nodes[i];
        float* b = boxes + i * 6;
        b[0] = n.box.mn.x; b[1] = n.box.mn.y; b[2] = n.box.mn.z; b[3] = n.box.mx.x; b[4] = n.box.mx.y; b[5] = n.box.mx.z;
        meta[i * 3] = n.kind; meta[i * 3 + 1] = n.a; meta[i * 3 + 2] = n.b;
    }
    for (size_t i = 0; i < t.leaf_prims.size(); ++i) leaf_prims[i] = t.leaf_prims[i];
}
static const Tlas& pick(const po_scene* s, int which) { return which ? s->sc.lights : s->sc.world; }
static const std::vector<const Blas*>& pickb(const po_scene* s, int which) { return which ? s->sc.light_blases : s->sc.world_blases; }
void po_tlas_counts(const po_scene* s, int which, uint32_t* node_count, uint32_t* root, uint32_t* nblas)
{
    const Tlas& t = pick(s, which);
    *node_count = (uint32_t)t.nodes.size();
    *root = t.root;
    *nblas = (uint32_t)t.blas_ids.size();
}
// boxes 6/node, meta 3/node (kind,a,b), matrices 24/node (matrix[12], inv_matrix[12]; zeros for branches), blas_ids nblas
void po_tlas_nodes(const po_scene* s, int which, float* boxes, uint32_t* meta, float* matrices, uint32_t* blas_ids)
{
    const Tlas& t = pick(s, which);
    for (size_t i = 0; i < t.nodes.size(); ++i)
    {
        const TNode& n = t.nodes[i];
        float* b = boxes + i * 6;
        b[0] = n.box.mn.x; b[1] = n.box.mn.y; b[2] = n.box.mn.z; b[3] = n.box.mx.x; b[4] = n.box.mx.y; b[5] = n.box.mx.z;
        meta[i * 3] = n.kind; meta[i * 3 + 1] = n.a; meta[i * 3 + 2] = n.b;
        float* m = matrices + i * 24;
        if (n.kind == TNODE_LEAF) { aff_to_cols12(n.matrix, m); aff_to_cols12(n.inv_matrix, m + 12); }
        else for (int k = 0; k < 24; ++k) m[k] = 0.0f;
    }
    for (size_t i = 0; i < t.blas_ids.size(); ++i) blas_ids[i] = t.blas_ids[i];
}
uint32_t po_light_count(const po_scene* s) { return (uint32_t)s->sc.sampler.lights.size(); }
// items: 2 u32 per light (lights-TLAS blas index, prim); pdf, cdf: float per light; max: 1 float
void po_light_sampler(const po_scene* s, uint32_t* items, float* pdf, float* cdf, float* max)
{
    const LightSampler& ls = s->sc.sampler;
    for (size_t i = 0; i < ls.lights.size(); ++i) { items[i * 2] = ls.lights[i].blas; items[i * 2 + 1] = ls.lights[i].prim; pdf[i] = ls.lights[i].pdf; cdf[i] = ls.cdf[i]; }
    *max = ls.max;
}

// Re-run generate_blas on a caller-supplied span (boxes 6 floats + prim id, in the given order, with the
// parent's split axis): lets tests compare sub-trees of a huge product build against the restatement.
// Returns node count; call twice (first with null outputs) to size buffers.
uint32_t po_generate_blas_span(const float* boxes, const uint32_t* prims, uint64_t n, uint32_t last_axis, float* out_boxes, uint32_t* out_meta, uint32_t* out_leaf_prims, uint32_t* leaf_prim_count)
{
    std::vector<PrimInfo> info(n);
    for (uint64_t i = 0; i < n; ++i) info[i] = {{{boxes[i * 6], boxes[i * 6 + 1], boxes[i * 6 + 2]}, {boxes[i * 6 + 3], boxes[i * 6 + 4], boxes[i * 6 + 5]}}, prims[i]};
    BlasTree t;
    t.root = generate_blas(t, info.data(), n, (uint8_t)last_axis);
    if (out_boxes)
    {
        for (size_t i = 0; i < t.nodes.size(); ++i)
        {
            const BNode& nd = t.nodes[i];
            float* b = out_boxes + i * 6;
            b[0] = nd.box.mn.x; b[1] = nd.box.mn.y; b[2] = nd.box.mn.z; b[3] = nd.box.mx.x; b[4] = nd.box.mx.y; b[5] = nd.box.mx.z;
            out_meta[i * 3] = nd.kind; out_meta[i * 3 + 1] = nd.a; out_meta[i * 3 + 2] = nd.b;
        }
        for (size_t i = 0; i < t.leaf_prims.size(); ++i) out_leaf_prims[i] = t.leaf_prims[i];
    }
    *leaf_prim_count = (uint32_t)t.leaf_prims.size();
    return (uint32_t)t.nodes.size();
}

}  // extern "C"

// ---- parallel-for helper (main.rs:72: num_cpus - 1 rayon workers, work stealing) ---------------------
static int resolve_threads(int nthreads)
{
    if (nthreads > 0) return nthreads;
    int hc = (int)std::thread::hardware_concurrency();
    return hc > 1 ? hc - 1 : 1;
}
template <class F>
static void parallel_chunks(uint64_t n, uint64_t chunk, int nthreads, F&& f)
{
    nthreads = resolve_threads(nthreads);
    std::atomic<uint64_t> next{0};
    auto worker = [&](int tid) {
        for (;;)
        {
            uint64_t b = next.fetch_add(chunk);
            if (b >= n) break;
            f(b, std::min(n, b + chunk), tid);
        }
    };
    if (nthreads == 1) { worker(0); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(worker, t);
    for (auto& t : th) t.join();
}

extern "C" {
int po_default_threads() { return resolve_threads(0); }

// ---- fixed ray batches --------------------------------------------------------------------------------
// rays: o[3n], d[3n] (d must be normalised), tmax[n].  Outputs (all optional except hit): hit u8, t,u,v f32,
// prim/blas/inst u32 (blas = scene blas index), front u8, normal[3n].  counters[3] += (boxes, tris, insts).
void po_trace_closest(const po_scene* s, int which, uint64_t n, const float* o, const float* d, const float* tmax, uint8_t* hit, float* t, float* u, float* v,
                      uint32_t* prim, uint32_t* blas, uint32_t* inst, uint8_t* front, float* normal, uint64_t* counters, int nthreads)
{
    const Tlas& tl = pick(s, which);
    const auto& bl = pickb(s, which);
    std::vector<Counters> cs(resolve_threads(nthreads));
    parallel_chunks(n, 4096, nthreads, [&](uint64_t b, uint64_t e, int tid) {
        for (uint64_t i = b; i < e; ++i)
        {
            Ray r = make_ray({o[i * 3], o[i * 3 + 1], o[i * 3 + 2]}, {d[i * 3], d[i * 3 + 1], d[i * 3 + 2]});
            TlasHit th;
            bool h = tlas_intersect(tl, bl, r, tmax[i], &th, &cs[tid]);
            hit[i] = h;
            if (t) t[i] = h ? th.hi.t : INF;
            if (u) u[i] = h ? th.hi.u : 0.0f;
            if (v) v[i] = h ? th.hi.v : 0.0f;
            if (prim) prim[i] = h ? th.prim : 0xffffffffu;
            if (blas) blas[i] = h ? tl.blas_ids[th.blas] : 0xffffffffu;
            if (inst) inst[i] = h ? th.instance : 0xffffffffu;
            if (front) front[i] = h ? th.hi.front_facing : 0;
            if (normal) { normal[i * 3] = h ? th.hi.normal.x : 0; normal[i * 3 + 1] = h ? th.hi.normal.y : 0; normal[i * 3 + 2] = h ? th.hi.normal.z : 0; }
        }
    });
    if (counters) for (auto& c : cs) { counters[0] += c.boxes; counters[1] += c.tris; counters[2] += c.insts; }
}
void po_trace_any(const po_scene* s, int which, uint64_t n, const float* o, const float* d, const float* tmax, uint8_t* hit, uint64_t* counters, int nthreads)
{
    const Tlas& tl = pick(s, which);
    const auto& bl = pickb(s, which);
    std::vector<Counters> cs(resolve_threads(nthreads));
    parallel_chunks(n, 4096, nthreads, [&](uint64_t b, uint64_t e, int tid) {
        for (uint64_t i = b; i < e; ++i)
        {
            Ray r = make_ray({o[i * 3], o[i * 3 + 1], o[i * 3 + 2]}, {d[i * 3], d[i * 3 + 1], d[i * 3 + 2]});
            hit[i] = tlas_any_intersect(tl, bl, r, tmax[i], &cs[tid]);
        }
    });
    if (counters) for (auto& c : cs) { counters[0] += c.boxes; counters[1] += c.tris; counters[2] += c.insts; }
}
// Brute force self-check: every TLAS leaf in arena order, every BLAS leaf node in arena order, with the
// reference's leaf semantics (exact leaf-box test with the running t_max, t_estimate = entry distance; a BLAS
// whose root is a leaf is tested with t_estimate 0 and no box test, blas.rs:217).  Independent of tree order.
void po_trace_brute_closest(const po_scene* s, int which, uint64_t n, const float* o, const float* d, const float* tmax, uint8_t* hit, float* t, uint32_t* prim, uint32_t* inst, int nthreads)
{
    const Tlas& tl = pick(s, which);
    const auto& bl = pickb(s, which);
    parallel_chunks(n, 256, nthreads, [&](uint64_t b, uint64_t e, int) {
        for (uint64_t i = b; i < e; ++i)
        {
            Ray r = make_ray({o[i * 3], o[i * 3 + 1], o[i * 3 + 2]}, {d[i * 3], d[i * 3 + 1], d[i * 3 + 2]});
            float tm = tmax[i];
            bool found = false;
            uint32_t bp = 0xffffffffu, bi = 0xffffffffu;
            for (const TNode& tn : tl.nodes)
            {
                if (tn.kind != TNODE_LEAF) continue;
                if (!aabb_intersect(tn.box, r, tm)) continue;
                Ray lr = ray_transform(r, tn.inv_matrix);
                const Blas& B = *bl[tn.a];
                for (uint32_t ni = 0; ni < B.tree.nodes.size(); ++ni)
                {
                    const BNode& bn = B.tree.nodes[ni];
                    if (bn.kind == NODE_BRANCH) continue;
                    float te = 0.0f;
                    if (ni != B.tree.root && !aabb_intersect_t(bn.box, lr, tm, &te)) continue;
                    uint32_t cnt = bn.kind == NODE_LEAF_SINGLE ? 1 : bn.b;
                    for (uint32_t k = 0; k < cnt; ++k)
                    {
                        uint32_t id = bn.kind == NODE_LEAF_SINGLE ? bn.a : B.tree.leaf_prims[bn.a + k];
                        HitInfo hi;
                        if (tri_intersect(B.prims[id], lr, tm, te, &hi)) { tm = hi.t; found = true; bp = id; bi = tn.b; }
                    }
                }
            }
            hit[i] = found;
            t[i] = found ? tm : INF;
            prim[i] = bp;
            inst[i] = bi;
        }
    });
}
// Evaluate ONE (instance, primitive) against a ray exactly as a leaf visit would: used to classify
// GPU/oracle index mismatches as ties.  t_enter is taken from the reference leaf that owns the primitive.
int po_eval_candidate(const po_scene* s, int which, uint32_t inst_ordinal, uint32_t prim_id, const float* o3, const float* d3, float t_max, float* t_out)
{
    const Tlas& tl = pick(s, which);
    const auto& bl = pickb(s, which);
    const TNode& tn = tl.nodes[inst_ordinal];  // leaves are allocated first, ordinal == arena index
    Ray r = make_ray({o3[0], o3[1], o3[2]}, {d3[0], d3[1], d3[2]});
    Ray lr = ray_transform(r, tn.inv_matrix);
    const Blas& B = *bl[tn.a];
    for (uint32_t ni = 0; ni < B.tree.nodes.size(); ++ni)
    {
        const BNode& bn = B.tree.nodes[ni];
        if (bn.kind == NODE_BRANCH) continue;
        bool owns = false;
        if (bn.kind == NODE_LEAF_SINGLE) owns = bn.a == prim_id;
        else for (uint32_t k = 0; k < bn.b; ++k) owns |= B.tree.leaf_prims[bn.a + k] == prim_id;
        if (!owns) continue;
        float te = 0.0f;
        if (ni != B.tree.root && !aabb_intersect_t(bn.box, lr, t_max, &te)) return 0;
        HitInfo hi;
        if (!tri_intersect(B.prims[prim_id], lr, t_max, te, &hi)) return 0;
        *t_out = hi.t;
        return 1;
    }
    return 0;
}

// ---- sampling / camera probes -------------------------------------------------------------------------
void po_sobol_table(uint32_t* out_2n)
{
    SobolSampler sb;
    for (uint32_t i = 0; i < SOBOL_N; ++i) { out_2n[i * 2] = sb.pts[i][0]; out_2n[i * 2 + 1] = sb.pts[i][1]; }
}
uint32_t po_lk_hash(uint32_t x, uint32_t seed) { return lk_hash(x, seed); }
uint32_t po_low_bias_hash(uint32_t x) { return low_bias_hash(x); }
uint32_t po_scramble_base2(uint32_t x, uint32_t seed) { return scramble_base2(x, seed); }
void po_ss_sobol(uint64_t n, const uint32_t* index, const uint32_t* seed, float* out_xy)
{
    SobolSampler sb;
    for (uint64_t i = 0; i < n; ++i) sb.get_ss_sobol(index[i], seed[i], &out_xy[i * 2], &out_xy[i * 2 + 1]);
}
void po_camera_rays(const float* origin, const float* target, float fov_deg, float aspect, uint64_t n, const float* st, float* out_o, float* out_d)
{
    Camera c = camera_new({origin[0], origin[1], origin[2]}, {target[0], target[1], target[2]}, fov_deg, aspect);
    for (uint64_t i = 0; i < n; ++i)
    {
        Ray r = camera_create_ray(c, st[i * 2], st[i * 2 + 1]);
        out_o[i * 3] = r.o.x; out_o[i * 3 + 1] = r.o.y; out_o[i * 3 + 2] = r.o.z;
        out_d[i * 3] = r.d.x; out_d[i * 3 + 1] = r.d.y; out_d[i * 3 + 2] = r.d.z;
    }
}
// RNG known-answer probes: first `n` u32 draws of the stream (seed, pixel, sample)
void po_rng_stream(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t n, uint32_t* out)
{
    WyRand r = stream_for(seed, pixel, sample);
    for (uint32_t i = 0; i < n; ++i) out[i] = r.u32();
}

// ---- output stage: ImageHelper::write_image (image_helper.rs:37-58) on the normalised accumulation buffer (shader.wgsl:63) ----
// gt_tonemap restated from image_helper/tonemapping.rs:2-96 (scalar, per channel).
static float gt_smoothstep_(float x, float e0, float e1)  // tonemapping.rs:39-56
{
    if (x < e0) return 0.0f;
    if (x > e1) return 1.0f;
    const float a = (x - e0) / (e1 - e0), b = 3.0f - 2.0f * a;
    return a * a * b;
}
static float gt_lerp_(float x, float e0, float e1)  // tonemapping.rs:21-35
{
    if (x < e0) return 0.0f;
    if (x > e1) return 1.0f;
    return (x - e0) / (e1 - e0);
}
static float gt_tonemap_(float x, float p, float a, float m, float l, float c, float b)  // tonemapping.rs:68-96
{
    if (x < 0.0f) return b;
    const float l0 = (p - m) * l / a;
    const float w0 = 1.0f - gt_smoothstep_(x, 0.0f, m);
    const float w2 = gt_lerp_(x, m + l0, m + l0);
    const float w1 = 1.0f - w0 - w2;
    const float t = (m * std::pow(x / m, c) + b) * w0;                                   // gt_toe :5
    const float u = (m + a * (x - m)) * w1;                                              // gt_linear :2
    const float s0 = m + l0, s1 = m + a * l0, c2 = a * p / (p - s1);
    const float v = (p - (p - s1) * std::exp(-c2 * (x - s0) / p)) * w2;                  // gt_shoulder :8-16
    return t + u + v;
}
static uint8_t as_u8(float v) { return !(v > 0.0f) ? 0 : (v >= 255.0f ? 255 : (uint8_t)v); }  // Rust `as u8`
void po_tonemap_rgb8(const float* accum_rgba, uint32_t W, uint32_t H, int flip_y, uint8_t* rgb)
{
    for (uint32_t y = 0; y < H; ++y)
        for (uint32_t x = 0; x < W; ++x)
        {
            const float* a = accum_rgba + ((size_t)y * W + x) * 4;
            uint8_t* o = rgb + ((size_t)(flip_y ? H - 1 - y : y) * W + x) * 3;
            for (int k = 0; k < 3; ++k) o[k] = as_u8(std::pow(gt_tonemap_(a[k] / a[3], 1.0f, 1.0f, 0.22f, 0.4f, 1.33f, 0.0f), 1.0f / 2.2f) * 255.0f);
        }
}

// ---- State::update (state.rs:505-579): accumulate.wgsl / velocity.wgsl + compute.wgsl on host arrays -------------
// accumulation[W*H*4] in/out, velocity[W*H*2] out (moving camera only).  Returns 1 when the TAA path ran.
int po_state_update(uint32_t W, uint32_t H, const float* data, const float* position, const uint32_t* id, const float* inv_projection, const float* last_inv_projection,
                    float* accumulation, float* velocity)
{
    return state_update((int)W, (int)H, data, position, id, inv_projection, last_inv_projection, accumulation, velocity);
}
// (cam.matrix * cam.inv_projection).inverse() (main.rs:209-216) for Camera::new(origin, target, fov, aspect): evaluated in
// double precision (exact view-projection rounded once), used by tests to bound the f32 cofactor inverse of the product path.
void po_view_projection(const float* origin, const float* target, float fov_deg, float aspect, double* m16)
{
    double e[3] = {origin[0], origin[1], origin[2]}, f[3] = {target[0] - e[0], target[1] - e[1], target[2] - e[2]};
    double fl = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
    for (double& v : f) v /= fl;
    double s[3] = {f[1] * 0.0 - f[2] * 1.0, f[2] * 0.0 - f[0] * 0.0, f[0] * 1.0 - f[1] * 0.0};  // f x (0,1,0)
    double sl = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
    for (double& v : s) v /= sl;
    double u[3] = {s[1] * f[2] - s[2] * f[1], s[2] * f[0] - s[0] * f[2], s[0] * f[1] - s[1] * f[0]};
    // view (look_at_rh), column-major
    double V[16] = {s[0], u[0], -f[0], 0, s[1], u[1], -f[1], 0, s[2], u[2], -f[2], 0, -(s[0] * e[0] + s[1] * e[1] + s[2] * e[2]), -(u[0] * e[0] + u[1] * e[1] + u[2] * e[2]),
                    (f[0] * e[0] + f[1] * e[1] + f[2] * e[2]), 1};
    // perspective_infinite_rh(fov, aspect, z_near = 1): cols (g/aspect,0,0,0) (0,g,0,0) (0,0,-1,-1) (0,0,-z_near,0)
    double g = 1.0 / std::tan(0.5 * (double)fov_deg * 3.14159265358979323846 / 180.0);
    double P[16] = {g / aspect, 0, 0, 0, 0, g, 0, 0, 0, 0, -1, -1, 0, 0, -1, 0};
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r)
        {
            double a = 0;
            for (int k = 0; k < 4; ++k) a += P[k * 4 + r] * V[c * 4 + k];
            m16[c * 4 + r] = a;
        }
}

// ---- frame driver: main.rs:181-207 + accumulate.wgsl:20-23 ------------------------------------------------
// accum[W*H*4] += (rgb, 1) per sample; position[W*H*4], id[W*H] as main.rs:204-206 (may be null).
// counters[14] += extend, shadow, probe, paths, bounces, extend{boxes,tris,insts}, lights{boxes,tris,insts}, shadow{boxes,tris,insts}.
// Pixel-major scheduling (each pixel's samples in order on one thread): sums are bit-identical to running
// `sample_count` successive frames.  Returns elapsed seconds of the per-pixel loop.
double po_render(const po_scene* s, const float* cam_origin, const float* cam_target, float fov_deg, float aspect, uint32_t W, uint32_t H, uint32_t sample_begin,
                 uint32_t sample_count, uint64_t seed, uint32_t max_bounces, int enable_nee, int nthreads, float* accum, float* accum_sq, float* position, uint32_t* id,
                 uint64_t* counters)
{
    const Scene& sc = s->sc;
    Camera cam = camera_new({cam_origin[0], cam_origin[1], cam_origin[2]}, {cam_target[0], cam_target[1], cam_target[2]}, fov_deg, aspect);
    int nt = resolve_threads(nthreads);
    std::vector<RayCounters> rcs(nt);
    auto t0 = std::chrono::steady_clock::now();
    parallel_chunks((uint64_t)W * H, 256, nthreads, [&](uint64_t b, uint64_t e, int tid) {
        RayCounters& rc = rcs[tid];
        for (uint64_t i = b; i < e; ++i)
        {
            uint32_t x = (uint32_t)(i % W), y = (uint32_t)(i / W);
            for (uint32_t k = 0; k < sample_count; ++k)
            {
                uint32_t smp = sample_begin + k;
                WyRand rng = stream_for(seed, (uint32_t)i, smp);
                uint32_t jseed = rng.u32();                                     // main.rs:193
                float jx, jy;
                sc.sobol.get_ss_sobol(smp == 0 ? 0u : 1u, jseed, &jx, &jy);     // main.rs:194: index = c.w as u32 (0 on frame 0, then 1)
                jx -= 0.5f;
                jy -= 0.5f;
                float u = ((float)x + jx) / (float)W;
                float v = ((float)y + jy) / (float)H;
                Ray ray = camera_create_ray(cam, u, v);
                PathResult pr = integrate(ray, sc, rng, max_bounces, enable_nee != 0, rc);
                float* a = accum + i * 4;
                a[0] += pr.colour.x; a[1] += pr.colour.y; a[2] += pr.colour.z; a[3] += 1.0f;  // accumulate.wgsl:23
                if (accum_sq) { float* q = accum_sq + i * 4; q[0] += pr.colour.x * pr.colour.x; q[1] += pr.colour.y * pr.colour.y; q[2] += pr.colour.z * pr.colour.z; q[3] += 1.0f; }
                if (position) { float* p = position + i * 4; p[0] = pr.position.x; p[1] = pr.position.y; p[2] = pr.position.z; p[3] = pr.position.w; }
                if (id) id[i] = (id[i] << 16) | (uint32_t)pr.first_id;           // main.rs:206
            }
        }
    });
    double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (counters)
        for (auto& rc : rcs)
        {
            counters[0] += rc.extend; counters[1] += rc.shadow; counters[2] += rc.probe; counters[3] += rc.paths; counters[4] += rc.bounces;
            counters[5] += rc.world.boxes; counters[6] += rc.world.tris; counters[7] += rc.world.insts;
            counters[8] += rc.lights.boxes; counters[9] += rc.lights.tris; counters[10] += rc.lights.insts;
            counters[11] += rc.shadow_w.boxes; counters[12] += rc.shadow_w.tris; counters[13] += rc.shadow_w.insts;
        }
    return el;
}

// load_obj (blas.rs:44-131) on a text image.  Two calls: with out == nullptr it only counts; otherwise it fills ntri*18 floats.
// Returns 0, or 1 where the reference would panic (message in po_last_error).
int po_load_obj_text(const char* text, uint64_t len, float* out, uint64_t* ntri)
{
    try
    {
        const std::vector<ObjVertex> v = load_obj_text(std::string(text, (size_t)len));
        *ntri = v.size() / 3;
        if (out)
            for (size_t i = 0; i < v.size(); ++i)
            {
                out[i * 6 + 0] = v[i].position.x; out[i * 6 + 1] = v[i].position.y; out[i * 6 + 2] = v[i].position.z;
                out[i * 6 + 3] = v[i].normal.x; out[i * 6 + 4] = v[i].normal.y; out[i * 6 + 5] = v[i].normal.z;
            }
        return 0;
    }
    catch (ObjError& e)
    {
        g_err = e.msg;
        return 1;
    }
}

}  // extern "C"
