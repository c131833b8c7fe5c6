// tests/hostemu — TEST-ONLY harness: compiles the product's __host__ __device__ traversal code (csrc/traverse.h)
// for the CPU so its logic can be checked against the oracle in the GPU-less container.  It is never linked into
// libptb.so and no C-ABI entry point routes here: the product has no CPU path (ptb_* compute calls return
// PTB_ERR_NO_DEVICE without a GPU).
#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../path_tracer_b200/csrc/camera_host.h"
#include "../../path_tracer_b200/csrc/host_scene.h"
#include "../../path_tracer_b200/csrc/traverse.h"
#include "../../path_tracer_b200/csrc/multiray.cuh"
#include "../../path_tracer_b200/csrc/temporal.cuh"
#include "../../path_tracer_b200/csrc/wavefront.h"

using namespace ptb;

extern "C" {

void* emu_create(const float* verts, uint64_t ntri, const ptb_blas_desc* blas, uint32_t nblas, const ptb_instance* inst, uint32_t ninst, int threads, char* err, int errlen)
{
    auto* sc = new HostScene();
    std::string e;
    int rc = sc->upload(verts, ntri, blas, nblas, inst, ninst, e);
    if (rc == PTB_OK) rc = sc->build(threads, e);
    if (rc != PTB_OK)
    {
        if (err && errlen > 0) { std::strncpy(err, e.c_str(), errlen - 1); err[errlen - 1] = 0; }
        delete sc;
        return nullptr;
    }
    return sc;
}
void emu_destroy(void* h) { delete (HostScene*)h; }

void emu_trace(void* h, int which, int kernel, int any, uint64_t n, const float* o, const float* d, const float* tmax, uint8_t* hit, float* t, float* u, float* v,
               uint32_t* prim, uint32_t* blas, uint32_t* inst, uint8_t* front, float* normal, int threads)
{
    const HostScene& sc = *(HostScene*)h;
    const SceneDev S = host_view(sc);
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint64_t> next{0};
    auto work = [&] {
        for (;;)
        {
            uint64_t b = next.fetch_add(1024);
            if (b >= n) break;
            uint64_t e = std::min(n, b + 1024);
            for (uint64_t i = b; i < e; ++i)
            {
                V3 ro = ex::v3(o[i * 3], o[i * 3 + 1], o[i * 3 + 2]), rd = ex::v3(d[i * 3], d[i * 3 + 1], d[i * 3 + 2]);
                Hit hr{};
                bool f;
                if (any) f = kernel ? bin_traverse<true>(S, which, ro, rd, tmax[i], &hr) : wide_traverse<true>(S, which, ro, rd, tmax[i], &hr);
                else f = kernel ? bin_traverse<false>(S, which, ro, rd, tmax[i], &hr) : wide_traverse<false>(S, which, ro, rd, tmax[i], &hr);
                hit[i] = f;
                if (any) continue;
                uint32_t p = 0xffffffffu, ib = 0xffffffffu, ii = 0xffffffffu;
                V3 nn = ex::v3(0, 0, 0);
                bool ff = false;
                if (f)
                {
                    p = kernel ? hr.slot_or_prim : S.tri_prim[hr.slot_or_prim];
                    ib = S.inst[hr.inst].blas;
                    ii = hr.inst - S.tlas[which].inst_base;
                    finish_hit(S, rd, hr.inst, p, hr.u, hr.v, &nn, &ff);
                }
                if (t) t[i] = f ? hr.t : ex::inf();
                if (u) u[i] = f ? hr.u : 0.0f;
                if (v) v[i] = f ? hr.v : 0.0f;
                if (prim) prim[i] = p;
                if (blas) blas[i] = ib;
                if (inst) inst[i] = ii;
                if (front) front[i] = ff;
                if (normal) { normal[i * 3] = nn.x; normal[i * 3 + 1] = nn.y; normal[i * 3 + 2] = nn.z; }
            }
        }
    };
    std::vector<std::thread> th;
    for (int k = 0; k < threads; ++k) th.emplace_back(work);
    for (auto& x : th) x.join();
}

// canonical dump, same shapes as ptb_bvh_* (lets the CPU tests check the builder before the C ABI exists on a box)
void emu_blas_counts(void* h, uint32_t b, uint32_t* nc, uint32_t* lc, uint32_t* root)
{
    const HostBlas& B = ((HostScene*)h)->blas[b];
    *nc = (uint32_t)B.nodes.size();
    *lc = (uint32_t)B.leaf_prims.size();
    *root = B.root;
}
void emu_stats(void* h, uint64_t* out8)
{
    const HostScene& sc = *(HostScene*)h;
    out8[0] = sc.wide.size();
    out8[1] = sc.leaf_headers.size();
    out8[2] = sc.tri_prim.size();
    out8[3] = sc.wide.size() * 80 + sc.leaf_headers.size() * sizeof(LeafHeader) + sc.tri_prim.size() * 52;
    out8[4] = 0;
    for (auto& B : sc.blas) out8[4] = std::max<uint64_t>(out8[4], B.wide_depth);
    out8[4] += sc.tlas[0].wide_depth;
    out8[5] = sc.bin.size();
    out8[6] = sc.inst_rows.size();
    out8[7] = sc.wide_stack_bound;
}

// Sequential emulation of the wavefront loop of render.cu using the very same per-slot bodies (wavefront.h).
// cam16: ray matrix (column-major) as ptb_set_camera computes it is NOT available here (render.cu is device code), so the
// caller passes the camera matrix + eye obtained from the oracle-independent helper emu_camera below.
void emu_render(void* h, const float* cam_m16, const float* cam_eye3, uint32_t W, uint32_t H, uint32_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_bounces,
                int enable_nee, int extend_kernel, float* accum, float* position, uint32_t* id, uint64_t* counters)
{
    const HostScene& sc = *(HostScene*)h;
    const uint32_t npix = W * H;
    RenderView V{};
    V.S = host_view(sc);
    std::memcpy(V.cam.m, cam_m16, 64);
    std::memcpy(V.cam.eye, cam_eye3, 12);
    V.W = W; V.H = H; V.npix = npix; V.K = 1; V.sample_begin = sample_begin; V.sample_count = sample_count; V.seed = seed; V.max_bounces = max_bounces;
    V.enable_nee = enable_nee ? 1u : 0u;
    V.extend_kernel = extend_kernel ? 1u : 0u;
    std::vector<RayRec> ray(npix);
    std::vector<PathRec> path(npix);
    std::vector<HitRec> hitrec(npix);
    std::vector<ShadowRec> shrec(npix);
    std::vector<ProbeRec> prrec(npix);
    std::vector<unsigned long long> volset(npix);
    V.accum = (float4*)accum; V.accum_sq = nullptr; V.position = (float4*)position; V.id = id;
    V.ray = ray.data(); V.path = path.data(); V.hit = hitrec.data(); V.shadow = shrec.data(); V.probe = prrec.data();
    V.volset = sc.volumes.empty() ? nullptr : volset.data();
    std::vector<uint32_t> ext, next, term, nee;
    for (uint32_t i = 0; i < npix; ++i) { ray[i].o.w = 0.0f; ray[i].d.w = 0.0f; term.push_back(i); }
    uint64_t n_extend = 0, n_shadow = 0, n_probe = 0, n_paths = 0, n_iters = 0;
    for (;;)
    {
        for (uint32_t s : term) if (regen_slot(V, s) & WF_NEXT) { next.push_back(s); ++n_paths; }
        term.clear();
        ext.swap(next);
        next.clear();
        if (ext.empty()) break;
        ++n_iters;
        // material-sorted shade order, as the device queues: kinds 0..6
        std::vector<uint32_t> mat[7];
        for (uint32_t s : ext) mat[extend_slot(V, s)].push_back(s);
        n_extend += ext.size();
        for (int k = 0; k < 7; ++k)
            for (uint32_t s : mat[k])
            {
                const uint32_t f = shade_slot(V, s);
                n_probe += (f & WF_PROBE_OUT) ? 1 : 0;
                if (f & (WF_SHADOW | WF_PROBE)) nee.push_back(s | ((f & WF_SHADOW) ? NEE_SHADOW_BIT : 0u) | ((f & WF_PROBE) ? NEE_PROBE_BIT : 0u));
                if (f & WF_NEXT) next.push_back(s);
                if (f & WF_TERM) term.push_back(s);
            }
        for (uint32_t e : nee)
        {
            const uint32_t f = ((e & NEE_SHADOW_BIT) ? WF_SHADOW : 0u) | ((e & NEE_PROBE_BIT) ? WF_PROBE : 0u);
            n_shadow += (f & WF_SHADOW) ? 1 : 0;
            n_probe += (f & WF_PROBE) ? 1 : 0;
            n_shadow += nee_slot(V, e & NEE_SLOT_MASK, f);
        }
        nee.clear();
    }
    if (counters) { counters[0] += n_extend; counters[1] += n_shadow; counters[2] += n_probe; counters[3] += n_paths; counters[4] += n_iters; }
}

void emu_ss_sobol(uint64_t n, const uint32_t* index, const uint32_t* seed, float* xy)
{
    for (uint64_t i = 0; i < n; ++i) ss_sobol(index[i], seed[i], &xy[i * 2], &xy[i * 2 + 1]);
}
void emu_rng_stream(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t n, uint32_t* out)
{
    Rng r = rng_stream(seed, pixel, sample);
    for (uint32_t i = 0; i < n; ++i) out[i] = rng_u32(r);
}
void emu_camera_rays(const float* cam_m16, const float* cam_eye3, uint64_t n, const float* st, float* o3, float* d3)
{
    CameraDev c;
    std::memcpy(c.m, cam_m16, 64);
    std::memcpy(c.eye, cam_eye3, 12);
    for (uint64_t i = 0; i < n; ++i)
    {
        V3 o, d;
        camera_ray(c, st[i * 2], st[i * 2 + 1], &o, &d);
        o3[i * 3] = o.x; o3[i * 3 + 1] = o.y; o3[i * 3 + 2] = o.z;
        d3[i * 3] = d.x; d3[i * 3 + 1] = d.y; d3[i * 3 + 2] = d.z;
    }
}
// the per-texel bodies of k_resolve / k_accumulate (temporal.cuh) looped over the image; history in, out separate
void emu_state_update(int W, int H, const float* data, const float* position, const uint32_t* id, const float* last16, int moved, const float* history, float* out, float* velocity)
{
    const tp::Image in{reinterpret_cast<const tp::F4*>(data), W, H, 0}, hist{reinterpret_cast<const tp::F4*>(history), W, H, 0};
    const tp::Image pos{reinterpret_cast<const tp::F4*>(position), W, H, 0};
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x)
        {
            const size_t i = (size_t)y * W + x;
            if (!moved)
            {
                for (int k = 0; k < 3; ++k) out[i * 4 + k] = history[i * 4 + k] + data[i * 4 + k];
                out[i * 4 + 3] = history[i * 4 + 3] + 1.0f;
                continue;
            }
            const tp::F2 v = tp::velocity_at(pos, last16, tp::F2{(float)W, (float)H}, x, y);
            const tp::F4 r = tp::resolve_at(in, hist, pos, id, last16, x, y);
            velocity[i * 2] = v.x; velocity[i * 2 + 1] = v.y;
            out[i * 4] = r.x; out[i * 4 + 1] = r.y; out[i * 4 + 2] = r.z; out[i * 4 + 3] = r.w;
        }
}
void emu_camera(const float* origin, const float* target, float fov_deg, float aspect, float* m16, float* eye3) { camera_setup(origin, target, fov_deg, aspect, m16, eye3); }
}

// K rays per lane with stage gating (multiray.cuh): one emulated warp of 32 lanes walks the batch with the loop structure of
// mr_persistent_trace — the stage functions are the product's own, the ballots are loops over the lanes.
template <int K>
static void emu_mr_warp(const SceneDev& S, int which, int any, uint64_t n, const float* o, const float* d, const float* tmax, int free_quorum, int qn, int ql, int qi,
                        uint8_t* hit, float* t, float* u, float* v, uint32_t* slot, uint32_t* inst, uint64_t* stats)
{
    constexpr int NW = MR_JOB + 1;
    using R = MrRay<1>;
    std::vector<uint32_t> mem((size_t)32 * K * NW, 0u);
    std::vector<uint2> stk((size_t)32 * K * WIDE_STACK_CAP);
    uint32_t tags[32] = {};
    auto ray = [&](int lane, int k) { return R{mem.data() + ((size_t)lane * K + k) * NW}; };
    auto stack = [&](int lane, int k) { return stk.data() + ((size_t)lane * K + k) * WIDE_STACK_CAP; };
    uint64_t next = 0;
    struct WorldRay
    {
        const float *o, *d;
        void world_ray(const R& r, V3* oo, V3* dd) const
        {
            const uint64_t i = r.ld(MR_JOB);
            *oo = ex::v3(o[i * 3], o[i * 3 + 1], o[i * 3 + 2]);
            *dd = ex::v3(d[i * 3], d[i * 3 + 1], d[i * 3 + 2]);
        }
    } job{o, d};
    for (;;)
    {
        int n_done = 0, n_idle = 0;
        for (int l = 0; l < 32; ++l) { n_done += mr_count2<K>(tags[l], MT_DONE, MT_DONE); n_idle += mr_count2<K>(tags[l], MT_IDLE, MT_IDLE); }
        const int n_act = 32 * K - n_done - n_idle;
        const bool more = next < n;
        if (more ? (n_done + n_idle >= free_quorum) : (n_done > 0 && (n_done >= 8 || n_act == 0)))
        {
            for (int l = 0; l < 32; ++l)
            {
                const int k = mr_first<K>(tags[l], MT_DONE);
                if (k < 0) continue;
                const R r = ray(l, k);
                const uint64_t i = r.ld(MR_JOB);
                const bool f = mr_hit(r);
                hit[i] = f;
                if (!any)
                {
                    t[i] = f ? r.ldf(MR_TMAX) : ex::inf();
                    u[i] = f ? mr_hit_u(r) : 0.0f;
                    v[i] = f ? mr_hit_v(r) : 0.0f;
                    slot[i] = f ? r.ld(MR_HSLOT) : 0xffffffffu;
                    inst[i] = f ? r.ld(MR_HINST) : 0xffffffffu;
                }
                tags[l] = mr_set(tags[l], k, MT_IDLE);
            }
            for (int pass = 0; pass < K && next < n; ++pass)
                for (int l = 0; l < 32 && next < n; ++l)
                {
                    const int k = mr_first<K>(tags[l], MT_IDLE);
                    if (k < 0) continue;
                    const uint64_t i = next++;
                    const R r = ray(l, k);
                    r.st(MR_JOB, (uint32_t)i);
                    tags[l] = mr_set(tags[l], k, mr_begin(r, S, which, ex::v3(o[i * 3], o[i * 3 + 1], o[i * 3 + 2]), ex::v3(d[i * 3], d[i * 3 + 1], d[i * 3 + 2]), tmax[i], any != 0));
                }
        }
        bool live = false;
        for (int l = 0; l < 32; ++l) live |= tags[l] != 0u;
        if (!live) { if (next >= n) break; continue; }
        if (stats) stats[0]++;
        for (int l = 0; l < 32; ++l)
        {
            const int k = mr_first<K>(tags[l], MT_P);
            if (k >= 0) tags[l] = mr_set(tags[l], k, mr_stage_p(ray(l, k), stack(l, k), job));
        }
        int cn = 0, cl = 0, ci = 0;
        for (int l = 0; l < 32; ++l) { cn += mr_has<K>(tags[l], MT_N); cl += mr_has<K>(tags[l], MT_L); ci += mr_has<K>(tags[l], MT_I); }
        bool run_n = cn >= qn, run_l = cl >= ql, run_i = ci >= qi;
        if (!(run_n && cn) && !(run_l && cl) && !(run_i && ci)) run_n = run_l = run_i = true;
        if (stats) { stats[1] += run_n && cn; stats[2] += run_n ? cn : 0; stats[3] += run_l && cl; stats[4] += run_l ? cl : 0; stats[5] += run_i && ci; stats[6] += run_i ? ci : 0; }
        for (int l = 0; l < 32; ++l)
        {
            if (run_i)
            {
                const int k = mr_first<K>(tags[l], MT_I);
                if (k >= 0) tags[l] = mr_set(tags[l], k, mr_stage_i(ray(l, k), S));
            }
            if (run_n)
            {
                const int k = mr_first<K>(tags[l], MT_N);
                if (k >= 0) tags[l] = mr_set(tags[l], k, mr_stage_n(ray(l, k), S, stack(l, k)));
            }
            if (run_l)
            {
                const int k = mr_first<K>(tags[l], MT_L);
                if (k >= 0) tags[l] = mr_set(tags[l], k, mr_stage_l<2>(ray(l, k), S));
            }
        }
    }
}

extern "C" void emu_trace_mr(void* h, int which, int any, uint64_t n, const float* o, const float* d, const float* tmax, int K, int free_quorum, int qn, int ql, int qi,
                             uint8_t* hit, float* t, float* u, float* v, uint32_t* prim, uint32_t* inst, uint64_t* stats)
{
    const HostScene& sc = *(HostScene*)h;
    const SceneDev S = host_view(sc);
    std::vector<uint32_t> slot(n), gi(n);
    if (K == 1) emu_mr_warp<1>(S, which, any, n, o, d, tmax, free_quorum, qn, ql, qi, hit, t, u, v, slot.data(), gi.data(), stats);
    else if (K == 2) emu_mr_warp<2>(S, which, any, n, o, d, tmax, free_quorum, qn, ql, qi, hit, t, u, v, slot.data(), gi.data(), stats);
    else if (K == 3) emu_mr_warp<3>(S, which, any, n, o, d, tmax, free_quorum, qn, ql, qi, hit, t, u, v, slot.data(), gi.data(), stats);
    else emu_mr_warp<4>(S, which, any, n, o, d, tmax, free_quorum, qn, ql, qi, hit, t, u, v, slot.data(), gi.data(), stats);
    if (any) return;
    for (uint64_t i = 0; i < n; ++i)
    {
        prim[i] = hit[i] ? S.tri_prim[slot[i]] : 0xffffffffu;
        inst[i] = hit[i] ? gi[i] - S.tlas[which].inst_base : 0xffffffffu;
    }
}

// candidate lists of camera-ray tiles (tile_list_build, wavefront.h) for a W x H frame: lists = n_tiles * TILE_LIST_STRIDE uint2;
// also the slot -> pixel map of the tiled layout for n slots of a layer
extern "C" void emu_tile_lists(void* h, const float* cam_m16, const float* cam_eye3, uint32_t W, uint32_t H, uint32_t* lists2, uint32_t* slot_to_pixel)
{
    const HostScene& sc = *(HostScene*)h;
    RenderView V{};
    V.S = host_view(sc);
    std::memcpy(V.cam.m, cam_m16, 64);
    std::memcpy(V.cam.eye, cam_eye3, 12);
    V.W = W; V.H = H; V.npix = W * H; V.K = 1;
    V.tiles_x = W / 8u;
    V.tile_nodes = 1u;
    const uint32_t n_tiles = V.tiles_x * (H / 4u);
    for (uint32_t t = 0; t < n_tiles; ++t) tile_list_build(V, t, reinterpret_cast<uint2*>(lists2) + (size_t)t * TILE_LIST_STRIDE);
    for (uint32_t j = 0; j < V.npix; ++j) slot_to_pixel[j] = slot_pixel(V, j);
}
// the TLAS leaf (leaf-header index) of an instance row: what a tile list names
extern "C" void emu_instance_leaf_headers(void* h, uint32_t* out, uint32_t n_inst)
{
    const HostScene& sc = *(HostScene*)h;
    for (uint32_t i = 0; i < n_inst; ++i) out[i] = 0xffffffffu;
    const uint32_t ib = sc.tlas[0].inst_base;
    for (size_t k = 0; k < sc.leaf_headers.size(); ++k)
    {
        // TLAS-level headers carry the instance row in `first` and the inv_matrix in the payload; a BLAS-level header's `first` is a triangle
        // slot, so identify TLAS leaves by walking the wide TLAS instead
        (void)k;
    }
    const SceneDev S = host_view(sc);
    std::vector<uint32_t> st{S.tlas[0].wide_root};
    while (!st.empty())
    {
        const uint32_t node = st.back(); st.pop_back();
        const uint4* np = S.wide + (size_t)node * 5;
        const uint32_t imask = np[0].w >> 24, lmask = np[1].z & 0xffu;
        for (uint32_t s8 = 0; s8 < 8; ++s8)
        {
            if ((imask >> s8) & 1u) st.push_back(np[1].x + (uint32_t)__builtin_popcount(imask & ((1u << s8) - 1u)));
            if ((lmask >> s8) & 1u)
            {
                const uint32_t hidx = np[1].y + (uint32_t)__builtin_popcount(lmask & ((1u << s8) - 1u));
                const float4* lp = S.leaf_headers + (size_t)hidx * LEAF_F4;
                const uint32_t row = ex::bits(lp[1].w);
                if (row >= ib && row - ib < n_inst) out[row - ib] = hidx;
            }
        }
    }
}

// closest hit of camera rays that start from their tile's list (wide_begin_list + wide_step_pop<true>), ray i belonging to tile tile_of[i]
extern "C" void emu_trace_list(void* h, const uint32_t* lists2, uint64_t n, const uint32_t* tile_of, const float* o, const float* d, uint8_t* hit, float* t, float* u, float* v,
                               uint32_t* prim, uint32_t* inst)
{
    const HostScene& sc = *(HostScene*)h;
    SceneDev S = host_view(sc);
    S.tile_lists = reinterpret_cast<const uint2*>(lists2);
    for (uint64_t i = 0; i < n; ++i)
    {
        const V3 ro = ex::v3(o[i * 3], o[i * 3 + 1], o[i * 3 + 2]), rd = ex::v3(d[i * 3], d[i * 3 + 1], d[i * 3 + 2]);
        uint2 stack[WIDE_STACK_CAP];
        WideTrav s;
        const uint32_t first = tile_of[i] * TILE_LIST_STRIDE, cnt = S.tile_lists[first].x;
        if (cnt == TILE_LIST_NONE) wide_begin(s, S, 0, ro, rd, ex::inf(), false);
        else wide_begin_list(s, S, ro, rd, ex::inf(), first + 1u, first + 1u + cnt);
        while (wide_step_pop<true>(s, S, stack) && wide_step_work<0>(s, S, stack, true)) {}
        const bool f = s.hit_inst != TRAV_NONE;
        hit[i] = f;
        t[i] = f ? s.t_max : ex::inf();
        u[i] = f ? wide_hit_u(s) : 0.0f;
        v[i] = f ? wide_hit_v(s) : 0.0f;
        prim[i] = f ? S.tri_prim[s.hit_slot] : 0xffffffffu;
        inst[i] = f ? s.hit_inst - S.tlas[0].inst_base : 0xffffffffu;
    }
}
