// Development tool (host only, never shipped): SIMT cost model of the persistent traversal loop (persist.cuh) with K rays per
// lane.  32 lanes x K ray slots per warp; every loop iteration runs the stages P, N, L, T once, each for at most one ray of a
// lane (the first of its rays that wants the stage).  A stage costs its instruction count when at least one lane of the warp runs
// it.  K = 1 is today's kernel; the model is calibrated against profiles/r2m_c3_extend_sass_blocks.txt.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>
#include "../tests/hostemu/hostemu.cpp"

namespace {

struct Slot
{
    WideTrav tr;
    uint2 stack[WIDE_STACK_CAP];
    int state = 0;  // 0 idle, 1 traversing, 2 finished (result not consumed)
    float4 pa, pb, pc;
    bool fresh = false;
};

bool wants_pop(const WideTrav& s) { return s.tri_i >= s.tri_end && s.lh == 0u && s.node == TRAV_NONE; }

// stage N of wide_step_work
void stage_n(WideTrav& s, const SceneDev& S, uint2* stack)
{
    const uint4* np = S.wide + (size_t)s.node * 5;
    const uint4 q0 = np[0], q1 = np[1], q2 = np[2], q3 = np[3], q4 = np[4];
    s.node = TRAV_NONE;
    const uint32_t hits = wide_node_test(q0, q2, q3, q4, s.ray, s.t_max);
    const uint32_t imask = q0.w >> 24, lmask = q1.z & 0xffu;
    const uint32_t perm = xor_permute16((hits & imask) | ((hits & lmask) << 8), s.ray.oi);
    const uint32_t ih = perm & 0xffu, lh = (perm >> 8) & 0xffu;
    s.G.x = q1.x;
    s.G.y = ih ? (ih | (imask << 8)) : 0u;
    if (s.cur_inst != TRAV_NONE)
    {
        s.lh = lh;
        s.lmask = lmask;
        s.leaf_base = q1.y;
    }
    else
    {
        if (s.G.y) { stack[s.sp++] = s.G; s.G.y = 0u; }
        uint32_t m = lh;
        while (m)
        {
            const int p = lo_bit(m);
            m &= m - 1u;
            const uint32_t sl = (uint32_t)p ^ s.ray.oi;
            uint2 e;
            e.x = q1.y + popc32(lmask & ((1u << sl) - 1u));
            e.y = 0x80000000u;
            stack[s.sp++] = e;
        }
    }
}
void stage_l(Slot& q, const SceneDev& S, bool any)
{
    WideTrav& s = q.tr;
    const int p = hi_bit(s.lh);
    s.lh &= ~(1u << p);
    const uint32_t sl = (uint32_t)p ^ s.ray.oi;
    const uint32_t hidx = s.leaf_base + popc32(s.lmask & ((1u << sl) - 1u));
    const float4* lp = S.leaf_headers + (size_t)hidx * LEAF_F4;
    const float4 h0 = lp[0], h1 = lp[1];
    q.pa = lp[2]; q.pb = lp[3]; q.pc = lp[4];
    const uint32_t cw = ex::bits(h0.w);
    float t_enter = 0.0f;
    bool go = true;
    if (any || !(cw & LEAF_ROOT_FLAG)) go = ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), s.ray.o, s.ray.inv, s.t_max, &t_enter);
    q.fresh = false;
    if (go)
    {
        s.tri_i = ex::bits(h1.w);
        s.tri_end = s.tri_i + (cw & ~LEAF_ROOT_FLAG);
        s.t_enter = t_enter;
        q.fresh = true;
    }
}
// returns false when an any-hit query is decided
bool stage_t(Slot& q, const SceneDev& S, bool any)
{
    WideTrav& s = q.tr;
    float4 a = q.pa, b = q.pb, c = q.pc;
    if (!q.fresh)
    {
        const float4* pp = S.tri_planes + (size_t)s.tri_i * 3;
        a = pp[0]; b = pp[1]; c = pp[2];
    }
    q.fresh = false;
    const float n0[4] = {a.x, a.y, a.z, a.w}, n1[4] = {b.x, b.y, b.z, b.w}, n2[4] = {c.x, c.y, c.z, c.w};
    if (any)
    {
        if (ex::tri_hit_bool(n0, n1, n2, s.ray.o, s.ray.d, s.t_max, s.t_enter))
        {
            s.hit_inst = s.cur_inst;
            return false;
        }
    }
    else
    {
        float t, raw[3];
        if (ex::tri_hit_raw(n0, n1, n2, s.ray.o, s.ray.d, s.t_max, s.t_enter, &t, raw))
        {
            s.t_max = t;
            s.hit_ud = raw[0]; s.hit_vd = raw[1]; s.hit_det = raw[2]; s.hit_slot = s.tri_i; s.hit_inst = s.cur_inst;
        }
    }
    ++s.tri_i;
    return true;
}

struct Warp
{
    std::vector<Slot> slots;  // [lane * K + k]
    uint64_t wbase = 0, wend = 0;
};

}  // namespace

extern "C" {

// out: [0] iterations, [1..2] N executions / lanes, [3..4] L, [5..6] T, [7..8] I (instance entry: slab test), [9..10] P (pop), [11..12] refill events / slots
// started, [13] sum over iterations of active rays, [14] checksum of hits, [15..16] finish events / slots
void mrsim_run(void* h, int which, int any, uint64_t n, const float* o, const float* d, const float* tmax, int K, int n_warps, int refill_below, int node_quorum,
               int policy, int q_leaf, int q_inst, uint64_t* out, uint8_t* hit_out, float* t_out)
{
    const HostScene& sc = *(HostScene*)h;
    const SceneDev S = host_view(sc);
    std::vector<Warp> warps(n_warps);
    for (auto& w : warps) w.slots.resize(32 * K);
    uint64_t cursor = 0;
    const uint64_t BATCH = 64;
    for (int i = 0; i < 17; ++i) out[i] = 0;
    std::vector<uint32_t> ray_of(n_warps * 32 * K, 0);
    bool any_alive = true;
    auto take = [&](Warp& w) -> bool {
        if (w.wbase < w.wend) return true;
        if (cursor >= n) return false;
        w.wbase = cursor;
        w.wend = std::min(n, cursor + BATCH);
        cursor = w.wend;
        return true;
    };
    while (any_alive)
    {
        any_alive = false;
        for (int wi = 0; wi < n_warps; ++wi)
        {
            Warp& w = warps[wi];
            uint32_t* rid = &ray_of[(size_t)wi * 32 * K];
            // ---- refill: count busy rays; finish + start when few are left (K = 1: busy lanes < refill_below; K > 1: free slots >= 32 - refill_below + ...)
            int busy = 0, fin = 0, idle = 0;
            for (auto& q : w.slots) { busy += q.state == 1; fin += q.state == 2; idle += q.state == 0; }
            const bool more = w.wbase < w.wend || cursor < n;
            bool do_refill;
            if (K == 1) do_refill = busy < (more ? refill_below : 1);
            else do_refill = more ? (fin + idle) >= (32 - refill_below) : busy == 0;  // a full warp's worth of free slots (less the threshold)
            if (do_refill && (fin || (idle && more)))
            {
                // finish: each lane consumes at most one finished ray per event; start: each lane starts at most one ray per event
                int nf = 0, ns = 0;
                for (int lane = 0; lane < 32; ++lane)
                {
                    for (int k = 0; k < K; ++k)
                    {
                        Slot& q = w.slots[lane * K + k];
                        if (q.state == 2)
                        {
                            const uint32_t r = rid[lane * K + k];
                            const bool f = q.tr.hit_inst != TRAV_NONE;
                            hit_out[r] = f;
                            t_out[r] = f && !any ? q.tr.t_max : 0.0f;
                            out[14] += f ? (uint64_t)(r + 1) : 0;
                            q.state = 0;
                            ++nf;
                            break;
                        }
                    }
                }
                for (int lane = 0; lane < 32; ++lane)
                {
                    for (int k = 0; k < K; ++k)
                    {
                        Slot& q = w.slots[lane * K + k];
                        if (q.state == 0 && take(w))
                        {
                            const uint64_t r = w.wbase++;
                            rid[lane * K + k] = (uint32_t)r;
                            wide_begin(q.tr, S, which, ex::v3(o[r * 3], o[r * 3 + 1], o[r * 3 + 2]), ex::v3(d[r * 3], d[r * 3 + 1], d[r * 3 + 2]), tmax[r], any != 0);
                            q.state = 1;
                            q.fresh = false;
                            ++ns;
                            break;
                        }
                    }
                }
                if (nf) { out[15]++; out[16] += nf; }
                if (ns) { out[11]++; out[12] += ns; }
            }
            busy = 0;
            for (auto& q : w.slots) busy += q.state == 1;
            int left = 0;
            for (auto& q : w.slots) left += q.state != 0;
            if (left || more) any_alive = true;
            if (!busy) continue;
            // ---- one loop iteration ---------------------------------------------------------------------------------------
            out[0]++;
            out[13] += busy;
            int pl = 0, il = 0, nl = 0, ll = 0, tl = 0;
            // P (pop / descend; an instance entry on top of the stack is stage I, gated separately with policy 2)
            auto wants_inst = [&](const Slot& q) { return wants_pop(q.tr) && q.tr.G.y == 0u && q.tr.sp > 0 && (q.stack[q.tr.sp - 1].y & 0x80000000u); };
            for (int lane = 0; lane < 32; ++lane)
                for (int k = 0; k < K; ++k)
                {
                    Slot& q = w.slots[lane * K + k];
                    if (q.state != 1 || !wants_pop(q.tr) || wants_inst(q)) continue;
                    if (!wide_step_pop(q.tr, S, q.stack)) q.state = 2;
                    ++pl;
                    break;
                }
            // vote
            int wn = 0, wl = 0, wq = 0;
            for (int lane = 0; lane < 32; ++lane)
            {
                bool a = false, b = false, c = false;
                for (int k = 0; k < K; ++k)
                {
                    const Slot& q = w.slots[lane * K + k];
                    if (q.state != 1) continue;
                    a |= wide_wants_node(q.tr);
                    b |= wide_has_leaf_work(q.tr);
                    c |= wants_inst(q);
                }
                wn += a; wl += b; wq += c;
            }
            bool allow = wn >= node_quorum || wl == 0;
            bool allow_l = true, allow_i = true;
            if (policy == 1) allow = wn > 0 && (wn >= node_quorum || wn >= wl);  // run N unless leaf work clearly dominates
            if (policy == 2)
            {
                allow = wn >= node_quorum; allow_l = wl >= q_leaf; allow_i = wq >= q_inst;
                if (!(allow && wn) && !(allow_l && wl) && !(allow_i && wq)) allow = allow_l = allow_i = true;
            }
            // I
            if (allow_i)
                for (int lane = 0; lane < 32; ++lane)
                    for (int k = 0; k < K; ++k)
                    {
                        Slot& q = w.slots[lane * K + k];
                        if (q.state != 1 || !wants_inst(q)) continue;
                        if (!wide_step_pop(q.tr, S, q.stack)) q.state = 2;
                        ++il;
                        break;
                    }
            // N
            if (allow)
                for (int lane = 0; lane < 32; ++lane)
                    for (int k = 0; k < K; ++k)
                    {
                        Slot& q = w.slots[lane * K + k];
                        if (q.state != 1 || !wide_wants_node(q.tr)) continue;
                        stage_n(q.tr, S, q.stack);
                        ++nl;
                        break;
                    }
            // L then T (T on the ray whose leaf was just opened, else on any ray with a pending triangle)
            for (int lane = 0; lane < 32 && allow_l; ++lane)
            {
                int kl = -1;
                for (int k = 0; k < K; ++k)
                {
                    Slot& q = w.slots[lane * K + k];
                    if (q.state == 1 && q.tr.tri_i >= q.tr.tri_end && q.tr.lh) { kl = k; break; }
                }
                int kt = -1;
                if (kl >= 0)
                {
                    Slot& q = w.slots[lane * K + kl];
                    stage_l(q, S, any != 0);
                    ++ll;
                    if (q.fresh) kt = kl;
                }
                if (kt < 0)
                    for (int k = 0; k < K; ++k)
                    {
                        Slot& q = w.slots[lane * K + k];
                        if (q.state == 1 && q.tr.tri_i < q.tr.tri_end) { kt = k; break; }
                    }
                if (kt >= 0)
                {
                    Slot& q = w.slots[lane * K + kt];
                    if (!stage_t(q, S, any != 0)) q.state = 2;
                    ++tl;
                }
            }
            if (pl) { out[9]++; out[10] += pl; }
            if (il) { out[7]++; out[8] += il; }
            if (nl) { out[1]++; out[2] += nl; }
            if (ll) { out[3]++; out[4] += ll; }
            if (tl) { out[5]++; out[6] += tl; }
        }
    }
}
}
