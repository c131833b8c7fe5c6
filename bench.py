#!/usr/bin/env python
"""bench.py — headline benchmark of the render hot path (BASELINE.json: Mrays/s, primary + shadow, 1080p).

One "step" = one full render of the workload configuration by the wavefront kernels through the C ABI (libptb.so).  The
default workload is BASELINE config 2, the scene the north-star target is stated on: the procedural 1 M-triangle instanced
scene (1026 instances under a TLAS), 1920x1080, 64 spp.  `value` is whole-job Mrays/s (world-TLAS closest-hit + any-hit rays of
all ranks / max-over-ranks time) with the scene, camera and accumulation buffer resident in HBM; `e2e` times the same step as a
caller of the reference's seam sees it: camera upload, render, and the read-back of the three per-frame arrays of
main.rs:209-215 (accumulated colour, position, id) into pinned host memory.

N > 1 (torchrun, one process per GPU): every rank joins a ptb_group (libptb.so's own NCCL communicator; torch.distributed only
ships the 128-byte id and the statistics), renders the full spp of a disjoint global sample range (weak scaling) and the group
reduces the accumulation buffers onto rank 0 with one ncclReduce per step, inside the timed step.  A `strong` sub-record times
BASELINE config 4's scene (the same scene at 4K) with a FIXED number of samples per step split over the ranks.

At N = 1 the same JSON line carries `workloads` (C2 and C4, each with its own value / e2e / roofline), `pipelined` (two
contexts rendering alternate frames) and `strong`.

`--impl reference`: the reference cannot be built here (Rust nightly + un-vendored crates, no toolchain), so this arm times the
CPU oracle's restatement of main.rs:181-207 ("kind": "port") compiled on this host with the reference's release settings, on
num_cpus - 1 threads as main.rs:72, over a bounded sample (a few spp) of the same frame.  Both arms print the same `config`.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
OUT = sys.stdout  # main() replaces it by a private copy of the original stdout
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c2": dict(desc="synthetic Cornell box (36 tris, diffuse + area light) 1920x1080, 256 spp, max depth 8"),
    "c3": dict(desc="procedural 1M-triangle instanced scene (1K BLAS instances under a TLAS) 1920x1080, 64 spp, mixed diffuse/metal"),
    "c4": dict(desc="procedural 10M-triangle displaced terrain + dielectric/glossy, 3840x2160, 16 spp, max depth 16 + RR"),
    # BASELINE config 4: the frame's 4096 samples are split over the ranks (strong scaling); opt-in, a step is ~60 s / N on B200
    "c5": dict(desc="4K 4096-spp sample-partitioned render of the 1M-triangle instanced scene, NCCL reduce of the accumulation buffers", strong=True),
}
STRONG_SPP = 512  # `strong` sub-record: samples per step of the 4K frame, split over the ranks


def get_scene(name):
    from path_tracer_b200 import scenes
    if name == "c5":
        return scenes.scene_c3(width=3840, height=2160, spp=4096)
    return {"c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()


def make_config(name, sd, spp, world, strong, ref_spp):
    """The `config` object — IDENTICAL in both arms (ours and --impl reference) for the same command line."""
    W, H = sd.width, sd.height
    return dict(workload=WORKLOADS[name]["desc"], width=W, height=H, spp=spp, max_bounces=sd.max_bounces, nee=True, n_gpus=world,
                scaling="strong: the step's spp are split over the GPUs" if strong else "weak: every GPU renders the full spp of its own sample range",
                partition="disjoint global sample ranges per GPU, one ncclReduce of the accumulation buffer per step (ptb_group_*)" if world > 1 else "single GPU",
                l2="no flush needed: the per-step working set (path state: 16 slots/pixel x (196 B of records + 52 B of queue and flag entries) = %.0f MB) exceeds L2 (126 MB)" % (W * H * 16 * 248 / 1e6),
                cpu_arm="Mrays/s is a rate: the CPU arms (reference arm, cpu_baseline) render %d spp of this same frame per step on num_cpus - 1 threads (main.rs:72), "
                        "not the full spp; rays are counted by the same rule on both sides" % ref_spp)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None, reasons=reasons,
                    samples=len(self.rows))


def alg_bytes_per_ray(counters):
    """bytes the *reference algorithm* touches per ray, from the oracle's work counters."""
    ne, ns = max(counters["extend"], 1), max(counters["shadow"], 1)
    ext = 32 + (32 * counters["extend_boxes"] + 48 * counters["extend_tris"] + 48 * counters["extend_insts"]) / ne + 48 + 32
    sha = 32 + (32 * counters["shadow_boxes"] + 48 * counters["shadow_tris"] + 48 * counters["shadow_insts"]) / ns + 4
    return ext, sha


def cpu_threads():
    return max(1, (os.cpu_count() or 1) - 1)  # rayon pool of the reference: num_cpus::get() - 1 (main.rs:72)


def cpu_arm(sd, spp, seed, steps, warmup, want_counters=False):
    """The CPU side of the comparison: the oracle's restatement of the frame loop (main.rs:181-207) on this host's cores."""
    from oracle import po  # the checker — executed here only as the timed CPU baseline / for the algorithmic-bytes counters
    flags = po.use_native_build()
    sc = po.OracleScene(sd)
    threads = cpu_threads()
    times, rays, last = [], 0, None
    for s in range(warmup + steps):
        r = sc.render(sample_begin=s * spp, sample_count=spp, seed=seed, threads=threads)
        if s >= warmup:
            times.append(r["seconds"])
            rays += r["counters"]["extend"] + r["counters"]["shadow"]
            last = r["counters"]
    total = sum(times)
    # one thread: a smaller sample of the same frame (1 spp; 4K frames at half resolution)
    div = 1 if sd.width <= 1920 else 2
    r1 = sc.render(width=sd.width // div, height=sd.height // div, sample_begin=0, sample_count=1, seed=seed, threads=1)
    one = (r1["counters"]["extend"] + r1["counters"]["shadow"]) / r1["seconds"] / 1e6
    out = dict(value=rays / total / 1e6, unit="Mrays/s", cores=threads, kind="port", flags=flags,
               threads_rule="num_cpus - 1 = %d of %d (main.rs:72)" % (threads, os.cpu_count() or 1),
               one_thread=dict(value=one, unit="Mrays/s", sample="1 spp of the frame at %dx%d, %.1f s" % (sd.width // div, sd.height // div, r1["seconds"])),
               sample="%d step(s) of %d spp of the %dx%d frame (%.1f s of CPU work); oracle restatement of main.rs:181-207 — the reference itself is Rust nightly + "
                      "un-vendored crates and cannot be built here" % (steps, spp, sd.width, sd.height, total),
               ms_per_step=total / max(steps, 1) * 1e3)
    return (out, last) if want_counters else out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    sd = get_scene(args.workload)
    strong = bool(WORKLOADS[args.workload].get("strong"))
    spp = args.spp or sd.spp
    cpu = cpu_arm(sd, args.ref_spp, args.seed, args.steps, args.warmup)
    v = cpu["value"]
    line = dict(impl="reference", metric="Mrays/s (primary+shadow)", value=v, unit="Mrays/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=cpu.pop("ms_per_step"), higher_is_better=True, scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=make_config(args.workload, sd, spp, world, strong, args.ref_spp), cpu_baseline=cpu,
                e2e=dict(value=v, unit="Mrays/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), file=OUT)
    OUT.flush()


class Runner:
    """One workload on this rank's GPU (+ the rank's member of the ptb_group when world > 1)."""

    def __init__(self, name, args, rank, world, local, dist, spp=None, strong=None):
        import torch
        import path_tracer_b200 as ptb
        from path_tracer_b200 import multi
        self.torch, self.dist, self.multi = torch, dist, multi
        self.name, self.args, self.rank, self.world, self.local = name, args, rank, world, local
        self.sd = sd = get_scene(name)
        self.W, self.H = sd.width, sd.height
        self.spp = spp or args.spp or sd.spp
        self.strong = bool(WORKLOADS[name].get("strong")) if strong is None else strong
        self.per_step = self.spp if self.strong else world * self.spp  # global samples per step
        self.dev = torch.device("cuda", local)
        self.ctx = ptb.Context(local)  # raises without a GPU / without libptb.so: no fallback
        t0 = time.time()
        self.ctx.upload(sd)
        self.build_s = time.time() - t0
        self.ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
        self.ctx.frame_begin(self.W, self.H)
        self.group = None
        if world > 1:
            uid = [multi.unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)  # plumbing: ships the 128-byte NCCL id; the collective itself is libptb's
            self.group = multi.Group.join(self.ctx, world, rank, uid[0])
        self.h_acc = torch.empty((self.H, self.W, 4), dtype=torch.float32).pin_memory().numpy()
        self.h_pos = torch.empty((self.H, self.W, 4), dtype=torch.float32).pin_memory().numpy()
        self.h_id = torch.empty((self.H, self.W), dtype=torch.int32).pin_memory().numpy().view(np.uint32)

    def close(self):
        if self.group:
            self.group.close()
        self.ctx.close()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def step(self, index, e2e):
        sd, ctx = self.sd, self.ctx
        if e2e:
            ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)  # the per-frame input of the seam (main.rs:199)
        ctx.frame_begin(self.W, self.H)  # clears the accumulation buffer
        # step `index` covers the global samples [index*per_step, +per_step); the group partitions them over the ranks (spp each under
        # weak scaling, spp / world under strong) and sums the accumulation buffers onto rank 0 with ONE ncclReduce (NVLink)
        if self.group:
            self.group.render(index * self.per_step, self.per_step, seed=self.args.seed, max_bounces=sd.max_bounces)
            self.group.reduce(0)
            ctx.synchronize()
        else:
            ctx.render(index * self.per_step, self.per_step, seed=self.args.seed, max_bounces=sd.max_bounces)
        if e2e and self.rank == 0:
            ctx.read_accum(self.h_acc)
            ctx.read_aux(self.h_pos, self.h_id)

    def timed(self, first, n, e2e):
        torch = self.torch
        self.ctx.reset_counters()
        self.barrier()
        t = time.perf_counter()
        for i in range(n):
            self.step(first + i, e2e)
        self.barrier()
        el = time.perf_counter() - t
        c = self.ctx.counters()
        dev_ms = self.ctx.timing()["render"][0]
        rays = torch.tensor([c["extend"] + c["settled"] + c["shadow"], c["probe"], c["launches"], c["paths"], c["settled"], c["extend"], c["shadow"]], dtype=torch.float64, device=self.dev)
        tt = torch.tensor([el, dev_ms], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(rays)
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
        return rays.tolist(), tt.tolist()

    def ray_record(self, rays):
        # `world` = every world-TLAS ray the reference would count (primary + bounce closest-hit, shadow any-hit); of these,
        # `settled_without_traversal` are primary rays of a camera outside the scene bounds that the bounds pre-test ends before any
        # traversal state is built (the reference ends them at its root-box test, tlas.rs:68)
        return dict(world=rays[0], traversed=rays[5] + rays[6], settled_without_traversal=rays[4], light_probes=rays[1], paths=rays[3])


def load_json(rel):
    p = os.path.join(ROOT, rel)
    return json.load(open(p)) if os.path.exists(p) else {}


def measure(R, steps, warmup, with_cpu, sampler=None):
    """value / e2e / roofline (+ cpu_baseline) of one workload; every rank runs it, rank 0 keeps the record."""
    args, ctx, sd, name = R.args, R.ctx, R.sd, R.name
    for i in range(warmup):
        R.step(i, False)
    if sampler:
        sampler.start()
    rays, (el, dev_ms) = R.timed(warmup, steps, False)
    if sampler:
        sampler.stop_flag = True
    value = rays[0] / el / 1e6
    rays_e, (el_e, _) = R.timed(warmup, steps, True)
    e2e_value = rays_e[0] / el_e / 1e6

    # roofline of the dominant kernel (extend): a few steps again with per-launch CUDA events on the context stream (one wavefront group)
    prof_steps = min(steps, 5)
    ctx.set_profiling(True)
    ctx.reset_counters()
    R.barrier()
    for i in range(prof_steps):
        R.step(warmup + i, False)
    R.barrier()
    tm, cn = ctx.timing(), ctx.counters()
    st, us = ctx.launch_log()
    act = ctx.iteration_log()
    ctx.set_profiling(False)
    if R.rank != 0:
        return None

    peaks = load_json("MEASURED_PEAKS.json")
    peak, peak_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)") if "hbm_gbs" in peaks else (6650.0, "B200_PROFILING.md fallback")
    cpu = alg_ext = alg_sha = None
    if with_cpu:
        cpu, cr = cpu_arm(sd, args.ref_spp, args.seed, 1, 0, want_counters=True)
        cpu.pop("ms_per_step")
        alg_ext, alg_sha = alg_bytes_per_ray(cr)

    rec = dict(value=value, unit="Mrays/s", ms_per_step=el / steps * 1e3, device_ms_per_step=dev_ms / steps, steps=steps, warmup=warmup,
               e2e=dict(value=e2e_value, unit="Mrays/s", h2d_bytes_per_step=76 + 24, d2h_bytes_per_step=R.W * R.H * 36 if R.rank == 0 else 0, ms_per_step=el_e / steps * 1e3),
               gpu_launches=int(rays[2]), rays=R.ray_record(rays), build_seconds=round(R.build_s, 2),
               value_traversed_only=(rays[5] + rays[6]) / el / 1e6)
    ext_ms, ext_launches = tm["extend"]
    ext_rays = cn["extend"] - cn["extend_path"]  # rays traced by k_extend launches (the tail kernel's rays are not in ext_ms either)
    if alg_ext is not None and ext_ms > 0:
        achieved = ext_rays * alg_ext / (ext_ms * 1e-3) / 1e9
        # the launches of the last profiled step, split at 4 rays per resident lane: below that a launch is floored by the latency of
        # its slowest ray, not by throughput
        n_slots = R.W * R.H * min(16, R.per_step // R.world if R.strong else R.spp)
        ext_us = us[st == 2]
        rays_in = np.concatenate([[n_slots], act[: max(len(ext_us) - 1, 0)]]).astype(np.float64)[: len(ext_us)]
        dense = rays_in >= 4 * 148 * 7 * 128
        dense_rec = None
        if dense.any() and len(rays_in) == len(ext_us):
            dense_rec = dict(launches=int(dense.sum()), of=int(len(ext_us)), ms=float(ext_us[dense].sum() * 1e-3), thin_ms=float(ext_us[~dense].sum() * 1e-3),
                             share_of_queue_entries=float(rays_in[dense].sum() / rays_in.sum()),
                             note="k_extend launches of one step whose queue holds >= 4 entries per resident lane (entries = continuing paths + slots awaiting regeneration) "
                                  "against the thin ones, each floored by the latency of its slowest ray (DESIGN.md section 5)")
        ncu = load_json("profiles/ncu_counters.json").get(name)
        roof = dict(bound="issue", kernel="k_extend (8-wide persistent closest-hit; also finalises/accumulates/raygens regenerated slots)", achieved=achieved, peak=peak, unit="GB/s",
                    frac=achieved / peak, traffic=load_json("profiles/traffic.json").get(name, {}).get("k_extend_dram_bytes_per_launch"),
                    peak_source=peak_src, alg_bytes_per_ray=alg_ext, rays_per_launch=ext_rays / max(ext_launches, 1), launch_ms=ext_ms / max(ext_launches, 1),
                    grays_per_s=ext_rays / (ext_ms * 1e-3) / 1e9, dense_launches=dense_rec, ncu=ncu,
                    note="`frac` = ALGORITHMIC bytes of the reference traversal (oracle-counted slab / triangle / instance work on its binary BVH) per second of k_extend time over the "
                         "measured HBM copy bandwidth (the figure SURVEY 8(d) defines).  It is not DRAM traffic: the kernel is bound by instruction issue — the ALU pipe first — and "
                         "by lanes per instruction, with DRAM at a few percent of peak (`ncu`, from the committed capture named there; DESIGN.md section 5).",
                    stage_ms={k: round(v[0] / prof_steps, 3) for k, v in tm.items()},
                    shadow=dict(alg_bytes_per_ray=alg_sha, achieved=(cn["shadow"] * alg_sha / (tm["nee"][0] * 1e-3) / 1e9) if tm["nee"][0] > 0 else None,
                                note="k_nee also traces the light probes (lights-TLAS closest hit), not counted here"))
        rec["roofline"] = roof
    if cpu:
        rec["cpu_baseline"] = cpu
    return rec


def measure_pipelined(name, args, local, steps, warmup):
    """Two contexts on one GPU rendering alternate frames (ptb_render_async / ptb_wait), results read back every frame."""
    import torch
    import path_tracer_b200 as ptb
    sd = get_scene(name)
    W, H, spp = sd.width, sd.height, args.spp or sd.spp
    ctxs = []
    for _ in range(2):
        c = ptb.Context(local)
        c.upload(sd)
        c.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
        c.frame_begin(W, H)
        ctxs.append(c)
    h_acc = torch.empty((H, W, 4), dtype=torch.float32).pin_memory().numpy()
    h_pos = torch.empty((H, W, 4), dtype=torch.float32).pin_memory().numpy()
    h_id = torch.empty((H, W), dtype=torch.int32).pin_memory().numpy().view(np.uint32)

    def run(first, n):
        for i in range(n + 2):
            c = ctxs[i % 2]
            if i >= 2:
                c.wait()
                c.read_accum(h_acc)
                c.read_aux(h_pos, h_id)
            if i < n:
                c.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
                c.frame_begin(W, H)
                c.render_async((first + i) * spp, spp, seed=args.seed, max_bounces=sd.max_bounces)

    run(0, max(warmup, 2))
    for c in ctxs:
        c.reset_counters()
    torch.cuda.synchronize()
    t = time.perf_counter()
    run(warmup, steps)
    torch.cuda.synchronize()
    el = time.perf_counter() - t
    rays = 0
    for c in ctxs:
        k = c.counters()
        rays += k["extend"] + k["settled"] + k["shadow"]
        c.close()
    return dict(frames_in_flight=2, value=rays / el / 1e6, unit="Mrays/s", ms_per_step=el / steps * 1e3, steps=steps,
                note="e2e-style: camera upload, ptb_render_async, ptb_wait and the read-back of the three frame arrays every frame; two contexts alternate, so the thin, "
                     "latency-bound end of one frame's wavefront overlaps the dense start of the next")


def multi_gpu_parity(R):
    """On the box: the N-GPU reduced image equals the 1-GPU image of the same global sample range (<= 1e-6 relative),
    w and the ray counters identical."""
    W, H, begin, count = 480, 270, 7, 64
    sd, ctx, g = R.sd, R.ctx, R.group
    ctx.frame_begin(W, H)
    ctx.reset_counters()
    g.render(begin, count, seed=R.args.seed, max_bounces=sd.max_bounces)
    g.reduce(0)
    ctx.synchronize()
    c = ctx.counters()
    t = R.torch.tensor([c["extend"] + c["settled"], c["shadow"], c["probe"], c["paths"]], dtype=R.torch.float64, device=R.dev)
    R.dist.all_reduce(t)
    out = None
    if R.rank == 0:
        a = ctx.read_accum().copy()
        ctx.frame_begin(W, H)
        ctx.reset_counters()
        ctx.render(begin, count, seed=R.args.seed, max_bounces=sd.max_bounces)
        b = ctx.read_accum()
        c1 = ctx.counters()
        one = [c1["extend"] + c1["settled"], c1["shadow"], c1["probe"], c1["paths"]]
        err = float((np.abs(a[..., :3] - b[..., :3]).max(axis=2) / (np.abs(b[..., :3]).max(axis=2) + 1e-9)).max())
        out = dict(ok=bool(err <= 1e-6 and np.array_equal(a[..., 3], b[..., 3]) and [int(x) for x in t.tolist()] == one), max_rel_err=err,
                   counters_equal=[int(x) for x in t.tolist()] == one, sample="%dx%d, samples [%d, %d) of the workload scene: %d-GPU group + ncclReduce vs one GPU" % (W, H, begin, begin + count, R.world))
    R.dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=list(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="samples per pixel per step (0 = the workload's own)")
    ap.add_argument("--ref-spp", type=int, default=4, help="spp per step of the bounded CPU sample (--impl reference and cpu_baseline)")
    ap.add_argument("--seed", type=int, default=20261104)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--only", action="store_true", help="the primary workload only: no `workloads`, `pipelined`, `strong` sub-records")
    args = ap.parse_args()
    # stdout carries the ONE JSON line and nothing else: file descriptor 1 is pointed at stderr for the run (libraries write banners
    # there, e.g. "NCCL version ..." on the first collective) and the line goes to the saved descriptor
    global OUT
    sys.stdout.flush()
    OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    name = args.workload
    R = Runner(name, args, rank, world, local, dist)
    sampler = ClockSampler(local)
    rec = measure(R, args.steps, args.warmup, not args.no_cpu_baseline, sampler)
    sampler.join(timeout=2)
    parity = multi_gpu_parity(R) if world > 1 else None
    sd, spp, strong = R.sd, R.spp, R.strong
    R.close()

    extra = {}
    sub_steps, sub_warm = min(args.steps, 5), 3
    if not args.only and world == 1 and name == "c3":
        extra["pipelined"] = measure_pipelined(name, args, local, args.steps, args.warmup)
        subs = {}
        for w in ("c2", "c4"):
            Rw = Runner(w, args, rank, world, local, dist)
            r = measure(Rw, sub_steps, sub_warm, not args.no_cpu_baseline)
            r["config"] = make_config(w, Rw.sd, Rw.spp, world, False, args.ref_spp)
            subs[w] = r
            Rw.close()
        extra["workloads"] = subs
    if not args.only and name == "c3":
        # strong scaling: BASELINE config 4's scene and resolution with a fixed number of samples per step, split over the ranks
        Rs = Runner("c5", args, rank, world, local, dist, spp=STRONG_SPP, strong=True)
        Rs.step(0, False)
        rays_s, (el_s, _) = Rs.timed(1, 2, True)
        red_ms = Rs.group.reduce_ms() if Rs.group else 0.0
        if rank == 0:
            extra["strong"] = dict(value=rays_s[0] / el_s / 1e6, unit="Mrays/s", ms_per_step=el_s / 2 * 1e3, steps=2, warmup=1, n_gpus=world, scaling="strong",
                                   config=dict(workload="the 1M-triangle instanced scene at 3840x2160 (BASELINE config 4's frame), %d spp per step split over the GPUs" % STRONG_SPP,
                                               spp_per_step=STRONG_SPP, spp_per_gpu=STRONG_SPP / world), reduce_ms_last=red_ms, reduce_bytes=Rs.W * Rs.H * 16,
                                   rays=Rs.ray_record(rays_s), note="e2e-style step (read-back on rank 0 included); value(N) / value(1) of this record is the strong-scaling speed-up")
        Rs.close()

    if rank == 0:
        line = dict(metric="Mrays/s (primary+shadow)", value=rec["value"], unit="Mrays/s", n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=rec["ms_per_step"],
                    higher_is_better=True, scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32", data="synthetic",
                    config=make_config(name, sd, spp, world, strong, args.ref_spp), e2e=rec["e2e"], gpu_launches=rec["gpu_launches"], clocks=sampler.summary(), rays=rec["rays"],
                    value_traversed_only=rec["value_traversed_only"],
                    run=dict(build_seconds=rec["build_seconds"], device_ms_per_step=rec["device_ms_per_step"], frames_in_flight=1))
        for k in ("roofline", "cpu_baseline"):
            if k in rec:
                line[k] = rec[k]
        if parity:
            line["multi_gpu_parity"] = parity
        line.update(extra)
        print(json.dumps(line), file=OUT)
        OUT.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
