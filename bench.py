#!/usr/bin/env python
"""bench.py -- 15-view depth frames/s of the B200 depth pipeline (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (CUDA, one process per GPU)
  python bench.py --impl reference [--steps K] [--warmup W]      the CPU path on the box's host cores

A "step" is one pass of the whole per-view depth pipeline (SLIC -> boundary -> block matching ->
flatness -> init state -> no_prop x propagate -> fusion) over one synthetic 5x3 camera-array frame
at 1920x1080 (config C2 of SURVEY.md 8(d), shipped literals of clMVDE.cpp:14-36).

N > 1 (torchrun): frames are independent, so the headline `value` is frame-parallel (each rank
runs its own frames, no data-path collective, "weak").  The same run also measures ONE frame
sharded by camera view over all ranks with the NCCL halo exchange (north_star) and reports it
under "view_sharded".
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
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (W, H, aw, ah)
    "15view_1080p": (1920, 1080, 5, 3),
    "4x4_1080p": (1920, 1080, 4, 4),
    "5x5_4k": (3840, 2160, 5, 5),
    "tiny": (192, 128, 3, 2),
    # the reference's own data (tests/data/_ref, copied from /root/reference/Images by tests/data/fetch_bundled.py)
    "bundled_15view": (1920, 1080, 5, 3),          # c0f1..c14f1 (backup/data-bar.txt), shipped literals
    "bundled_15view_legacy": (1920, 1080, 5, 3),   # same views, the legacy 15-view settings (levels 10..100 step 2, bl 0.625)
    "beer_garden_3x3": (1920, 1080, 3, 3),         # Images/Beer-Garden/img0..8 (data.txt), shipped literals
    # no structure at all: independent uniform noise in every view (block matching returns arbitrary levels, so
    # neighbouring superpixels rarely share a plane: the candidate dedup of k_propagate has nothing to skip)
    "noise_15view_1080p": (1920, 1080, 5, 3),
}
BUNDLED = {"bundled_15view": "bar15", "bundled_15view_legacy": "bar15", "beer_garden_3x3": "beer9"}


def settings_for(name):
    import cl_multiview_stereo_b200 as mvs
    W, H, aw, ah = WORKLOADS[name]
    if name == "tiny":
        return mvs.SystemSettings(array_width=aw, array_height=ah, min_disp=0, max_disp=16, inc=1, bl_ratio=1.0,
                                  kernel_size=128), W, H
    if name == "bundled_15view_legacy":   # SURVEY 8(d) settings B (clMVDE.cpp:26-29 comments, photo_consistency.cpp:325-326)
        return mvs.SystemSettings(array_width=aw, array_height=ah, min_disp=10, max_disp=100, inc=2, bl_ratio=0.625), W, H
    # shipped literals (clMVDE.cpp:14-36) with the array geometry of the workload
    return mvs.SystemSettings(array_width=aw, array_height=ah), W, H


def data_kind(name):
    return "bundled" if name in BUNDLED else "synthetic"


def make_input(name, seed=0xC2):
    from cl_multiview_stereo_b200 import synth
    s, W, H = settings_for(name)
    if name in BUNDLED:
        from cl_multiview_stereo_b200 import datasets
        if not datasets.available(BUNDLED[name]):
            raise SystemExit(f"workload {name}: the reference's images are not under tests/data/_ref "
                             "(tests/data/fetch_bundled.py copies them where /root/reference exists)")
        rgbx, aw, ah = datasets.load_set(BUNDLED[name])
        assert (aw, ah) == (s.array_width, s.array_height) and rgbx.shape[1:3] == (H, W)
    elif name == "tiny":
        rgbx, gt = synth.make_frame(7, W, H, s.array_width, s.array_height, 1.0, 5, 4, 14)
    elif name == "noise_15view_1080p":
        rng = np.random.default_rng(seed)
        rgbx = rng.integers(0, 256, size=(s.array_width * s.array_height, H, W, 4), dtype=np.uint8)
        rgbx[..., 3] = 0
    else:
        rgbx, gt = synth.make_frame(seed, W, H, s.array_width, s.array_height, s.bl_ratio, 12, 32.0, 58.0)
    return s, W, H, rgbx


# algorithmic bytes per view and stage (API data types, every array once per pass)
def stage_bytes(W, H, S, no_iter, no_prop):
    N = W * H
    mw, mh = -(-W // S), -(-H // S)
    M = mw * mh
    assoc = N * (16 + 4) + M * 32
    return {
        "slic": N * (4 + 16) + M * (16 + 32) + (no_iter + 1) * assoc + no_iter * assoc,
        "boundary": N * 4 + M * (32 + 8),
        "block_match": N * 16 + M * (32 + 8 + 4),
        "flatness": M * (32 + 8),
        "init_state": N * 4 + M * (32 + 8 + 8 + 24),
        "propagate": no_prop * (N * 4 + M * (32 + 8 + 8 + 24 + 24)),
        "fusion": N * (4 + 4) + M * (32 + 24),
    }


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 9:
                for k, n in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU path: WHOLE frames through the reference's own kernel source on the host cores (no sampling, no model)
# ---------------------------------------------------------------------------------------------
class CpuPath:
    """One step = one whole frame, stage by stage as pipeline::exe_pipeline runs it (SLIC of every view, boundary,
    block matching, flatness, init state, no_prop x propagate, spixl_to_image), every kernel from
    oracle/_ref/libclref.so -- the reference's clcode.cl compiled as C++ over a work-item shim, OpenMP over
    work-items -- where that library was built (kind "reference"); the C port of the same kernels otherwise
    (kind "port").  One exception: update_cluster_center needs 256-work-item groups with barriers, which the
    shim can only run as 256 real threads per group (minutes per launch at this size); its C port -- pinned bit
    for bit to the reference kernel on small images, tests/test_cpu_reference_pin.py -- is timed in its place."""

    def __init__(self, name):
        from tests import orc, clref
        self.o = orc.load()
        self.r = clref.load() if clref.available() else None
        self.kind = "reference" if self.r is not None else "port"
        self.name = name
        self.s, self.W, self.H, self.rgbx = make_input(name)
        s = self.s
        self.V = s.array_width * s.array_height
        self.levels = self.o.make_levels(s.min_disp, s.max_disp, s.inc)
        self.sub, self.num = self.o.make_subsets(s.array_width, s.array_height, s.neib_hor, s.neib_ver)
        self.cores = self.o.num_threads()
        self.last = None

    def step(self, k=0):
        """-> (seconds of this frame, per-stage seconds); nothing is extrapolated"""
        o, r, s, V = self.o, self.r, self.s, self.V
        S, aw = s.spixl_size, s.array_width
        T = {}
        t_frame = time.perf_counter()
        t = time.perf_counter()
        labs, maps, idxs = [], [], []
        mxy, mcol = o.slic_params(S)
        for v in range(V):
            if r is None:
                lab, mp, ix = o.slic(self.rgbx[v], S, s.slic_color_weight, s.no_iter, s.enforce_connectivity)
            else:   # clSLIC::do_super_pixel_seg (clSLIC.cpp:67-122) kernel by kernel
                lab = r.cvt(self.rgbx[v])
                mp = r.init_centers(lab, S)
                ix = r.assoc(lab, mp, S, s.slic_color_weight, mxy, mcol)
                for _ in range(s.no_iter):
                    mp = o.update_centers(lab, ix, mp, S)
                    ix = r.assoc(lab, mp, S, s.slic_color_weight, mxy, mcol)
                if s.enforce_connectivity:
                    ix = r.suppress(r.suppress(ix))
            labs.append(lab); maps.append(mp); idxs.append(ix)
        lab, mp, idx = np.stack(labs), np.stack(maps), np.stack(idxs)
        del labs, maps, idxs
        T["slic"] = time.perf_counter() - t
        k_ = r or o
        t = time.perf_counter()
        rep = k_.boundary(mp, idx, S)
        T["boundary"] = time.perf_counter() - t
        t = time.perf_counter()
        mp = (r.block_match(lab, mp, rep, idx, self.levels, self.sub, self.num, aw, s.bl_ratio, S) if r is not None
              else o.block_match(lab, mp, rep, self.levels, self.sub, self.num, aw, s.bl_ratio))
        T["block_match"] = time.perf_counter() - t
        g_, a_ = np.float32(2 * float(s.gamma) ** 2), np.float32(2 * float(s.alpha) ** 2)
        gam, alp, fus, kss = o.refine_params(g_, a_, s.fuse, s.kernel_step, s.kernel_size // 2, S)
        t = time.perf_counter()
        flat = k_.flatness(mp, np.float32(1.0 / float(g_)))
        T["flatness"] = time.perf_counter() - t
        t = time.perf_counter()
        st = [k_.init_state(mp, idx, rep, flat, self.sub, self.num, aw, s.bl_ratio, gam, alp, fus, s.kernel_step, kss), None]
        T["init_state"] = time.perf_counter() - t
        t = time.perf_counter()
        for it in range(s.no_prop):
            src, dst = it % 2, 1 - it % 2
            st[dst] = k_.propagate(st[src], mp, idx, rep, flat, self.sub, self.num, it, aw, s.bl_ratio, gam, alp, fus,
                                   s.kernel_step // (it + 1), np.float32(kss) / np.float32(it + 1))
        T["propagate"] = time.perf_counter() - t
        t = time.perf_counter()
        disp = k_.spixl_to_image(mp, idx, st[0])
        T["fusion"] = time.perf_counter() - t
        self.last = disp
        return time.perf_counter() - t_frame, T

    def sample_desc(self):
        who = ("the reference's own kernel source (clcode.cl compiled as C++ over a work-item shim, oracle/_ref/libclref.so; "
               "update_cluster_center alone runs its bit-identical C port: the kernel needs 256-work-item barrier groups)"
               if self.r is not None else "CPU oracle (C port of clcode.cl)")
        return (who + ", OpenMP over work-items, on WHOLE frames of the workload: every stage, every view, every "
                "superpixel (SLIC -> boundary -> block matching -> flatness -> init state -> %d x propagate -> "
                "spixl_to_image); one step = one frame, measured, nothing sampled or extrapolated" % self.s.no_prop)


def run_reference(args, rank, world):
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is one process that should use all
    # host cores (the OpenMP runtime reads this when the oracle libraries are loaded, below)
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    cp = CpuPath(args.workload)
    budget = float(os.environ.get("MVS_REF_BUDGET_S", "1500"))
    t_all = time.perf_counter()
    warm = min(args.warmup, 1)          # a CPU frame takes seconds: one untimed frame warms pages and the thread pool
    for k in range(warm):
        cp.step(k)
    secs = []
    stages = {}
    for k in range(args.steps):
        sec, T = cp.step(k)
        secs.append(sec)
        for a, b in T.items():
            stages[a] = stages.get(a, 0.0) + b * 1e3
        # never extrapolate: if the wall budget would be overrun, stop and report the steps that were timed
        if k + 1 < args.steps and (time.perf_counter() - t_all) + 1.5 * sec > budget:
            break
    done = len(secs)
    frame_s = float(np.mean(secs))
    value = 1.0 / frame_s
    s = cp.s
    line = {
        "impl": "reference", "metric": "15-view depth frames/s", "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": done, "warmup": warm, "ms_per_step": frame_s * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": data_kind(args.workload),
        "config": dict(workload_config(args.workload, s, cp.W, cp.H), parallelism=f"frame-parallel x{world}"),
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cp.cores, "kind": cp.kind,
                         "sample": cp.sample_desc()},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "stage_ms": {a: b / done for a, b in stages.items()},
        "steps_asked": args.steps, "warmup_asked": args.warmup, "timed_wall_s": float(sum(secs)),
        "ms_per_step_minmax": [min(secs) * 1e3, max(secs) * 1e3], "gpu_launches": 0,
    }
    emit(line)


_REAL_STDOUT = None


def _stdout_to_stderr():
    """stdout carries exactly one JSON line: libraries that write to fd 1 on their own (NCCL prints its version
    banner there when the first communicator is created) are sent to stderr for the rest of the run"""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def workload_config(name, s, W, H):
    return {"workload": name, "views": s.array_width * s.array_height, "array": [s.array_width, s.array_height],
            "image": [W, H], "spixl_size": s.spixl_size, "no_iter": s.no_iter,
            "disp_levels": (s.max_disp - s.min_disp) // s.inc + 1, "no_prop": s.no_prop, "bl_ratio": s.bl_ratio,
            "l2": "per-frame working set (Lab 16 B/px + labels + RGBX + disparity over all views, >1 GB at 1080p) "
                  "exceeds the 126 MB L2; no flush needed"}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def measure_resident(mvs, ctx, frame, steps, warmup, barrier=lambda: None):
    """device-resident frames: (ms per frame by CUDA events around `steps` runs, launches per frame).  The timed
    loop carries no per-stage events (mvs_frame_set_timing is off)."""
    frame.set_timing(False)
    for _ in range(warmup):
        frame.run()
    ctx.sync()
    barrier()
    l0 = ctx.launches
    ctx.event_record(0)
    for _ in range(steps):
        frame.run()
    ctx.event_record(1)
    ctx.sync()
    return ctx.event_elapsed_ms(0, 1) / steps, (ctx.launches - l0) / steps


def stage_split(frame, steps):
    """per-stage CUDA-event times, in a pass of its own"""
    frame.set_timing(True)
    frame.stage_ms()
    for _ in range(steps):
        frame.run()
    st = {k: v / steps for k, v in frame.stage_ms().items()}
    frame.set_timing(False)
    return st


def propagate_work(frame, s):
    """candidate planes and bit-identical-unique planes per superpixel and iteration (one untimed diagnostic pass):
    k_propagate evaluates only the unique ones, so its time depends on how diverse the scene's planes are"""
    frame.slic(); frame.block_match(); frame.flatness(); frame.init_state()
    st = [frame.propagate_stats(i) for i in range(s.no_prop)]
    return {"candidates_per_superpixel": [round(x["candidates"] / max(x["superpixels"], 1), 2) for x in st],
            "unique_planes_per_superpixel": [round(x["unique_planes"] / max(x["superpixels"], 1), 2) for x in st],
            "plane_view_items": [x["items"] for x in st],
            "samples_inside": [x["samples_inside"] for x in st]}


def run_ours(args, rank, world, local_rank):
    dist = None
    if world > 1:
        _stdout_to_stderr()
        import torch                      # plumbing only: rendezvous, barrier, max over ranks
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import cl_multiview_stereo_b200 as mvs

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    s, W, H, rgbx = make_input(args.workload)
    V = s.array_width * s.array_height
    ctx = mvs.Context(local_rank)
    frame = mvs.Frame(ctx, s, W, H)
    frame.upload(mvs.BUF_RGBX, rgbx)

    # ---- device-resident throughput (inputs already in HBM) -------------------------------
    sampler = ClockSampler(local_rank)
    frame.set_timing(False)
    for _ in range(args.warmup):
        frame.run()
    ctx.sync()
    barrier()
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ctx.event_record(0)
    for _ in range(args.steps):
        frame.run()
    ctx.event_record(1)
    ctx.sync()
    ms_total = ctx.event_elapsed_ms(0, 1)
    launches = ctx.launches - l0
    barrier()
    ms_total = max_over_ranks(ms_total)
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = world * 1e3 / ms_step          # frames/s over all ranks (frame-parallel)
    stage = stage_split(frame, min(args.steps, 5))

    # ---- end to end through the public API with HOST buffers ---------------------------------
    # NCTX contexts/frames take turns, so that the copies of one frame overlap the kernels of another.  Two: with three
    # (upload, kernels and download of three frames at once) the kernels of more frames interleave and evict each other
    # from L2 -- measured 52.3 against 53.9 frames/s (MVS_E2E_CONTEXTS=3); the copies are 5 ms of an 18 ms frame anyway
    NCTX = max(1, int(os.environ.get("MVS_E2E_CONTEXTS", "2")))
    ctxs, frames = [ctx], [frame]
    for _ in range(NCTX - 1):
        ctxs.append(mvs.Context(local_rank))
        frames.append(mvs.Frame(ctxs[-1], s, W, H))
    pin_in = mvs.PinnedBuffer(rgbx.shape, np.uint8)
    pin_in.array[...] = rgbx
    pin_out = [mvs.PinnedBuffer((V, H, W), np.float32) for _ in range(NCTX)]

    # The streams interleave freely.  MVS_E2E_SERIAL=1 runs the contexts' KERNELS one frame after the other
    # instead (mvs_ctx_wait_event on the other context's "run queued" event), so that only copies overlap kernels:
    # measured slower (19.10 against 18.91 ms per frame), the tails of one frame's kernels fill with the other's.
    serial = bool(os.environ.get("MVS_E2E_SERIAL"))
    ran = [False] * NCTX

    def e2e_step(i):
        k = i % NCTX
        ctxs[k].sync()                      # the previous frame on this context has fully landed
        frames[k].upload(mvs.BUF_RGBX, pin_in.array, sync=False)
        if serial and NCTX > 1 and ran[(k - 1) % NCTX]:
            ctxs[k].wait_event(ctxs[(k - 1) % NCTX], 8)
        frames[k].run()
        ctxs[k].event_record(8)
        ran[k] = True
        frames[k].download(mvs.BUF_DISP, out=pin_out[k].array, sync=False)

    for i in range(max(args.warmup, NCTX)):
        e2e_step(i)
    for c in ctxs:
        c.sync()
    barrier()
    ctx.event_record(2)
    for i in range(args.steps):
        e2e_step(i)
    for c in reversed(ctxs):
        c.sync()
    ctx.event_record(3)
    ctx.sync()
    e2e_ms = max_over_ranks(ctx.event_elapsed_ms(2, 3)) / args.steps
    checksum = float(np.nansum(pin_out[(args.steps - 1) % NCTX].array[0, ::64, ::64]))
    barrier()
    for f_, c_ in zip(frames[1:], ctxs[1:]):
        f_.close(); c_.close()
    for pb in pin_out:
        pb.free()
    pin_in.free()

    work = propagate_work(frame, s) if rank == 0 else None

    # ---- one frame sharded by view over all ranks (NCCL halo exchange) ----------------------------
    view_sharded = None
    if world > 1:
        uid = [mvs.Context.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(uid[0], world, rank)
        _, sub_num = mvs.make_subsets(s.array_width, s.array_height, s.neib_hor, s.neib_ver)
        tables = {"even": [mvs.view_partition(V, world, r) for r in range(world)],
                  "weighted": mvs.view_partition_weighted(mvs.view_weights(sub_num), world)}
        if tables["weighted"] == tables["even"]:
            del tables["weighted"]
        # any SET of cameras per rank (a slot order of the cameras + contiguous slot ranges), balanced by measured work
        orders = {}
        bal_order, bal_table = mvs.view_order_balanced(s.array_width, s.array_height, world, s.neib_hor, s.neib_ver,
                                                       w_halo=int(os.environ.get("MVS_W_HALO", "4")))
        if bal_order != list(range(V)):
            tables["balanced"] = bal_table
            orders["balanced"] = bal_order
        frame.set_timing(False)
        frame.upload(mvs.BUF_RGBX, rgbx); frame.run(); frame.view_fusion()     # the single-GPU result to compare with
        runs = {}
        for name, table in tables.items():
            lo, hi = table[rank]
            order = orders.get(name, list(range(V)))      # slot -> camera
            fs = mvs.Frame(ctx, s, W, H, lo, hi)
            if name in orders:
                fs.set_view_order(order)
            fs.set_partition(table)
            if hi > lo:
                fs.upload(mvs.BUF_RGBX, np.ascontiguousarray(rgbx[order[lo:hi]]), lo, hi)
            # a rank ingests the camera images of the views its own views read as well (host data, like its own):
            # no image crosses between ranks
            halo = fs.halo_views()
            for v in halo:
                fs.upload(mvs.BUF_RGBX, rgbx[order[v]:order[v] + 1], v, v + 1)
            fs.set_halo_inputs(True)
            for _ in range(args.warmup):
                fs.run()
            ctx.sync(); barrier()
            ctx.event_record(4)
            for _ in range(args.steps):
                fs.run()
            ctx.event_record(5)
            ctx.sync()
            vs_ms = max_over_ranks(ctx.event_elapsed_ms(4, 5)) / args.steps
            barrier()
            fs.set_timing(True); fs.stage_ms()
            for _ in range(min(args.steps, 5)):
                fs.run()
            st = {k: v / min(args.steps, 5) for k, v in fs.stage_ms().items()}
            fs.set_timing(False)
            per_rank = [None] * world
            dist.all_gather_object(per_rank, {k: round(v, 3) for k, v in st.items()})
            # sharded result must equal the single-GPU result bit for bit on the views this rank owns
            same = True
            checked = [mvs.BUF_DISP]
            if name not in orders:             # multi-view fusion wants the cameras in row-major order
                fs.view_fusion()               # every rank takes part in the exchange of the disparity images
                checked.append(mvs.BUF_DISP_FUSED)
            if hi > lo:
                for which in checked:
                    a = fs.download(which, lo, hi)
                    b = np.stack([frame.download(which, c, c + 1)[0] for c in order[lo:hi]])
                    same = same and bool(np.array_equal(a.view(np.uint32), b.view(np.uint32)))
            same = max_over_ranks(0.0 if same else 1.0) == 0.0
            runs[name] = {"value": 1e3 / vs_ms, "unit": "frames/s", "ms_per_step": vs_ms, "scaling": "strong",
                          "partition": name, "views_per_rank": [b - a for a, b in table],
                          "cameras_per_rank": [order[a:b] for a, b in table], "halo_views_rank0": len(halo),
                          "halo_inputs": "uploaded by each rank with its own views (no image exchange)",
                          "exchange_ms_rank0": st["exchange"], "stage_ms_rank0": st,
                          # per rank: time in its own kernels, and in the exchanges (transfers + waiting for its peers)
                          "compute_ms_per_rank": [round(sum(v for k, v in r.items() if k != "exchange"), 3) for r in per_rank],
                          "exchange_ms_per_rank": [r["exchange"] for r in per_rank],
                          "bit_identical_to_1gpu": same}
            fs.close()
            barrier()
        best = min(runs, key=lambda k: runs[k]["ms_per_step"])
        view_sharded = dict(runs[best])
        view_sharded["other_partition"] = {k: v for k, v in runs.items() if k != best}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        sb = stage_bytes(W, H, s.spixl_size, s.no_iter, s.no_prop)
        stage_hbm = {}
        for k, b in sb.items():
            if stage.get(k, 0) > 0:
                gbs = b * V / (stage[k] * 1e-3) / 1e9
                stage_hbm[k] = {"ms": stage[k], "alg_MB": b * V / 1e6, "GBps": gbs, "frac": gbs / peak}
        # dominant kernel: propagate, one launch per iteration covering all views
        prop_bytes = sb["propagate"] / s.no_prop * V
        prop_ms = stage["propagate"] / s.no_prop
        prof = {}
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            pass
        roofline = {"kernel": "k_propagate", "bound": "hbm", "achieved": prop_bytes / (prop_ms * 1e-3) / 1e9,
                    "peak": peak, "unit": "GB/s", "frac": prop_bytes / (prop_ms * 1e-3) / 1e9 / peak,
                    "traffic": prof.get("k_propagate"), "peak_source": peak_src, "alg_bytes_per_launch": prop_bytes,
                    "avg_launch_ms": prop_ms,
                    # what actually binds it (from the committed ncu capture, not from this run)
                    "issue": prof.get("k_propagate_issue"),
                    "note": "propagate is bound by instruction issue and dependent-gather latency (about 2e5 instructions "
                            "per superpixel-iteration against 350 B of unique traffic), so its HBM fraction is tiny by "
                            "construction; avg_launch_ms includes the k_pixel_plane pass that feeds it; HBM-bound "
                            "stages are listed in stage_hbm"}
        cpu = None
        if world == 1 and not args.no_cpu:
            cp = CpuPath(args.workload)
            sec, T = cp.step(0)
            cpu = {"value": 1.0 / sec, "unit": "frames/s", "cores": cp.cores, "kind": cp.kind,
                   "sample": cp.sample_desc() + "; here: one frame", "stage_ms": {k: v * 1e3 for k, v in T.items()}}
        other = None
        if world == 1 and not args.no_extra and args.workload == "15view_1080p":
            other = {}
            for name in ("bundled_15view", "noise_15view_1080p"):
                try:
                    s2, W2, H2, rgbx2 = make_input(name)
                except SystemExit as e:
                    other[name] = {"unavailable": str(e)}
                    continue
                f2 = mvs.Frame(ctx, s2, W2, H2)
                f2.upload(mvs.BUF_RGBX, rgbx2)
                ms2, _ = measure_resident(mvs, ctx, f2, max(3, args.steps // 2), 3)
                st2 = stage_split(f2, 3)
                other[name] = {"value": 1e3 / ms2, "unit": "frames/s", "ms_per_step": ms2, "data": data_kind(name),
                               "stage_ms": st2, "propagate_work": propagate_work(f2, s2)}
                f2.close()
        line = {
            "metric": "15-view depth frames/s", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": data_kind(args.workload),
            "config": dict(workload_config(args.workload, s, W, H), parallelism=f"frame-parallel x{world}"),
            "clocks": clocks,
            "e2e": {"value": world * 1e3 / e2e_ms, "unit": "frames/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(rgbx.nbytes), "d2h_bytes_per_step": int(V * H * W * 4),
                    "result_checksum": checksum},
            "gpu_launches": int(launches),
            "roofline": roofline, "stage_ms": stage, "stage_hbm": stage_hbm, "propagate_work": work,
            "cpu_baseline": cpu, "view_sharded": view_sharded, "other_data": other,
        }
        emit(line)
    frame.close()
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# SLIC-only workload (BASELINE.json configs[2]): batches of 3840x2160 images, S = 8, 5 iterations
# ---------------------------------------------------------------------------------------------
def run_slic_batch(args, rank, world, local_rank):
    dist = None
    if world > 1:
        _stdout_to_stderr()
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import cl_multiview_stereo_b200 as mvs
    from cl_multiview_stereo_b200 import synth

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    W, H, B = 3840, 2160, 16
    s = mvs.SystemSettings(array_width=B, array_height=1)      # B independent images = a B x 1 "array"
    imgs = np.stack([synth.voronoi_image(0x5EED0000 + rank * B + k, W, H) for k in range(B)])
    ctx = mvs.Context(local_rank)
    frame = mvs.Frame(ctx, s, W, H)
    frame.upload(mvs.BUF_RGBX, imgs)
    frame.set_timing(False)
    for _ in range(args.warmup):
        frame.slic()
    ctx.sync()
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ctx.event_record(0)
    for _ in range(args.steps):
        frame.slic()
    ctx.event_record(1)
    ctx.sync()
    ms = max_over_ranks(ctx.event_elapsed_ms(0, 1)) / args.steps
    launches = ctx.launches - l0
    clocks = sampler.stop() if rank == 0 else None
    # end to end: pinned host images in, labels + centres out, every step; NCTX frames on as many streams take
    # turns so that the upload of one batch, the kernels of another and the download of a third overlap: here copies
    # and kernels take about 10 ms each per batch, and three contexts give 1104 images/s against 934 with two
    NCTX = max(1, int(os.environ.get("MVS_E2E_CONTEXTS", "3")))
    ctxs, frames = [ctx], [frame]
    for _ in range(NCTX - 1):
        ctxs.append(mvs.Context(local_rank))
        frames.append(mvs.Frame(ctxs[-1], s, W, H))
    for f in frames:
        f.set_timing(False)
    pin_in = mvs.PinnedBuffer(imgs.shape, np.uint8)
    pin_in.array[...] = imgs
    mw, mh = mvs.map_size(W, H, s.spixl_size)
    pin_idx = [mvs.PinnedBuffer((B, H, W), np.uint32) for _ in range(NCTX)]
    pin_map = [mvs.PinnedBuffer((B, mh, mw, 8), np.float32) for _ in range(NCTX)]

    def e2e_step(i):
        k = i % NCTX
        ctxs[k].sync()
        frames[k].upload(mvs.BUF_RGBX, pin_in.array, sync=False)
        frames[k].slic()
        frames[k].download(mvs.BUF_IDX, out=pin_idx[k].array, sync=False)
        frames[k].download(mvs.BUF_MAP, out=pin_map[k].array, sync=False)

    for i in range(NCTX):
        e2e_step(i)
    for c in ctxs:
        c.sync()
    ctx.event_record(2)
    for i in range(args.steps):
        e2e_step(i)
    for c in reversed(ctxs):
        c.sync()
    ctx.event_record(3)
    ctx.sync()
    e2e_ms = max_over_ranks(ctx.event_elapsed_ms(2, 3)) / args.steps
    pin_idx, pin_map = pin_idx[(args.steps - 1) % NCTX], pin_map[(args.steps - 1) % NCTX]
    for f_, c_ in zip(frames[1:], ctxs[1:]):
        f_.close(); c_.close()
    if rank == 0:
        peak = 6650.0
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            pass
        alg = stage_bytes(W, H, s.spixl_size, s.no_iter, s.no_prop)["slic"] * B
        gbs = alg / (ms * 1e-3) / 1e9
        emit({
            "metric": "SLIC images/s", "value": world * B * 1e3 / ms, "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "slic_4k", "image": [W, H], "batch_per_step": B, "spixl_size": s.spixl_size,
                       "no_iter": s.no_iter, "color_weight": s.slic_color_weight,
                       "l2": "one batch (16 x 33 MB RGBX, 16 x 133 MB Lab) exceeds the 126 MB L2"},
            "clocks": clocks,
            "e2e": {"value": world * B * 1e3 / e2e_ms, "unit": "images/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(imgs.nbytes), "d2h_bytes_per_step": int(pin_idx.array.nbytes + pin_map.array.nbytes),
                    "result_checksum": int(pin_idx.array[0, ::97, ::89].sum())},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "slic (13 launches per batch)", "bound": "hbm", "achieved": gbs, "peak": peak,
                         "unit": "GB/s", "frac": gbs / peak, "traffic": None, "alg_bytes_per_step": alg},
            "cpu_baseline": None})
    frame.close(); ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="15view_1080p", choices=sorted(WORKLOADS) + ["slic_4k"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra lines on the bundled / noise frames (other_data)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        args.warmup = max(args.warmup, 3)
        if args.workload == "slic_4k":
            run_slic_batch(args, rank, world, local_rank)
        else:
            run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
