#!/usr/bin/env python
"""bench.py — headline benchmark of the tiled-splatting hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path through the C ABI
  python bench.py --impl reference --gpus N --steps K ...   # the CPU oracle (the reference is Metal-only) on host cores
  python bench.py --workload C4 --render-only               # BASELINE config 4: render-only frames/s, views sharded over N GPUs

One "step" = one fwd+bwd pass of the hot path over one synthetic view of config C2 (1M Gaussians, 1920x1080).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.  The driver keeps the base keys of the
line (metric ... config, roofline, cpu_baseline, clocks, e2e), so the numbers that must survive in its records — oracle
parity of the measured frame, BASELINE configs 4 and 5, the gradient all-reduce — ride compactly inside `config`.
"""
import argparse
import csv
import glob
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from gaussiansplatting_b200 import synth  # noqa: E402

METRIC = "fwd+bwd ms/frame @1M Gaussians 1080p"
LRS = [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3]          # mtl_engine.mm:1063-1068


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


def host_threads():
    """host cores this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1, which must not handicap the CPU arm:
    the oracle is given an explicit thread count (its OpenMP regions take num_threads(threads))."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def workload_string(name):
    n = synth.CONFIG_N[name]
    W, H = synth.CONFIG_RES[name]
    return f"{name}: {n} Gaussians, {W}x{H}, 1 view per GPU per step"


def spread(xs):
    a = np.asarray(xs, np.float64)
    return {"median": float(np.median(a)), "p10": float(np.percentile(a, 10)), "p90": float(np.percentile(a, 90)),
            "mean": float(a.mean()), "n": int(a.size)}


def ncu_traffic(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (bytes) of a kernel from the newest committed `ncu --set full`
    counter summary under profiles/ (r*_blend_counters.csv); (None, None) if no capture holds it"""
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_blend_counters.csv")), key=lambda p: os.path.basename(p))
    for path in reversed(files):
        try:
            rows = list(csv.reader(open(path)))
        except OSError:
            continue
        if len(rows) < 3 or "dram__bytes_read.sum" not in rows[0]:
            continue
        ir, iw = rows[0].index("dram__bytes_read.sum"), rows[0].index("dram__bytes_write.sum")
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        it = rows[0].index("gpu__time_duration.sum") if "gpu__time_duration.sum" in rows[0] else None
        best = None                                   # several kernels may match (the flagged-tile pass): the longest launch
        for r in rows[2:]:
            if r and kernel_substr in r[0]:
                try:
                    t = float(r[it]) if it is not None else 0.0
                    b = float(r[ir]) * unit.get(rows[1][ir], 1.0) + float(r[iw]) * unit.get(rows[1][iw], 1.0)
                except (ValueError, IndexError):
                    continue
                if best is None or t > best[0]:
                    best = (t, b)
        if best is not None:
            return best[1], os.path.relpath(path, ROOT)
    return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.p = index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.rows.append(l) for l in self.p.stdout], daemon=True).start()
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        for l in self.rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload(name):
    n = synth.CONFIG_N[name]
    W, H = synth.CONFIG_RES[name]
    return n, W, H, synth.CONFIG_S0[name], synth.CONFIG_SEED[name], synth.CONFIG_ORBIT[name]


# ---------------------------------------------------------------------------------------------------------
# reference arm: the reference's own algorithm on the host cores (oracle port; Metal cannot run here)
# ---------------------------------------------------------------------------------------------------------
def oracle_frame(orc, g, u, gt, threads, timing=None, keep=None, render_only=False):
    t0 = time.perf_counter()
    r = orc.render(g, u, threads)
    t1 = time.perf_counter()
    gr = None
    if not render_only:
        if gt is None:
            gt = np.full_like(r["image"], 0.5)
        gr = orc.backward(g, r["proj"], r["svals"], r["ranges"], u, r["last"], r["image"], gt, threads=threads)
    t2 = time.perf_counter()
    if timing is not None:
        timing.append((t1 - t0, t2 - t1))
    if keep is not None:
        keep["render"], keep["grads"], keep["gt"] = r, gr, gt
    return t2 - t0


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import orc
    orc.build()
    n, W, H, s0, seed, orbit = workload(args.workload)
    threads = host_threads()
    g = synth.make_scene(n, seed, s0)
    steps, warm = max(1, args.steps), max(0, min(args.warmup, 1))
    # a full 1M-Gaussian 1080p frame costs about a second of CPU on 16-32 cores: bound the run to ~a few minutes
    steps = min(steps, 5 if args.workload != "C4" else 3)
    times = []
    for i in range(warm + steps):
        u = synth.make_camera(i, 64, W, H, n, orbit)
        t = oracle_frame(orc, g, u, None, threads, render_only=args.render_only)
        if i >= warm:
            times.append(t)
    ms = 1e3 * float(np.mean(times))
    what = "project+pairs+CPU LSD sort+ranges+forward" + ("" if args.render_only else "+backward")
    sample = f"{len(times)} full frames ({what}), {threads} OpenMP threads"
    metric, unit, value, hib = METRIC, "ms/frame", ms, False
    if args.render_only:
        metric, unit, value, hib = render_metric(args.workload), "frames/s", 1e3 / ms, True
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": len(times),
        "warmup": warm, "ms_per_step": ms, "higher_is_better": hib, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": workload_string(args.workload)},
        "cpu_baseline": {"value": value, "unit": unit, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference is Metal/Objective-C++ and cannot run on Linux: this arm times oracle/ (scalar C++ restatement of "
                "its kernels + its own CPU radix sort) on the host cores; thread count = this process's CPU affinity, passed "
                "explicitly so that torchrun's OMP_NUM_THREADS=1 does not apply"}))


def render_metric(wl):
    n = synth.CONFIG_N[wl]
    W, H = synth.CONFIG_RES[wl]
    return f"render-only frames/s @{n // 1000000}M Gaussians {W}x{H}"


# ---------------------------------------------------------------------------------------------------------
# this repo's arm
# ---------------------------------------------------------------------------------------------------------
class Bench:
    def __init__(self, args):
        import torch
        from gaussiansplatting_b200 import capi as gs
        self.torch, self.gs, self.args = torch, gs, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        torch.cuda.set_device(self.local)
        if self.world > 1:
            import datetime
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local), timeout=datetime.timedelta(seconds=180))
            self.dist = dist
        self.ctx = gs.Context(self.local)
        self.stream = torch.cuda.ExternalStream(self.ctx.stream, device=self.local)
        self.K, self.Wm = max(1, args.steps), max(3, args.warmup)
        self.hbm_peak, self.sm_max_mhz, self.peak_kind = measured_peaks()

    # ---- timing helpers -----------------------------------------------------------------------------------
    def barrier(self):
        self.ctx.sync()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def allmax(self, vals):
        if self.dist is None:
            return [float(v) for v in vals]
        t = self.torch.tensor([float(v) for v in vals], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]

    def timed(self, fn, steps, collective=True, per_step=None):
        """device time of `steps` calls (ms), CUDA events on the launching stream, max over ranks; per_step (a list) receives
        this rank's per-step times.  collective=False: this rank only (rank-0-only diagnostics)"""
        torch = self.torch
        if collective:
            self.barrier()
        else:
            self.ctx.sync()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1 if per_step is not None else 2)]
        evs[0].record(self.stream)
        for i in range(steps):
            fn(i)
            if per_step is not None:
                evs[i + 1].record(self.stream)
        if per_step is None:
            evs[1].record(self.stream)
        evs[-1].synchronize()
        ms = evs[0].elapsed_time(evs[-1])
        if per_step is not None:
            per_step.extend(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
        if collective:
            ms = self.allmax([ms])[0]
            self.barrier()
        return ms

    # ---- BASELINE config 4: render-only, views dealt v mod g, no communication -----------------------------
    def render_only(self, wl, num_views=64, reps=1, try_two_in_flight=False):
        """forward-only frames/s for `wl`, 64 views sharded over the ranks (dp.render_views_for_rank), every frame's RGBA8 read
        back to pinned host memory (e2e); mtl_engine.mm:1224-1306 (exportTrainingViews, forward at :1285)"""
        torch, gs, ctx = self.torch, self.gs, self.ctx
        from gaussiansplatting_b200 import dp
        n, W, H, s0, seed, orbit = workload(wl)
        G = gs.Gaussians(ctx, n); G.upload(synth.make_scene(n, seed, s0))
        R = gs.Rasterizer(ctx, n)
        imgs = [gs.Image(ctx, W, H), gs.Image(ctx, W, H)]
        outs = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
        mine = dp.render_views_for_rank(num_views, self.world, self.rank) * reps
        cams = {v: synth.make_camera(v, num_views, W, H, n, orbit) for v in set(mine)}
        # Views are independent, so TWO rasterizers alternate (frame i on rasterizer i mod 2) with gs_rast_set_overlap: the
        # latency-bound stages before the blend of frame i+1 run on a high-priority side stream under the issue-bound blend of
        # frame i.  The same frames on ONE rasterizer (every kernel in one stream) are timed beside it.
        Rb = gs.Rasterizer(ctx, n)
        Rs = [R, Rb]
        for Rx in Rs:
            for v in sorted(set(mine)) + mine[:3]:    # one untimed forward per distinct view sizes the pair buffers (see run())
                Rx.forward(G, cams[v], imgs[0])
        ctx.sync()
        ms_one = self.timed(lambda i: R.forward(G, cams[mine[i]], imgs[i % 2]), len(mine))
        for Rx in Rs:
            Rx.set_overlap(True)
        for i in range(4):
            Rs[i % 2].forward(G, cams[mine[i % len(mine)]], imgs[i % 2])
        ms_dev = self.timed(lambda i: Rs[i % 2].forward(G, cams[mine[i]], imgs[i % 2]), len(mine))
        for Rx in Rs:
            Rx.set_overlap(False)
        Rb.close()
        chk = [0]

        def e2e():     # one rasterizer: with every frame's 33 MB read back the run is bound by the D2H copies, not by the kernels
            for i, v in enumerate(mine):
                if i >= 2:
                    imgs[i % 2].wait(); chk[0] += int(outs[i % 2][0, 0, 0])
                R.forward(G, cams[v], imgs[i % 2])
                imgs[i % 2].download_rgba8_async(outs[i % 2].data_ptr())
            for im in imgs:
                im.wait()
            ctx.sync()
        self.barrier()
        t0 = time.perf_counter(); e2e(); ms_e2e = (time.perf_counter() - t0) * 1e3
        ms_e2e = self.allmax([ms_e2e])[0]
        self.barrier()
        frames = num_views * reps
        # two frames in flight (--render-only runs only): views are independent, so a second context (own stream, rasterizer,
        # images) on the same GPU could let one frame's latency-bound ordering stage run under the other frame's blend.  Wall
        # clock between full syncs.  MEASURED: no gain (613.8 against 613.5 frames/s at C4) -- the block scheduler dispatches
        # the 32,400 CTAs of the earlier blend grid before the other stream's small kernels get SM slots, only tails overlap.
        two = None
        try:
            if not try_two_in_flight:
                raise StopIteration
            ctx2 = gs.Context(self.local)
            G2 = gs.Gaussians(ctx2, n); G2.upload(synth.make_scene(n, seed, s0))
            R2 = gs.Rasterizer(ctx2, n)
            imgs_b = [gs.Image(ctx2, W, H), gs.Image(ctx2, W, H)]
            outs_b = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]
            lanes = [(ctx, G, R, imgs, outs), (ctx2, G2, R2, imgs_b, outs_b)]

            def run2(readback):
                for i, v in enumerate(mine):
                    c_, G_, R_, im_, ou_ = lanes[i % 2]
                    k = (i // 2) % 2
                    if readback and i >= 4:
                        im_[k].wait(); chk[0] += int(ou_[k][0, 0, 0])
                    R_.forward(G_, cams[v], im_[k])
                    if readback:
                        im_[k].download_rgba8_async(ou_[k].data_ptr())
                for c_, G_, R_, im_, ou_ in lanes:
                    for im in im_:
                        im.wait()
                    c_.sync()
            for v in mine[:2]:
                R2.forward(G2, cams[v], imgs_b[0])
            ctx2.sync()
            t2 = []
            for readback in (False, True):
                self.barrier(); ctx2.sync()
                t0 = time.perf_counter(); run2(readback); t2.append((time.perf_counter() - t0) * 1e3)
            t2 = self.allmax(t2)
            two = {"frames_per_s": 1e3 * frames / t2[0], "e2e_frames_per_s": 1e3 * frames / t2[1]}
            for o in imgs_b + [R2, G2]:
                o.close()
            ctx2.close()
        except StopIteration:
            two = None
        except Exception as e:
            two = {"error": repr(e)[:200]}
        self.barrier()
        res = {"frames_per_s": 1e3 * frames / ms_dev, "frames_per_s_one_rasterizer": 1e3 * frames / ms_one,
               "how": "two rasterizers alternate, the stages before the blend on a high-priority side stream (gs_rast_set_overlap); "
                      "frames_per_s_one_rasterizer = the same frames on one rasterizer, every kernel in one stream",
               "two_in_flight": two, "e2e_frames_per_s": 1e3 * frames / ms_e2e, "ms_per_frame_per_gpu": ms_dev / max(len(mine), 1),
               "views": frames, "pairs": R.pair_count, "sort_passes": R.sort_passes, "d2h_bytes_per_frame": W * H * 4,
               "workload": f"{wl}: {n} Gaussians, {W}x{H}, {num_views} views dealt v mod {self.world}, forward only, RGBA8 read-back per frame"}
        stage = None
        if self.rank == 0:
            R.set_profiling(True)
            acc = None
            for i in range(min(8, len(mine))):
                R.forward(G, cams[mine[i]], imgs[0])
                t = R.timings(False)
                acc = t if acc is None else {k: acc[k] + t[k] for k in t}
            R.set_profiling(False)
            stage = {k: v / min(8, len(mine)) for k, v in acc.items() if k not in ("blend_bwd", "preprocess_bwd")}
            R.forward(G, cams[mine[0]], imgs[0])
            E, C = R.count_evals()
            fp32_peak = 2 * 128 * 148 * self.sm_max_mhz * 1e6 / 1e12
            a = (25.0 * C + 14.0 * (E - C)) / (stage["blend_fwd"] * 1e-3) / 1e12
            res["stage_ms"] = stage
            res["blend_fwd_frac_fp32"] = a / fp32_peak
            res["evals"] = {"E": E, "C": C}
        for o in imgs + [R, G]:
            o.close()
        return res

    # ---- oracle parity of the measured frame (N = 1, rank 0) -------------------------------------------------
    def parity_leg(self, G, R, img, gti, D, g_host, u, gt_f, out):
        """the CPU baseline frame (timed) and, from ITS outputs, the parity of this library's frame for the same view:
        sorted keys / per-tile lists / ranges / lastContribIdx bit-exact, image max-abs, gradients (oracle/parity.py)"""
        gs = self.gs
        from oracle import orc, parity
        orc.build()
        th = host_threads()
        tm, keep = [], {}
        t = oracle_frame(orc, g_host, u, gt_f, th, tm, keep)
        out["cpu_baseline"] = {"value": t * 1e3, "unit": "ms/frame", "cores": th, "kind": "port",
                               "sample": f"1 full frame of the same workload (forward {tm[0][0]:.2f} s incl. the reference's "
                                         f"CPU LSD sort, backward {tm[0][1]:.2f} s)"}
        try:
            ref = keep["render"]
            # the graded backward takes the ORACLE's image and lastContribIdx; ours takes its own (bit-equal when parity holds)
            ref_gr, ab, cnt = orc.backward(g_host, ref["proj"], ref["svals"], ref["ranges"], u, ref["last"], ref["image"], gt_f,
                                           threads=th, want_abs=True)
            gti.upload(gt_f)
            R.forward(G, u, img)
            R.backward(G, D, u, img, gti)
            got = D.download(len(g_host))
            p = parity.frame_parity(gs, R, img.download(), ref, got, ref_gr, ab, cnt,
                                    what=f"{self.args.workload}, view 0 of the timed loop, ground truth = the bench's RGBA8 target")
            out["parity"] = p
            out["config"]["parity"] = {k: p[k] for k in ("ok", "keys_equal", "lists_equal", "ranges_equal", "last_equal", "image_max_abs",
                                                         "grad_worst_pure_rel", "grad_frac_over_1e-3_pure", "grad_frac_need_band",
                                                         "grad_n_fail", "grad_components")}
        except Exception as e:  # a parity failure must not hide the measurement; it is reported as such
            out["parity"] = {"ok": False, "error": repr(e)}
            out["config"]["parity"] = {"ok": False, "error": repr(e)[:200]}

    # ---- the main run ------------------------------------------------------------------------------------------
    def run(self):
        torch, gs, ctx, args = self.torch, self.gs, self.ctx, self.args
        rank, world, local, dist = self.rank, self.world, self.local, self.dist
        K, Wm = self.K, self.Wm
        n, W, H, s0, seed, orbit = workload(args.workload)
        hbm_peak, sm_max_mhz, peak_kind = self.hbm_peak, self.sm_max_mhz, self.peak_kind

        g_host = synth.make_scene(n, seed, s0)
        G = gs.Gaussians(ctx, n); G.upload(g_host)
        R = gs.Rasterizer(ctx, n)
        img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
        D = gs.Grads(ctx, n)
        A = gs.Adam(ctx, n)
        Dn = gs.Density(ctx); Dn.set_scene_extent(synth.scene_extent(orbit))
        NV = 64
        cams = [synth.make_camera(v, NV, W, H, n, orbit) for v in range(NV)]
        # ground truth: render of a different scene seed through this same path, kept as RGBA8 in pinned memory
        Ggt = gs.Gaussians(ctx, n); Ggt.upload(synth.make_scene(n, seed ^ 0xA5A5, s0))
        gt_pinned = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
        out_pinned = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
        R.forward(Ggt, cams[rank % NV], gti)
        gti.download_rgba8_ptr(gt_pinned.data_ptr())
        Ggt.close()
        gti.upload_rgba8_ptr(gt_pinned.data_ptr()); ctx.sync()

        def view(i):
            return cams[(i * world + rank) % NV]

        def fwd_bwd(i):
            R.forward(G, view(i), img)
            R.backward(G, D, view(i), img, gti)

        # The pair buffers (and the visit words behind them) grow on demand, as the reference's do (ensurePairsCapacity,
        # tiled_rasterizer.mm:242-272): a view with more pairs than any before it costs a reallocation and a redone frame, once.
        # One untimed forward per distinct view of the timed loop takes those one-off events out of the timed region.
        for v in sorted({(i * world + rank) % NV for i in range(K)}):
            R.forward(G, cams[v], img)
        for i in range(Wm):
            fwd_bwd(i)
        sampler = ClockSampler(local)       # rank 0 only: eight nvidia-smi pollers on one box disturb the launch path of every rank
        if rank == 0:
            sampler.start()
        per_step = []
        ms_total = self.timed(fwd_bwd, K, per_step=per_step)
        ms_step = ms_total / K                       # one view per rank per step
        value = ms_total / (K * world)               # whole-job ms per frame
        # the same frames with TWO rasterizers alternating and gs_rast_set_overlap (frame i+1's stages before the blend on a
        # high-priority side stream under frame i's blend kernels): a throughput figure for independent frames of one parameter
        # set; a training loop (Adam between frames) cannot use it, so the headline stays the one-rasterizer number.
        ms_overlap = None
        try:
            R2 = gs.Rasterizer(ctx, n)
            for v in sorted({(i * world + rank) % NV for i in range(K)}):
                R2.forward(G, cams[v], img)
            R2.forward(G, view(0), img); R2.backward(G, D, view(0), img, gti)      # its visit words exist from here on
            Rs = [R, R2]
            for Rx in Rs:
                Rx.set_overlap(True)

            def fwd_bwd2(i):
                Rs[i % 2].forward(G, view(i), img)
                Rs[i % 2].backward(G, D, view(i), img, gti)
            for i in range(Wm):
                fwd_bwd2(i)
            # every rank measures on its own (no barrier inside the try: a rank that fails here must not leave the others waiting)
            ms_overlap = self.timed(fwd_bwd2, K, collective=False) / (K * world)
            for Rx in Rs:
                Rx.set_overlap(False)
            R2.close()
        except Exception as e:                           # diagnostic only
            ms_overlap = "failed: " + repr(e)[:120]
        pairs, passes = R.pair_count, R.sort_passes

        out = {"metric": METRIC, "value": value, "unit": "ms/frame", "n_gpus": world, "steps": K, "warmup": Wm,
               "ms_per_step": ms_step, "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
               "data": "synthetic",
               "config": {"workload": workload_string(args.workload), "pairs": pairs,
                          "parallelism": f"views sharded over {world} GPU(s), no data-path collective for fwd+bwd (config.c5 holds the step with the gradient all-reduce)",
                          "l2": "per-step working set (~0.6 GB: params, records, pair buffers, images) exceeds the 126 MB L2",
                          "sizing": "one untimed forward per distinct view before the warm-up: on-demand growth of the pair buffers is a one-off reallocation, not a step",
                          "ms_per_frame_two_rasterizers_overlapped": ms_overlap},
               "ms_per_step_stats": spread(per_step)}

        # ---- end to end through the C ABI with host buffers: every step copies that step's ground truth (pinned RGBA8) and
        # TiledUniforms H2D and reads the step's rendered RGBA8 back D2H.  The transfers use the library's copy streams
        # (gs_image_*_async): the NEXT view's ground truth is prefetched and the render is read back while the blend kernels
        # run; the host consumes step i's pixels before it enqueues step i+2 (two images each way).
        imgs2, gts2 = [img, gs.Image(ctx, W, H)], [gti, gs.Image(ctx, W, H)]
        outs2 = [out_pinned, torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()]
        e2e_sum = [0]

        def e2e_run(steps):
            gts2[0].upload_rgba8_async(gt_pinned.data_ptr())
            for i in range(steps):
                if i + 1 < steps:
                    gts2[(i + 1) % 2].upload_rgba8_async(gt_pinned.data_ptr())
                if i >= 2:
                    imgs2[i % 2].wait()                                   # step i-2's pixels are on the host: consume them
                    e2e_sum[0] += int(outs2[i % 2][0, 0, 0])
                R.forward(G, view(i), imgs2[i % 2])
                imgs2[i % 2].download_rgba8_async(outs2[i % 2].data_ptr())
                R.backward(G, D, view(i), imgs2[i % 2], gts2[i % 2])
            for j in range(2):
                imgs2[j].wait()
            ctx.sync()

        def e2e_step_serial(i):
            gti.upload_rgba8_ptr(gt_pinned.data_ptr())
            R.forward(G, view(i), img)
            R.backward(G, D, view(i), img, gti)
            img.download_rgba8_ptr(out_pinned.data_ptr())      # synchronises
        e2e_run(3)
        self.barrier()
        t0 = time.perf_counter()
        e2e_run(K)
        e2e_ms = (time.perf_counter() - t0) * 1e3
        self.barrier()
        t0 = time.perf_counter()
        for i in range(K):
            e2e_step_serial(i)
        ctx.sync()
        e2e_serial_ms = (time.perf_counter() - t0) * 1e3
        e2e_ms, e2e_serial_ms = self.allmax([e2e_ms, e2e_serial_ms])
        out["e2e"] = {"value": e2e_ms / (K * world), "unit": "ms/frame", "h2d_bytes_per_step": W * H * 4 + 240,
                      "d2h_bytes_per_step": W * H * 4, "serial_ms_per_frame": e2e_serial_ms / (K * world),
                      "what": "per step: pinned RGBA8 ground truth + TiledUniforms H2D, forward, backward, rendered RGBA8 D2H "
                              "(every step's pixels are read on the host); transfers on the library's copy streams overlap the "
                              "kernels, wall clock incl. all syncs; serial_ms_per_frame = the same with blocking transfers; "
                              "Gaussians/gradients stay device-resident (model state)"}
        gts2[1].close(); imgs2[1].close()
        gti.upload_rgba8_ptr(gt_pinned.data_ptr()); ctx.sync()

        # ---- training iterations (fwd + bwd + accumulateGradients + Adam), with ONE gradient all-reduce per step when N > 1.
        # Adam clamps log-scale to [-4, 4] (shaders.metal:55,655), so the C2 scene (scales ~e^-5.1) would jump to e^-4 on the
        # first step; training runs on the C5 scene (same 1M Gaussians / 1080p, scales inside the clamp, orbit 12).
        from gaussiansplatting_b200 import dp
        n5, (W5, H5) = synth.CONFIG_N["C5"], synth.CONFIG_RES["C5"]
        s05, seed5, orbit5 = synth.CONFIG_S0["C5"], synth.CONFIG_SEED["C5"], synth.CONFIG_ORBIT["C5"]
        NVT = 16
        g5_host = synth.make_scene(n5, seed5, s05)
        G5 = gs.Gaussians(ctx, n5); G5.upload(g5_host)
        Gt5 = gs.Gaussians(ctx, n5); Gt5.upload(synth.make_scene(n5, seed5 ^ 0xA5A5, s05))
        cams5 = [synth.make_camera(v, NVT, W5, H5, n5, orbit5) for v in range(NVT)]
        gts5 = [gs.Image(ctx, W5, H5) for _ in range(NVT)]
        for v in range(NVT):
            R.forward(Gt5, cams5[v], gts5[v])
        ctx.sync(); Gt5.close()
        A5 = gs.Adam(ctx, n5)
        Dn.set_scene_extent(synth.scene_extent(orbit5))
        D.set_stat_mode(True)                            # per-view densify statistics ride in columns 14-15 (survive sums)
        vp = dp.ViewParallel(dist)
        gbuf = vp.device_tensor(D) if dist is not None else None
        # the gradient sum: this library's own kernels over NVLink peer memory (gs_dp_*, csrc/peer_reduce.cu); NCCL's all-reduce on
        # the same buffer is timed beside it (--allreduce nccl makes it the headline instead)
        PR = SG = None
        if dist is not None:
            PR = gs.PeerReduce(ctx, rank, world)
            blobs = [None] * world
            dist.all_gather_object(blobs, PR.export(D))
            PR.connect(blobs)
            # the same payload in symmetric, multicast-mapped memory: the sum runs in the NVSwitch (multimem.ld_reduce / multimem.st)
            try:
                SG = dp.SymmetricGrads(gs, ctx, dist, n5, local)
                SG.grads.set_stat_mode(True)
                if not SG.multicast:
                    SG.close(); SG = None
            except Exception as e:
                out.setdefault("notes", []).append("no symmetric / multicast memory on this box: " + repr(e)[:160])
                SG = None
            ok = torch.tensor([1 if SG is not None else 0], device="cuda"); dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if not int(ok.item()) and SG is not None:
                SG.close(); SG = None

        def make_train_step(views_per_step, reduce_impl):
            Dg = SG.grads if reduce_impl == "nvls" else D
            def step(i):
                mine = vp.views(i, views_per_step, NVT)
                for k, v in enumerate(mine):
                    R.forward(G5, cams5[v], img)
                    R.backward(G5, Dg, cams5[v], img, gts5[v], gs.GS_BWD_ACCUMULATE if k else 0)
                if dist is not None:
                    if reduce_impl == "peer":
                        PR.allreduce(D, n5)              # enqueued on the ctx stream behind the backward kernels
                    elif reduce_impl == "nvls":
                        SG.reduce.allreduce(Dg, n5)
                    elif reduce_impl == "nccl":
                        with torch.cuda.stream(self.stream):  # NCCL orders itself after / before the ctx stream: no host sync
                            dist.all_reduce(gbuf)
                Dn.accumulate(Dg, n5)
                A5.step(G5, Dg, LRS)
            return step

        main = args.allreduce
        if main == "auto":
            # measured on this pool (64 MB of gradients): 2 GPUs peer 125 us / in-switch 184 us; 4 GPUs 181 / 161; 8 GPUs 213 / 156
            main = "nvls" if (SG is not None and world > 2) else "peer"
        if main == "nvls" and SG is None:
            main = "peer"
        alts = [a for a in ("nvls", "peer", "nccl") if a != main and (a != "nvls" or SG is not None)]
        runs = [("train", world, main), ("c5", 8, main)]
        if dist is not None:
            runs += [("c5_noreduce", 8, "none")] + [(f"c5_alt_{a}", 8, a) for a in alts] + [(f"train_alt_{a}", world, a) for a in alts[:1]]
        c5 = {}
        for name, vps, impl in runs:
            if vps % world:
                continue
            stepf = make_train_step(vps, impl)
            G5.upload(g5_host); A5.reset(); Dn.reset_accumulator(n5)
            for i in range(3):
                stepf(i)
            ps = []
            ms = self.timed(stepf, K, per_step=ps)
            if name == "train":
                out["train_iters_per_s"] = 1e3 * K / ms
                out["train_views_per_s"] = 1e3 * K * world / ms
                out["train_step_ms_stats"] = spread(ps)
                out["train_config"] = (f"C5 scene: {n5} Gaussians, {W5}x{H5}, 1 view per GPU per step, {R.pair_count} tile pairs, "
                                       "fwd + bwd + accumulateGradients + Adam" + (f", {main} all-reduce of the 64 B/Gaussian packed gradients" if world > 1 else ""))
                c5["train_views_per_s"] = out["train_views_per_s"]
            elif name == "c5":
                out["c5_steps_per_s"] = 1e3 * K / ms
                out["c5_config"] = f"BASELINE config 5: 8 views per step over {world} GPU(s) (strong scaling), one all-reduce per step"
                c5.update({"steps_per_s": 1e3 * K / ms, "views_per_s": 8e3 * K / ms, "ms_per_step": ms / K})
                # bit-equality of the replicas after the timed steps: every rank's parameters hashed, compared on rank 0
                if dist is not None:
                    ctx.sync()
                    hsh = hashlib.sha1(G5.download().tobytes()).hexdigest()
                    allh = [None] * world
                    dist.all_gather_object(allh, hsh)
                    c5["replicas_bit_equal"] = bool(all(h == allh[0] for h in allh))
            elif name == "c5_noreduce":
                c5["allreduce_us_exposed"] = 1e3 * (c5["ms_per_step"] - ms / K)
            else:
                kind, a = name.split("_alt_")
                out[f"{kind}_{'iters' if kind == 'train' else 'steps'}_per_s_{a}_allreduce"] = 1e3 * K / ms
                if kind == "c5":
                    c5[f"steps_per_s_{a}"] = 1e3 * K / ms
        # One GPU only: the 8 views of a step share one parameter set, so two rasterizers can alternate over them with
        # gs_rast_set_overlap (view k+1's stages before the blend under view k's blend kernels), re-armed after every Adam step.
        # Reported beside the plain figure; the multi-GPU runs keep one rasterizer per rank (one view per rank at 8 GPUs).
        if dist is None:
            try:
                Rb = gs.Rasterizer(ctx, n5)
                for v in range(NVT):
                    Rb.forward(G5, cams5[v], img)                         # sizes its pair buffers
                Rb.forward(G5, cams5[0], img); Rb.backward(G5, D, cams5[0], img, gts5[0])   # its visit words exist from here on
                Rs5 = [R, Rb]

                def step_ov(i):
                    for k, v in enumerate(vp.views(i, 8, NVT)):
                        Rs5[k % 2].forward(G5, cams5[v], img)
                        Rs5[k % 2].backward(G5, D, cams5[v], img, gts5[v], gs.GS_BWD_ACCUMULATE if k else 0)
                    Dn.accumulate(D, n5)
                    A5.step(G5, D, LRS)
                    for Rx in Rs5:
                        Rx.set_overlap(True)                              # re-arm: the next forwards wait for this Adam step
                G5.upload(g5_host); A5.reset(); Dn.reset_accumulator(n5)
                for Rx in Rs5:
                    Rx.set_overlap(True)
                for i in range(3):
                    step_ov(i)
                ms = self.timed(step_ov, K, collective=False)
                c5["steps_per_s_two_rasterizers_overlapped"] = 1e3 * K / ms
                for Rx in Rs5:
                    Rx.set_overlap(False)
                Rb.close()
            except Exception as e:
                c5["steps_per_s_two_rasterizers_overlapped"] = "failed: " + repr(e)[:120]
        if dist is not None:
            # correctness of this library's sums: the same per-rank inputs summed by NCCL, compared element for element
            G5.upload(g5_host)
            v0 = vp.views(0, world, NVT)[0]
            par = {}
            for impl in [main] + [a for a in alts if a != "nccl"]:
                Dg = SG.grads if impl == "nvls" else D
                gb = vp.device_tensor(Dg)
                R.forward(G5, cams5[v0], img); R.backward(G5, Dg, cams5[v0], img, gts5[v0])
                ctx.sync()
                want = gb.clone()
                dist.all_reduce(want)
                torch.cuda.synchronize()
                (SG.reduce if impl == "nvls" else PR).allreduce(Dg, n5); ctx.sync()
                err = (gb - want).abs().max().item()
                scale = want.abs().max().item()
                t = torch.tensor([err / max(scale, 1e-30)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
                # every rank must hold the same bits
                hs = [None] * world
                dist.all_gather_object(hs, hashlib.sha1(gb.cpu().numpy().tobytes()).hexdigest())
                par[impl] = {"max_abs_diff_vs_nccl_over_max": float(t.item()), "all_ranks_bit_equal": bool(all(h == hs[0] for h in hs))}
                (SG.reduce if impl == "nvls" else PR).check()
            # the collective alone: 20 back-to-back sums of the 64 MB payload behind a barrier (no compute in between), max over ranks
            pure = {}
            for impl in [main] + alts:
                Dg = SG.grads if impl == "nvls" else D
                gb = vp.device_tensor(Dg)
                gb.zero_()

                def one(i, impl=impl, Dg=Dg, gb=gb):
                    if impl == "nccl":
                        with torch.cuda.stream(self.stream):
                            dist.all_reduce(gb)
                    else:
                        (SG.reduce if impl == "nvls" else PR).allreduce(Dg, n5)
                for i in range(3):
                    one(i)
                pure[impl] = 1e3 * self.timed(one, 20) / 20
            c5["allreduce_us"] = pure
            out["allreduce"] = {"impl": main, "payload_bytes": 64 * n5, "parity": par, "us_per_call": pure,
                                "what": "nvls = gs_dp_allreduce_grads on a symmetric multicast-mapped buffer (own kernel: multimem.ld_reduce in the "
                                        "NVSwitch + multimem.st broadcast); peer = the same entry point over plain NVLink peer loads / stores, sums "
                                        "in rank order; nccl = torch.distributed.all_reduce on the same buffer; parity = one view per rank, each "
                                        "sum against NCCL's of the same inputs"}
            c5["allreduce_parity_max_rel"] = par[main]["max_abs_diff_vs_nccl_over_max"]
            c5["allreduce_ranks_bit_equal"] = par[main]["all_ranks_bit_equal"]
            c5["allreduce_impl"] = main
        out["config"]["c5"] = c5
        for o in gts5 + [A5, G5] + ([PR] if PR is not None else []) + ([SG] if SG is not None else []):
            o.close()
        Dn.set_scene_extent(synth.scene_extent(orbit))
        D.set_stat_mode(False)

        # ---- BASELINE config 4: render-only 3M Gaussians at 3840x2160, 64 views sharded over the ranks
        if not args.no_c4:
            try:
                c4 = self.render_only("C4", 64, reps=max(1, world // 2))
                out["c4_render"] = c4
                out["config"]["c4"] = {k: c4[k] for k in ("frames_per_s", "frames_per_s_one_rasterizer", "e2e_frames_per_s", "ms_per_frame_per_gpu", "pairs", "views", "two_in_flight")}
            except Exception as e:
                out["config"]["c4"] = {"error": repr(e)[:200]}
        clocks = sampler.stop() if rank == 0 else {}    # sampled over every timed region above (fwd+bwd, e2e, training, C4)
        out["clocks"] = clocks
        # kernels per fwd+bwd frame (binned pipeline): preprocess (it also fills the depth sort's histograms), 4 depth passes, scan +
        # coarse emit + bin histogram (one kernel), bins from the histogram (side stream), the bin passes (1; with two passes the
        # run boundaries take 2 more kernels), chunk counts, bin scan + tile scan (one kernel), fill (its CTA 0 writes the blend
        # tile order), blend fwd, blend bwd (two launches: the walk on the forward's visit words, then the pinned walk for the
        # tiles the forward flagged), per-Gaussian chain (memsets not counted): 15 at C2, profiles/r2_launches_C2.md
        launches_fb = 10 + passes + (2 if passes > 5 else 0)
        out["gpu_launches"] = K * launches_fb

        if rank == 0:
            self.diagnostics(out, G, R, img, gti, D, g_host, view, fwd_bwd, clocks, gt_pinned)
            print(json.dumps(out))
        self.barrier()
        for o in (Dn, A, D, gti, img, R, G):
            o.close()
        ctx.close()
        if dist is not None:
            dist.destroy_process_group()

    # ---- rank-0 diagnostics: per-kernel times and rooflines, sort micro-bench, viewer, C3 loop, CPU baseline + parity ----
    def diagnostics(self, out, G, R, img, gti, D, g_host, view, fwd_bwd, clocks, gt_pinned):
        torch, gs, ctx, args = self.torch, self.gs, self.ctx, self.args
        K, world, local = self.K, self.world, self.local
        n, W, H, s0, seed, orbit = workload(args.workload)
        hbm_peak, sm_max_mhz, peak_kind = self.hbm_peak, self.sm_max_mhz, self.peak_kind
        passes = R.sort_passes
        timed = lambda fn, steps: self.timed(fn, steps, collective=False)   # noqa: E731
        R.set_profiling(True)
        acc = None
        for i in range(K):
            fwd_bwd(i)
            t = R.timings()
            acc = t if acc is None else {k: acc[k] + t[k] for k in t}
        R.set_profiling(False)
        stage = {k: v / K for k, v in acc.items()}
        R.forward(G, view(0), img)
        E, C = R.count_evals()
        last = R.get(gs.RAST_LAST_IDX); rng = R.get(gs.RAST_TILE_RANGES)
        P = R.pair_count
        tiles = len(rng)
        # E' = entries the backward re-walks: [start, min(lastContribIdx + 1, start + count)) for every pixel that has a
        # contributor (tiled_shaders.metal:404-407, 427)
        ntx = (W + 15) // 16
        ys, xs = np.nonzero(last != 0xFFFFFFFF)
        tl = (ys // 16) * ntx + xs // 16
        st_, ct_ = rng["start"][tl].astype(np.int64), rng["count"][tl].astype(np.int64)
        Eb = int((np.minimum(last[ys, xs].astype(np.int64) + 1, st_ + ct_) - st_).sum())
        fp32_peak = 2 * 128 * 148 * sm_max_mhz * 1e6 / 1e12          # TFLOP/s at max clock
        dpasses, tpasses = 4, passes - 4
        Pc = R.coarse_pair_count
        bytes_alg = {"depth_sort": n * (4.0 + 16.0 * dpasses), "scan": 12.0 * n,
                     "emit": 8.0 * Pc + 24.0 * n, "sort": Pc * (4.0 + 16.0 * tpasses), "ranges": 4.0 * P + 8.0 * tiles + 72.0 * Pc,
                     "preprocess_bwd": 196.0 * n}
        if stage["scan"] < 0.004:       # fused: the scan runs inside the coarse-emit kernel (look-back over its blocks), no kernel of its own
            bytes_alg.pop("scan")
        flops_alg = {"blend_fwd": 25.0 * C + 14.0 * (E - C), "blend_bwd": 70.0 * C + 14.0 * (Eb - C)}
        roofs = {}
        for k, b in bytes_alg.items():
            a = b / (stage[k] * 1e-3) / 1e9 if stage[k] > 0 else 0.0
            roofs[k] = {"bound": "hbm", "ms": stage[k], "achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak}
        for k, f in flops_alg.items():
            a = f / (stage[k] * 1e-3) / 1e12 if stage[k] > 0 else 0.0
            roofs[k] = {"bound": "fp32", "ms": stage[k], "achieved": a, "peak": fp32_peak, "unit": "TFLOP/s",
                        "frac": a / fp32_peak}
        # K1 is ISSUE bound (about 750 warp instructions per Gaussian of pinned IEEE divides / square roots, issue active 60 %:
        # profiles/r1zf_order_counters.csv), not HBM bound; its 136 B/Gaussian are reported as achieved GB/s for reference only
        a = 136.0 * n / (stage["preprocess"] * 1e-3) / 1e9 if stage["preprocess"] > 0 else 0.0
        roofs["preprocess"] = {"bound": "issue", "ms": stage["preprocess"], "achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak,
                               "note": "instruction-issue bound (pinned IEEE arithmetic); frac is against HBM and is not its roofline"}
        dom = max((k for k in roofs), key=lambda k: stage[k])
        traffic, traffic_src = ncu_traffic({"blend_bwd": "blend_backward_kernel", "blend_fwd": "blend_forward_kernel"}.get(dom, dom)) \
            if args.workload == "C2" else (None, None)
        out["roofline"] = dict(roofs[dom], kernel=dom, traffic=traffic, peak_kind=peak_kind,
                               note="algorithmic flops per SURVEY 8(d): 70*C + 14*(E' - C), E' = entries the backward re-walks; fp32 peak = "
                                    f"2*128*148*sm_max_mhz; traffic = DRAM bytes per launch read from {traffic_src} (ncu --set full)")
        out["rooflines"] = roofs
        out["stage_ms"] = stage
        out["evals"] = {"E": E, "C": C, "E_backward": Eb, "pairs": P, "sort_passes": passes}

        # ---- radix sort Gkeys/s: standalone GPURadixSort64 on full-entropy keys and the real tile|depth keys
        sort_res = {}
        for logn in (20, 23, 24, 26):
            m = 1 << logn
            keys = synth.splitmix64(1234 + logn, m, 5)
            dk, dv = ctx.alloc(m * 8), ctx.alloc(m * 4)
            ctx.upload(dk, keys); ctx.upload(dv, np.arange(m, dtype=np.uint32))
            del keys
            S = gs.Sort64(ctx, m)
            for _ in range(3):
                S.sort(dk, dv, m)
            reps = 20 if logn <= 24 else 5
            ms = timed(lambda i: S.sort(dk, dv, m), reps) / reps
            hybrid = S.last_plan_hybrid
            gs.tune("sort_hybrid", 0)                     # the plain 8-pass LSD plan on the same keys, for comparison
            for _ in range(2):
                S.sort(dk, dv, m)
            ms_lsd = timed(lambda i: S.sort(dk, dv, m), reps) / reps
            gs.tune("sort_hybrid", 1)
            gk = m / (ms * 1e-3) / 1e9
            # P * (8 + 24 p) bytes with p = the digit passes ACTUALLY run (hybrid: the prefix passes + the local pass)
            pre = 3 if logn > 25 else 2
            p_run = (pre + 1) if hybrid else 8
            moved = m * (8.0 + 24.0 * p_run)
            sort_res[f"random_u64_2^{logn}"] = {"gkeys_per_s": gk, "ms": ms, "passes": p_run, "GB/s": moved / (ms * 1e-3) / 1e9,
                                                "frac_of_hbm": moved / (ms * 1e-3) / 1e9 / hbm_peak,
                                                "plan": f"hybrid: {pre} onesweep passes over the top bits + 1 local pass" if hybrid else "8 LSD passes",
                                                "bound_gkeys_per_s_for_this_plan": hbm_peak * 1e9 / (8.0 + 24.0 * p_run) / 1e9,
                                                "lsd_8_pass_gkeys_per_s": m / (ms_lsd * 1e-3) / 1e9,
                                                "lsd_8_pass_frac_of_hbm": m * 200.0 / (ms_lsd * 1e-3) / 1e9 / hbm_peak,
                                                "note": "GB/s and frac_of_hbm count the bytes of the passes that ran (SURVEY 8d, p = passes run); "
                                                        "the 8-pass LSD plan on the same keys is the side note"}
            S.close(); ctx.free(dk); ctx.free(dv)
        order_ms = stage["depth_sort"] + stage["scan"] + stage["emit"] + stage["sort"] + stage["ranges"]
        sort_res["tile_depth_order_C2_binned"] = {
            "gkeys_per_s": P / (order_ms * 1e-3) / 1e9, "ms": order_ms, "passes": passes, "n": P, "coarse_pairs": Pc,
            "what": "everything between the projection and the blend (per-tile lists in depth order + ranges) for the frame's P "
                    "pairs: 4 depth passes on the N Gaussians, coarse pairs, one pass by bin id, per-bin expansion"}
        # the reference's structure for comparison: full 64-bit pair sort of the same keys
        gs.tune("sort_mode", 1)
        R.set_profiling(True)
        acc1 = accall = 0.0
        for i in range(3 + 5):
            R.forward(G, view(i), img)
            if i >= 3:
                t1 = R.timings(False)
                acc1 += t1["sort"]
                accall += t1["depth_sort"] + t1["scan"] + t1["emit"] + t1["sort"] + t1["ranges"]
        R.set_profiling(False)
        p64 = R.sort_passes
        gs.tune("sort_mode", 0)
        R.forward(G, view(0), img)
        sort_res["tile_depth_keys_C2_pairs64"] = {"gkeys_per_s": P / (acc1 / 5 * 1e-3) / 1e9, "ms": acc1 / 5, "passes": p64, "n": P,
                                                  "GB/s": P * (8.0 + 24.0 * p64) / (acc1 / 5 * 1e-3) / 1e9,
                                                  "order_ms": accall / 5,
                                                  "what": "the reference's structure: 64-bit (tile|depth) keys of the P pairs, radix sort over the "
                                                          "varying digits; order_ms = scan + emit + sort + ranges, to compare with the binned line"}
        out["sort"] = sort_res
        out["sort_gkeys_per_s"] = sort_res["random_u64_2^24"]["gkeys_per_s"]
        out["config"]["sort_gkeys_per_s_2^24"] = out["sort_gkeys_per_s"]
        if "train_iters_per_s" in out:
            out["config"]["train_iters_per_s"] = out["train_iters_per_s"]

        # ---- the viewer frame (GPURadixSort32 draw order + the instanced-quad draw of MTLEngine::render as CUDA
        # compute into a headless image) on the same scene; CPU restatement of the same frame beside it
        if not args.no_viewer:
            from gaussiansplatting_b200.layouts import view_uniforms
            S32, Vw = gs.Sort32(ctx, n), gs.Viewer(ctx, n)
            dpos = ctx.alloc(n * 16)
            G.positions(dpos)

            def viewer_frame(i):
                u = view(i)
                S32.sort_depth(dpos, u["cameraPos"][0][:3], n)
                Vw.render(G, S32.sorted_indices, view_uniforms(u), img)
            for i in range(3):
                viewer_frame(i)
            reps = 10
            vms = timed(viewer_frame, reps) / reps
            Vw.render(G, S32.sorted_indices, view_uniforms(view(0)), img, gs.GS_VIEW_COUNT_FRAGMENTS)
            shaded, blended = Vw.fragments()
            out["viewer"] = {"ms_per_frame": vms, "frames_per_s": 1e3 / vms, "pairs": Vw.pair_count, "fragments_shaded": shaded,
                             "fragments_blended": blended, "gfragments_per_s": shaded / (vms * 1e-3) / 1e9,
                             "what": f"{args.workload} scene through gs_sort32_sort_depth + gs_viewer_render (back-to-front, no early "
                                     "exit: every quad fragment is shaded), device time incl. the pair-count read-back"}
            if world == 1 and not args.no_cpu:
                from oracle import orc
                orc.build()
                u0 = view(0)
                t0 = time.time()
                order = orc.viewer_order(g_host, u0["cameraPos"][0][:3])
                orc.viewer_render(g_host, order, orc.view_uniforms(u0), threads=host_threads())
                out["viewer"]["cpu_oracle_ms_per_frame"] = (time.time() - t0) * 1e3
                out["viewer"]["cpu_cores"] = host_threads()
            ctx.free(dpos); Vw.close(); S32.close()

        # ---- BASELINE config 3: training loop with Adam + densify/prune through the C++ host layer (Engine::train,
        # the reference's mtl_engine.mm:1047-1221 schedule time-compressed so that densify / opacity-reset events occur)
        if world == 1 and not args.no_c3:
            from gaussiansplatting_b200 import host
            n3, (W3, H3) = synth.CONFIG_N["C3"], synth.CONFIG_RES["C3"]
            o3, s3, seed3 = synth.CONFIG_ORBIT["C3"], synth.CONFIG_S0["C3"], synth.CONFIG_SEED["C3"]
            V3, ep3 = 16, args.c3_epochs
            poses = [synth.orbit_pose(v, V3, o3) for v in range(V3)]
            cfg = host.TrainConfig(width=W3, height=H3, num_views=V3, epochs=ep3, fx=0.9 * W3, fy=0.9 * W3, cx=W3 / 2, cy=H3 / 2,
                                   scene_extent=synth.scene_extent(o3), densify_from=100, densify_until=10 ** 6,
                                   densify_interval=100, opacity_reset_interval=600, seed=0x5EED, quantize_rgba8=1)
            res, _ = host.train(synth.make_scene(n3, seed3, s3), synth.make_scene(n3, seed3 ^ 0xA5A5, s3),
                                np.array([p[0] for p in poses], np.float32), np.array([p[1] for p in poses], np.float32), cfg, local)
            out["c3"] = {"iters_per_s": res["iterations"] / res["seconds"], "iterations": res["iterations"], "start_gaussians": n3,
                         "final_gaussians": res["final_count"], "densify_events": res["densify_events"],
                         "ms_per_densify_event": 1e3 * res["densify_seconds"] / max(res["densify_events"], 1),
                         "opacity_resets": res["opacity_resets"], "pruned": res["pruned"], "cloned": res["cloned"], "split": res["split"],
                         "what": "500k-Gaussian 1080p training loop (fwd, bwd, accumulateGradients, Adam, densify every 100 its, "
                                 "opacity reset every 600) through the C++ Engine::train; wall clock incl. densify read-backs"}
            out["config"]["c3"] = {k: out["c3"][k] for k in ("iters_per_s", "ms_per_densify_event", "densify_events", "final_gaussians")}

        # ---- CPU baseline beside it (the oracle on the host cores, one full frame of the same workload) and, from the same
        # oracle frame, the parity of the frame this bench measures
        if world == 1 and not args.no_cpu:
            gt_f = gt_pinned.numpy().astype(np.float32) / 255.0
            self.parity_leg(G, R, img, gti, D, g_host, view(0), gt_f, out)

    # ---- `--render-only`: BASELINE config 4 as the headline of its own invocation ---------------------------------
    def run_render_only(self):
        args = self.args
        sampler = ClockSampler(self.local)
        if self.rank == 0:
            sampler.start()
        # the 64-view pool is repeated so that every rank times at least 32 frames (8 frames = 13 ms would be all start-up jitter)
        c4 = self.render_only(args.workload, 64, reps=max(1, args.steps // 16, self.world // 2), try_two_in_flight=True)
        clocks = sampler.stop() if self.rank == 0 else {}
        n, W, H, *_ = workload(args.workload)
        out = {"metric": render_metric(args.workload), "value": c4["frames_per_s"], "unit": "frames/s", "n_gpus": self.world,
               "steps": c4["views"], "warmup": 3, "ms_per_step": c4["ms_per_frame_per_gpu"], "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": c4["workload"], "pairs": c4["pairs"], "two_in_flight": c4["two_in_flight"],
                          "how": c4["how"], "frames_per_s_one_rasterizer": c4["frames_per_s_one_rasterizer"],
                          "parallelism": f"views dealt v mod {self.world}, no communication (BASELINE config 4; mtl_engine.mm:1224-1306)"},
               "e2e": {"value": c4["e2e_frames_per_s"], "unit": "frames/s", "h2d_bytes_per_step": 240, "d2h_bytes_per_step": W * H * 4,
                       "what": "every frame's RGBA8 pixels read back to pinned host memory on the copy-out stream, wall clock"},
               "clocks": clocks,
               # forward-only launches per frame: projection, the depth and bin passes, emit, bins-from-histogram (one bin pass) or
               # run boundaries + chunk map (two), count / scan / fill, blend
               "gpu_launches": c4["views"] * (7 + c4["sort_passes"] + (2 if c4["sort_passes"] > 5 else 0))}
        if self.rank == 0:
            fp32_peak = 2 * 128 * 148 * self.sm_max_mhz * 1e6 / 1e12
            st = c4["stage_ms"]
            out["stage_ms"] = st
            out["roofline"] = {"bound": "fp32", "kernel": "blend_fwd", "ms": st["blend_fwd"], "achieved": c4["blend_fwd_frac_fp32"] * fp32_peak,
                               "peak": fp32_peak, "unit": "TFLOP/s", "frac": c4["blend_fwd_frac_fp32"], "traffic": None}
            out["evals"] = c4["evals"]
            if self.world == 1 and not args.no_cpu:
                from oracle import orc
                orc.build()
                n, W, H, s0, seed, orbit = workload(args.workload)
                g = synth.make_scene(n, seed, s0)
                th = host_threads()
                t = oracle_frame(orc, g, synth.make_camera(0, 64, W, H, n, orbit), None, th, render_only=True)
                out["cpu_baseline"] = {"value": 1.0 / t, "unit": "frames/s", "cores": th, "kind": "port", "sample": "1 full forward frame (project, pairs, CPU LSD sort, ranges, blend)"}
            print(json.dumps(out))
        self.barrier()
        self.ctx.close()
        if self.dist is not None:
            self.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=list(synth.CONFIG_N))
    ap.add_argument("--render-only", action="store_true", help="BASELINE config 4: forward-only frames/s, 64 views sharded over the GPUs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline + parity leg")
    ap.add_argument("--no-c3", action="store_true", help="skip the config-3 training loop")
    ap.add_argument("--no-c4", action="store_true", help="skip the config-4 render-only leg")
    ap.add_argument("--allreduce", default="auto", choices=["auto", "nvls", "peer", "nccl"],
                    help="gradient sum of view-parallel training (N > 1); auto = nvls on more than two GPUs where the box has multicast memory, else peer")
    ap.add_argument("--no-viewer", action="store_true", help="skip the viewer-frame line (SURVEY 8f-3)")
    ap.add_argument("--c3-epochs", type=int, default=100, help="epochs of 16 views for the config-3 loop")
    args = ap.parse_args()
    # the contract is ONE JSON line on stdout: library chatter on fd 1 (e.g. "NCCL version ...") is sent to stderr, and
    # print() keeps writing to the real stdout
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    elif args.render_only:
        Bench(args).run_render_only()
    else:
        Bench(args).run()
    sys.stdout.flush()


if __name__ == "__main__":
    main()
