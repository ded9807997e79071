#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200-native dense-stereo hot path.

Metric (BASELINE.json): Gpix-disp/s = H*W*D / s for SSD + SGM + WTA at 1920x1080, D=128.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--frames F] [--impl b200|reference]

A *step* is one pass of the hot path (SSD k=7 -> Semiglobal P1=.1 P2=.2 -> WTA, through the
reference-shaped Python API over the C ABI) over a batch of F distinct synthetic 1080p
stereo pairs per GPU.  Frames are sharded across ranks with no collective (weak scaling).
One JSON line is printed by rank 0:
  value     device-resident throughput (inputs already in HBM), CUDA events, max over ranks
  e2e       same metric with HOST buffers through stereomatch.streaming.HostFrameStreamer:
            pinned H2D of both images and D2H of the disparity map inside the timed
            region, every frame (copies overlap the previous/next frame's compute)
  roofline  dominant stage (SGM) against the measured HBM copy bandwidth: algorithmic bytes over
            the MEDIAN stand-alone stage time (each stage warmed up first), every pass listed
  stages    SSD / SGM / WTA stand-alone, median CUDA-event time, fraction of the roofline
  configs   BASELINE.json configs 2, 3, 4 (one frame each, per stage and whole frame)
  config5   the 64-frame 1080p D=192 batch sharded over the ranks of this run (strong scaling)
  cpu_baseline  the unmodified reference CPU ops (oracle/_ref) on a bounded sample
`--impl reference` times the reference's own CPU implementation (rank 0 only).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "stereo-matching_b200"))

H, W, D = 1080, 1920, 128
KERNEL_SIZE, P1, P2 = 7, 0.1, 0.2
METRIC = "Gpix-disp/s (H*W*D) SSD+SGM+WTA 1080p D=128"
# algorithmic (compulsory) bytes per pixel-disparity, SURVEY.md 8(d) / DESIGN.md
ALGO_BYTES = {"ssd": 4.0 + 8.0 / D, "sgm": 8.0 + 4.0 / D, "wta": 4.0 + 4.0 / D}
STAGE_WARM = 3   # untimed launches of every stand-alone stage before its timed ones


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if (t0 is None or t0 <= t <= t1)] or [r for _, r in self.rows]
        mhz, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                mhz.append(float(r[1]))
                mx = float(r[2])
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        mhz.sort()
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(mhz)}


def make_frames(seeds):
    from oracle import synth
    return [synth.random_dot_pair(H, W, D, seed=s, scale="raw")[:2] for s in seeds]


# ---------------------------------------------------------------------------------------
# reference CPU arm / cpu_baseline
# ---------------------------------------------------------------------------------------
def cpu_reference_sample(rows: int, seed: int = 1000):
    """SSD+SGM+WTA of the reference's own CPU ops on a `rows` x 1920 x 128 crop of the
    workload; returns (seconds, pix-disp, kind, cores)."""
    import numpy as np
    from oracle import ref_ops, oracle as O, synth
    left, right, _ = synth.random_dot_pair(rows, W, D, seed=seed, scale="raw")
    if ref_ops.available():
        import torch
        torch.set_num_threads(1)  # the reference CPU path is single-threaded as shipped
        ops, kind = ref_ops, "reference"
    else:
        ops, kind = None, "port"
    t0 = time.perf_counter()
    if ops is not None:
        cost = ops.ssd(left, right, D, KERNEL_SIZE)
        agg = ops.sgm(cost, left, P1, P2)
        disp = ops.wta(agg)
    else:
        cost = O.ssd(left, right, D, KERNEL_SIZE)
        agg = O.sgm(cost, left, P1, P2)
        disp = O.wta(agg)
    dt = time.perf_counter() - t0
    assert disp.shape == (rows, W) and np.isfinite(agg[:, D:, :]).all()
    return dt, rows * W * D, kind, 1


# stdout carries exactly ONE JSON line.  Native libraries print there too (NCCL's "NCCL version ..."
# banner when NCCL_DEBUG is set on the box), so file descriptor 1 is pointed at stderr for the
# whole run and the line is written to the saved descriptor.
_JSON_FD = None


def _claim_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    rows = args.ref_rows
    for _ in range(args.warmup):
        cpu_reference_sample(rows)
    t_total, px_total, kind, cores = 0.0, 0, "port", 1
    for s in range(args.steps):
        dt, px, kind, cores = cpu_reference_sample(rows, seed=1000 + s)
        t_total += dt
        px_total += px
    value = px_total / t_total / 1e9
    sample = f"{rows}x{W}x{D} crop (rows of the 1080p workload) per step, SSD k=7 + SGM + WTA"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Gpix-disp/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_total / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "1920x1080 D=128 SSD(k=7)+SGM(P1=.1,P2=.2)+WTA random-dot stereogram",
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": "Gpix-disp/s", "cores": cores, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "Gpix-disp/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)
    return 0


# ---------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    import stereomatch
    from stereomatch import _cstereomatch as C
    from stereomatch import sharding

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    F = args.frames
    # frame f of rank r: seed 1000 + r*F + f (frame-sharded batch, no collective)
    host = make_frames(sharding.frame_seeds(1000, F, rank))
    pinned = [(torch.from_numpy(l).pin_memory(), torch.from_numpy(r).pin_memory()) for l, r in host]
    dev_frames = [(l.to(dev), r.to(dev)) for l, r in pinned]
    out_host = [torch.empty(H, W, dtype=torch.int32).pin_memory() for _ in range(F)]

    ssd = stereomatch.cost.SSD(D, KERNEL_SIZE)
    sgm = stereomatch.aggregation.Semiglobal(P1, P2)
    wta = stereomatch.disparity_reduce.WinnerTakesAll()
    pipe = stereomatch.Pipeline(ssd, wta, sgm)

    # the library's batch API: consecutive frames alternate between two Pipelines on two streams
    # (each owns its volumes), which overlaps one frame's issue-bound cost stage with the other's
    # HBM-bound aggregation passes; every frame runs exactly Pipeline.estimate
    pipe2 = stereomatch.Pipeline(stereomatch.cost.SSD(D, KERNEL_SIZE), stereomatch.disparity_reduce.WinnerTakesAll(),
                                 stereomatch.aggregation.Semiglobal(P1, P2))
    batch = stereomatch.streaming.BatchEstimator([pipe, pipe2], dev)
    dev_out = [torch.empty(H, W, dtype=torch.int32, device=dev) for _ in range(F)]

    def step_resident():
        batch.estimate_batch(dev_frames, dev_out)

    streamer = stereomatch.streaming.HostFrameStreamer(pipe, dev, extra_pipelines=[pipe2])

    def step_e2e():
        # the call a user with host frames makes: pinned H2D of both images and D2H of the
        # disparity map every frame, inside the timed region (overlapped with compute)
        streamer.run(pinned, out_host)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        t1 = time.perf_counter()
        return sharding.max_over_ranks(e0.elapsed_time(e1), dev), t0, t1

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = C.launch_count()
    ms, t0, t1 = timed(step_resident, args.steps)
    launches = int(sharding.sum_over_ranks(C.launch_count() - launches0, dev))  # whole job
    clocks = sampler.stop(t0, t1) if rank == 0 else None

    for _ in range(max(1, args.warmup // 2)):
        step_e2e()
    ms_e2e, _, _ = timed(step_e2e, args.steps)

    # ---- per-stage device times (CUDA events on the launching stream), for the roofline ----
    # Every stand-alone stage is launched `STAGE_WARM` times untimed first (the variants timed
    # here -- unfused SGM, the WTA kernel -- are not the ones Pipeline.estimate ran above: their
    # first launch pays module loading and a new per-stream workspace), then timed `reps` times,
    # one event pair per launch, and the MEDIAN is reported.
    reps = max(5, min(args.steps, 20))

    def stage_ms(fn):
        for _ in range(STAGE_WARM):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        return ts[len(ts) // 2]

    def pass_times(sgm_op, cost, img, out):
        """Median CUDA-event time of every pass of one stand-alone aggregation (the library
        records the events on the launching stream: stm_sgm_pass_timing)."""
        acc = {}
        C.AggregationOps.pass_timing(True)
        try:
            for _ in range(STAGE_WARM + reps):
                sgm_op(cost, img, out)
                for k, (d, t) in enumerate(C.AggregationOps.last_pass_ms()):
                    acc.setdefault((k, d), []).append(t)
        finally:
            C.AggregationOps.pass_timing(False)
        return {kd: sorted(ts[STAGE_WARM:])[len(ts[STAGE_WARM:]) // 2] for kd, ts in acc.items()}

    peak, peak_kind = measured_peak()

    def stage_record(ms_, algo_bytes_per_pd, n_pd):
        gbs = algo_bytes_per_pd * n_pd / (ms_ * 1e-3) / 1e9
        return {"ms": round(ms_, 4), "algo_GBs": round(gbs, 1), "frac": round(gbs / peak, 4),
                "Gpix_disp_s": round(n_pd / (ms_ * 1e-3) / 1e9, 2)}

    N = H * W * D
    l, r = dev_frames[0]
    cost = ssd(l, r, cost_volume=pipe._cost_volume)
    vol = pipe._aggregation_volume
    ms_ssd = stage_ms(lambda: ssd(l, r, cost_volume=cost))
    ms_sgm = stage_ms(lambda: sgm(cost, l, vol))
    agg = sgm(cost, l, vol)
    ms_wta = stage_ms(lambda: wta(agg, dev_out[0]))
    # what Pipeline.estimate runs for SGM + WTA: one call, argmin in the last pass, no write-back
    ms_fused = stage_ms(lambda: sgm.aggregate_and_reduce(cost, l, vol, dev_out[0], keep_volume=False))
    sgm_launches0 = C.launch_count()
    sgm(cost, l, vol)
    sgm_launches = C.launch_count() - sgm_launches0
    pass_ms = pass_times(sgm, cost, l, vol)
    torch.cuda.synchronize()
    stage_t = {"ssd": ms_ssd, "sgm": ms_sgm, "wta": ms_wta}
    stages = {k: stage_record(v, ALGO_BYTES[k], N) for k, v in stage_t.items()}
    stages["sgm_wta_fused_call"] = {"ms": round(ms_fused, 4),
                                    "note": "stm_sgm_wta_f32 + STM_SGM_DISPARITY_ONLY, what Pipeline.estimate runs"}
    ms_single = stage_ms(lambda: pipe.estimate(l, r))
    stages["frame_single"] = {"ms": round(ms_single, 4), "Gpix_disp_s": round(N / (ms_single * 1e-3) / 1e9, 2),
                              "note": "Pipeline.estimate, ONE frame in flight, device-resident (latency of a frame)"}

    passes = []
    fused = len(pass_ms) == 4  # horizontal passes + two fused vertical sweeps (sgm_vsweep.cu)
    overlapped = len(pass_ms) == 2  # STM_SGM_OVERLAP: (-> <- on a helper stream || down sweep), up sweep
    for (k, d), t in sorted(pass_ms.items()):
        if overlapped:
            # phase 0: -> writes S (8 B), <- read-modify-writes S (12 B), down sweep reads C and
            # writes T (8 B); phase 1: up sweep reads C, S, T and writes S (16 B)
            b = H * W * D * 4 * (7 if k == 0 else 4) + H * W * 4 * (3 if k == 0 else 1)
            label = "0,1 || 2+4+6" if k == 0 else "3+5+7 (+T)"
            passes.append({"dir": label, "ms": round(t, 4), "bytes": b, "GBs": round(b / (t * 1e-3) / 1e9, 1),
                           "frac_of_peak": round(b / (t * 1e-3) / 1e9 / peak, 3)})
            continue
        b = sgm_pass_bytes(d, k == 0, H, W, D, fused_sweep=fused and k >= 2)
        label = {2: "2+4+6", 3: "3+5+7"}[d] if fused and k >= 2 else str(d)
        passes.append({"dir": label, "ms": round(t, 4), "bytes": b, "GBs": round(b / (t * 1e-3) / 1e9, 1),
                       "frac_of_peak": round(b / (t * 1e-3) / 1e9 / peak, 3)})
    sum_passes = sum(p_["ms"] for p_ in passes)
    # the stage time and the sum of its passes are two measurements of the same thing
    passes_agree = abs(sum_passes - ms_sgm) <= 0.10 * ms_sgm + 0.05

    # ---- the other BASELINE.json configs, one frame each (parity: tests/test_gpu_named_shapes.py) ----
    def run_config(Hc, Wc, Dc, cost_kind, reducer, label, l2_note=None):
        from oracle import synth
        lh, rh, _ = synth.random_dot_pair(Hc, Wc, Dc, seed=1000, scale="raw")
        lc, rc = torch.from_numpy(lh).to(dev), torch.from_numpy(rh).to(dev)
        n_pd = Hc * Wc * Dc
        cost_op = stereomatch.cost.SSD(Dc, KERNEL_SIZE) if cost_kind == "ssd" else stereomatch.cost.Birchfield(Dc)
        rec = {"shape": f"{Wc}x{Hc} D={Dc}", "stages": {}}
        cvol = cost_op(lc, rc)
        rec["stages"][cost_kind] = stage_record(stage_ms(lambda: cost_op(lc, rc, cost_volume=cvol)),
                                                4.0 + 8.0 / Dc, n_pd)
        if reducer == "wta":
            sg = stereomatch.aggregation.Semiglobal(P1, P2)
            wt = stereomatch.disparity_reduce.WinnerTakesAll()
            avol = sg(cvol, lc)
            rec["stages"]["sgm"] = stage_record(stage_ms(lambda: sg(cvol, lc, avol)), 8.0 + 4.0 / Dc, n_pd)
            dimg = wt(avol)
            rec["stages"]["wta"] = stage_record(stage_ms(lambda: wt(avol, dimg)), 4.0 + 4.0 / Dc, n_pd)
            p_ = stereomatch.Pipeline(cost_op, wt, sg)
            p_._cost_volume, p_._aggregation_volume, p_._disparity_image = cvol, avol, dimg
            algo = (4.0 + 8.0 / Dc) + (8.0 + 4.0 / Dc) + (4.0 + 4.0 / Dc)
        else:
            algo_dp = 5.0 + (8.0 * Hc * Dc + 5.0 * Hc * Wc) / n_pd
            for mode in ("cuda", "cpu_exact"):
                dpo = stereomatch.disparity_reduce.DynamicProgramming(mode=mode)
                dimg = dpo(cvol)
                rec["stages"]["dp" if mode == "cuda" else "dp_cpu_exact"] = stage_record(
                    stage_ms(lambda: dpo(cvol, dimg)), algo_dp, n_pd)
            p_ = stereomatch.Pipeline(cost_op, stereomatch.disparity_reduce.DynamicProgramming())
            p_._cost_volume = cvol
            algo = (4.0 + 8.0 / Dc) + algo_dp
        frame = stage_ms(lambda: p_.estimate(lc, rc))
        rec["frame"] = stage_record(frame, algo, n_pd)
        rec["frame"]["note"] = "Pipeline.estimate, one frame in flight, device-resident"
        if l2_note:
            rec["l2"] = l2_note
        rec["workload"] = label
        torch.cuda.synchronize()
        return rec

    def run_config1(Hc, Wc, Dc, label, l2_note=None):
        """BASELINE.json config 1 in its GPU form: SSD + WTA without aggregation -- two stages with the
        volume in HBM, and the single pass that never writes it (stm_ssd_wta_f32, what Pipeline runs)."""
        from oracle import synth
        lh, rh, _ = synth.random_dot_pair(Hc, Wc, Dc, seed=1000, scale="raw")
        lc, rc = torch.from_numpy(lh).to(dev), torch.from_numpy(rh).to(dev)
        n_pd = Hc * Wc * Dc
        cost_op = stereomatch.cost.SSD(Dc, KERNEL_SIZE)
        wt = stereomatch.disparity_reduce.WinnerTakesAll()
        cvol = cost_op(lc, rc)
        dimg = wt(cvol)
        rec = {"shape": f"{Wc}x{Hc} D={Dc}", "stages": {}}
        rec["stages"]["ssd"] = stage_record(stage_ms(lambda: cost_op(lc, rc, cost_volume=cvol)), 4.0 + 8.0 / Dc, n_pd)
        rec["stages"]["wta"] = stage_record(stage_ms(lambda: wt(cvol, dimg)), 4.0 + 4.0 / Dc, n_pd)
        p_ = stereomatch.Pipeline(cost_op, wt)
        p_._disparity_image = dimg
        p_.fuse_cost_and_reduce = "always"   # Pipeline's own rule folds above 48 MB only (L2-resident volumes lose)
        fused = stage_ms(lambda: p_.estimate(lc, rc))
        rec["frame"] = {"ms": round(fused, 4), "Gpix_disp_s": round(n_pd / (fused * 1e-3) / 1e9, 2),
                        "hbm_bytes_per_pixel": 28,
                        "note": "one pass, no volume (stm_ssd_wta_f32); Pipeline.estimate's default above 48 MB; device-resident"}
        p_.fuse_cost_and_reduce = False
        p_._cost_volume = cvol
        two = stage_ms(lambda: p_.estimate(lc, rc))
        rec["frame_two_stages"] = {"ms": round(two, 4), "Gpix_disp_s": round(n_pd / (two * 1e-3) / 1e9, 2)}
        if l2_note:
            rec["l2"] = l2_note
        rec["workload"] = label
        torch.cuda.synchronize()
        return rec

    configs = {}
    if not args.no_configs:
        configs["config1_gpu"] = run_config1(375, 450, 64, "450x375 D=64 SSD+WTA, no aggregation (GPU form of config 1)",
                                             l2_note="L2-resident (43 MB volume): not HBM-roofline evidence")
        configs["config1_gpu_1080p"] = run_config1(H, W, D, "1920x1080 D=128 SSD+WTA, no aggregation")
        configs["config2"] = run_config(375, 450, 64, "ssd", "wta", "450x375 D=64 SSD+SGM+WTA",
                                        l2_note="L2-resident (43 MB volumes): not HBM-roofline evidence")
        configs["config3"] = run_config(H, W, D, "birchfield", "wta", "1920x1080 D=128 Birchfield+SGM+WTA")
        configs["config4"] = run_config(H, W, D, "ssd", "dp", "1920x1080 D=128 SSD+DP (default and cpu_exact)")

    # ---- config 5: the 64-frame D=192 video batch, frame-sharded over the ranks of THIS run ----
    config5 = None
    if not args.no_configs:
        from oracle import synth
        D5, F5 = 192, args.config5_frames
        b5, e5 = sharding.shard_range(F5, rank, world)
        mine = [synth.random_dot_pair(H, W, D5, seed=1000 + f, scale="raw")[:2] for f in range(b5, e5)]
        dev5 = [(torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)) for a, b in mine]
        out5 = [torch.empty(H, W, dtype=torch.int32, device=dev) for _ in dev5]
        mk5 = lambda: stereomatch.Pipeline(stereomatch.cost.SSD(D5, KERNEL_SIZE),
                                           stereomatch.disparity_reduce.WinnerTakesAll(),
                                           stereomatch.aggregation.Semiglobal(P1, P2))
        batch5 = stereomatch.streaming.BatchEstimator([mk5(), mk5()], dev)
        for _ in range(2):
            batch5.estimate_batch(dev5[:4], out5[:4])
        ms5, _, _ = timed(lambda: batch5.estimate_batch(dev5, out5), 1)
        ms5b, _, _ = timed(lambda: batch5.estimate_batch(dev5, out5), 1)
        ms5 = min(ms5, ms5b)
        n5 = H * W * D5
        config5 = {"workload": f"{F5} x (1920x1080 D={D5}) SSD+SGM+WTA, frames [{b5},{e5}) on rank 0 of {world}",
                   "frames": F5, "n_gpus": world, "ms_per_batch": round(ms5, 3),
                   "ms_per_frame_per_gpu": round(ms5 / max(e5 - b5, 1), 4),
                   "Gpix_disp_s": round(F5 * n5 / (ms5 * 1e-3) / 1e9, 2),
                   "scaling": "strong (fixed 64-frame batch split over the ranks), no collective",
                   "timing": "CUDA events, barrier both sides, max over ranks, best of 2 batches"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    total_px = N * F * world * args.steps
    value = total_px / (ms * 1e-3) / 1e9
    e2e = total_px / (ms_e2e * 1e-3) / 1e9
    dom = max(stage_t, key=stage_t.get)
    achieved = ALGO_BYTES[dom] * N / (stage_t[dom] * 1e-3) / 1e9
    traffic = None
    tp = ROOT / "profiles" / "traffic.json"
    if tp.exists():
        try:
            traffic = json.loads(tp.read_text()).get(dom)
        except Exception:
            traffic = None

    cpu_baseline = None
    if not args.no_cpu_baseline:
        dt, px, kind, cores = cpu_reference_sample(args.cpu_rows)
        cpu_baseline = {"value": px / dt / 1e9, "unit": "Gpix-disp/s", "cores": cores, "kind": kind,
                        "sample": f"{args.cpu_rows}x{W}x{D} crop of frame 0, SSD+SGM+WTA, "
                                  f"{dt:.1f} s on 1 of {os.cpu_count()} host cores"}

    line = {
        "metric": METRIC, "value": value, "unit": "Gpix-disp/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "ms_per_frame": ms / args.steps / F, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "1920x1080 D=128 SSD(k=7)+SGM(P1=.1,P2=.2, reference path set; 2 horizontal "
                               "passes next to the down sweep, then the up sweep)+WTA, "
                               "random-dot stereograms", "frames_per_gpu_per_step": F,
                   "frames_in_flight": 2,
                   "sharding": f"frame-sharded x{world}, no collective",
                   "l2": "working set per frame (2 x 1.06 GB volumes) exceeds the 126 MB L2"},
        "e2e": {"value": e2e, "unit": "Gpix-disp/s",
                "h2d_bytes_per_step": 2 * H * W * 4 * F, "d2h_bytes_per_step": H * W * 4 * F,
                "ms_per_frame": ms_e2e / args.steps / F},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": f"{dom} stage ({sgm_launches} launches)" if dom == "sgm" else dom,
                     "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "stage_ms": round(stage_t[dom], 4), "algo_bytes": ALGO_BYTES[dom] * N,
                     "timing": f"median of {reps} launches after {STAGE_WARM} untimed, CUDA events",
                     # what each pass of the stage HAS to move (C read, S read-modify-write;
                     # the first pass only writes) over its own duration
                     "passes": passes, "sum_passes_ms": round(sum_passes, 4),
                     "passes_agree_with_stage": bool(passes_agree)},
        "stages": stages,
        "configs": configs,
        "config5": config5,
        "cpu_baseline": cpu_baseline,
    }
    _emit(line)
    if world > 1:
        dist.destroy_process_group()
    if not passes_agree:
        print(f"bench.py: sum of SGM passes {sum_passes:.3f} ms vs stage {ms_sgm:.3f} ms disagree",
              file=sys.stderr)
    return 0


def sgm_pass_bytes(direction, first, H, W, D, fused_sweep=False):
    """Bytes one pass must move: read C, read and write S (the first pass only writes S), read
    the intensity, for the pixels the pass visits (the defective up-diagonals 5 and 7 skip a
    band of a landscape image; a fused sweep of three directions visits every
    pixel once)."""
    import numpy as np
    if fused_sweep:
        return H * W * D * 4 * 3 + H * W * 4
    y, x = np.mgrid[0:H, 0:W]
    active = np.ones((H, W), dtype=bool)
    if direction == 5 and W != H:
        c, cs = x - y, W - H
        active = ~((c >= 0) & (c < cs)) if W > H else ~((c > cs) & (c <= -1))
    elif direction == 7 and W > H:
        s_ = x + y
        active = ~((s_ >= H) & (s_ <= W - 1))
    px = int(active.sum())
    return px * D * 4 * (2 if first else 3) + px * 4


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--frames", type=int, default=4, help="stereo pairs per GPU per step")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-rows", type=int, default=160, help="rows of the cpu_baseline crop")
    ap.add_argument("--ref-rows", type=int, default=64, help="rows per step of --impl reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs 2-5 block")
    ap.add_argument("--config5-frames", type=int, default=64, help="frames of the D=192 batch (config 5)")
    args = ap.parse_args()
    _claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
