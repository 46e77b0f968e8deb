#!/usr/bin/env python
"""bench.py -- events/s of the AutoBub3hs trigger hot path on B200 (BASELINE.json metric).

Workload (config[1] of BASELINE.json): "entropy-trigger only: 1000 events x 4 cameras x 40
frames" of 1050x1680 8-bit frames.  One *step* = one pass of the trigger path (K1
ProcessFrame+histogram over every frame pair, K2 significance state machine) over a batch
of `--events-per-step` events resident in HBM (default 125 -> 8 steps = 1000 events; the
batch is 35 GB, so every step re-reads far more than the 126 MB L2: no flush needed).

  value       events/s with the batch resident in HBM (CUDA events on the library stream)
  e2e         events/s through the C ABI from pinned HOST buffers: H2D of every frame +
              K1 + K2 + D2H of the trigger results inside the timed region
  roofline    K1's algorithmic bytes (41 planes per (event,camera)) / its
              measured launch time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline the cv2 oracle (oracle/cv2_oracle.py, a port: the reference cannot be built
              here) on a bounded sample of the same batch, one process per host core

`--impl reference` times the CPU oracle alone (all host cores) on the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS, NCAMS, NFRAMES = 1050, 1680, 4, 40
NPIX = ROWS * COLS
K1_BYTES_PER_EVENT_CAM = (NFRAMES + 1) * NPIX          # every frame once + the sigma plane once
K1_DRAM_TRAFFIC_125 = 37.733456e9 + 8.645e6             # bytes per 125-event launch, measured (profiles/r1_k1_v6_125ev_full.txt)
WORKLOAD = "trigger-only: 1000 events x 4 cams x 40 frames, 1050x1680 u8 (BASELINE.json configs[1])"


# --------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md recipe)
# --------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 7:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------
# synthetic events on the device (same pixel model as autobub3hs_b200.synth, torch RNG)
# --------------------------------------------------------------------------------------
class _DevPtr:
    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "|u1", "data": (ptr, False), "version": 2}


def slot_tensor(torch, ctx, slot):
    return torch.as_tensor(_DevPtr(ctx.event_device_ptr(slot), (NCAMS, NFRAMES, ROWS, COLS)), device="cuda")


def fill_slot(torch, t, bgs, specs_for_event, gen, noise_sigma=2.0):
    """t: u8 [cam, frame, rows, cols] view of a slot.  Renders background + N(0,sigma) noise +
    growing discs, rounds, clips (SURVEY.md 8(d) pixel model)."""
    for c in range(NCAMS):
        img = torch.randn((NFRAMES, ROWS, COLS), generator=gen, device="cuda", dtype=torch.float32)
        img.mul_(noise_sigma).add_(bgs[c])
        for b in specs_for_event.get(c, ()):
            for f in range(b.f_trig, NFRAMES):
                r = b.r0 + b.growth * (f - b.f_trig)
                x0, x1 = max(0, int(b.cx - r - 2)), min(COLS, int(b.cx + r + 3))
                y0, y1 = max(0, int(b.cy - r - 2)), min(ROWS, int(b.cy + r + 3))
                yy = torch.arange(y0, y1, device="cuda", dtype=torch.float32)[:, None]
                xx = torch.arange(x0, x1, device="cuda", dtype=torch.float32)[None, :]
                disc = ((xx - b.cx) ** 2 + (yy - b.cy) ** 2) <= r * r
                img[f, y0:y1, x0:x1] += disc * float(b.amp)
        t[c].copy_(img.round_().clamp_(0, 255).to(torch.uint8))


def event_classes(n, rank):
    out = []
    for i in range(n):
        k = (i + rank) % 20
        out.append("multi" if k == 7 else ("notrigger" if k == 13 else "normal"))
    return out


# --------------------------------------------------------------------------------------
# CPU oracle leg (also the --impl reference arm)
# --------------------------------------------------------------------------------------
_G = {}


def _cpu_one(idx):
    import oracle.cv2_oracle as o
    ev, cam = divmod(idx, NCAMS)
    fr = _G["events"][ev][cam]
    st = o.SignificanceState()
    r = o.find_trigger_frame(o.FrameSource(fr), _G["sigma"][cam], _G["tss"][cam], 1, st, 0)
    return (r.status, r.trig, r.loc_thres)


def cpu_trigger_pass(events, sigmas, tss, cores):
    """One pass of the reference trigger path (early exit at the trigger kept) over `events`
    (list of [cam][N,rows,cols] arrays) with `cores` worker processes.  Returns (seconds, results)."""
    import multiprocessing as mp
    _G["events"], _G["sigma"], _G["tss"] = events, sigmas, tss
    n = len(events) * NCAMS
    ctx = mp.get_context("fork")
    with ctx.Pool(min(cores, n)) as pool:
        pool.map(_cpu_noop, range(min(cores, n)))       # spin the workers up outside the timing
        t0 = time.perf_counter()
        res = pool.map(_cpu_one, range(n), chunksize=1)
        dt = time.perf_counter() - t0
    return dt, res


def _cpu_noop(i):
    import oracle.cv2_oracle  # noqa: F401  (import cost outside the timed region)
    return i


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """--impl reference: the CPU oracle with every host core, bounded sample per step."""
    if rank != 0:
        return
    from autobub3hs_b200 import synth
    import oracle.cv2_oracle as o
    cores = host_cores()
    n_ev = max(1, min(args.ref_events, max(1, cores // NCAMS)))
    cam_masks, bellows = synth.make_masks(ROWS, COLS, NCAMS)
    bgs = [synth.make_background(c, ROWS, COLS) for c in range(NCAMS)]
    classes = event_classes(n_ev, 0)
    events = []
    for ev in range(n_ev):
        es = synth.make_event_spec(ev, classes[ev], cam_masks, bellows, seed=11)
        events.append([synth.render_camera(es, c, bgs[c], seed=11) for c in range(NCAMS)])
    trained = o.train_all(events, NCAMS)
    sig = [t.sigma for t in trained]
    tss = [max(t.training_set_size, 6) for t in trained]
    times = []
    for s in range(args.warmup + args.steps):
        dt, _ = cpu_trigger_pass(events, sig, tss, cores)
        if s >= args.warmup:
            times.append(dt)
    total = sum(times)
    v = n_ev * len(times) / total
    line = {"metric": "events_per_sec", "value": v, "unit": "events/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOAD, "events_per_step": n_ev, "note": "cv2 oracle port of the reference CPU path"},
            "cpu_baseline": {"value": v, "unit": "events/s", "cores": cores, "kind": "port",
                             "sample": f"{n_ev} events x 4 cams x 40 frames per step, early exit at trigger kept"},
            "e2e": {"value": v, "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
def bind_to_gpu_numa_node(local_rank):
    """Pins this process to the CPUs of the NUMA node its GPU hangs off, so that the pinned staging buffers of the
    end-to-end leg are allocated (first touch) in memory local to that GPU's PCIe root.  Best effort: returns the
    node number or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dev = "/sys/bus/pci/devices/" + bus.lower()[-12:]
        node = int(open(dev + "/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from autobub3hs_b200 import capi, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    numa_node = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    B = args.events_per_step
    ctx = capi.Context(ROWS, COLS, NCAMS, NFRAMES, max_slots=B, device=local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)

    # ---- synthetic batch, generated on the device ----
    cam_masks, bellows = synth.make_masks(ROWS, COLS, NCAMS)
    bgs_np = [synth.make_background(c, ROWS, COLS) for c in range(NCAMS)]
    bgs = [torch.from_numpy(b.astype(np.float32)).cuda() for b in bgs_np]
    gen = torch.Generator(device="cuda")
    gen.manual_seed(7919 + rank)
    classes = event_classes(B, rank)
    specs = []
    t_gen = time.perf_counter()
    for s in range(B):
        es = synth.make_event_spec(rank * 100000 + s, classes[s], cam_masks, bellows, seed=11)
        specs.append(es)
        fill_slot(torch, slot_tensor(torch, ctx, s), bgs, es.bubbles, gen, args.noise_sigma)
        ctx.event_set_nframes(s, [NFRAMES] * NCAMS)
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t_gen

    # ---- Trainer (K3 gate + K4 Welford) on frames {0,1} of the batch; not part of the timed step ----
    n_train = min(B, args.train_events)
    kept, sig_np, tss_np = [], [], []
    for c in range(NCAMS):
        ctx.train_begin(c)
        k = 0
        for s in range(n_train):
            _, kp = ctx.train_add_slot(c, s)
            k += int(kp)
        kept.append(k)
        _, sg_plane, tss, status = ctx.train_finish(c)
        sig_np.append(sg_plane)
        tss_np.append(tss)
        if status != 0:
            raise SystemExit(f"training failed for camera {c} (status {status})")
        ctx.set_background_from_training(c)

    def step():
        ctx.trigger_hist(0, B)
        ctx.trigger_scan(0, B)      # synchronises the library stream (start-frame upload)

    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    k1a = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    k1b = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]

    for _ in range(max(args.warmup, 0)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    with ClockSampler(local_rank) as clk:
        torch.cuda.synchronize()
        ev0.record(stream)
        for i in range(args.steps):
            k1a[i].record(stream)
            ctx.trigger_hist(0, B)
            k1b[i].record(stream)
            ctx.trigger_scan(0, B)
        ev1.record(stream)
        torch.cuda.synchronize()
    ms_total = ev0.elapsed_time(ev1)
    k1_ms = float(np.mean([a.elapsed_time(b) for a, b in zip(k1a, k1b)]))
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    trig, status, lt, _ = ctx.trigger_results(0, B, want_sig=False)
    value = world * B * args.steps / (ms_total * 1e-3)

    # ---- the same batch with the reference's early exit (abub_trigger_early): extra key, not the headline ----
    early = None
    if args.early_chunk > 0:
        ctx.trigger_early(0, B, args.early_chunk)
        t_e, s_e, l_e, _ = ctx.trigger_results(0, B, want_sig=False)
        ok = status == 0
        if not (np.array_equal(s_e, status) and np.array_equal(t_e[ok], trig[ok]) and np.array_equal(l_e[status != -9], lt[status != -9])):
            raise SystemExit("abub_trigger_early disagrees with the full-table trigger path")
        n_e = max(1, min(args.steps, 8))
        ee0 = torch.cuda.Event(enable_timing=True)
        ee1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ee0.record(stream)
        for _ in range(n_e):
            ctx.trigger_early(0, B, args.early_chunk)
        ee1.record(stream)
        torch.cuda.synchronize()
        ms_e = ee0.elapsed_time(ee1) / n_e
        ch = args.early_chunk
        need = np.where(ok, np.minimum(trig + 3, NFRAMES), NFRAMES)
        bounds = np.minimum(NFRAMES, np.maximum(5, -(-need // ch) * ch))
        early = {"value": B / (ms_e * 1e-3), "unit": "events/s (this rank)", "ms_per_step": ms_e, "chunk_frames": ch,
                 "frame_pairs_computed_frac": float((bounds - 1).sum() / (B * NCAMS * (NFRAMES - 1))),
                 "note": "K1 in rounds of chunk_frames frame pairs, K2 after every round, decided cameras drop out "
                         "(AnalyzerUnit.cpp:297-300); results equal to the full-table path"}

    # ---- end-to-end through the C ABI from pinned host buffers ----
    Be = min(args.e2e_events, B)
    nf = [NFRAMES] * NCAMS
    host_res = None
    for s in range(Be):
        pin = ctx.pinned(s)
        pin_t = torch.from_numpy(pin)
        pin_t.copy_(slot_tensor(torch, ctx, s).cpu())
    torch.cuda.synchronize()

    nf_arr = np.full((Be, NCAMS), NFRAMES, np.int32)
    sent_total = [0]

    def e2e_step():
        if args.e2e_mode == "stream":
            # frames leave the pinned buffers chunk by chunk; a camera whose trigger search is decided gets no
            # more of them (the reference reads frame by frame and returns at the trigger, AnalyzerUnit.cpp:191-300)
            t_, s_, l_, sent = ctx.trigger_stream(0, Be, nf_arr, args.e2e_chunk)
            sent_total[0] += int(sent.sum())
            return t_, s_, l_, None
        for s in range(Be):
            ctx.event_submit(s, nf)                 # cudaMemcpyAsync H2D of every frame from pinned staging
        ctx.trigger_hist(0, Be)
        ctx.trigger_scan(0, Be)
        sent_total[0] += Be * NCAMS * NFRAMES
        return ctx.trigger_results(0, Be, want_sig=False)   # D2H of the step's result

    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    for _ in range(1):
        e2e_step()
    sent_total[0] = 0
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    ee0 = torch.cuda.Event(enable_timing=True)
    ee1 = torch.cuda.Event(enable_timing=True)
    ee0.record(stream)
    for _ in range(e2e_steps):
        host_res = e2e_step()
    ee1.record(stream)
    torch.cuda.synchronize()
    e2e_wall = time.perf_counter() - t0
    e2e_ms = max(ee0.elapsed_time(ee1), e2e_wall * 1e3)     # host-side D2H waits are part of the path
    if world > 1:
        t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e_value = world * Be * e2e_steps / (e2e_ms * 1e-3)
    # what the link does on one large pinned copy (context for e2e, which is bound by it)
    link_gbs = None
    try:
        pin0 = torch.from_numpy(ctx.pinned(0))
        dst0 = slot_tensor(torch, ctx, 0)
        lk0 = torch.cuda.Event(enable_timing=True)
        lk1 = torch.cuda.Event(enable_timing=True)
        cur = torch.cuda.current_stream()
        dst0.copy_(pin0, non_blocking=True)
        lk0.record(cur)
        for _ in range(4):
            dst0.copy_(pin0, non_blocking=True)
        lk1.record(cur)
        torch.cuda.synchronize()
        link_gbs = 4 * pin0.numel() / (lk0.elapsed_time(lk1) * 1e-3) / 1e9
    except Exception:
        pass
    assert np.array_equal(host_res[1], status[:Be]) and np.array_equal(host_res[0], trig[:Be])

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- CPU baseline on a bounded sample of the same batch + parity check of that sample ----
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores = host_cores()
        n_s = max(1, min(args.cpu_events, B, max(1, cores // NCAMS)))
        sample = [[slot_tensor(torch, ctx, s)[c].cpu().numpy() for c in range(NCAMS)] for s in range(n_s)]
        dt, res = cpu_trigger_pass(sample, sig_np, tss_np, cores)
        mism = 0
        for i, (st_, tr_, lt_) in enumerate(res):
            s, c = divmod(i, NCAMS)
            if st_ != status[s, c] or (st_ == 0 and (tr_ != trig[s, c] or lt_ != lt[s, c])):
                mism += 1
        cpu = {"value": n_s / dt, "unit": "events/s", "cores": cores, "kind": "port",
               "sample": f"{n_s} events x 4 cams x 40 frames of the benchmark batch (D2H), cv2 oracle, "
                         f"one process per (event,camera), early exit at trigger kept; parity mismatches vs GPU: {mism}"}
        if mism:
            raise SystemExit(f"PARITY FAILURE: {mism} (event,camera) results differ from the oracle")
        # the cv2-free C restatement (oracle/abub_ref.c), one thread: ProcessFrame + calcHist of a few frame pairs of
        # the same sample, checked against the GPU's histograms of those pairs
        try:
            from oracle import cref
            ctx.trigger_hist(0, 1)               # the streamed / early-exit passes above leave partial tables
            H_gpu = ctx.trigger_hist_read(0, 0)
            fr = sample[0][0]
            n_pairs = 4
            t0 = time.perf_counter()
            hs = [cref.hist256(cref.process_frame(fr[j], fr[j - 2], sig_np[0])) for j in range(2, 2 + n_pairs)]
            dt_c = (time.perf_counter() - t0) / n_pairs
            bad = sum(0 if np.array_equal(h, H_gpu[2 + i]) else 1 for i, h in enumerate(hs))
            if bad:
                raise SystemExit(f"PARITY FAILURE: {bad} histograms differ from the C restatement")
            cpu["c_port_one_thread"] = {"ms_per_frame_pair": dt_c * 1e3,
                                        "events_per_s_without_early_exit": 1.0 / (dt_c * NCAMS * (NFRAMES - 1)),
                                        "sample": f"{n_pairs} frame pairs of event 0, camera 0; histograms equal to the GPU's"}
        except SystemExit:
            raise
        except Exception as e:
            cpu["c_port_one_thread"] = {"unavailable": repr(e)[:120]}

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    k1_bytes = B * NCAMS * K1_BYTES_PER_EVENT_CAM
    achieved = k1_bytes / (k1_ms * 1e-3) / 1e9
    line = {
        "metric": "events_per_sec", "value": value, "unit": "events/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "events_per_step": B, "l2": "inputs (35 GB/step at 125 events) exceed L2; no flush",
                   "event_mix": f"90% normal (1-3 bubbles), 5% multi (>=20), 5% no-trigger; noise sigma {args.noise_sigma}",
                   "train_events": n_train, "generated_on": "device (torch RNG), seconds=%.1f" % t_gen},
        "e2e": {"value": e2e_value, "unit": "events/s", "h2d_bytes_per_step": sent_total[0] * NPIX // e2e_steps,
                "d2h_bytes_per_step": Be * NCAMS * 16 * (1 if args.e2e_mode == "full" else (NFRAMES + args.e2e_chunk - 1) // args.e2e_chunk),
                "events_per_step": Be, "steps": e2e_steps, "mode": args.e2e_mode, "rank0_numa_node": numa_node,
                "h2d_gbs": sent_total[0] * NPIX / e2e_steps / (e2e_ms / e2e_steps * 1e-3) / 1e9 * 1.0,
                "h2d_link_gbs_one_large_pinned_copy": link_gbs,
                "note": ("abub_trigger_stream: host frames in pinned memory, sent %d frames at a time, a camera stops "
                         "receiving frames once its trigger search is decided (the reference's early exit); "
                         "frames_sent/frames_total = %.3f" % (args.e2e_chunk, sent_total[0] / max(1, e2e_steps * Be * NCAMS * NFRAMES)))
                        if args.e2e_mode == "stream" else "every frame of every event copied H2D, then K1+K2"},
        "gpu_launches": args.steps * 4,      # k1_trigger_tma + k1_trigger (border rows) + k1_fill_bin0 + k2_trigger_scan
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": K1_DRAM_TRAFFIC_125 if B == 125 else None,
                     "traffic_source": "profiles/r1_k1_v6_125ev_full.txt (ncu --set full, dram read+write of one 125-event launch)",
                     "kernel": "k1_trigger_tma_kernel (+ k1_trigger_kernel on the 4 border rows, k1_fill_bin0_kernel)",
                     "ms_per_launch": k1_ms,
                     "algorithmic_bytes_per_launch": k1_bytes,
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s"},
        "cpu_baseline": cpu,
        "clocks": clk.summary(),
        "triggered_fraction": float((status == 0).mean()),
        "early_exit": early,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--events-per-step", type=int, default=125)
    ap.add_argument("--train-events", type=int, default=25)
    ap.add_argument("--e2e-events", type=int, default=8)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-mode", default="stream", choices=["stream", "full"])
    ap.add_argument("--e2e-chunk", type=int, default=4)
    ap.add_argument("--early-chunk", type=int, default=6, help="frame pairs per round of abub_trigger_early (0 = skip)")
    ap.add_argument("--cpu-events", type=int, default=8)
    ap.add_argument("--ref-events", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--noise-sigma", type=float, default=2.0,
                    help="camera noise of the synthetic frames (grey levels); 2.0 is the workload of the headline number")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
