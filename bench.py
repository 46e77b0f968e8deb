#!/usr/bin/env python
"""bench.py -- events/s of the AutoBub3hs per-event image path on B200 (BASELINE.json metric).

Workload = BASELINE.json configs[3]: "full recon run (trigger + L3Localizer + CCL bubble boxes) of 10,000 synthetic
events ... with PICOFormatWriterV4 output", 4 cameras x 40 frames of 1050x1680 8-bit frames per event.  10,000
distinct events are 2.8 TB, so a RING of `--ring-events` (32) distinct synthetic events is rendered
once on the host (numpy PCG64, autobub3hs_b200.synth -- the same pixels in both arms) and reused: the `--events-per-step`
(125) HBM slots of a set hold ring event (slot % ring), two such sets are used in turn (--slot-sets), and each step
announces the other set so that its trigger search is queued inside the current step's localizer call
(abubhost_recon_hint); 10,000 events = 80 steps, the default run times 20 of them.

One *step* = the whole reference event loop (AutoBubStart3.cpp:342-388) over the resident slots through the host
library (libabubhost.so = the abub3hs detect loop): K1+K2 trigger search with the reference's early exit, K5-K7
localizer for every triggered camera, isInMask / tracking / bubble bookkeeping on the host cores, OutputWriter rows
appended to abub3hs_<run>.txt and flushed per event.  Every step re-reads 35 GB of frames (>> 126 MB L2: no flush).

  value        events/s, frames resident in HBM when the timed region starts (wall clock between synchronisations:
               the host part of the path is inside)
  e2e          the same from pinned HOST buffers: frames cross PCIe inside the timed region (cudaMemcpyAsync, per camera
               only as far as the trigger search and the localizer read), results and blob records come back
  roofline     dominant kernel (K1) + `kernels`: device time per kernel group (CUDA events on the library stream)
               against the algorithmic bytes of SURVEY.md 8(d) and MEASURED_PEAKS.json hbm_gbs
  cpu_baseline the cv2 oracle (a port; the reference's own sources compiled against API stand-ins, oracle/_ref, are much
               slower and are timed from files in the reference arm) on a bounded sample of the same ring, every host core
               busy; its output text is the parity check of the GPU's text
  trainer      (N > 1) BASELINE.json configs[2]: 200 training events sharded over the ranks, exact integer partials,
               ncclAllReduce + all-gather fix-up (abub_train_allreduce / abub_train_finish_allreduced), planes compared
               bit for bit with the single-GPU Trainer
  trigger_only BASELINE.json configs[1] as a named extra (K1 full table + K2, round-1 headline)

`--impl reference` times the CPU oracle alone (all host cores) on the same workload, plus oracle/_ref/abub3hs_ref from
BMP files (reference_binary_from_files).
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS, COLS, NCAMS, NFRAMES = 1050, 1680, 4, 40
NPIX = ROWS * COLS
K1_BYTES_PER_EVENT_CAM = (NFRAMES + 1) * NPIX          # full table: every frame once + the sigma plane once
LOC_PLANES_MAX = 15                                     # localizer: 3 genesis planes + (10 + 2) post-trigger planes
RUN_ID = "20210101_0"
SEED = 11
WORKLOAD = ("full recon: trigger + L3Localizer + PICOFormatWriterV4 rows, 4 cams x 40 frames of 1050x1680 u8 per event "
            "(BASELINE.json configs[3]; ring of distinct synthetic events reused)")


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


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# --------------------------------------------------------------------------------------
# the ring of synthetic events (host, numpy PCG64): identical in both arms and on every rank
# --------------------------------------------------------------------------------------
def ring_classes(n):
    out = []
    for i in range(n):
        k = i % 20
        out.append("multi" if k == 7 else ("notrigger" if k == 13 else "normal"))
    return out


def render_ring(first, count, total, workers):
    """Events [first, first+count) of the ring of `total` events: SyntheticRun (events[i][cam] = [40,1050,1680] u8)."""
    from autobub3hs_b200 import synth
    return synth.make_run(count, ring_classes(total), rows=ROWS, cols=COLS, seed=SEED, run_id=RUN_ID, workers=workers,
                          first_event=first)


def write_masks(run, mask_dir):
    from autobub3hs_b200 import synth
    os.makedirs(mask_dir, exist_ok=True)
    for c in range(NCAMS):
        synth.write_bmp8(os.path.join(mask_dir, f"cam{c}_mask.bmp"), run.cam_masks[c])
        synth.write_bmp8(os.path.join(mask_dir, f"cam{c}_bellows_mask.bmp"), run.bellows_masks[c])


# --------------------------------------------------------------------------------------
# CPU oracle leg (cpu_baseline of the GPU arm and the --impl reference arm)
# --------------------------------------------------------------------------------------
_G = {}


def _cpu_noop(i):
    import cv2
    import oracle.cv2_oracle  # noqa: F401  (import cost outside the timed region)
    cv2.setNumThreads(1)
    return i


def _cpu_task(idx):
    """AnyCamAnalysis (AutoBubStart3.cpp:67-125) of one (event, camera): trigger search with the early exit, localizer,
    tracking.  Returns (CameraOutput, seconds)."""
    import oracle.cv2_oracle as o
    t0 = time.perf_counter()
    slot, cam = divmod(idx, NCAMS)
    ev = _G["order"][slot]
    a = o.any_cam_analysis(o.FrameSource(_G["events"][ev][cam]), _G["trained"][cam], _G["cam_masks"][cam], _G["bellows"][cam],
                           _G["numbers"][slot], True, None)
    return a.out, time.perf_counter() - t0


class CpuRecon:
    """The reference's event loop on the host cores: one worker process per core (cv2 single-threaded inside a worker,
    like one OpenMP thread per event in AutoBubStart3.cpp:342), tasks = (event, camera) handed out one at a time."""

    def __init__(self, run, trained, cores):
        import multiprocessing as mp
        self.cores = cores
        _G.update(events=run.events, trained=trained, cam_masks=run.cam_masks, bellows=run.bellows_masks)
        self.n_ring = len(run.events)
        self.pool = None
        self.mp = mp.get_context("fork")

    def start(self, order, numbers):
        _G["order"], _G["numbers"] = list(order), list(numbers)
        self.pool = self.mp.Pool(self.cores)                 # forked after _G is complete
        self.pool.map(_cpu_noop, range(self.cores * 2))

    def one_pass(self):
        """One pass over the sample.  Returns (seconds, utilisation, text)."""
        import oracle.cv2_oracle as o
        n = len(_G["order"]) * NCAMS
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_task, range(n), chunksize=1)
        outs = [r[0] for r in res]
        text = "".join(o.write_camera_output(RUN_ID, outs[s * NCAMS:(s + 1) * NCAMS], 30) for s in range(len(_G["order"])))
        dt = time.perf_counter() - t0
        util = sum(r[1] for r in res) / (self.cores * dt)
        return dt, util, text

    def close(self):
        if self.pool:
            self.pool.close()
            self.pool.join()
            self.pool = None


def check_timed_rows(out_path, cpu_text, B, R, base_no):
    """Every row the warm-up and timed steps wrote, against the oracle's rows: the event with number base_no + step*B + s was
    read from slot s of a set = ring event s % R, and the oracle's sample event i (i < R) is ring event i -- so apart from
    the event number a step's rows must be the oracle's rows of that ring event, in order.  Returns (events, rows) checked."""
    exp = {}
    for line in cpu_text.splitlines():
        p = line.split(None, 2)
        if len(p) == 3 and int(p[1]) < R:
            exp.setdefault(int(p[1]), []).append(p[2])
    got, order = {}, []
    with open(out_path) as f:
        for k, line in enumerate(f):
            if k < 6:
                continue
            p = line.rstrip("\n").split(None, 2)
            ev = int(p[1])
            if ev not in got:
                got[ev] = []
                order.append(ev)
            got[ev].append(p[2])
    if order != sorted(order):
        raise SystemExit("PARITY FAILURE: the rows of the timed steps are not in event order")
    for ev in order:
        ring = ((ev - base_no) % B) % R
        if ring in exp and got[ev] != exp[ring]:
            raise SystemExit(f"PARITY FAILURE: rows of timed event {ev} (ring event {ring}) differ from the oracle's: "
                             f"{got[ev][:1]} != {exp[ring][:1]} ({len(got[ev])} vs {len(exp[ring])} rows)")
    return len(order), sum(len(v) for v in got.values())


def cpu_sample_size(cores, ring):
    """>= 2 events (8 tasks) per worker, at least the whole ring (the arm must not be starved)."""
    return max(ring, 2 * cores)


def run_reference(args, rank, world):
    """--impl reference: the CPU oracle with every host core on the same workload; rank 0 alone."""
    if rank != 0:
        return
    import oracle.cv2_oracle as o
    cores = host_cores()
    R = args.ring_events
    t0 = time.perf_counter()
    run = render_ring(0, R, R, cores)
    t_render = time.perf_counter() - t0
    trained = o.train_all(run.events[:min(R, args.train_events)], NCAMS)
    n_s = cpu_sample_size(cores, R)
    cpu = CpuRecon(run, trained, cores)
    cpu.start([i % R for i in range(n_s)], list(range(n_s)))
    times, utils = [], []
    for s in range(args.warmup + args.steps):
        dt, util, _ = cpu.one_pass()
        if s >= args.warmup:
            times.append(dt)
            utils.append(util)
    cpu.close()
    total = sum(times)
    v = n_s * len(times) / total
    line = {"metric": "events_per_sec", "value": v, "unit": "events/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOAD, "events_per_step": n_s, "ring_events": R, "train_events": min(R, args.train_events),
                       "note": "cv2 oracle port of the reference CPU path (trigger with early exit + localizer + tracking + rows); "
                               "ring rendered in %.1f s" % t_render},
            "cpu_baseline": {"value": v, "unit": "events/s", "cores": cores, "kind": "port",
                             "worker_utilisation": float(np.mean(utils)),
                             "sample": f"{n_s} events x 4 cams x 40 frames per step ({n_s * NCAMS} (event,camera) tasks on {cores} "
                                       f"worker processes, handed out one at a time), early exit at the trigger kept"},
            "e2e": {"value": v, "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    rb = reference_binary_leg(run, cores, args.ref_binary_events)
    if rb:
        line["reference_binary_from_files"] = rb
    print(json.dumps(line), flush=True)


def reference_binary_leg(run, cores, n):
    """The reference's own compiled code from image FILES: oracle/_ref/abub3hs_ref (the unmodified first-party sources
    built -O0 -fopenmp as CMakeLists.txt:35 does, against oracle/refshim) on the first n ring events written as 8-bit BMP:
    file read + decode + training pass + detect loop + output file, OMP_NUM_THREADS = the host cores.  The pixel routines
    behind its cv:: calls are the plain C of oracle/abub_ref.c, not OpenCV's SIMD code, which is why this figure is far
    below the cv2 port above and is reported beside it, not as the baseline."""
    exe = os.path.join(ROOT, "oracle", "_ref", "abub3hs_ref")
    if n <= 0 or not os.path.exists(exe):
        return None
    import dataclasses
    from autobub3hs_b200 import synth
    n = min(n, len(run.events))
    sub = dataclasses.replace(run, events=run.events[:n], event_numbers=run.event_numbers[:n], specs=run.specs[:n])
    tmp = tempfile.mkdtemp(prefix="abub_refbin_")
    try:
        synth.write_run(sub, os.path.join(tmp, "data"), os.path.join(tmp, "masks"))
        os.makedirs(os.path.join(tmp, "out"))
        nbytes = sum(os.path.getsize(os.path.join(dp, f)) for dp, _, fs in os.walk(os.path.join(tmp, "data")) for f in fs)
        t0 = time.perf_counter()
        r = subprocess.run([exe, "-D", "40l-19", "-d", os.path.join(tmp, "data"), "-r", run.run_id, "-o", os.path.join(tmp, "out"),
                            "-c", os.path.join(tmp, "masks")], capture_output=True, text=True, timeout=1200,
                           env=dict(os.environ, OMP_NUM_THREADS=str(cores)))
        dt = time.perf_counter() - t0
        if r.returncode != 0:
            return {"error": (r.stdout + r.stderr)[-300:]}
        rows = sum(1 for _ in open(os.path.join(tmp, "out", f"abub3hs_{run.run_id}.txt"))) - 6
        return {"value": n / dt, "unit": "events/s", "events": n, "seconds": dt, "cores": cores, "kind": "reference",
                "rows_written": rows, "bytes_read": nbytes,
                "what": "oracle/_ref/abub3hs_ref from 8-bit BMP files: read + decode + Trainer pass + detect loop + output file"}
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
class _DevPtr:
    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "|u1", "data": (ptr, False), "version": 2}


def slot_tensor(torch, ctx, slot):
    return torch.as_tensor(_DevPtr(ctx.event_device_ptr(slot), (NCAMS, NFRAMES, ROWS, COLS)), device="cuda")


def bind_to_gpu_numa_node(local_rank):
    """Pins this process to the CPUs of the NUMA node its GPU hangs off (first-touch placement of the pinned staging
    buffers).  Best effort: returns (node or None, why)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dev = "/sys/bus/pci/devices/" + bus.lower()[-12:]
        node = int(open(dev + "/numa_node").read())
        if node < 0:
            return None, "sysfs reports numa_node -1 for " + bus
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None, f"no allowed CPU on node {node}"
        n_nodes = len([d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")])
        if n_nodes > 1:
            os.sched_setaffinity(0, cpus)
        return node, f"{n_nodes} NUMA node(s) visible"
    except Exception as e:
        return None, repr(e)[:100]


def load_traffic():
    """DRAM bytes per launch of the profiled kernels, written by tools/ncu_summary.py from the committed ncu captures."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
    except Exception:
        return {}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from autobub3hs_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    cores = host_cores()
    R, B = args.ring_events, args.events_per_step
    if R % world:
        raise SystemExit("--ring-events must be a multiple of the number of ranks")
    # ---- ring: every rank renders its share on its share of the host cores (before CUDA is touched: fork) ----
    share = R // world
    t0 = time.perf_counter()
    run = render_ring(rank * share, share, R, max(1, cores // world))
    t_render = time.perf_counter() - t0

    # ---- CPU baseline first (rank 0, N = 1): the oracle trains and detects on the same ring ----
    cpu, cpu_text, n_s = None, None, 0
    if world == 1 and not args.no_cpu_baseline:
        import oracle.cv2_oracle as o
        trained_cpu = o.train_all(run.events[:min(R, args.train_events)], NCAMS)
        n_s = cpu_sample_size(cores, R)
        cr = CpuRecon(run, trained_cpu, cores)
        cr.start([i % R for i in range(n_s)], list(range(n_s)))
        cr.one_pass()                                        # warm-up
        dt, util, cpu_text = cr.one_pass()
        cr.close()
        cpu = {"value": n_s / dt, "unit": "events/s", "cores": cores, "kind": "port", "worker_utilisation": util,
               "sample": f"{n_s} events x 4 cams x 40 frames of the ring ({n_s * NCAMS} (event,camera) tasks on {cores} worker "
                         f"processes, one at a time), cv2 oracle: trigger with early exit + localizer + tracking + rows"}

    torch.cuda.set_device(local_rank)
    numa_node, numa_why = bind_to_gpu_numa_node(local_rank) if world > 1 else (None, "single rank")
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    # --slot-sets sets of B slots: step i works on set i % sets, so the host stage of one step's last sub-batch never holds
    # the slots the next step needs (abub3hs does the same: 2 x batch slots)
    SETS = max(1, args.slot_sets)
    hr = capi.HostRecon(ROWS, COLS, NCAMS, NFRAMES, max_slots=SETS * B, device=local_rank, early_chunk=args.early_chunk)
    ctx = hr.ctx
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)
    tmp = tempfile.mkdtemp(prefix=f"abub_bench_r{rank}_")
    mask_dir = os.path.join(tmp, "masks")
    write_masks(run, mask_dir)

    # ---- slots: ring event (slot % R); the other ranks' shares arrive over NCCL ----
    for i in range(share):
        for c in range(NCAMS):
            ctx.event_upload(rank * share + i, c, run.events[i][c])
    if world > 1:
        mine = torch.stack([slot_tensor(torch, ctx, rank * share + i) for i in range(share)])
        full = torch.empty((world,) + tuple(mine.shape), dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(full, mine)
        for r in range(world):
            if r != rank:
                for i in range(share):
                    slot_tensor(torch, ctx, r * share + i).copy_(full[r, i])
        del full, mine
    for s in range(R, SETS * B):
        slot_tensor(torch, ctx, s).copy_(slot_tensor(torch, ctx, (s % B) % R))
    for s in range(SETS * B):
        ctx.event_set_nframes(s, [NFRAMES] * NCAMS)
    torch.cuda.synchronize()

    # ---- Trainer (K3 gate + K4 Welford) on frames {0,1} of the first --train-events ring events; not in the timed step ----
    n_train = min(R, args.train_events)
    tss = []
    for c in range(NCAMS):
        ctx.train_begin(c)
        ctx.train_add_slots(c, list(range(n_train)))
        _, _, n, status = ctx.train_finish(c)
        if status != 0:
            raise SystemExit(f"training failed for camera {c} (status {status})")
        ctx.set_background_from_training(c)
        tss.append(n)

    out_dir = os.path.join(tmp, "out")
    os.makedirs(out_dir)
    host_threads = args.host_threads or max(1, cores // world - 1)   # one core stays with the thread that drives the device
    hr.set_threads(host_threads, args.sub_batch)                     # (torchrun exports OMP_NUM_THREADS=1)
    out_path = hr.begin_run(RUN_ID, out_dir, mask_dir, 30, tss)
    nf_all = np.full((B, NCAMS), NFRAMES, np.int32)
    base_no = rank * 10_000_000
    step_no = [0]

    def step():
        numbers = [base_no + step_no[0] * B + s for s in range(B)]
        first = (step_no[0] % SETS) * B
        step_no[0] += 1
        if SETS > 1 and not args.no_hint:
            # the set the NEXT step works on: its trigger search is queued inside this step's localizer call, so the GPU has work
            # during the localizer's host round trips and between two steps (every timed step launches exactly one such
            # search -- the next step's -- and finds its own already launched by the step before)
            hr.recon_hint((step_no[0] % SETS) * B, B, nf_all)
        hr.recon(first, numbers, nf_all)

    for _ in range(max(args.warmup, 0)):
        step()
    hr.read_stats()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        torch.cuda.synchronize()
        l0 = ctx.launch_count()
        t0 = time.perf_counter()
        ev0.record(stream)
        for i in range(args.steps):
            step()
        hr.drain()                                        # the rows of the last events are in the file
        ev1.record(stream)
        torch.cuda.synchronize()
        wall_ms = (time.perf_counter() - t0) * 1e3
        launches = ctx.launch_count() - l0
    dev_ms = ev0.elapsed_time(ev1)
    ms_total = max(wall_ms, dev_ms)                      # host analysis and the file appends are inside the path
    stage = hr.read_stats()
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    value = world * B * args.steps / (ms_total * 1e-3)
    rows_written = sum(1 for _ in open(out_path)) - 6

    # ---- parity of the output text: ring events 0..n_s-1 (slot s holds ring event s % R) against the oracle's text ----
    parity = None
    if cpu_text is not None:
        hr.end_run()
        p_dir = os.path.join(tmp, "parity")
        os.makedirs(p_dir)
        p_path = hr.begin_run(RUN_ID, p_dir, mask_dir, 30, tss)
        if n_s > B:
            raise SystemExit("--events-per-step is smaller than the CPU sample")
        hr.recon(0, list(range(n_s)), nf_all[:n_s])
        hr.drain()
        hr.end_run()
        import oracle.cv2_oracle as o
        got = open(p_path).read()
        want = o.write_header() + cpu_text
        if got != want:
            gl, wl = got.splitlines(), want.splitlines()
            bad = next((i for i, (a, b) in enumerate(zip(gl, wl)) if a != b), min(len(gl), len(wl)))
            raise SystemExit(f"PARITY FAILURE: abub3hs text differs from the oracle's at line {bad}: {gl[bad:bad+1]} != {wl[bad:bad+1]}")
        parity = {"events": n_s, "rows": len(want.splitlines()) - 6, "text_identical": True}
        # ... and every row of the warm-up and timed steps themselves (the path with the trigger searches queued ahead)
        ev_chk, rows_chk = check_timed_rows(out_path, cpu_text, B, R, base_no)
        parity["timed_steps_events_checked"] = ev_chk
        parity["timed_steps_rows_checked"] = rows_chk
        cpu["sample"] += f"; GPU output text of the same {n_s} events identical ({parity['rows']} rows)"
        hr.begin_run(RUN_ID, out_dir, mask_dir, 30, tss)
    hr.read_stats()

    # ---- per-kernel device time (two more steps with CUDA events around every kernel group) ----
    trig, status, lt, _ = ctx.trigger_results(0, B, want_sig=False)
    ok = status == 0
    npost = np.array([min(10 if t < 29 else 39 - t, NFRAMES - 1 - t) for t in trig.reshape(-1)]).reshape(trig.shape)
    n_trig = int(ok.sum())
    loc_frames = int((npost[ok] + 1).sum())
    ch = max(args.early_chunk, 1)
    need = np.where(ok, np.minimum(trig + 3, NFRAMES), NFRAMES)
    bounds = np.minimum(NFRAMES, np.maximum(5, -(-need // ch) * ch)) if args.early_chunk > 0 else np.full_like(need, NFRAMES)
    pairs = int((bounds - 1).sum())
    ctx.profile(True)
    nprof = 2
    for _ in range(nprof):
        step()
    prof = ctx.profile_read()
    ctx.profile(False)
    hr.read_stats()
    if args.profile_step:
        # one more step between cudaProfilerStart / Stop: `ncu --profile-from-start off` sees exactly one step's launches
        # (tools/round_profile.sh -> tools/ncu_traffic.py -> profiles/r2_traffic.json)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step()
        hr.drain()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        hr.read_stats()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = load_traffic()
    # algorithmic bytes per step of every kernel group; groups without their own algorithmic bytes
    # (candidate scan, sort, border tracing: scratch traffic) count against the localizer stage
    alg = {
        "K1": (pairs + B * NCAMS * 2) * NPIX,                 # every frame read once (pairs + 1) + the sigma plane, per camera
        "K5a": 3 * n_trig * NPIX,                             # trigger frame, pre-trigger frame, sigma
        "K5b": (loc_frames - n_trig + 2 * n_trig) * NPIX,     # post frames + mean + sigma per triggered camera
    }
    kernels, loc_ms = [], 0.0
    for name, ms, calls in prof:
        key = name.split()[0]
        ms_step = ms / nprof
        if key in ("K5a", "K5b", "K5c", "K6", "cub", "K7a", "K7b", "K7c"):
            loc_ms += ms_step
        if calls == 0:
            continue
        a = alg.get(key, 0)
        ent = {"kernel": name, "ms_per_step": ms_step, "launch_groups_per_step": calls / nprof,
               "algorithmic_bytes_per_step": a, "achieved": (a / (ms_step * 1e-3) / 1e9) if a and ms_step > 0 else None,
               "frac": (a / (ms_step * 1e-3) / 1e9 / peak) if a and ms_step > 0 else None,
               "traffic": traffic.get(key)}
        kernels.append(ent)
    loc_bytes = (3 * n_trig + (loc_frames - n_trig) + 2 * n_trig) * NPIX
    kernels.append({"kernel": "localizer stage (K5a+K5b+K5c+K6+sort+K7, device time)", "ms_per_step": loc_ms,
                    "algorithmic_bytes_per_step": loc_bytes,
                    "achieved": loc_bytes / (loc_ms * 1e-3) / 1e9 if loc_ms > 0 else None,
                    "frac": loc_bytes / (loc_ms * 1e-3) / 1e9 / peak if loc_ms > 0 else None, "traffic": None})
    k1 = next((k for k in kernels if k["kernel"].startswith("K1")), None)

    # ---- configs[1] as a named extra: K1 over every frame pair + K2 (the round-1 headline) ----
    trig_only = None
    if args.trigger_steps > 0:
        ctx.trigger_hist(0, B)
        ctx.trigger_scan(0, B)
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k1a = [torch.cuda.Event(enable_timing=True) for _ in range(args.trigger_steps)]
        k1b = [torch.cuda.Event(enable_timing=True) for _ in range(args.trigger_steps)]
        a0.record(stream)
        for i in range(args.trigger_steps):
            k1a[i].record(stream)
            ctx.trigger_hist(0, B)
            k1b[i].record(stream)
            ctx.trigger_scan(0, B)
        a1.record(stream)
        torch.cuda.synchronize()
        t_ms = a0.elapsed_time(a1) / args.trigger_steps
        k1_ms = float(np.mean([a.elapsed_time(b) for a, b in zip(k1a, k1b)]))
        k1_bytes = B * NCAMS * K1_BYTES_PER_EVENT_CAM
        t2, s2, l2, _ = ctx.trigger_results(0, B, want_sig=False)
        if not (np.array_equal(s2, status) and np.array_equal(t2[ok], trig[ok]) and np.array_equal(l2[status != -9], lt[status != -9])):
            raise SystemExit("PARITY FAILURE: the early-exit trigger search disagrees with the full-table search")
        trig_only = {"workload": "trigger-only: 1000 events x 4 cams x 40 frames (BASELINE.json configs[1]), all 39 frame pairs",
                     "value": B / (t_ms * 1e-3), "unit": "events/s (this rank)", "ms_per_step": t_ms,
                     "k1_full_table": {"ms_per_launch": k1_ms, "algorithmic_bytes_per_launch": k1_bytes,
                                       "achieved": k1_bytes / (k1_ms * 1e-3) / 1e9, "frac": k1_bytes / (k1_ms * 1e-3) / 1e9 / peak,
                                       "traffic": traffic.get("K1_full")},
                     "results_equal_to_early_exit": True}

    # ---- end to end from pinned host buffers ----
    Be = min(args.e2e_events, B)
    for s in range(Be):
        pin = torch.from_numpy(ctx.pinned(s))
        pin.copy_(slot_tensor(torch, ctx, s).cpu())
    torch.cuda.synchronize()
    nf_e = nf_all[:Be]
    hr.set_threads(host_threads, args.e2e_sub_batch)

    def e2e_step():
        numbers = [base_no + 5_000_000 + step_no[0] * Be + s for s in range(Be)]
        step_no[0] += 1
        hr.recon(0, numbers, nf_e, pinned=True, chunk_frames=args.e2e_chunk)

    for _ in range(2):
        e2e_step()
    hr.read_stats()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    hr.drain()
    torch.cuda.synchronize()
    e2e_ms_local = (time.perf_counter() - t0) * 1e3
    e2e_stage = hr.read_stats()
    e2e_ms = e2e_ms_local
    h2d_bytes_step = e2e_stage["frames_sent"] * NPIX / args.e2e_steps
    h2d_gbs_local = h2d_bytes_step * args.e2e_steps / (e2e_ms_local * 1e-3) / 1e9
    h2d_ranks = [h2d_gbs_local]
    if world > 1:
        t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
        g = torch.zeros(world, device="cuda", dtype=torch.float64)
        g[rank] = h2d_gbs_local
        dist.all_reduce(g)
        h2d_ranks = [float(v) for v in g.tolist()]
    e2e_value = world * Be * args.e2e_steps / (e2e_ms * 1e-3)
    # what the link does on one large pinned copy (context for e2e, which is bound by it)
    link_gbs = None
    try:
        pin0 = torch.from_numpy(ctx.pinned(0))
        dst0 = slot_tensor(torch, ctx, 0)
        lk0, lk1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
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
    # the same copy by every rank at once: what the box's host memory system gives N GPUs together (the ceiling of e2e)
    link_all = None
    if world > 1 and link_gbs is not None:
        try:
            dist.barrier()
            torch.cuda.synchronize()
            lk0.record(cur)
            for _ in range(8):
                dst0.copy_(pin0, non_blocking=True)
            lk1.record(cur)
            torch.cuda.synchronize()
            g = torch.zeros(world, device="cuda", dtype=torch.float64)
            g[rank] = 8 * pin0.numel() / (lk0.elapsed_time(lk1) * 1e-3) / 1e9
            dist.all_reduce(g)
            per = [float(v) for v in g.tolist()]
            link_all = {"per_rank_min": min(per), "per_rank_max": max(per), "aggregate": float(sum(per)),
                        "what": "every rank copies one 282 MB pinned event buffer to its GPU 8 times, all ranks at once"}
        except Exception:
            pass
    hr.end_run()

    # ---- configs[2]: the Trainer sharded over the ranks, one NCCL all-reduce + all-gather fix-up (last: it replaces the planes) ----
    trainer = None
    if world > 1 and args.trainer_events > 0:
        trainer = trainer_leg(args, ctx, torch, dist, rank, world, R)

    if rank == 0:
        line = {
            "metric": "events_per_sec", "value": value, "unit": "events/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "events_per_step": B, "ring_events": R,
                       "events_timed": world * B * args.steps, "note_10000": "10,000 events = %d steps of %d events" % (-(-10000 // B), B),
                       "l2": "inputs (35 GB/step at 125 events) exceed L2; no flush",
                       "event_mix": "90% normal (1-3 bubbles), 5% multi (>=20), 5% no-trigger; noise sigma 2.0",
                       "train_events": n_train, "early_chunk": args.early_chunk, "sub_batch": args.sub_batch, "slot_sets": SETS, "host_threads": host_threads,
                       "rendered_on": "host (numpy PCG64, %d of %d ring events on this rank, %.1f s)" % (share, R, t_render)},
            "stage_seconds_per_step": {k: v / args.steps for k, v in stage.items() if k != "frames_sent"},
            "device_ms_per_step": dev_ms / args.steps,
            "rows_written": rows_written,
            "triggered_fraction": float(ok.mean()),
            "e2e": {"value": e2e_value, "unit": "events/s", "h2d_bytes_per_step": int(h2d_bytes_step),
                    "d2h_bytes_per_step": int(Be * NCAMS * 16 * (-(-NFRAMES // args.e2e_chunk)) + 88 * rows_written / max(1, args.steps * B) * Be * 11),
                    "events_per_step": Be, "steps": args.e2e_steps, "seconds_timed": e2e_ms * 1e-3,
                    "frames_sent_frac": e2e_stage["frames_sent"] / (args.e2e_steps * Be * NCAMS * NFRAMES),
                    "h2d_gbs_per_rank": {"min": min(h2d_ranks), "median": float(np.median(h2d_ranks)), "max": max(h2d_ranks)},
                    "h2d_gbs_aggregate": float(sum(h2d_ranks)),
                    "h2d_link_gbs_one_large_pinned_copy": link_gbs, "h2d_all_ranks_copying_at_once_gbs": link_all, "rank0_numa_node": numa_node, "numa_note": numa_why,
                    "stage_seconds_per_step": {k: v / args.e2e_steps for k, v in e2e_stage.items() if k != "frames_sent"},
                    "note": "abubhost_recon mode 1: frames wait in pinned host memory and are sent %d at a time; a camera stops "
                            "receiving frames at its trigger decision, the localizer's frames (up to trig+10) follow; trigger "
                            "results and blob records are read back; rows appended per event" % args.e2e_chunk},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": k1["achieved"] if k1 else None, "peak": peak, "unit": "GB/s",
                         "frac": k1["frac"] if k1 else None, "traffic": traffic.get("K1"),
                         "traffic_source": traffic.get("source"),
                         "kernel": "k1_trigger_tma_kernel (+ border rows, bin 0) in rounds of %d frame pairs (early exit)" % args.early_chunk,
                         "ms_per_step": k1["ms_per_step"] if k1 else None,
                         "algorithmic_bytes_per_step": alg["K1"],
                         "frame_pairs_computed_frac": pairs / (B * NCAMS * (NFRAMES - 1)),
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                         "kernels": kernels},
            "cpu_baseline": cpu,
            "parity": parity,
            "clocks": clk.summary(),
            "trigger_only": trig_only,
            "trainer": trainer,
        }
        print(json.dumps(line), flush=True)
    hr.close()
    shutil.rmtree(tmp, ignore_errors=True)
    if world > 1:
        dist.destroy_process_group()


def trainer_leg(args, ctx, torch, dist, rank, world, R):
    """BASELINE.json configs[2]: `--trainer-events` training events (event i = ring event i % R) sharded by contiguous
    range; per camera abub_train_partials -> abub_train_allreduce -> abub_train_finish_allreduced, compared bit for
    bit with rank 0's single-GPU abub_train_finish over all the events."""
    E = args.trainer_events
    lo, hi = E * rank // world, E * (rank + 1) // world
    idfile = os.path.join(tempfile.gettempdir(), "abub_bench_%s.ncclid" % os.environ.get("MASTER_PORT", "0"))
    comm = ctx.nccl_comm_create(idfile, rank, world)
    ctx.nccl_barrier(comm)                     # first collective on the communicator: connection set-up, not timed
    # pass 0 is the warm-up (partial-sum planes, gather scratch and NCCL channels are set up on first use), pass 1 is timed
    for timed in (False, True):
        t_gate = t_part = t_ar = t_fix = 0.0
        flagged, planes = [], []
        dist.barrier()
        for c in range(NCAMS):
            ctx.train_begin(c)
            t0 = time.perf_counter()
            ctx.train_add_slots(c, [i % R for i in range(lo, hi)])
            ctx.sync()
            t1 = time.perf_counter()
            ctx.train_partials(c)
            ctx.sync()
            t2 = time.perf_counter()
            ctx.train_allreduce(c, comm)
            ctx.sync()
            t3 = time.perf_counter()
            mean, sigma, n, st, nfl = ctx.train_finish_allreduced(c, comm)
            t4 = time.perf_counter()
            t_gate += t1 - t0; t_part += t2 - t1; t_ar += t3 - t2; t_fix += t4 - t3
            flagged.append(nfl)
            planes.append((mean, sigma, n, st))
    tt = torch.tensor([t_gate, t_part, t_ar, t_fix], device="cuda", dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_gate, t_part, t_ar, t_fix = [float(v) for v in tt.tolist()]
    identical, t_single = None, None
    if rank == 0:
        identical = True
        t_single = 0.0
        for c in range(NCAMS):
            t0 = time.perf_counter()
            ctx.train_begin(c)
            ctx.train_add_slots(c, [i % R for i in range(E)])
            m1, s1, n1, st1 = ctx.train_finish(c)
            t_single += time.perf_counter() - t0                 # like the sharded form: planes read back to the host
            m, s, n, st = planes[c]
            if not (n1 == n and st1 == st and np.array_equal(m1, m) and np.array_equal(s1, s)):
                identical = False
    ok = torch.tensor([1 if (identical is None or identical) else 0], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    ctx.nccl_barrier(comm)
    ctx.nccl_comm_destroy(comm)
    if rank != 0:
        return None
    if not bool(ok.item()):
        raise SystemExit("PARITY FAILURE: the all-reduced Trainer planes differ from the single-GPU Trainer's")
    nkept = planes[0][2]
    return {"workload": "Trainer background statistics over %d training events sharded over %d ranks (BASELINE.json configs[2])" % (E, world),
            "planes_identical": True, "kept_frames_per_cam": nkept,
            "gate_keep_ms": 1e3 * t_gate, "partials_ms": 1e3 * t_part, "allreduce_ms": 1e3 * t_ar, "fixup_ms": 1e3 * t_fix,
            "train_ms": 1e3 * (t_gate + t_part + t_ar + t_fix),
            "single_gpu_train_ms_rank0": 1e3 * t_single,
            "allreduce_bytes_per_cam": NPIX * 12 + 4,
            "flagged_pixel_fraction": float(np.mean(flagged)) / NPIX,
            "allgather_bytes_per_cam": float(np.mean(flagged)) * nkept,
            "collective": "ncclAllReduce(sum) of u32 sum / u64 sum-of-squares / count + ncclAllGather of the flagged pixels' samples "
                          "(csrc/nccl_glue.cu), communicator from abub_nccl_comm_create (ncclCommInitRank)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--events-per-step", type=int, default=125)
    ap.add_argument("--ring-events", type=int, default=32)
    ap.add_argument("--train-events", type=int, default=25)
    ap.add_argument("--early-chunk", type=int, default=5, help="frame pairs per round of the trigger search's early exit (0 = full table)")
    ap.add_argument("--host-threads", type=int, default=0, help="threads of the host event loop (0 = cores per rank - 1)")
    ap.add_argument("--slot-sets", type=int, default=2, help="sets of --events-per-step slots the steps alternate between")
    ap.add_argument("--sub-batch", type=int, default=125, help="events handed from the device stage to the host stage at a time")
    ap.add_argument("--e2e-events", type=int, default=32)
    ap.add_argument("--e2e-sub-batch", type=int, default=16)
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--e2e-chunk", type=int, default=4)
    ap.add_argument("--trigger-steps", type=int, default=8, help="steps of the trigger-only extra (0 = skip)")
    ap.add_argument("--trainer-events", type=int, default=200, help="training events of the sharded Trainer leg at N > 1 (0 = skip)")
    ap.add_argument("--ref-binary-events", type=int, default=8, help="--impl reference: events the compiled reference binary runs from BMP files (0 = skip)")
    ap.add_argument("--no-hint", action="store_true", help="do not announce the next step's slots (abubhost_recon_hint)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-step", action="store_true", help="run one extra step between cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    args = ap.parse_args()
    args.e2e_steps = max(1, args.e2e_steps)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
