#!/usr/bin/env python
"""Per-stage timings of the path on one B200 (not the headline bench: bench.py is):
  trainer    K3 gate + K4 Welford over `--train-events` events (BASELINE.json configs[2] shape on 1 GPU)
  localizer  K5-K7 over every triggered (event, camera) of a resident batch
  recon      K1+K2 followed by the localizer = the device part of a full recon (configs[3])
Each line carries the algorithmic bytes of SURVEY.md 8(d) and the GB/s they imply.  Wall-clock
around synchronising C-ABI calls (the localizer has host round trips inside, they are part of it)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402

NCAMS, NFRAMES, ROWS, COLS = B.NCAMS, B.NFRAMES, B.ROWS, B.COLS


# synthetic events rendered on the device (torch RNG; same pixel model as autobub3hs_b200.synth): fast set-up for stage timings
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





def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--events", type=int, default=32)
    ap.add_argument("--train-events", type=int, default=200)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--png-events", type=int, default=8, help="events of the PNG decode stage (0 = skip)")
    ap.add_argument("--only-png", action="store_true", help="only the PNG decode stage")
    ap.add_argument("--kernels", action="store_true", help="device time per kernel group of one localizer pass and one trigger pass (abub_profile)")
    args = ap.parse_args()
    import torch
    from autobub3hs_b200 import capi, synth
    from tests import glue

    E = args.events
    ctx = capi.Context(B.ROWS, B.COLS, B.NCAMS, B.NFRAMES, max_slots=E)
    cam_masks, bellows = synth.make_masks(B.ROWS, B.COLS, B.NCAMS)
    bgs = [torch.from_numpy(synth.make_background(c, B.ROWS, B.COLS).astype(np.float32)).cuda() for c in range(B.NCAMS)]
    gen = torch.Generator(device="cuda")
    gen.manual_seed(7919)
    classes = event_classes(E, 0)
    for s in range(E):
        es = synth.make_event_spec(s, classes[s], cam_masks, bellows, seed=11)
        fill_slot(torch, B.slot_tensor(torch, ctx, s), bgs, es.bubbles, gen)
        ctx.event_set_nframes(s, [B.NFRAMES] * B.NCAMS)
    torch.cuda.synchronize()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0

    if not args.only_png:
        # ---- trainer: the ring of E resident events is reused until --train-events have been added ----
        best = None
        t_gate = t_fin = 0.0
        for _ in range(args.reps):
            tg = tf = 0.0
            for c in range(B.NCAMS):
                ctx.train_begin(c)
                t0 = time.perf_counter()
                for i0 in range(0, args.train_events, E):
                    ctx.train_add_slots(c, [i % E for i in range(i0, min(i0 + E, args.train_events))])
                t1 = time.perf_counter()
                _, _, n, st = ctx.train_finish(c, read=False)     # the planes stay in HBM: set_background_from_training installs them
                t2 = time.perf_counter()
                assert st == 0
                ctx.set_background_from_training(c)
                tg += t1 - t0
                tf += t2 - t1
            if best is None or tg + tf < best:
                best, t_gate, t_fin = tg + tf, tg, tf
        byt_gate = B.NCAMS * args.train_events * 2 * B.NPIX
        byt_fin = B.NCAMS * (n * B.NPIX + 2 * B.NPIX)
        byt = byt_gate + byt_fin                                    # gate reads + Welford reads + planes
        print(json.dumps({"stage": "trainer", "train_events": args.train_events, "kept_frames_per_cam": n, "seconds": best,
                          "algorithmic_GB": byt / 1e9, "GBps": byt / best / 1e9, "frac_of_peak": byt / best / 1e9 / peak,
                          "gate_seconds": t_gate, "gate_GBps": byt_gate / t_gate / 1e9,
                          "welford_seconds": t_fin, "welford_GBps": byt_fin / t_fin / 1e9,
                          "note": "gate: batched K3 over resident slots, kept frames stay in place; "
                                  "welford: K4, planes left in HBM (abub_set_background_from_training installs them from there)"}))

        # ---- trigger + localizer ----
        ctx.trigger_hist(0, E)
        ctx.trigger_scan(0, E)
        trig, status, lt, _ = ctx.trigger_results(0, E, want_sig=False)
        items = []
        for s in range(E):
            for c in range(B.NCAMS):
                if status[s, c] == 0:
                    t = int(trig[s, c])
                    items.append((s, c, t, max(0, t - 2), int(lt[s, c]), glue.n_post_frames(t, B.NFRAMES)))
        nplanes = sum(3 + it[5] + 2 for it in items)
        best = None
        _, _, nblobs, bufs = ctx.localize_batch_raw(items, cap=1 << 18)
        for _ in range(args.reps):
            t0 = time.perf_counter()
            ctx.localize_batch_raw(items, cap=1 << 18, buffers=bufs)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        byt = nplanes * B.NPIX
        print(json.dumps({"stage": "localizer", "items": len(items), "frames": sum(it[5] + 1 for it in items), "blobs": nblobs,
                          "seconds": best, "items_per_s": len(items) / best, "algorithmic_GB": byt / 1e9,
                          "GBps": byt / best / 1e9, "frac_of_peak": byt / best / 1e9 / peak}))

        if args.kernels:
            ctx.profile(True)
            for _ in range(args.reps):
                ctx.localize_batch_raw(items, cap=1 << 18, buffers=bufs)
                ctx.trigger_hist(0, E)
                ctx.trigger_scan(0, E)
            prof = ctx.profile_read()
            ctx.profile(False)
            print(json.dumps({"stage": "kernel groups (device ms per pass)", "items": len(items), "frames": sum(it[5] + 1 for it in items),
                              "kernels": {n: round(ms / args.reps, 4) for n, ms, c in prof if c}}))
        best = None
        for _ in range(args.reps):
            t0 = time.perf_counter()
            ctx.trigger_hist(0, E)
            ctx.trigger_scan(0, E)
            ctx.trigger_results(0, E, want_sig=False)
            ctx.localize_batch_raw(items, cap=1 << 18, buffers=bufs)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        byt = E * B.NCAMS * B.K1_BYTES_PER_EVENT_CAM + nplanes * B.NPIX
        print(json.dumps({"stage": "recon (K1+K2+K5-K7, resident)", "events": E, "seconds": best, "events_per_s": E / best,
                          "algorithmic_GB": byt / 1e9, "GBps": byt / best / 1e9, "frac_of_peak": byt / best / 1e9 / peak}))
        # ---- bellows veto: TrackAFeature at full size (a14; rare branch, timed for completeness) ----
        pat, _ = synth.make_bellows_object(1, B.ROWS, B.COLS)
        ctx.set_bellows_template(1, pat)
        ctx.bellows_track(0, 1, 0)
        best = None
        for _ in range(args.reps):
            t0 = time.perf_counter()
            ctx.bellows_track(0, 1, 0)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        macs = (B.ROWS - pat.shape[0] + 1) * (B.COLS - pat.shape[1] + 1) * pat.shape[0] * pat.shape[1]
        print(json.dumps({"stage": "bellows TrackAFeature (K8, one frame)", "template": list(pat.shape), "seconds": best,
                          "u8_MACs": macs, "TMAC_per_s": macs / best / 1e12,
                          "note": "exact integer CCORR_NORMED numerators with dp4a; includes window sums, min/max/argmax and the read-back"}))
    # ---- K9: on-device PNG decode of a staged batch (8(f)-2) ----
    if args.png_events > 0:
        import cv2
        import struct

        def idat(png):
            pos, z = 8, b""
            while pos + 12 <= len(png):
                n = struct.unpack(">I", png[pos:pos + 4])[0]
                if png[pos + 4:pos + 8] == b"IDAT":
                    z += png[pos + 8:pos + 8 + n]
                pos += 12 + n
            return z

        pe = min(args.png_events, E)
        ring = []
        t_cpu = []
        for i in range(8):
            frame = B.slot_tensor(torch, ctx, i % E)[i % B.NCAMS, (5 * i) % B.NFRAMES].cpu().numpy()
            enc = cv2.imencode(".png", frame)[1]
            t0 = time.perf_counter()
            dec = cv2.imdecode(enc, cv2.IMREAD_GRAYSCALE)
            t_cpu.append(time.perf_counter() - t0)
            assert np.array_equal(dec, frame)
            ring.append((idat(enc.tobytes()), frame))
        jobs = [(s, c, f, ring[(s + c + f) % 8][0]) for s in range(pe) for c in range(B.NCAMS) for f in range(B.NFRAMES)]
        zbytes = sum(len(j[3]) for j in jobs)
        okv = ctx.decode_png(jobs)                       # warm-up (allocations), also fills the pinned regions
        assert okv.all()
        dev = ctx.download_frames(0)
        assert np.array_equal(dev[1, 3], ring[(0 + 1 + 3) % 8][1])
        arr = (capi.PngJob * len(jobs))(*[capi.PngJob(s, c, f, len(z)) for (s, c, f, z) in jobs])
        okc = (capi.C.c_int * len(jobs))()
        best = None
        for _ in range(args.reps):
            t0 = time.perf_counter()
            ctx._ck(ctx.lib.abub_event_decode_png(ctx.h, arr, len(jobs), okc), "abub_event_decode_png")
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        cpu_ms = 1e3 * float(np.median(t_cpu))
        ncores = len(os.sched_getaffinity(0))
        print(json.dumps({"stage": "PNG decode on the device (K9: H2D of the compressed bytes + inflate + unfilter)", "events": pe,
                          "frames": len(jobs), "seconds": best, "events_per_s": pe / best, "frames_per_s": len(jobs) / best,
                          "compressed_GB": zbytes / 1e9, "pixels_GB": len(jobs) * B.NPIX / 1e9,
                          "cv2_imdecode_ms_per_frame_one_core": cpu_ms, "host_cores": ncores,
                          "host_events_per_s_all_cores": ncores / (cpu_ms * 1e-3) / (B.NCAMS * B.NFRAMES)}))
    ctx.close()


if __name__ == "__main__":
    main()
