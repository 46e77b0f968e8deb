"""Test-side host glue: drives the C ABI the way the C++ host (AnyCamAnalysis / LocalizeOMatic)
does and turns abub_blob records into the oracle's BubbleImageFrame / Bubble bookkeeping, so
that the final abub3hs text can be compared byte for byte with the oracle's."""
import numpy as np

import oracle.cv2_oracle as o

F32, F64 = np.float32, np.float64


def blob_to_frame(b, genesis: bool) -> o.BubbleImageFrame:
    with np.errstate(all="ignore"):
        radius = float(np.sqrt(F64(b.area) / F64(3.14159)))
        if (not genesis) or b.m00 > 0:
            cx = F32(F64(b.m10) / F64(b.m00))
            cy = F32(F64(b.m01) / F64(b.m00))
        else:
            cx = F32(F64(b.sum_x) / F64(b.npts))
            cy = F32(F64(b.sum_y) / F64(b.npts))
    return o.BubbleImageFrame(b.x, b.y, b.w, b.h, b.area, radius, b.m00, b.m10, b.m01, cx, cy)


def n_post_frames(trig: int, nframes: int) -> int:
    """L3Localizer.cpp:944-954."""
    kmax = o.NUM_FRAMES_BUBBLE_TRACK if trig < 29 else 39 - trig
    n = 0
    for k in range(1, kmax + 1):
        if trig + k >= nframes:
            break
        n += 1
    return n


def genesis_from_blobs(blobs, cam_mask, bellows_mask, bellows_veto=None):
    """CalculateInitialBubbleParams bookkeeping (L3Localizer.cpp:273-290, 392-444).  `bellows_veto` is called when
    every contour lies in the bellows mask (:292-390) and returns the blobs of the re-binarised frame, or None when
    the camera has no template (the reference then re-reads the contours it already had, :374)."""
    kept, largest = [], 0
    all_in_bellows = len(blobs) > 0
    for b in blobs:
        if not o.is_in_mask(b.x, b.y, b.w, b.h, bellows_mask, True):
            all_in_bellows = False
            largest = max(largest, b.w * b.h)
            kept.append(b)
    if all_in_bellows:
        again = bellows_veto() if bellows_veto is not None else None
        kept = list(blobs) if again is None else list(again)
        largest = 0
        for b in kept:                                               # :383-389
            largest = max(largest, b.w * b.h)
    out = []
    for b in kept:
        if b.w * b.h > 10 or b.w * b.h >= largest:
            f = blob_to_frame(b, True)
            if o.is_in_mask(f.x, f.y, f.w, f.h, cam_mask, False):
                out.append(f)
    return out


def post_from_blobs(blobs, cam_mask):
    out = []
    for b in blobs:
        if b.w * b.h > 10:
            f = blob_to_frame(b, False)
            if o.is_in_mask(f.x, f.y, f.w, f.h, cam_mask, False):
                out.append(f)
    return out


class _Minus6(Exception):
    """Where the reference throws inside AnyCamAnalysis (caught at AutoBubStart3.cpp:114 -> -6)."""


def gpu_any_cam_analysis(ctx, slot, cam, nframes, tss, cam_mask, bellows_mask, event_number, has_template=False,
                         log=None):
    """AnyCamAnalysis (AutoBubStart3.cpp:67-125) with trigger + localizer on the device.  has_template: the context
    holds this camera's bellows template (abub_set_bellows_template).  log (list) receives the bellows positions."""
    bubbles, mat_trig, loc_thres = [], 0, 3
    out = o.CameraOutput()
    try:
        while True:
            t, status, lt, _ = ctx.trigger(slot, cam, mat_trig + 1)
            if status == 0:
                mat_trig = t
            if lt >= 0:
                loc_thres = lt
            if status != 0:
                return o.CameraOutput(status, 0, event_number)
            if nframes <= 5:
                return o.CameraOutput(-8, 0, event_number)
            pre = max(0, mat_trig - (1 if tss < 6 else 2))
            npost = n_post_frames(mat_trig, nframes)
            frames = ctx.localize_batch([(slot, cam, mat_trig, pre, loc_thres, npost)])[0]

            def veto():
                if not has_template:
                    return None
                try:
                    pt, _ = ctx.bellows_track(slot, cam, mat_trig)           # :313
                    pp, _ = ctx.bellows_track(slot, cam, pre)                # :314
                except RuntimeError as e:
                    raise _Minus6(str(e))
                blobs, st, roi = ctx.localize_bellows(slot, cam, mat_trig, pre, loc_thres, pt, pp)
                if log is not None:
                    log.append((pt, pp, roi, st))
                if st != 0:
                    raise _Minus6("Rect outside the frame")
                return blobs

            for g in genesis_from_blobs(frames[0], cam_mask, bellows_mask, veto):
                bubbles.append(o.Bubble(g))
            for k in range(1, npost + 1):
                o.track_candidates(bubbles, post_from_blobs(frames[k], cam_mask))
            out = o.stage_camera_output(bubbles, mat_trig, event_number)
            if bubbles:
                return out
    except _Minus6:
        return o.CameraOutput(-6, 0, event_number)


def host_recon_text(run, trained, tmp_path, tag, pinned, early_chunk, batch=None, sub_batch=None, hint=False):
    """Trains on the resident events through the C ABI, then runs the full recon through libabubhost.so."""
    import os
    from autobub3hs_b200 import capi, synth
    nev = len(run.events)
    mask_dir = str(tmp_path / ("masks_" + tag))
    os.makedirs(mask_dir, exist_ok=True)
    for c in range(4):
        synth.write_bmp8(os.path.join(mask_dir, f"cam{c}_mask.bmp"), run.cam_masks[c])
        synth.write_bmp8(os.path.join(mask_dir, f"cam{c}_bellows_mask.bmp"), run.bellows_masks[c])
        if run.bellows_templates is not None and run.bellows_templates[c] is not None:
            synth.write_png8(os.path.join(mask_dir, f"cam{c}_bellows_template.png"), run.bellows_templates[c])
    out_dir = str(tmp_path / ("out_" + tag))
    os.makedirs(out_dir, exist_ok=True)
    rows, cols = run.cam_masks[0].shape
    nfmax = max(len(st) for ev in run.events for st in ev)
    with capi.HostRecon(rows, cols, 4, max(nfmax, 5), max_slots=nev, early_chunk=early_chunk) as hr:
        ctx = hr.ctx
        nframes = np.zeros((nev, 4), np.int32)
        for ev in range(nev):
            for c in range(4):
                ctx.event_upload(ev, c, run.events[ev][c])
                nframes[ev, c] = len(run.events[ev][c])
        tss = []
        for c in range(4):
            ctx.train_begin(c)
            ctx.train_add_slots(c, list(range(nev)))
            mean, sigma, n, st = ctx.train_finish(c)
            assert st == trained[c].status and n == trained[c].training_set_size
            assert np.array_equal(mean, trained[c].avg) and np.array_equal(sigma, trained[c].sigma), f"trained planes of camera {c}"
            ctx.set_background_from_training(c)
            tss.append(n)
        if pinned:
            for ev in range(nev):
                pin = ctx.pinned(ev)
                for c in range(4):
                    pin[c, :nframes[ev, c]] = run.events[ev][c]
                # what is in HBM must not help: poison the slot so that only frames really sent can be used
                for c in range(4):
                    ctx.event_upload(ev, c, np.full_like(run.events[ev][c], 77))
        if sub_batch:
            hr.set_threads(0, sub_batch)       # several sub-batches per call: the next one's trigger search is queued ahead
        path = hr.begin_run(run.run_id, out_dir, mask_dir, 30, tss)
        b = batch or nev
        for e0 in range(0, nev, b):
            e1 = min(nev, e0 + b)
            if hint and e1 < nev:              # announce the next call's slots (abubhost_recon_hint)
                e2 = min(nev, e1 + b)
                # hint == "wrong": the announcement names fewer frames than the call will bring -- the call must not use it
                hr.recon_hint(e1, e2 - e1, nframes[e1:e2] if hint is True else np.maximum(nframes[e1:e2] - 7, 0))
            hr.recon(e0, run.event_numbers[e0:e1], nframes[e0:e1], pinned=pinned, chunk_frames=4)
        hr.drain()
        stats = hr.read_stats()
        hr.end_run()
        text = open(path).read()
    return text, stats


