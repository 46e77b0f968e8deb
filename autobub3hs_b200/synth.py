"""Synthetic 40l-19-layout PICO runs.

Pure numpy; no dependency on the CUDA library or on ``oracle/``.  The reference ships
no data besides four sample frames and the camera masks, which are not copied here:
backgrounds and masks are procedural look-alikes (dark surround, bright jar region with
fixed-pattern texture, bellows strip on cameras 1 and 3, cf. reference
cam_masks/40l-19/cam{c}_mask.bmp / cam{c}_bellows_mask.bmp).

Layout written by :func:`write_run`: ``<data>/<run>/<ev>/Images/cam{c}_image{30..}.bmp``
(AutoBubStart3.cpp:226-233: 40l-19 series, frameOffset 30, 4 cameras for run ids
>= "20200713_7"; RawParser.cpp:146 accepts ``png|bmp``).
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

ROWS, COLS = 1050, 1680
NCAMS = 4
NFRAMES = 40
FRAME_OFFSET = 30
RUN_ID = "20210101_0"

EVENT_CLASSES = ("normal", "multi", "notrigger", "late", "short4", "short5", "flicker", "highnoise")
# classes that need a run made with bellows_object=True (cameras 1 and 3 show a moving bellows, L3Localizer.cpp:292-390)
BELLOWS_CLASSES = ("bellows", "bellows_bubble")


@dataclass
class BubbleSpec:
    cx: float
    cy: float
    r0: float
    growth: float
    amp: int
    f_trig: int


@dataclass
class EventSpec:
    event: int
    klass: str = "normal"
    nframes: int = NFRAMES
    noise_sigma: float = 2.0
    bubbles: Dict[int, List[BubbleSpec]] = field(default_factory=dict)   # per camera
    flicker_frame: int = -1
    bellows_move: Dict[int, Tuple[int, int, int]] = field(default_factory=dict)   # per camera (f_move, dx, dy) per step


def _rng(*key: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(list(key)))


def make_masks(rows: int = ROWS, cols: int = COLS, ncams: int = NCAMS):
    """(cam_masks, bellows_masks): u8 planes, 255/0 (bellows: graded non-zero values)."""
    yy, xx = np.mgrid[0:rows, 0:cols]
    cam_masks, bellows = [], []
    for c in range(ncams):
        cx = cols * (0.50 + 0.02 * (c - 1.5))
        cy = rows * (0.53 + 0.01 * (c % 2))
        ax, ay = cols * 0.43, rows * 0.40
        m = (((xx - cx) / ax) ** 2 + ((yy - cy) / ay) ** 2 <= 1.0)
        cam_masks.append((m * 255).astype(np.uint8))
        b = np.zeros((rows, cols), np.uint8)
        if c % 2 == 1:        # bellows strip only on cams 1 and 3 (reference masks: cam0/2 all zero)
            x0, x1 = int(cols * 0.80), cols - 1
            y0, y1 = int(rows * 0.11), int(rows * 0.87)
            ramp = np.linspace(1, 255, x1 - x0, dtype=np.float64).astype(np.uint8)
            b[y0:y1, x0:x1] = ramp[None, :]
        bellows.append(b)
    return cam_masks, bellows


def make_background(cam: int, rows: int = ROWS, cols: int = COLS) -> np.ndarray:
    """Fixed scene of one camera: dark surround, bright elliptical jar, texture."""
    rng = _rng(4242, cam)
    yy, xx = np.mgrid[0:rows, 0:cols].astype(np.float32)
    cx = cols * (0.50 + 0.02 * (cam - 1.5))
    cy = rows * 0.53
    r2 = ((xx - cx) / (cols * 0.47)) ** 2 + ((yy - cy) / (rows * 0.46)) ** 2
    jar = np.clip(1.15 - r2, 0.0, 1.0)
    base = 8.0 + 150.0 * jar ** 0.7 + 18.0 * np.sin(xx / cols * 9.0 + cam) * jar
    # fixed-pattern texture, low-pass filtered white noise
    t = rng.standard_normal((rows // 4 + 2, cols // 4 + 2)).astype(np.float32)
    t = np.kron(t, np.ones((4, 4), np.float32))[:rows, :cols]
    base = base + 6.0 * t * jar
    base[r2 > 1.12] = 0.0            # saturated-black surround as in the real frames
    return np.clip(np.rint(base), 0, 255).astype(np.uint8)


def make_bellows_object(cam: int, rows: int = ROWS, cols: int = COLS):
    """(pattern u8 [th, tw], (x, y) rest position) of the bellows seen by an odd camera, None for the others.
    Sizes follow the reference templates (cam1 557x178, cam3 509x152 at 1050x1680); horizontal ridges + fixed texture."""
    if cam % 2 == 0:
        return None
    th = int(rows * (0.53 if cam == 1 else 0.485))
    tw = int(cols * (0.106 if cam == 1 else 0.0905))
    rng = _rng(777, cam)
    yy, xx = np.mgrid[0:th, 0:tw].astype(np.float32)
    period = max(6.0, rows / 40.0)
    pat = 62.0 + 50.0 * np.sin(2 * np.pi * yy / period + 0.8 * np.sin(xx / tw * 3.0)) + 5.0 * rng.standard_normal((th, tw))
    pat *= (0.75 + 0.25 * np.cos((xx / tw - 0.5) * 2.2))
    pat = np.clip(np.rint(pat), 4, 123).astype(np.uint8)
    sx0, sx1 = int(cols * 0.80), cols - 1
    sy0, sy1 = int(rows * 0.11), int(rows * 0.87)
    x = sx0 + (sx1 - sx0 - tw) // 2
    y = sy0 + (sy1 - sy0 - th) // 2
    return pat, (x, y)


def _pick_centres(rng, n, cam_mask, bellows_mask, margin=40):
    rows, cols = cam_mask.shape
    ok = (cam_mask > 0) & (bellows_mask == 0)
    out = []
    guard = 0
    while len(out) < n and guard < 100000:
        guard += 1
        x = float(rng.uniform(margin, cols - margin))
        y = float(rng.uniform(margin, rows - margin))
        if not ok[int(y), int(x)]:
            continue
        if all((x - a) ** 2 + (y - b) ** 2 > 60 ** 2 for a, b in out):
            out.append((x, y))
    return out


def make_event_spec(event: int, klass: str, cam_masks, bellows_masks, seed: int = 0,
                    nframes: int = NFRAMES) -> EventSpec:
    rng = _rng(7919, seed, event)
    es = EventSpec(event, klass, nframes, 2.0)
    ncams = len(cam_masks)
    rows, cols = cam_masks[0].shape
    scale = min(1.0, rows / ROWS * 2.0)
    lo = min(5, max(2, nframes // 8))
    hi = max(lo + 1, nframes - 6)
    if klass == "highnoise":
        es.noise_sigma = 6.0
    if klass == "short4":
        es.nframes = 4
    if klass == "short5":
        es.nframes = 5
    if klass == "flicker":
        es.flicker_frame = int(rng.integers(lo, hi))
    if klass in BELLOWS_CLASSES:
        f_move = int(rng.integers(lo, hi))
        for c in range(ncams):
            if c % 2 == 1:
                es.bellows_move[c] = (f_move, int(rng.integers(0, 2)), int(rng.integers(1, 3)))
                if klass == "bellows_bubble":
                    # a bubble beside the bellows, inside both the camera mask and the bellows mask
                    ok = (cam_masks[c] > 0) & (bellows_masks[c] > 0)
                    obj = make_bellows_object(c, rows, cols)
                    ok[:, obj[1][0] - 6:] = False
                    ys, xs = np.nonzero(ok)
                    if len(xs):
                        i = int(rng.integers(0, len(xs)))
                        es.bubbles[c] = [BubbleSpec(float(xs[i]), float(ys[i]), 2.0, 1.0, 120, f_move)]
    nb = {"normal": (1, 4), "multi": (20, 27), "highnoise": (1, 4), "late": (1, 3)}.get(klass)
    if nb is not None:
        f_trig = 0 if klass == "late" else int(rng.integers(lo, hi))
        for c in range(ncams):
            k = int(rng.integers(nb[0], nb[1]))
            specs = []
            for (x, y) in _pick_centres(rng, k, cam_masks[c], bellows_masks[c],
                                        margin=max(8, int(40 * scale))):
                specs.append(BubbleSpec(x, y, float(rng.uniform(1.0, 3.0)),
                                        float(rng.uniform(0.3, 1.5)),
                                        int(rng.integers(60, 151)), f_trig))
            es.bubbles[c] = specs
    return es


def render_camera(es: EventSpec, cam: int, background: np.ndarray, seed: int = 0, bellows_object=None) -> np.ndarray:
    """[nframes, rows, cols] u8 stack of one camera of one event.  `bellows_object` = make_bellows_object(cam, ...):
    pasted (opaque) at its rest position, displaced by up to 3 steps from the event's f_move on."""
    rows, cols = background.shape
    rng = _rng(101, seed, es.event, cam)
    out = np.empty((es.nframes, rows, cols), np.uint8)
    bg = background.astype(np.float32)
    for f in range(es.nframes):
        scene = bg
        if bellows_object is not None:
            pat, (bx, by) = bellows_object
            mv = es.bellows_move.get(cam)
            if mv is not None and f >= mv[0]:
                k = min(f - mv[0] + 1, 3)
                bx, by = bx + k * mv[1], by + k * mv[2]
            scene = bg.copy()
            scene[by:by + pat.shape[0], bx:bx + pat.shape[1]] = pat
        img = rng.standard_normal((rows, cols), dtype=np.float32)         # in place from here on: same values as
        img *= np.float32(es.noise_sigma)                                  # scene + noise * sigma, fewer passes over memory
        img += scene
        if f == es.flicker_frame:
            img += 8.0
        for b in es.bubbles.get(cam, ()):
            if f < b.f_trig:
                continue
            r = b.r0 + b.growth * (f - b.f_trig)
            x0, x1 = max(0, int(b.cx - r - 2)), min(cols, int(b.cx + r + 3))
            y0, y1 = max(0, int(b.cy - r - 2)), min(rows, int(b.cy + r + 3))
            yy, xx = np.mgrid[y0:y1, x0:x1]
            disc = (xx - b.cx) ** 2 + (yy - b.cy) ** 2 <= r * r
            img[y0:y1, x0:x1][disc] += b.amp
        np.rint(img, out=img)
        np.clip(img, 0, 255, out=img)
        out[f] = img                                                       # exact: integral values in [0, 255]
    return out


@dataclass
class SyntheticRun:
    run_id: str
    event_numbers: List[int]
    specs: List[EventSpec]
    events: List[List[np.ndarray]]          # events[ev][cam] = [N,rows,cols] u8
    cam_masks: List[np.ndarray]
    bellows_masks: List[np.ndarray]
    backgrounds: List[np.ndarray]
    frame_offset: int = FRAME_OFFSET
    bellows_templates: Optional[List[Optional[np.ndarray]]] = None   # cam<c>_bellows_template.png (bellows_object runs)


def _render_job(job):
    es, c, bg, seed, obj = job
    return render_camera(es, c, bg, seed, obj)


def make_run(n_events: int, classes: Optional[Sequence[str]] = None, rows: int = ROWS, cols: int = COLS,
             ncams: int = NCAMS, nframes: int = NFRAMES, seed: int = 0, run_id: str = RUN_ID,
             bellows_object: bool = False, workers: int = 1, first_event: int = 0) -> SyntheticRun:
    """`classes[i]` = class of event i (cycled); default all "normal".  bellows_object=True puts the bellows of
    make_bellows_object into every frame of cameras 1 and 3 and provides their templates (BELLOWS_CLASSES move it)."""
    cam_masks, bellows = make_masks(rows, cols, ncams)
    bgs = [make_background(c, rows, cols) for c in range(ncams)]
    objs = [make_bellows_object(c, rows, cols) if bellows_object else None for c in range(ncams)]
    classes = list(classes) if classes else ["normal"]
    specs, events = [], []
    for ev in range(first_event, first_event + n_events):
        specs.append(make_event_spec(ev, classes[ev % len(classes)], cam_masks, bellows, seed, nframes))
    if workers > 1 and n_events * ncams > 1:
        # full-size frames cost about a second per camera to render: spread the cameras over processes (same pixels,
        # every camera has its own generator keyed by (seed, event, camera))
        import multiprocessing as mp
        jobs = [(es, c, bgs[c], seed, objs[c]) for es in specs for c in range(ncams)]
        with mp.get_context("fork").Pool(min(workers, len(jobs))) as pool:
            flat = pool.map(_render_job, jobs, chunksize=1)
        events = [flat[i * ncams:(i + 1) * ncams] for i in range(n_events)]
    else:
        for es in specs:
            events.append([render_camera(es, c, bgs[c], seed, objs[c]) for c in range(ncams)])
    templates = [None if o is None else o[0] for o in objs] if bellows_object else None
    return SyntheticRun(run_id, list(range(first_event, first_event + n_events)), specs, events, cam_masks, bellows, bgs, FRAME_OFFSET,
                        templates)


# ---------------------------------------------------------------------------------------
# 8-bit BMP writer (the format the reference's RawParser regex accepts next to png)
# ---------------------------------------------------------------------------------------
def write_bmp8(path: str, img: np.ndarray) -> None:
    rows, cols = img.shape
    stride = (cols + 3) & ~3
    pal = b"".join(struct.pack("<BBBB", i, i, i, 0) for i in range(256))
    off = 14 + 40 + len(pal)
    size = off + stride * rows
    hdr = struct.pack("<2sIHHI", b"BM", size, 0, 0, off)
    dib = struct.pack("<IiiHHIIiiII", 40, cols, rows, 1, 8, 0, stride * rows, 2835, 2835, 256, 0)
    body = np.zeros((rows, stride), np.uint8)
    body[:, :cols] = img[::-1]            # bottom-up rows
    with open(path, "wb") as f:
        f.write(hdr + dib + pal + body.tobytes())


def write_png8(path: str, img: np.ndarray) -> None:
    """8-bit greyscale PNG (filter 0 on every row) -- for cam<c>_bellows_template.png (L3Localizer.cpp:294)."""
    import zlib
    rows, cols = img.shape

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    raw = b"".join(b"\x00" + np.ascontiguousarray(img[r]).tobytes() for r in range(rows))
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", cols, rows, 8, 0, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def write_run(run: SyntheticRun, data_dir: str, mask_dir: Optional[str] = None, fmt: str = "bmp") -> str:
    """Writes ``<data_dir>/<run>/<ev>/Images/cam{c}_image{30+f}.bmp`` (+ masks) and returns the run dir.
    ``fmt="png"``: 8-bit grey PNG frames instead (adaptive row filters when cv2 is there to encode them)."""
    rdir = os.path.join(data_dir, run.run_id)
    encode = None
    if fmt == "png":
        try:
            import cv2
            encode = lambda img: cv2.imencode(".png", img)[1].tobytes()
        except ImportError:
            encode = None
    for evn, ev in zip(run.event_numbers, run.events):
        idir = os.path.join(rdir, str(evn), "Images")
        os.makedirs(idir, exist_ok=True)
        for c, stack in enumerate(ev):
            for f in range(len(stack)):
                if fmt == "png":
                    path = os.path.join(idir, f"cam{c}_image{run.frame_offset + f}.png")
                    if encode:
                        with open(path, "wb") as fh:
                            fh.write(encode(np.ascontiguousarray(stack[f])))
                    else:
                        write_png8(path, stack[f])
                else:
                    write_bmp8(os.path.join(idir, f"cam{c}_image{run.frame_offset + f}.bmp"), stack[f])
    if mask_dir:
        os.makedirs(mask_dir, exist_ok=True)
        for c in range(len(run.cam_masks)):
            write_bmp8(os.path.join(mask_dir, f"cam{c}_mask.bmp"), run.cam_masks[c])
            write_bmp8(os.path.join(mask_dir, f"cam{c}_bellows_mask.bmp"), run.bellows_masks[c])
            if run.bellows_templates is not None and run.bellows_templates[c] is not None:
                write_png8(os.path.join(mask_dir, f"cam{c}_bellows_template.png"), run.bellows_templates[c])
    return rdir
