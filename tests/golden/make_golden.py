#!/usr/bin/env python
"""Generates the golden vectors under tests/golden/ from the cv2-backed oracle
(oracle/cv2_oracle.py driving cv2 4.13.0 exactly as the reference drives OpenCV).

The reference ships no test vectors, so these pin *this* oracle: the
C restatement, the host-side contour code and the CUDA path are all checked against them.
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import cv2  # noqa: E402

import oracle.cv2_oracle as o  # noqa: E402
from autobub3hs_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def frame_pairs(seed, n, rows, cols):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        base = rng.integers(0, 256, (rows, cols)).astype(np.float32)
        sd = [2, 6, 25, 80][i % 4]
        a = np.clip(base + rng.normal(0, sd, (rows, cols)), 0, 255).astype(np.uint8)
        b = np.clip(base + rng.normal(0, sd, (rows, cols)), 0, 255).astype(np.uint8)
        s = rng.integers(0, [2, 3, 6, 43][i % 4], (rows, cols)).astype(np.uint8)
        out.append((a, b, s))
    return out


def hist_tables(seed, n):
    """Synthetic H tables incl. >46340 counts in bins >= 2 (int32 wrap), zero-variance and NaN/inf cases."""
    rng = np.random.default_rng(seed)
    total = 1764000
    tabs = []
    for i in range(n):
        nf = int(rng.integers(5, 41))
        H = np.zeros((nf, 256), np.uint32)
        kind = i % 6
        for f in range(1, nf):
            row = np.zeros(256, np.int64)
            maxbin = int(rng.integers(0, 12))
            if kind == 0:                       # quiet, then a growing bubble
                grow = max(0, f - nf // 2)
                for b in range(1, min(255, 2 + grow * 3)):
                    row[b] = int(rng.integers(0, 40 + 30 * grow))
            elif kind == 1:                     # big counts -> val*val overflows int32
                for b in range(1, maxbin + 2):
                    row[b] = int(rng.integers(40000, 90000))
            elif kind == 2:                     # constant history -> sigma == 0 -> inf / skipped terms
                row[1] = 100
                row[2] = 50
                row[3] = 7
                if f > nf - 4:
                    row[2] = 51 + f
            elif kind == 3:                     # all zero frames, then flicker
                if f == nf // 2:
                    row[1:6] = 1000
            elif kind == 4:                     # random sparse
                for b in rng.integers(1, 256, int(rng.integers(0, 6))):
                    row[b] = int(rng.integers(1, 5000))
            else:                               # overflow that makes the radicand negative -> NaN
                row[2] = 70000 if f % 2 else 65536
                row[3] = int(rng.integers(0, 3))
                if f > nf - 5:
                    row[4] = 5 * f
            row[0] = total - row[1:].sum()
            H[f] = row.astype(np.uint32)
        tabs.append(H)
    return tabs, total


BELLOWS_CLASSES = ["normal", "bellows", "bellows_bubble", "notrigger", "bellows", "normal"]


def bellows_golden():
    """6. a run whose cameras 1 and 3 show a moving bellows (a14, L3Localizer.cpp:292-390): text, and for every time
    the veto branch was taken the two TrackAFeature positions after the nudge (float32, hex) and the ROI."""
    run = synth.make_run(6, BELLOWS_CLASSES, rows=264, cols=420, seed=5, bellows_object=True)
    trained = o.train_all(run.events, 4)
    text, recs = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks, 30,
                              run.bellows_templates)
    open(os.path.join(OUT, "bellows_run.txt"), "w").write(o.write_header() + text)
    branch = []
    for ev, rr in enumerate(recs):
        for cam, a in enumerate(rr):
            for info in a.bellows:
                if info.get("bellows_branch"):
                    pt, pp, roi = info["bellows_positions"]
                    branch.append({"event": ev, "cam": cam, "pos_trig": [float(v).hex() for v in pt],
                                   "pos_pre": [float(v).hex() for v in pp], "roi": [int(v) for v in roi]})
    json.dump({"rows": 264, "cols": 420, "seed": 5, "classes": BELLOWS_CLASSES, "branch": branch},
              open(os.path.join(OUT, "bellows_run.json"), "w"))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--only-bellows":
        bellows_golden()
        return
    # 1. ProcessFrame + histogram
    pf = {}
    for i, (a, b, s) in enumerate(frame_pairs(101, 8, 61, 133)):
        d = o.process_frame(a, b, s)
        pf[f"diff{i}"] = d
        pf[f"hist{i}"] = o.hist256(d).astype(np.uint32)
    np.savez_compressed(os.path.join(OUT, "process_frame.npz"), **pf)

    # 2. significance / trigger on H tables
    tabs, total = hist_tables(202, 48)
    trig_out = []
    for H in tabs:
        for tss in (12, 4):
            st = o.SignificanceState()
            src = o.TableSource(len(H), total)
            hp = o.TableHistProvider(H)
            mat, rec = 0, []
            for attempt in range(3):
                r = o.find_trigger_frame(src, None, tss, mat + 1, st, mat, hp)
                rec.append({"start": mat + 1, "status": r.status, "trig": r.trig, "loc_thres": r.loc_thres,
                            "sig": [[i, float(v).hex()] for i, v in r.sig]})
                if r.status != 0:
                    break
                mat = r.trig
            trig_out.append({"tss": tss, "runs": rec})
    np.savez_compressed(os.path.join(OUT, "trigger_tables.npz"), **{f"H{i}": H for i, H in enumerate(tabs)})
    json.dump({"total": total, "cases": trig_out}, open(os.path.join(OUT, "trigger_expected.json"), "w"))

    # 3. Trainer
    tr = {}
    rng = np.random.default_rng(303)
    base = rng.integers(0, 256, (40, 56)).astype(np.float32)
    stack = [np.clip(base + rng.normal(0, 2.5, base.shape), 0, 255).astype(np.uint8) for _ in range(40)]
    for n in (1, 2, 5, 6, 40):
        m, s = o.welford_mean_sigma(stack[:n])
        tr[f"mean{n}"], tr[f"sigma{n}"] = m, s
    ents = [float(o.training_gate_entropy(stack[0], stack[i])) for i in range(1, 6)]
    ents.append(float(o.training_gate_entropy(stack[0], rng.integers(0, 256, base.shape).astype(np.uint8))))
    tr["entropies"] = np.array(ents, np.float64)
    np.savez_compressed(os.path.join(OUT, "trainer.npz"), **tr)

    # 4. contours: vertices, boxes, areas, moments
    rng = np.random.default_rng(404)
    cont = []
    imgs = {}
    for i in range(24):
        h, w = int(rng.integers(20, 70)), int(rng.integers(20, 90))
        if i % 3 == 0:
            img = ((rng.random((h, w)) < rng.uniform(0.1, 0.5)) * 255).astype(np.uint8)
        elif i % 3 == 1:
            img = np.zeros((h, w), np.uint8)
            for _ in range(rng.integers(1, 6)):
                cv2.ellipse(img, (int(rng.integers(0, w)), int(rng.integers(0, h))), (int(rng.integers(1, 18)), int(rng.integers(1, 18))),
                            float(rng.uniform(0, 180)), 0, 360, 255, int(rng.choice([-1, 1])))
        else:
            img = np.zeros((h, w), np.uint8)
            c = (w // 2, h // 2)
            for r in range(min(h, w) // 2 - 1, 2, -4):
                cv2.circle(img, c, r, 255, 1)
            img |= ((rng.random((h, w)) < 0.03) * 255).astype(np.uint8)
        imgs[f"img{i}"] = img
        cs = o.find_external_contours(img)
        rec = []
        for c in cs:
            m = cv2.moments(c)
            rec.append({"pts": c.reshape(-1, 2).tolist(), "rect": list(cv2.boundingRect(c)),
                        "area": float(cv2.contourArea(c)).hex(), "m00": float(m["m00"]).hex(),
                        "m10": float(m["m10"]).hex(), "m01": float(m["m01"]).hex()})
        cont.append(rec)
    np.savez_compressed(os.path.join(OUT, "contour_images.npz"), **imgs)
    json.dump(cont, open(os.path.join(OUT, "contours_expected.json"), "w"))

    # 5. end-to-end text of a small synthetic run (all event classes)
    run = synth.make_run(8, list(synth.EVENT_CLASSES), rows=132, cols=212, seed=9)
    trained = o.train_all(run.events, 4)
    text, recs = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks)
    open(os.path.join(OUT, "e2e_small_run.txt"), "w").write(o.write_header() + text)
    trig = [[[(t.status, t.trig, t.loc_thres) for t in a.triggers] for a in rr] for rr in recs]
    json.dump({"rows": 132, "cols": 212, "seed": 9, "classes": list(synth.EVENT_CLASSES), "triggers": trig,
               "training": [[t.status, t.training_set_size, t.kept_events] for t in trained]},
              open(os.path.join(OUT, "e2e_small_run.json"), "w"))
    bellows_golden()
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
