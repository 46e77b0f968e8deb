"""CPU suite (-m "not gpu"): the cv2 oracle, the C restatement and the host build of the device
contour code against the committed golden vectors and against cv2 itself."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import oracle.cv2_oracle as o
from oracle import cref

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pairs():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLD, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    return mg


def test_process_frame_golden_cv2_and_c():
    mg = _pairs()
    g = np.load(os.path.join(GOLD, "process_frame.npz"))
    for i, (a, b, s) in enumerate(mg.frame_pairs(101, 8, 61, 133)):
        d = o.process_frame(a, b, s)
        assert np.array_equal(d, g[f"diff{i}"])
        assert np.array_equal(o.hist256(d).astype(np.uint32), g[f"hist{i}"])
        dc = cref.process_frame(a, b, s)
        assert np.array_equal(dc, g[f"diff{i}"])
        assert np.array_equal(cref.hist256(dc), g[f"hist{i}"])


def test_trigger_tables_golden():
    exp = json.load(open(os.path.join(GOLD, "trigger_expected.json")))
    tabs = np.load(os.path.join(GOLD, "trigger_tables.npz"))
    total, ci = exp["total"], 0
    statuses = set()
    for i in range(len(tabs.files)):
        H = tabs[f"H{i}"]
        for tss in (12, 4):
            case = exp["cases"][ci]
            ci += 1
            assert case["tss"] == tss
            st = o.SignificanceState()
            mat = 0
            for run in case["runs"]:
                r = o.find_trigger_frame(o.TableSource(len(H), total), None, tss, mat + 1, st, mat, o.TableHistProvider(H))
                assert (r.status, r.trig, r.loc_thres) == (run["status"], run["trig"], run["loc_thres"])
                for (fi, hx), (fj, v) in zip(run["sig"], r.sig):
                    e = float.fromhex(hx)
                    assert fi == fj and (e == v or (np.isnan(e) and np.isnan(v)))
                statuses.add(r.status)
                mat = r.trig
    assert statuses == {0, -3}
    nan_seen = any(np.isnan(float.fromhex(hx)) or np.isinf(float.fromhex(hx))
                   for c in exp["cases"] for run in c["runs"] for _, hx in run["sig"])
    assert nan_seen, "golden set is meant to contain NaN/inf significance values"


def test_trainer_golden_cv2_and_c():
    g = np.load(os.path.join(GOLD, "trainer.npz"))
    rng = np.random.default_rng(303)
    base = rng.integers(0, 256, (40, 56)).astype(np.float32)
    stack = [np.clip(base + rng.normal(0, 2.5, base.shape), 0, 255).astype(np.uint8) for _ in range(40)]
    for n in (1, 2, 5, 6, 40):
        m, s = o.welford_mean_sigma(stack[:n])
        assert np.array_equal(m, g[f"mean{n}"]) and np.array_equal(s, g[f"sigma{n}"])
        mc, sc = cref.welford(stack[:n])
        assert np.array_equal(mc, g[f"mean{n}"]) and np.array_equal(sc, g[f"sigma{n}"])
    ents = [float(o.training_gate_entropy(stack[0], stack[i])) for i in range(1, 6)]
    assert np.allclose(ents, g["entropies"][:5], atol=1e-7, rtol=0)


def _host_driver():
    so = os.path.join(ROOT, "tests", "_build", "libcontour_host.so")
    src = os.path.join(ROOT, "tests", "host_contour_driver.cu")
    core = os.path.join(ROOT, "autobub3hs_b200", "csrc", "contour_core.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(core)):
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-O2", "-std=c++17", "-Xcompiler", "-fPIC", "--fmad=false",
                               "-Wno-deprecated-gpu-targets", "-shared", "-o", so, src])
    from autobub3hs_b200 import capi
    lib = C.CDLL(so)
    lib.host_find_contours.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(capi.Blob), C.c_int]
    lib.host_find_contours.restype = C.c_int

    def run(img):
        img = np.ascontiguousarray(img, np.uint8)
        blobs = (capi.Blob * 20000)()
        n = lib.host_find_contours(img.ctypes.data, 0, img.shape[0], img.shape[1], blobs, 20000)
        return [blobs[i] for i in range(n)]
    return run


def test_contours_golden_c_and_device_code_on_host():
    imgs = np.load(os.path.join(GOLD, "contour_images.npz"))
    exp = json.load(open(os.path.join(GOLD, "contours_expected.json")))
    host = _host_driver()
    total = 0
    for i, rec in enumerate(exp):
        img = imgs[f"img{i}"]
        blobs, polys = cref.find_contours(img)
        hb = host(img)
        assert len(blobs) == len(rec) == len(hb)
        for b, p, h, r in zip(blobs, polys, hb, rec):
            assert p.tolist() == r["pts"]
            for q in (b, h):
                assert [q.x, q.y, q.w, q.h] == r["rect"]
                assert q.area == float.fromhex(r["area"])
                assert (q.m00, q.m10, q.m01) == tuple(float.fromhex(r[k]) for k in ("m00", "m10", "m01"))
                assert q.npts == len(r["pts"])
            total += 1
    assert total > 300


def test_contours_fuzz_vs_cv2():
    """C restatement and the device code path (host build) vs cv2.findContours on random shapes,
    nested rings included (RETR_EXTERNAL drops components inside holes)."""
    import cv2
    host = _host_driver()
    rng = np.random.default_rng(2024)
    ncont = 0
    for it in range(300):
        kind = it % 5
        h, w = int(rng.integers(6, 90)), int(rng.integers(6, 120))
        img = np.zeros((h, w), np.uint8)
        if kind == 0:
            img = ((rng.random((h, w)) < rng.uniform(0.03, 0.7)) * 255).astype(np.uint8)
        elif kind == 1:
            f = cv2.GaussianBlur(rng.random((h, w)).astype(np.float32), (0, 0), rng.uniform(1, 4))
            img = ((f > np.quantile(f, rng.uniform(0.3, 0.8))) * 255).astype(np.uint8)
        elif kind == 2:
            for _ in range(rng.integers(1, 8)):
                cv2.ellipse(img, (int(rng.integers(0, w)), int(rng.integers(0, h))), (int(rng.integers(1, 25)), int(rng.integers(1, 25))),
                            float(rng.uniform(0, 180)), 0, 360, 255, int(rng.choice([-1, 1, 2])))
        elif kind == 3:
            for _ in range(rng.integers(1, 6)):
                x0, y0 = int(rng.integers(0, w - 3)), int(rng.integers(0, h - 3))
                x1, y1 = int(rng.integers(x0 + 1, w)), int(rng.integers(y0 + 1, h))
                cv2.rectangle(img, (x0, y0), (x1, y1), 255, int(rng.choice([-1, 1])))
                if x1 - x0 > 4 and y1 - y0 > 4:
                    cv2.circle(img, ((x0 + x1) // 2, (y0 + y1) // 2), int(rng.integers(0, 3)), 255, -1)
        else:
            for _ in range(rng.integers(1, 4)):
                c = (int(rng.integers(0, w)), int(rng.integers(0, h)))
                for r in range(int(rng.integers(3, 30)), 2, -int(rng.integers(2, 6))):
                    cv2.circle(img, c, r, 255, 1)
            img |= ((rng.random((h, w)) < 0.03) * 255).astype(np.uint8)
        cs, _ = cv2.findContours(img, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_TC89_L1)
        blobs, polys = cref.find_contours(img)
        hb = host(img)
        assert len(cs) == len(blobs) == len(hb), (it, kind)
        for c, b, p, q in zip(cs, blobs, polys, hb):
            assert np.array_equal(c.reshape(-1, 2), p), (it, kind)
            m = cv2.moments(c)
            for z in (b, q):
                assert cv2.boundingRect(c) == (z.x, z.y, z.w, z.h)
                assert cv2.contourArea(c) == z.area
                assert (m["m00"], m["m10"], m["m01"]) == (z.m00, z.m10, z.m01)
        ncont += len(cs)
    assert ncont > 5000


def test_pixel_ops_c_vs_cv2():
    import cv2
    rng = np.random.default_rng(77)
    for it in range(25):
        h, w = int(rng.integers(5, 150)), int(rng.integers(5, 200))
        base = rng.integers(0, 256, (h, w)).astype(np.float32)
        sd = [2, 10, 60][it % 3]
        a = np.clip(base + rng.normal(0, sd, (h, w)), 0, 255).astype(np.uint8)
        avg = np.clip(base, 0, 255).astype(np.uint8)
        s = rng.integers(0, [2, 4, 50][it % 3], (h, w)).astype(np.uint8)
        over = cv2.blur(cv2.addWeighted(cv2.absdiff(a, avg), 1.0, s, -6.0, 0.0), (3, 3))
        assert np.array_equal(cref.post_over(a, avg, s), over)
        for thr in (2, 3):
            _, tz = cv2.threshold(over, thr, 255, cv2.THRESH_TOZERO)
            t, bn = cv2.threshold(tz, 0, 255, cv2.THRESH_BINARY | cv2.THRESH_OTSU)
            mine, tm = cref.tozero_otsu_binary(over, thr)
            assert tm == int(t) and np.array_equal(mine, bn)


def test_e2e_golden_text_and_statuses():
    from autobub3hs_b200 import synth
    meta = json.load(open(os.path.join(GOLD, "e2e_small_run.json")))
    run = synth.make_run(8, meta["classes"], rows=meta["rows"], cols=meta["cols"], seed=meta["seed"])
    trained = o.train_all(run.events, 4)
    assert [[t.status, t.training_set_size, t.kept_events] for t in trained] == meta["training"]
    text, recs = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks)
    assert o.write_header() + text == open(os.path.join(GOLD, "e2e_small_run.txt")).read()
    codes = {int(l.split()[5]) for l in text.splitlines() if int(l.split()[2]) == 0}
    assert {-3, -9} <= codes


def test_writer_formats():
    """nan / -nan and the padded track blocks (PICOFormatWriterV4.cpp:132-283, bubble.cpp:101-118)."""
    g = o.BubbleImageFrame(10, 20, 4, 5, 12.0, 1.95, 12.0, 150.0, 270.0, np.float32(12.5), np.float32(22.5))
    b = o.Bubble(g)
    txt, nxt = o.form_each_bubble_output("20210101_0", 2, o.stage_camera_output([b], 7, 3), 1, 1, 30)
    f = txt.split("  ")
    assert f[:12] == ["20210101_0", "3", "1", "1", "2", "37", "12.50", "22.50", "4", "5", "-nan", "-nan"]
    assert f[12:22] == [str(37 + j) for j in range(10)] and f[22:72] == ["-1"] * 50 and txt.endswith("1  \n")
    err, _ = o.form_each_bubble_output("r", 1, o.CameraOutput(-3, 0, 9), 1, 0, 30)
    assert err.startswith("r  9  0  0  1  -3  0.00  0.00  0  0  0.00  0.00  0  0")
    assert err.count("0.00") == 2 + 2 + 50


def test_bellows_golden_run():
    """a14: the oracle's bellows veto (cv2.matchTemplate positions, ROI, final text) against the committed vectors."""
    from autobub3hs_b200 import synth
    meta = json.load(open(os.path.join(GOLD, "bellows_run.json")))
    run = synth.make_run(6, meta["classes"], rows=meta["rows"], cols=meta["cols"], seed=meta["seed"], bellows_object=True)
    trained = o.train_all(run.events, 4)
    text, recs = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks, 30,
                              run.bellows_templates)
    assert o.write_header() + text == open(os.path.join(GOLD, "bellows_run.txt")).read()
    got = []
    for ev, rr in enumerate(recs):
        for cam, a in enumerate(rr):
            for info in a.bellows:
                if info.get("bellows_branch"):
                    pt, pp, roi = info["bellows_positions"]
                    got.append((ev, cam, [float(v).hex() for v in pt], [float(v).hex() for v in pp], [int(v) for v in roi]))
    want = [(b["event"], b["cam"], b["pos_trig"], b["pos_pre"], b["roi"]) for b in meta["branch"]]
    assert got == want and len(got) >= 6
    assert {b["cam"] for b in meta["branch"]} == {1, 3}


def test_bellows_helpers():
    """GetDiffROI / nudge arithmetic (L3Localizer.cpp:316-323, :462-470) and the ROI ProcessFrame."""
    F = np.float32
    (t, p) = o.nudge_apart((F(10.25), F(7.5)), (F(10.25), F(9.0)))
    assert t == (F(11.25), F(7.5)) and p == (F(9.25), F(9.0))                     # >= : trig moves right
    (t, p) = o.nudge_apart((F(3.0), F(1.0)), (F(4.5), F(1.0)))
    assert t == (F(2.0), F(1.0)) and p == (F(5.5), F(1.0))
    assert o.get_diff_roi((F(11.25), F(7.5)), (F(9.25), F(9.0)), (30, 20)) == (11, 9, 18, 28)
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (40, 50), dtype=np.uint8)
    b = rng.integers(0, 256, (40, 50), dtype=np.uint8)
    s = rng.integers(0, 4, (40, 50), dtype=np.uint8)
    roi = (7, 5, 20, 11)
    d = o.process_frame_roi(a, b, s, roi)
    assert np.array_equal(d[5:16, 7:27], o.process_frame(a[5:16, 7:27].copy(), b[5:16, 7:27].copy(), s[5:16, 7:27].copy()))
    d[5:16, 7:27] = 0
    assert not d.any()
    for bad in ((45, 5, 20, 11), (-1, 0, 5, 5), (0, 0, 0, 5)):
        with pytest.raises(ValueError):
            o.process_frame_roi(a, b, s, bad)


def test_entropy_trigger_variants_known_answers():
    """calculateEntropyFrame (AnalyzerUnit.cpp:386-423) on histograms with closed-form entropies, the pairing of the
    256-bin table into 128 bins, and calculateEntropySignificance (:425-433) on a hand-computed history."""
    import oracle.cv2_oracle as o
    h = np.zeros(256, np.int64)
    h[0] = 1 << 20
    assert o.calculate_entropy_frame128(h, 1 << 20) == 0.0                 # one bin: p = 1
    h[:] = 0
    h[0] = h[1] = 1 << 19                                                   # grey 0 and 1 share bin 0
    assert o.calculate_entropy_frame128(h, 1 << 20) == 0.0
    h[:] = 0
    h[0] = h[2] = 1 << 19                                                   # two bins of 1/2
    assert o.calculate_entropy_frame128(h, 1 << 20) == 1.0
    h[:] = 1 << 12                                                          # uniform: log2(128)
    assert o.calculate_entropy_frame128(h, 256 << 12) == 7.0
    # against cv2.calcHist with 128 bins on an image
    import cv2
    rng = np.random.default_rng(3)
    img = rng.integers(0, 40, (120, 200)).astype(np.uint8)
    h128 = cv2.calcHist([img], [0], None, [128], [0, 256]).reshape(128)
    assert np.array_equal(h128, o.hist256(img).reshape(128, 2).sum(axis=1))
    # z-score: entropies 0, 1, 1 -> mean 2/3, sigma sqrt(2/3 - 4/9)
    st = o.SignificanceState()
    one = np.zeros(256, np.int64); one[0] = 1 << 20
    two = np.zeros(256, np.int64); two[0] = two[2] = 1 << 19
    with np.errstate(all="ignore"):
        assert np.isnan(o.calculate_entropy_significance(one, 1 << 20, True, st))      # 0/0 on the first frame
    o.calculate_entropy_significance(two, 1 << 20, True, st)
    z = o.calculate_entropy_significance(two, 1 << 20, True, st)
    assert abs(z - (1 - 2 / 3) / np.sqrt(2 / 3 - 4 / 9)) < 1e-12
    z2 = o.calculate_entropy_significance(one, 1 << 20, False, st)                      # store = false: history unchanged
    assert len(st.entropies) == 3 and abs(z2 - (0 - 2 / 3) / np.sqrt(2 / 3 - 4 / 9)) < 1e-12
