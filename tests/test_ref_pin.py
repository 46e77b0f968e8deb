"""The oracle pinned by the reference's OWN code (oracle/_ref, built by oracle/Makefile from the unmodified sources under
/root/reference against oracle/refshim -- stand-ins for the OpenCV / Boost APIs only):

  * oracle/_ref/abub3hs_ref  = AutoBubStart3.cpp main + AnalyzerUnit + Trainer + L3Localizer + bubble + OutputWriter +
    RawParser: run on synthetic BMP runs, its abub3hs_<run>.txt must equal the committed golden texts and the cv2
    oracle's text byte for byte (trigger state machine, significance, Welford, gate, LocalizeOMatic bookkeeping,
    tracking, dZdT / dRdT, -nan printing, error rows);
  * oracle/_ref/ref_harness  = single functions of the same object code: calculateSignificanceFrame / CalcMean /
    CalcStdDev replayed on the 96 golden trigger tables (int32 wrap of val*val, NaN / inf), CalculateMeanSigmaImageVector
    on the golden Trainer stacks, calculateEntropyFrame of the gate.

What is NOT pinned by the reference: the pixel arithmetic inside OpenCV (GaussianBlur, blur, threshold, findContours,
moments ...), which the reference does not own or pin; there the pin is cv2 4.13.0 (tests/test_oracle_cpu.py).
The binaries are prebuilt here (where /root/reference exists) and travel with the repository snapshot."""
import json
import os
import struct
import subprocess

import numpy as np
import pytest

import oracle.cv2_oracle as o
from autobub3hs_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(HERE, "golden")
REFDIR = os.path.join(ROOT, "oracle", "_ref")


@pytest.fixture(scope="module")
def ref_bins():
    if os.path.exists("/root/reference/AutoBubStart3.cpp"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    exe, har = os.path.join(REFDIR, "abub3hs_ref"), os.path.join(REFDIR, "ref_harness")
    if not (os.path.exists(exe) and os.path.exists(har)):
        pytest.skip("oracle/_ref is not built (the reference sources are not on this machine)")
    return exe, har


def run_reference(exe, run, tmp, extra=()):
    data, masks, out = os.path.join(tmp, "data"), os.path.join(tmp, "masks"), os.path.join(tmp, "out")
    synth.write_run(run, data, masks)
    os.makedirs(out, exist_ok=True)
    r = subprocess.run([exe, "-D", "40l-19", "-d", data, "-r", run.run_id, "-o", out, "-c", masks, *extra], capture_output=True,
                       text=True, timeout=1800)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return open(os.path.join(out, f"abub3hs_{run.run_id}.txt")).read()


def test_reference_binary_reproduces_golden_text(ref_bins, tmp_path):
    """The committed end-to-end golden (every event class: normal, multi, no-trigger, late, 4 and 5 frames, flicker,
    high noise) was generated from the cv2 oracle; the reference's own code gives the same bytes."""
    exe, _ = ref_bins
    meta = json.load(open(os.path.join(GOLD, "e2e_small_run.json")))
    run = synth.make_run(8, meta["classes"], rows=meta["rows"], cols=meta["cols"], seed=meta["seed"])
    gold = open(os.path.join(GOLD, "e2e_small_run.txt")).read()
    assert run_reference(exe, run, str(tmp_path)) == gold
    statuses = {int(l.split()[5]) for l in gold.splitlines()[6:]}
    assert {-3, -9}.issubset(statuses) and any(s > 0 for s in statuses)


def test_reference_binary_equals_oracle_on_other_runs(ref_bins, tmp_path):
    """Fresh runs (another pitch, retrigger-prone classes, TrainingSetSize < 6 with its one-frame offset and loc_thres
    rule): cv2 oracle text == reference text."""
    exe, _ = ref_bins
    cases = [
        (6, ["normal", "multi", "notrigger", "flicker", "short5", "highnoise"], 136, 208, 21),
        (2, ["normal", "flicker"], 120, 160, 5),                    # 2 events -> TrainingSetSize 4 < 6
        (7, ["late", "late", "normal", "multi", "highnoise", "flicker", "normal"], 96, 160, 33),
    ]
    for k, (n, classes, rows, cols, seed) in enumerate(cases):
        run = synth.make_run(n, classes, rows=rows, cols=cols, seed=seed)
        trained = o.train_all(run.events, 4)
        want, _ = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks)
        got = run_reference(exe, run, str(tmp_path / f"c{k}"))
        assert got == o.write_header() + want, f"case {k}"


def test_reference_binary_bellows_branch(ref_bins, tmp_path):
    """The bellows veto (L3Localizer.cpp:292-390) through the reference's own TrackAFeature / GetDiffROI / paste code.
    The correlation map behind cv::matchTemplate is exact here and a float32 DFT in OpenCV, so the sub-pixel positions
    can differ in the last float bits; the text only sees them through integer truncation."""
    exe, _ = ref_bins
    meta = json.load(open(os.path.join(GOLD, "bellows_run.json")))
    run = synth.make_run(6, meta["classes"], rows=meta["rows"], cols=meta["cols"], seed=meta["seed"], bellows_object=True)
    gold = open(os.path.join(GOLD, "bellows_run.txt")).read()
    assert len(meta["branch"]) >= 6
    assert run_reference(exe, run, str(tmp_path)) == gold


def _record_significance_calls():
    """Every calculateSignificanceFrame call the oracle makes on the golden trigger tables: (tss, [(store, total, hist)]),
    with the oracle's result (double, loc_thres after the call)."""
    exp = json.load(open(os.path.join(GOLD, "trigger_expected.json")))
    tabs = np.load(os.path.join(GOLD, "trigger_tables.npz"))
    total = exp["total"]
    cases, log = [], []
    orig = o.calculate_significance

    def spy(hist, tot, store, st, tss, *a, **k):
        v = orig(hist, tot, store, st, tss, *a, **k)
        log.append((bool(store), int(tot), np.asarray(hist, np.uint32).copy(), float(v), int(st.loc_thres)))
        return v

    o.calculate_significance = spy
    try:
        for i in range(len(tabs.files)):
            H = tabs[f"H{i}"]
            for tss in (12, 4):
                log.clear()
                st = o.SignificanceState()
                # start frame 1 only: a re-trigger continues with the history of the first search, which a fresh
                # reference object does not have
                o.find_trigger_frame(o.TableSource(len(H), total), None, tss, 1, st, 0, o.TableHistProvider(H))
                cases.append((tss, list(log)))
    finally:
        o.calculate_significance = orig
    return cases


def test_reference_significance_on_golden_tables(ref_bins, tmp_path):
    _, har = ref_bins
    cases = _record_significance_calls()
    fin, fout = str(tmp_path / "sig.in"), str(tmp_path / "sig.out")
    with open(fin, "wb") as f:
        f.write(struct.pack("<i", len(cases)))
        for tss, calls in cases:
            f.write(struct.pack("<ii", tss, len(calls)))
            for store, total, hist, _, _ in calls:
                f.write(struct.pack("<ii", int(store), total))
                f.write(hist.astype("<u4").tobytes())
    subprocess.check_call([har, "significance", fin, fout])
    raw = open(fout, "rb").read()
    ncalls = sum(len(c) for _, c in cases)
    assert len(raw) == 12 * ncalls
    k, n_nan, n_inf, n_wrap = 0, 0, 0, 0
    for tss, calls in cases:
        for store, total, hist, want, want_lt in calls:
            got, lt = struct.unpack_from("<di", raw, 12 * k)
            k += 1
            assert (got == want) or (np.isnan(got) and np.isnan(want)), (tss, k, got, want)
            assert lt == want_lt
            n_nan += int(np.isnan(want))
            n_inf += int(np.isinf(want))
            n_wrap += int((hist[2:] > 46340).any())
    assert ncalls > 1000 and n_nan > 0 and n_inf > 0 and n_wrap > 0, (ncalls, n_nan, n_inf, n_wrap)


def test_reference_welford_and_gate_on_golden_stacks(ref_bins, tmp_path):
    _, har = ref_bins
    g = np.load(os.path.join(GOLD, "trainer.npz"))
    rng = np.random.default_rng(303)
    base = rng.integers(0, 256, (40, 56)).astype(np.float32)
    stack = [np.clip(base + rng.normal(0, 2.5, base.shape), 0, 255).astype(np.uint8) for _ in range(40)]
    for n in (1, 2, 5, 6, 40):
        fin, fout = str(tmp_path / f"w{n}.in"), str(tmp_path / f"w{n}.out")
        with open(fin, "wb") as f:
            f.write(struct.pack("<iii", n, 40, 56))
            for fr in stack[:n]:
                f.write(fr.tobytes())
        subprocess.check_call([har, "welford", fin, fout])
        raw = open(fout, "rb").read()
        mean = np.frombuffer(raw, np.uint8, 40 * 56, 0).reshape(40, 56)
        sigma = np.frombuffer(raw, np.uint8, 40 * 56, 40 * 56).reshape(40, 56)
        assert np.array_equal(mean, g[f"mean{n}"]) and np.array_equal(sigma, g[f"sigma{n}"]), n
        assert struct.unpack_from("<i", raw, 2 * 40 * 56)[0] == n
    # gate entropies: float32, same operations (calcHist, * (float)(1/N), log2f) -> equal to the oracle's float32 value
    noise = rng.integers(0, 256, base.shape).astype(np.uint8)
    pairs = [(stack[0], stack[i]) for i in range(1, 6)] + [(stack[0], noise)]
    fin, fout = str(tmp_path / "g.in"), str(tmp_path / "g.out")
    with open(fin, "wb") as f:
        f.write(struct.pack("<iii", len(pairs), 40, 56))
        for a, b in pairs:
            f.write(a.tobytes())
            f.write(b.tobytes())
    subprocess.check_call([har, "gate", fin, fout])
    got = np.frombuffer(open(fout, "rb").read(), np.float32)
    want = np.array([o.training_gate_entropy(a, b) for a, b in pairs], np.float32)
    assert np.allclose(got, want, atol=1e-6, rtol=0), (got, want)
    assert np.allclose(got.astype(np.float64), g["entropies"], atol=1e-6, rtol=0)
    assert ((got <= 0.0005) == (want <= 0.0005)).all()
