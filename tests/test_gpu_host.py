"""The drop-in binary: abub3hs (host C++ class surface over libabub_cuda.so) on synthetic runs written
to disk, its abub3hs_<run>.txt compared byte for byte with the oracle's text."""
import json
import os
import subprocess
import zipfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
HOST = os.path.join(ROOT, "autobub3hs_b200", "host")
GOLD = os.path.join(HERE, "golden")


@pytest.fixture(scope="module")
def abub3hs():
    from autobub3hs_b200 import build
    build.build()
    subprocess.check_call(["make", "-s", "-C", HOST])
    return os.path.join(HOST, "abub3hs")


def _run(abub3hs, data_dir, out_dir, mask_dir, extra=()):
    os.makedirs(out_dir, exist_ok=True)
    cmd = [abub3hs, "-D", "40l-19", "-d", data_dir, "-r", "20210101_0", "-o", out_dir, "-c", mask_dir, *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return open(os.path.join(out_dir, "abub3hs_20210101_0.txt")).read()


def test_golden_run_raw(abub3hs, tmp_path):
    """The committed golden run (all event classes, 132x212: generic-pitch kernels) from BMP files."""
    from autobub3hs_b200 import synth
    meta = json.load(open(os.path.join(GOLD, "e2e_small_run.json")))
    run = synth.make_run(8, meta["classes"], rows=meta["rows"], cols=meta["cols"], seed=meta["seed"])
    synth.write_run(run, str(tmp_path / "data"), str(tmp_path / "masks"))
    gold = open(os.path.join(GOLD, "e2e_small_run.txt")).read()
    for batch in ("1", "3", "8"):
        text = _run(abub3hs, str(tmp_path / "data"), str(tmp_path / ("out" + batch)), str(tmp_path / "masks"), ["--batch", batch])
        assert text == gold, "batch " + batch
    # -e: a single event index (AutoBubStart3.cpp:350)
    text = _run(abub3hs, str(tmp_path / "data"), str(tmp_path / "out_e"), str(tmp_path / "masks"), ["-e", "1"])
    import oracle.cv2_oracle as o
    hdr = o.write_header()
    rows = [l for l in gold[len(hdr):].splitlines(True) if l.split("  ")[1] == "1"]
    assert text == hdr + "".join(rows)


@pytest.fixture(scope="module")
def tma_run():
    """208 columns (multiple of 16): the TMA-fed trigger kernel is on the path."""
    from autobub3hs_b200 import synth
    import oracle.cv2_oracle as o
    run = synth.make_run(6, ["normal", "multi", "notrigger", "flicker", "short5", "highnoise"], rows=136, cols=208, seed=21)
    trained = o.train_all(run.events, 4)
    text, _ = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks)
    return run, o.write_header() + text


def test_tma_pitch_raw_png_and_zip(abub3hs, tma_run, tmp_path):
    import cv2
    from autobub3hs_b200 import synth
    run, gold = tma_run
    synth.write_run(run, str(tmp_path / "data"), str(tmp_path / "masks"))
    assert _run(abub3hs, str(tmp_path / "data"), str(tmp_path / "out_bmp"), str(tmp_path / "masks")) == gold
    # the same run as PNG files inside a zip archive (-z)
    zdir = tmp_path / "zdata"
    zdir.mkdir()
    with zipfile.ZipFile(str(zdir / "20210101_0.zip"), "w", zipfile.ZIP_STORED) as z:
        z.writestr("20210101_0/", "")
        for evn, ev in zip(run.event_numbers, run.events):
            z.writestr(f"20210101_0/{evn}/", "")
            z.writestr(f"20210101_0/{evn}/Images/", "")
            for c, stack in enumerate(ev):
                for f in range(len(stack)):
                    ok, buf = cv2.imencode(".png", stack[f])
                    z.writestr(f"20210101_0/{evn}/Images/cam{c}_image{run.frame_offset + f}.png", buf.tobytes())
    assert _run(abub3hs, str(zdir), str(tmp_path / "out_zip"), str(tmp_path / "masks"), ["-z", "--batch", "4"]) == gold


def test_png_frames_decoded_on_device(abub3hs, tma_run, tmp_path):
    """Row f-2: 8-bit grey PNG frames are inflated and unfiltered on the device (K9); frames the device path does not
    take (RGB PNG, a BMP among PNGs) or rejects (truncated stream) go through the host decoder, with the same result as
    ABUB_HOST_DECODE=1 and as the oracle on the decoded pixels."""
    import cv2
    from autobub3hs_b200 import synth
    run, gold = tma_run
    synth.write_run(run, str(tmp_path / "data"), str(tmp_path / "masks"), fmt="png")

    def go(out, env=None):
        os.makedirs(out, exist_ok=True)
        r = subprocess.run([abub3hs, "-D", "40l-19", "-d", str(tmp_path / "data"), "-r", run.run_id, "-o", out, "-c", str(tmp_path / "masks"),
                            "--batch", "4"], capture_output=True, text=True, env=dict(os.environ, ABUB_TIMING="1", **(env or {})))
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        dec = [l for l in r.stderr.splitlines() if l.startswith("ABUB_DECODE")][-1]
        n_dev, n_host = int(dec.split()[6].rstrip(",")), int(dec.split()[-1])
        return open(os.path.join(out, f"abub3hs_{run.run_id}.txt")).read(), n_dev, n_host

    text, n_dev, n_host = go(str(tmp_path / "o1"))
    assert text == gold
    nframes = sum(len(stack) for ev in run.events for stack in ev)
    assert n_dev == nframes and n_host == 0
    text, n_dev, n_host = go(str(tmp_path / "o2"), {"ABUB_HOST_DECODE": "1"})
    assert text == gold and n_dev == 0
    # event 0: cam 1 frame 3 as an RGB PNG (same grey values), cam 2 frame 5 as the tail of a truncated file
    idir = tmp_path / "data" / run.run_id / str(run.event_numbers[0]) / "Images"
    f3 = idir / f"cam1_image{run.frame_offset + 3}.png"
    cv2.imwrite(str(f3), cv2.cvtColor(run.events[0][1][3], cv2.COLOR_GRAY2BGR))
    f5 = idir / f"cam2_image{run.frame_offset + 5}.png"
    blob = f5.read_bytes()
    f5.write_bytes(blob[: len(blob) * 2 // 3])
    # event 1: a flipped byte in the compressed data of cam 0 frame 2 (chunk CRC fails -> unreadable, as with libpng)
    f2 = tmp_path / "data" / run.run_id / str(run.event_numbers[1]) / "Images" / f"cam0_image{run.frame_offset + 2}.png"
    blob = bytearray(f2.read_bytes())
    blob[len(blob) // 2] ^= 0x10
    f2.write_bytes(bytes(blob))
    t_dev, n_dev, n_host = go(str(tmp_path / "o3"))
    t_host, _, _ = go(str(tmp_path / "o4"), {"ABUB_HOST_DECODE": "1"})
    assert t_dev == t_host
    assert n_dev == nframes - 3 and n_host >= 1
    # events 2.. are untouched
    touched = {str(run.event_numbers[0]), str(run.event_numbers[1])}
    rest = lambda t: [l for l in t.splitlines() if l.split()[:1] == [run.run_id] and l.split()[1] not in touched]
    assert rest(t_dev) == rest(gold)


@pytest.mark.parametrize("metric,thr", [("entropy_significance", None), ("entropy", 0.02)])
def test_trigger_metric_option(abub3hs, tma_run, tmp_path, metric, thr):
    """8(f)-4 through the binary: --trigger_metric selects calculateEntropyFrame / calculateEntropySignificance
    (AnalyzerUnit.cpp:226-227) for the whole run; output text as the oracle's with the same line compiled in."""
    from autobub3hs_b200 import synth
    import oracle.cv2_oracle as o
    run, _ = tma_run
    trained = o.train_all(run.events, 4)
    text, _ = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks,
                           metric=metric, threshold=thr)
    synth.write_run(run, str(tmp_path / "data"), str(tmp_path / "masks"))
    extra = ["--trigger_metric", metric] + ([] if thr is None else ["--trigger_threshold", str(thr)])
    assert _run(abub3hs, str(tmp_path / "data"), str(tmp_path / "out"), str(tmp_path / "masks"), extra) == o.write_header() + text


def test_unreadable_run_is_minus5_and_bad_training_is_minus7(abub3hs, tmp_path):
    from autobub3hs_b200 import synth
    import oracle.cv2_oracle as o
    out = tmp_path / "out"
    out.mkdir()
    r = subprocess.run([abub3hs, "-D", "40l-19", "-z", "-d", str(tmp_path / "nowhere"), "-r", "20210101_0", "-o", str(out)],
                       capture_output=True, text=True)
    assert r.returncode == (-5) % 256
    text = open(out / "abub3hs_20210101_0.txt").read()
    exp = o.write_header() + o.write_camera_output("20210101_0", [o.CameraOutput(-5, 0, -1) for _ in range(4)], 30)
    assert text == exp
    # every event fails the Trainer gate (frame 1 differs wildly from frame 0) -> -7 rows for all events
    rng = np.random.default_rng(0)
    data = tmp_path / "data"
    for ev in range(2):
        d = data / "20210101_0" / str(ev) / "Images"
        d.mkdir(parents=True)
        for c in range(4):
            for f in range(6):
                synth.write_bmp8(str(d / f"cam{c}_image{30 + f}.bmp"), rng.integers(0, 256, (48, 64)).astype(np.uint8))
    out7 = tmp_path / "out7"
    out7.mkdir()
    r = subprocess.run([abub3hs, "-D", "40l-19", "-d", str(data), "-r", "20210101_0", "-o", str(out7)], capture_output=True, text=True)
    assert r.returncode == (-7) % 256, r.stdout[-1500:]
    exp = o.write_header()
    for ev in range(2):
        exp += o.write_camera_output("20210101_0", [o.CameraOutput(-7, 0, ev) for _ in range(4)], 30)
    assert open(out7 / "abub3hs_20210101_0.txt").read() == exp


def _ngpus():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


@pytest.mark.skipif(_ngpus() < 2, reason="needs 2 GPUs")
def test_two_ranks_same_text(abub3hs, tma_run, tmp_path):
    """One process per GPU (--rank/--world): contiguous event ranges, Trainer partials all-reduced over
    NCCL with the exact fix-up, rank-ordered concatenation of the rows == the single-process file."""
    from autobub3hs_b200 import synth
    run, gold = tma_run
    synth.write_run(run, str(tmp_path / "data"), str(tmp_path / "masks"))
    out = tmp_path / "out2"
    out.mkdir()
    procs = []
    for r in range(2):
        cmd = [abub3hs, "-D", "40l-19", "-d", str(tmp_path / "data"), "-r", "20210101_0", "-o", str(out), "-c",
               str(tmp_path / "masks"), "--rank", str(r), "--world", "2", "--batch", "2"]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    logs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(l[-1500:] for l in logs)
    assert open(out / "abub3hs_20210101_0.txt").read() == gold
    assert not [f for f in os.listdir(out) if "rank" in f or "ncclid" in f]


def test_two_ranks_one_cannot_train_both_end_minus7(abub3hs, tmp_path):
    """Multi-rank failure path (both processes on GPU 0: no communicator is ever created): rank 1's event range holds no
    readable frame, so it cannot take part in the training.  The ranks agree on that through the status files BEFORE
    anyone enters ncclCommInitRank: both end with -7, rank 0 writes the -7 rows of every event (AutoBubStart3.cpp:315-325),
    nobody blocks, and no part / status / id file is left behind."""
    from autobub3hs_b200 import synth
    import oracle.cv2_oracle as o
    rng = np.random.default_rng(1)
    data = tmp_path / "data"
    quiet = rng.integers(60, 120, (48, 64)).astype(np.uint8)
    for ev in range(2):
        d = data / "20210101_0" / str(ev) / "Images"
        d.mkdir(parents=True)
        if ev == 1:
            continue                                     # rank 1's only event: the folder is there, the frames are not
        for c in range(4):
            for f in range(6):
                synth.write_bmp8(str(d / f"cam{c}_image{30 + f}.bmp"), quiet)
    out = tmp_path / "out"
    out.mkdir()
    procs = []
    for r in range(2):
        cmd = [abub3hs, "-D", "40l-19", "-d", str(data), "-r", "20210101_0", "-o", str(out), "--rank", str(r), "--world", "2",
               "--device", "0"]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                      env=dict(os.environ, ABUB_RANK_TIMEOUT="120")))
    logs = [p.communicate(timeout=300)[0] for p in procs]
    assert [p.returncode for p in procs] == [(-7) % 256] * 2, "\n".join(l[-1500:] for l in logs)
    exp = o.write_header()
    for ev in range(2):
        exp += o.write_camera_output("20210101_0", [o.CameraOutput(-7, 0, ev) for _ in range(4)], 30)
    assert open(out / "abub3hs_20210101_0.txt").read() == exp
    assert sorted(os.listdir(out)) == ["abub3hs_20210101_0.txt"], os.listdir(out)


def test_bellows_run_with_and_without_template(abub3hs, tmp_path):
    """Cameras 1 and 3 show a moving bellows: every genesis contour lies in the bellows mask and the veto branch
    (L3Localizer.cpp:292-390) runs -- with the template PNG in the mask directory, and with it missing."""
    from autobub3hs_b200 import synth
    import oracle.cv2_oracle as o
    run = synth.make_run(6, ["normal", "bellows", "bellows_bubble", "notrigger", "bellows", "normal"], rows=264, cols=420, seed=5,
                         bellows_object=True)
    trained = o.train_all(run.events, 4)
    synth.write_run(run, str(tmp_path / "data"), str(tmp_path / "masks"))
    want, recs = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks, 30,
                              run.bellows_templates)
    assert sum(1 for rr in recs for a in rr for i in a.bellows if i.get("bellows_branch")) >= 6
    assert _run(abub3hs, str(tmp_path / "data"), str(tmp_path / "out"), str(tmp_path / "masks"), ["--batch", "3"]) == o.write_header() + want
    for c in (1, 3):
        os.remove(str(tmp_path / "masks" / f"cam{c}_bellows_template.png"))
    want2, _ = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks, 30, None)
    assert want2 != want
    assert _run(abub3hs, str(tmp_path / "data"), str(tmp_path / "out2"), str(tmp_path / "masks")) == o.write_header() + want2


@pytest.mark.parametrize("pinned", [False, True])
def test_host_library_recon_in_memory(small_run, small_trained, tmp_path, pinned):
    """libabubhost.so (include/abub_host.h): the abub3hs detect loop over events placed in HBM slots / pinned staging
    buffers -- every event class at 264x420 (generic-pitch kernels), one batch and ragged batches."""
    import oracle.cv2_oracle as o
    from tests import glue
    run, trained = small_run, small_trained
    want, _ = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks)
    want = o.write_header() + want
    for batch, early in ((None, 6), (3, 0), (1, 4)):
        text, stats = glue.host_recon_text(run, trained, tmp_path, f"b{batch}e{early}", pinned=pinned, early_chunk=early, batch=batch)
        assert text == want, (batch, early)
    # trigger searches queued ahead of time (abub_trigger_early_deferred): the next sub-batch of a call, the announced
    # next call (abubhost_recon_hint), and a hint that is wrong about the next call
    for tag, kw in (("sub3", dict(sub_batch=3)), ("hint3", dict(batch=3, hint=True)), ("hint2sub1", dict(batch=2, sub_batch=1, hint=True)),
                    ("wronghint", dict(batch=3, hint="wrong"))):
        text, stats = glue.host_recon_text(run, trained, tmp_path, tag, pinned=pinned, early_chunk=5, **kw)
        assert text == want, tag


def test_host_library_recon_bellows_and_tma_pitch(tma_run, tmp_path):
    import oracle.cv2_oracle as o
    from autobub3hs_b200 import synth
    from tests import glue
    run, gold = tma_run
    trained = o.train_all(run.events, 4)
    for pinned in (False, True):
        text, _ = glue.host_recon_text(run, trained, tmp_path, f"tma{int(pinned)}", pinned=pinned, early_chunk=6, batch=4)
        assert text == gold
    brun = synth.make_run(6, ["normal", "bellows", "bellows_bubble", "notrigger", "bellows", "normal"], rows=264, cols=420, seed=5,
                          bellows_object=True)
    btr = o.train_all(brun.events, 4)
    want, recs = o.run_detect(brun.run_id, brun.event_numbers, brun.events, btr, brun.cam_masks, brun.bellows_masks, 30,
                              brun.bellows_templates)
    for pinned in (False, True):
        text, _ = glue.host_recon_text(brun, btr, tmp_path, f"bel{int(pinned)}", pinned=pinned, early_chunk=6, batch=3)
        assert text == o.write_header() + want
