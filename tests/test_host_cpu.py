"""Host C++ pieces that need no GPU (autobub3hs_b200/host): image codecs against cv2.imdecode,
RawParser / ZipParser listings, OutputWriter text against the oracle's writer restatement."""
import os
import subprocess
import zipfile

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
HOST = os.path.join(ROOT, "autobub3hs_b200", "host")
TOOL = os.path.join(HOST, "host_tool")


@pytest.fixture(scope="module")
def tool():
    subprocess.check_call(["make", "-s", "-C", HOST, "host_tool"])
    return TOOL


def _decode(tool, path, tmp_path):
    out = str(tmp_path / "dec.raw")
    r = subprocess.run([tool, "decode", path, out], capture_output=True, text=True)
    head = r.stdout.strip()
    if head == "FAIL":
        return None
    rows, cols = map(int, head.split())
    return np.fromfile(out, np.uint8).reshape(rows, cols)


def test_codecs_match_cv2(tool, tmp_path):
    import cv2
    from autobub3hs_b200 import synth
    rng = np.random.default_rng(1)
    gray = rng.integers(0, 256, (37, 53)).astype(np.uint8)          # odd width: BMP row padding
    color = rng.integers(0, 256, (29, 41, 3)).astype(np.uint8)
    cases = {}
    p = str(tmp_path / "g8.bmp"); synth.write_bmp8(p, gray); cases["bmp8"] = p
    p = str(tmp_path / "c24.bmp"); cv2.imwrite(p, color); cases["bmp24"] = p
    p = str(tmp_path / "g8.png"); cv2.imwrite(p, gray); cases["png_gray"] = p
    p = str(tmp_path / "c24.png"); cv2.imwrite(p, color); cases["png_rgb"] = p
    p = str(tmp_path / "c32.png"); cv2.imwrite(p, np.dstack([color, gray[:29, :41]])); cases["png_rgba"] = p
    p = str(tmp_path / "g16.png"); cv2.imwrite(p, (gray.astype(np.uint16) << 8) | 0x55); cases["png_gray16"] = p
    p = str(tmp_path / "grayrgb.png"); cv2.imwrite(p, np.dstack([gray] * 3)); cases["png_rgb_gray_valued"] = p
    try:
        from PIL import Image
        p = str(tmp_path / "pal.png"); Image.fromarray(gray).convert("P").save(p); cases["png_palette"] = p
        p = str(tmp_path / "bit1.png"); Image.fromarray((gray > 127).astype(np.uint8) * 255).convert("1").save(p); cases["png_1bit"] = p
    except ImportError:
        pass
    for name, path in cases.items():
        ref = cv2.imdecode(np.fromfile(path, np.uint8), cv2.IMREAD_GRAYSCALE)
        got = _decode(tool, path, tmp_path)
        assert got is not None, name
        assert got.shape == ref.shape, name
        if name in ("png_rgb", "png_rgba"):
            # colour PNG frames do not occur in detector data; the gray weights differ between libpng
            # builds by at most one count
            assert np.abs(got.astype(int) - ref.astype(int)).max() <= 1, name
        else:
            assert np.array_equal(got, ref), name
    bad = str(tmp_path / "bad.png")
    open(bad, "wb").write(open(cases["png_gray"], "rb").read()[:200])
    assert _decode(tool, bad, tmp_path) is None
    assert _decode(tool, str(tmp_path / "missing.bmp"), tmp_path) is None


def _make_tree(tmp_path):
    from autobub3hs_b200 import synth
    run = tmp_path / "20210101_0"
    names = {}
    for ev in (0, 2, 10):
        d = run / str(ev) / "Images"
        d.mkdir(parents=True)
        for cam in range(2):
            for f in (30, 31, 100, 9):
                img = np.full((8, 12), (ev * 7 + cam * 3 + f) % 256, np.uint8)
                synth.write_bmp8(str(d / f"cam{cam}_image{f}.bmp"), img)
        (d / "notes.txt").write_text("x")
        names[ev] = sorted(f"cam1_image{f}.bmp" for f in (30, 31, 100, 9))
    (run / ".hidden").mkdir()
    (run / "readme.txt").write_text("x")
    return run, names


def test_raw_parser(tool, tmp_path):
    run, names = _make_tree(tmp_path)
    ev = subprocess.run([tool, "raw-events", str(run) + "/"], capture_output=True, text=True).stdout.split()
    assert sorted(ev, key=int) == ["0", "2", "10"]
    got = subprocess.run([tool, "raw-list", str(run) + "/", "/Images/", "10", "1"], capture_output=True, text=True).stdout.split()
    assert got == names[10]            # lexicographic order (RawParser.cpp:155), only camera 1, only png|bmp
    got = subprocess.run([tool, "raw-list", str(run) + "/", "/Images/", "5", "1"], capture_output=True, text=True).stdout.split()
    assert got == []


@pytest.mark.parametrize("compression", [zipfile.ZIP_STORED, zipfile.ZIP_DEFLATED])
def test_zip_parser(tool, tmp_path, compression):
    run, names = _make_tree(tmp_path)
    zp = str(tmp_path / "20210101_0.zip")
    with zipfile.ZipFile(zp, "w", compression) as z:
        z.writestr("20210101_0/", "")
        for ev in (0, 2, 10):
            z.writestr(f"20210101_0/{ev}/", "")
            z.writestr(f"20210101_0/{ev}/Images/", "")
            d = run / str(ev) / "Images"
            for fn in sorted(os.listdir(d)):
                z.write(str(d / fn), f"20210101_0/{ev}/Images/{fn}")
    ev = subprocess.run([tool, "zip-events", zp], capture_output=True, text=True).stdout.split()
    assert ev == ["0", "2", "10"]
    got = subprocess.run([tool, "zip-list", zp, "2", "1"], capture_output=True, text=True).stdout.split()
    assert got == names[2]
    out = str(tmp_path / "z.raw")
    r = subprocess.run([tool, "zip-get", zp, "2", "cam1_image100.bmp", out], capture_output=True, text=True)
    assert r.stdout.split() == ["8", "12"]
    assert np.all(np.fromfile(out, np.uint8) == (2 * 7 + 3 + 100) % 256)
    r = subprocess.run([tool, "zip-get", zp, "2", "cam1_image77.bmp", out], capture_output=True, text=True)
    assert r.stdout.strip() == "MISSING"           # -1, ZipParser.cpp:198
    # the ".zip" extension is added when absent, a missing archive throws -10 (ZipParser.cpp:41-45, :66-70)
    assert subprocess.run([tool, "zip-events", zp[:-4] + "/"], capture_output=True, text=True).stdout.split() == ["0", "2", "10"]
    assert subprocess.run([tool, "zip-events", str(tmp_path / "nope")], capture_output=True, text=True).stdout.strip() == "THROW -10"


def test_frame_path_matcher_equals_the_reference_regex(tool):
    """RawParser::ParseAndSortFramesInFolder selects frames with regex_match(path, "^.*cam<N>.*(png|bmp)")
    (reference RawParser.cpp:145-150); the written-out matcher must select exactly the same paths."""
    r = subprocess.run([tool, "frame-pattern-fuzz", "100000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1500:]
    assert " 0 mismatches" in r.stdout and " 0 matches" not in r.stdout


def test_batch_formatter_prints_what_the_stream_prints(tool):
    """the to_chars formatter of OutputWriter::appendCameraOutput against `stream << value` under
    fixed / precision(2) on 2M values (random floats of every class, ties at the last digit, non-finite)."""
    r = subprocess.run([tool, "format-fuzz", "500000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1500:]
    assert "0 mismatches" in r.stdout


def test_output_writer_matches_oracle(tool, tmp_path):
    import oracle.cv2_oracle as o
    F32 = np.float32
    r = subprocess.run([tool, "writer-selftest", str(tmp_path) + "/"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1500:]          # includes: batch formatter == stream formatter on these rows
    text = open(r.stdout.strip()).read()

    def frame(x, y, w, h, area, cx, cy):
        return o.BubbleImageFrame(x, y, w, h, area, float(np.sqrt(np.float64(area) / np.float64(3.14159))), 0, 0, 0, F32(cx), F32(cy))

    b0 = o.Bubble(frame(100, 50, 4, 3, 7.5, 101.256, 51.5))
    b1 = o.Bubble(frame(10, 20, 6, 6, 20.0, 12.994999, 23.005))
    for k in range(1, 13):
        b1.lock = False
        b1.push(frame(10 - k, 20 + k, 6 + k, 6 + 2 * k, 20.0 + 3.3 * k, F32(12.25) - F32(k), F32(23.125) + F32(1.5) * F32(k)))
    b2 = o.Bubble(frame(500, 400, 2, 9, 0.0, 500.5, 404.0))
    for k in range(1, 4):
        b2.lock = False
        with np.errstate(all="ignore"):
            nan = F32(0) / F32(0)
        b2.push(frame(500, 400 + k, 3, 9, 0.5 * k, nan if k == 2 else 500.125, 404.0 + k))
    cams = [o.stage_camera_output([b0, b1], 17, 3), o.stage_camera_output([b2], 33, 3), o.CameraOutput(-3, 0, 3),
            o.stage_camera_output([], 5, 3)]
    exp = o.write_header() + o.write_camera_output("20210101_0", cams, 30)
    exp += o.write_camera_output("20210101_0", [o.CameraOutput(-7, 0, 4) for _ in range(4)], 30)
    assert text == exp


def test_zip_parser_rejects_sizes_a_corrupt_directory_claims(tool, tmp_path):
    """Entry sizes come from the archive's central directory; a corrupt one must end as "frame does not decode" (the
    per-camera -9 / -6 row), not as a bad_alloc inside a parallel region.  The compressed / uncompressed sizes of one
    entry are overwritten in the central directory with values far beyond the file: that frame reads as MISSING, its
    neighbours still decode, and a directory whose own size field points outside the file throws -10 like an unreadable
    archive (ZipParser.cpp:66-70)."""
    import struct
    run, names = _make_tree(tmp_path)
    zp = str(tmp_path / "20210101_0.zip")
    with zipfile.ZipFile(zp, "w", zipfile.ZIP_DEFLATED) as z:
        for ev in (0, 2):
            d = run / str(ev) / "Images"
            for fn in sorted(os.listdir(d)):
                z.write(str(d / fn), f"20210101_0/{ev}/Images/{fn}")
    raw = bytearray(open(zp, "rb").read())
    target = b"20210101_0/2/Images/cam1_image100.bmp"
    pos, hits = 0, []
    while True:
        pos = raw.find(b"PK\x01\x02", pos)
        if pos < 0:
            break
        nlen = struct.unpack_from("<H", raw, pos + 28)[0]
        if bytes(raw[pos + 46:pos + 46 + nlen]) == target:
            hits.append(pos)
        pos += 4
    assert len(hits) == 1
    out = str(tmp_path / "z.raw")
    for comp, uncomp in ((0xFFFFFFF0, 0x1000), (0x20, 0xFFFFFFF0)):       # compressed bytes past the file / a 4 GB frame
        bad = bytearray(raw)
        struct.pack_into("<II", bad, hits[0] + 20, comp, uncomp)
        zb = str(tmp_path / "bad" / "20210101_0.zip")
        os.makedirs(os.path.dirname(zb), exist_ok=True)
        open(zb, "wb").write(bad)
        r = subprocess.run([tool, "zip-get", zb, "2", "cam1_image100.bmp", out], capture_output=True, text=True)
        assert r.returncode == 0 and r.stdout.strip() == "MISSING", (r.returncode, r.stdout, r.stderr[-300:])
        r = subprocess.run([tool, "zip-get", zb, "2", "cam1_image31.bmp", out], capture_output=True, text=True)
        assert r.stdout.split() == ["8", "12"]
    eocd = raw.rfind(b"PK\x05\x06")
    bad = bytearray(raw)
    struct.pack_into("<I", bad, eocd + 12, 0x7FFFFFF0)                     # central directory "larger than the file"
    open(zb, "wb").write(bad)
    assert subprocess.run([tool, "zip-events", zb], capture_output=True, text=True).stdout.split()[-2:] == ["THROW", "-10"]


def test_rank_status_rendezvous(tool, tmp_path):
    """RankSync (host/RankSync.hpp), the status exchange `abub3hs --world N` makes before every phase that ends in a
    collective: every rank sees the minimum status, a failing rank takes everybody out of the next phase, a rank that
    never shows up costs a bounded wait instead of a hang, and leftovers of an earlier run are ignored."""
    import time
    base = str(tmp_path / ".abub3hs_run.status")

    def launch(rank, world, phases, start=0, timeout="30"):
        return subprocess.Popen([tool, "rank-sync", base, str(rank), str(world), str(start)] + phases, stdout=subprocess.PIPE, text=True,
                                env=dict(os.environ, ABUB_RANK_TIMEOUT=timeout))

    # all fine through two phases
    ps = [launch(r, 3, ["context=0", "trained=0"]) for r in range(3)]
    outs = [p.communicate(timeout=60)[0].split() for p in ps]
    assert outs == [["context", "0", "trained", "0"]] * 3
    assert not [f for f in os.listdir(tmp_path) if "status" in f]              # every rank removed its own files
    # rank 1 fails in the first phase: everybody stops there
    ps = [launch(r, 3, [f"context={-1 if r == 1 else 0}", "trained=0"]) for r in range(3)]
    outs = [p.communicate(timeout=60)[0].split() for p in ps]
    assert outs == [["context", "-1"]] * 3
    # rank 2 never arrives: bounded wait, -1; a status file left by a run of an hour ago does not count as its arrival
    open(base + ".context.2", "w").write(f"{int(time.time()) - 3600} 0\n")
    t0 = time.time()
    ps = [launch(r, 3, ["context=0", "trained=0"], timeout="2") for r in range(2)]
    outs = [p.communicate(timeout=60)[0].split() for p in ps]
    assert [o[-2:] for o in outs] == [["context", "-1"]] * 2 and time.time() - t0 < 20      # after the "did not reach" notice
    assert all("did not reach phase" in " ".join(o) for o in outs)
