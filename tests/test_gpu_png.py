"""GPU parity of the on-device PNG decode (K9): frames encoded by cv2 / by hand-made encoders with
every row filter and every deflate block type, decoded through abub_event_decode_png and compared byte for byte with
cv2.imdecode (what cv::imread returns in the reference, RawParser.cpp:39)."""
import struct
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def capi():
    from autobub3hs_b200 import capi
    capi.load()
    return capi


def idat_stream(png: bytes) -> bytes:
    assert png[:8] == b"\x89PNG\r\n\x1a\n"
    pos, z = 8, b""
    while pos + 12 <= len(png):
        n = struct.unpack(">I", png[pos:pos + 4])[0]
        if png[pos + 4:pos + 8] == b"IDAT":
            z += png[pos + 8:pos + 8 + n]
        pos += 12 + n
    return z


def filter_rows(img: np.ndarray, ftypes) -> bytes:
    """PNG filtering of an 8-bit grey image with the given filter type per row."""
    rows, cols = img.shape
    out = bytearray()
    prev = np.zeros(cols, np.int32)
    for y in range(rows):
        cur = img[y].astype(np.int32)
        ft = ftypes[y % len(ftypes)]
        left = np.concatenate([[0], cur[:-1]])
        ul = np.concatenate([[0], prev[:-1]])
        if ft == 0:
            f = cur
        elif ft == 1:
            f = cur - left
        elif ft == 2:
            f = cur - prev
        elif ft == 3:
            f = cur - ((left + prev) >> 1)
        else:
            p = left + prev - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
            f = cur - pred
        out.append(ft)
        out += (f & 255).astype(np.uint8).tobytes()
        prev = cur
    return bytes(out)


def synth_frame(rows, cols, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:rows, 0:cols]
    base = 90 + 60 * np.sin(xx / 37.0) * np.cos(yy / 23.0)
    base[rows // 4: rows // 2, cols // 5: cols // 3] = 0            # flat black block (long matches)
    return np.clip(np.rint(base + rng.normal(0, 2.0, (rows, cols))), 0, 255).astype(np.uint8)


def test_cv2_encoded_frames(capi):
    import cv2
    rows, cols = 264, 420
    frames = [synth_frame(rows, cols, s) for s in range(6)]
    jobs = []
    for f, img in enumerate(frames):
        for level in ((1, 9, 3, 6, 4, 2)[f],):
            ok, enc = cv2.imencode(".png", img, [cv2.IMWRITE_PNG_COMPRESSION, level])
            assert ok and np.array_equal(cv2.imdecode(enc, cv2.IMREAD_GRAYSCALE), img)
            jobs.append((0, f % 4, f, idat_stream(enc.tobytes())))
    with capi.Context(rows, cols, 4, 8, max_slots=1) as ctx:
        okv = ctx.decode_png(jobs)
        assert okv.all()
        ctx.event_set_nframes(0, [8, 8, 8, 8])
        dev = ctx.download_frames(0)
        for (s, c, f, _), img in zip(jobs, frames):
            assert np.array_equal(dev[c, f], img), (c, f)


@pytest.mark.parametrize("ftypes", [(0,), (1,), (2,), (3,), (4,), (0, 1, 2, 3, 4), (4, 3, 1)])
@pytest.mark.parametrize("mode", ["default", "stored", "fixed", "rle"])
def test_filters_and_block_types(capi, ftypes, mode):
    rows, cols = 70, 134                     # scanlines of 135 bytes: no row starts aligned
    img = synth_frame(rows, cols, 11 + len(ftypes))
    raw = filter_rows(img, ftypes)
    if mode == "stored":
        # a stored block with the first 3000 bytes, an empty stored block (sync flush), then compressed data
        k = 3000
        co = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = co.compress(raw[k:k + 2000]) + co.flush(zlib.Z_SYNC_FLUSH) + co.compress(raw[k + 2000:]) + co.flush()
        z = b"\x78\x01" + b"\x00" + struct.pack("<HH", k, k ^ 0xFFFF) + raw[:k] + body + struct.pack(">I", zlib.adler32(raw))
        assert zlib.decompress(z) == raw
    else:
        strat = {"default": zlib.Z_DEFAULT_STRATEGY, "fixed": zlib.Z_FIXED, "rle": zlib.Z_RLE}[mode]
        co = zlib.compressobj(6, zlib.DEFLATED, 15, 8, strat)
        z = co.compress(raw) + co.flush()
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        okv = ctx.decode_png([(0, 0, 2, z)])
        assert okv.all()
        ctx.event_set_nframes(0, [4])
        assert np.array_equal(ctx.download_frames(0)[0, 2], img)


def test_full_size_frames_and_corrupt_stream(capi):
    import cv2
    from autobub3hs_b200 import synth
    rows, cols = synth.ROWS, synth.COLS
    bg = synth.make_background(1, rows, cols).astype(np.float32)
    rng = np.random.default_rng(4)
    frames = [np.clip(np.rint(bg + 2.0 * rng.standard_normal((rows, cols), dtype=np.float32)), 0, 255).astype(np.uint8) for _ in range(3)]
    streams = [idat_stream(cv2.imencode(".png", f)[1].tobytes()) for f in frames]
    bad = bytearray(streams[1])
    bad[len(bad) // 2] ^= 0x5A                # flips bits in the middle of the deflate data
    trunc = streams[2][: len(streams[2]) // 2]
    with capi.Context(rows, cols, 1, 8, max_slots=1) as ctx:
        okv = ctx.decode_png([(0, 0, 0, streams[0]), (0, 0, 1, bytes(bad)), (0, 0, 2, trunc), (0, 0, 3, streams[1]), (0, 0, 4, streams[2])])
        ctx.event_set_nframes(0, [8])
        dev = ctx.download_frames(0)[0]
        assert okv[0] and okv[3] and okv[4] and not okv[2]
        assert np.array_equal(dev[0], frames[0]) and np.array_equal(dev[3], frames[1]) and np.array_equal(dev[4], frames[2])
        if not okv[1]:
            assert not dev[1].any()            # undecodable frames are zero-filled
        assert not dev[2].any()


@pytest.mark.parametrize("shape", [(7, 12), (70, 136), (1050, 1680)])
def test_bmp_pixel_arrays_turned_over_on_device(capi, shape):
    """abub_event_flip_rows: frames uploaded in BMP file order (bottom-up) come out as cv::imread returns them; frames
    not named keep their rows (odd and even row counts, the full-size frame)."""
    rows, cols = shape
    rng = np.random.default_rng(rows + cols)
    fr = rng.integers(0, 256, (2, 5, rows, cols)).astype(np.uint8)
    with capi.Context(rows, cols, 2, 5, max_slots=2) as ctx:
        flipped = [(1, 0, 0), (1, 1, 3), (1, 1, 4), (0, 0, 2)]
        for slot in range(2):
            for c in range(2):
                up = fr[c].copy()
                for (s_, c_, f_) in flipped:
                    if s_ == slot and c_ == c:
                        up[f_] = fr[c, f_, ::-1]
                ctx.event_upload(slot, c, up)
        ctx.event_flip_rows(flipped)
        for slot in range(2):
            assert np.array_equal(ctx.download_frames(slot), fr), slot
