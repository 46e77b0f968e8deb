"""GPU parity: K3/K4 Trainer statistics and K5-K7 localizer through the C ABI vs the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def capi():
    from autobub3hs_b200 import capi
    capi.load()
    return capi


def test_trainer_matches_oracle(capi, small_run, small_trained):
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        for c in range(4):
            ctx.train_begin(c)
            for ev in range(len(run.events)):
                fr = run.events[ev][c]
                e, kept = ctx.train_add_event(c, fr[0], fr[1])
                assert abs(e - trained[c].entropies[ev]) <= 1e-6          # float32 entropy, abs tolerance 1e-6
                assert kept == (ev in trained[c].kept_events)
            mean, sigma, n, st = ctx.train_finish(c)
            assert (n, st) == (trained[c].training_set_size, trained[c].status)
            assert np.array_equal(mean, trained[c].avg)
            assert np.array_equal(sigma, trained[c].sigma)


@pytest.mark.parametrize("n", [1, 2, 3, 5, 6, 33])
def test_welford_sizes(capi, n):
    import oracle.cv2_oracle as o
    rng = np.random.default_rng(n)
    rows, cols = 50, 68
    base = rng.integers(0, 256, (rows, cols))
    frames = [np.clip(base + rng.normal(0, 3, (rows, cols)).round(), 0, 255).astype(np.uint8) for _ in range(n + (n & 1))]
    frames = frames[: n + (n & 1)]
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        ctx.train_begin(0)
        kept = []
        for i in range(0, len(frames), 2):
            e, k = ctx.train_add_event(0, frames[i], frames[i])      # identical pair: entropy 0 -> always kept
            assert k and abs(e) <= 1e-6
            kept += [frames[i], frames[i]]
        mean, sigma, cnt, st = ctx.train_finish(0)
        m_ref, s_ref = o.welford_mean_sigma(kept)
        assert st == 0 and cnt == len(kept)
        assert np.array_equal(mean, m_ref) and np.array_equal(sigma, s_ref)


def test_train_nothing_kept_is_minus7(capi):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 40, (40, 52)).astype(np.uint8)
    b = rng.integers(0, 256, (40, 52)).astype(np.uint8)       # difference spread over many bins
    with capi.Context(40, 52, 1, 4, max_slots=1) as ctx:
        ctx.train_begin(0)
        e, k = ctx.train_add_event(0, a, b)
        assert not k and e > 0.0005
        _, _, n, st = ctx.train_finish(0)
        assert (n, st) == (0, -7)


def test_train_partials_single_rank_equals_sequential(capi, small_run, small_trained):
    """Exact integer partials + boundary fix-up reproduce the sequential float32 Welford."""
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        ctx.train_begin(1)
        for ev in range(len(run.events)):
            ctx.train_add_event(1, run.events[ev][1][0], run.events[ev][1][1])
        ctx.train_partials(1)
        mean, sigma, n, st, nflag = ctx.train_finish_partials(1)
        assert st == 0 and n == trained[1].training_set_size
        assert np.array_equal(mean, trained[1].avg) and np.array_equal(sigma, trained[1].sigma)
        assert nflag < mean.size * 0.2


def _check_blobs(o, blobs, binary):
    import cv2
    cs, _ = cv2.findContours(binary, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_TC89_L1)
    assert len(cs) == len(blobs)
    for c, b in zip(cs, blobs):
        assert cv2.boundingRect(c) == (b.x, b.y, b.w, b.h)
        assert len(c) == b.npts
        assert cv2.contourArea(c) == b.area
        m = cv2.moments(c)
        assert (m["m00"], m["m10"], m["m01"]) == (b.m00, b.m10, b.m01)
        assert int(c[:, 0, 0].sum()) == b.sum_x and int(c[:, 0, 1].sum()) == b.sum_y


def test_localizer_blobs_vs_cv2(capi, small_run, small_trained):
    import oracle.cv2_oracle as o
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    with capi.Context(rows, cols, 4, 40, max_slots=nev) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
        items = []
        for ev in range(nev):
            for c in range(4):
                ctx.event_upload(ev, c, run.events[ev][c])
                n = len(run.events[ev][c])
                if n > 12:
                    for trig, lt in ((n // 2, 3), (n - 3, 2), (2, 3)):
                        items.append((ev, c, trig, max(0, trig - 2), lt, min(10, n - 1 - trig)))
        res = ctx.localize_batch(items)
        nblobs = 0
        for it, frames in zip(items, res):
            ev, c, trig, pre, lt, npost = it
            fr = run.events[ev][c]
            _check_blobs(o, frames[0], o.genesis_binary(fr[trig], fr[pre], trained[c].sigma, lt))
            for k in range(1, npost + 1):
                _check_blobs(o, frames[k], o.post_trigger_binary(fr[trig + k], trained[c].avg, trained[c].sigma))
                nblobs += len(frames[k])
        assert nblobs > 50
        # single-item entry point + binary tap
        ev, c, trig, pre, lt, npost = items[0]
        blobs, counts = ctx.localize(ev, c, trig, pre, lt, npost)
        assert counts == [len(f) for f in res[0]]
        fr = run.events[ev][c]
        assert np.array_equal(ctx.localize_binary(0), o.genesis_binary(fr[trig], fr[pre], trained[c].sigma, lt))
        assert np.array_equal(ctx.localize_binary(1), o.post_trigger_binary(fr[trig + 1], trained[c].avg, trained[c].sigma))


def test_contours_fuzz_through_device(capi):
    """Random binary shapes pushed through the device contour path (as a post-trigger frame with
    avg = 0, sigma = 0 the binarised image is controlled by the input) vs cv2."""
    import cv2
    import oracle.cv2_oracle as o
    rng = np.random.default_rng(11)
    rows, cols = 96, 160
    zero = np.zeros((rows, cols), np.uint8)
    with capi.Context(rows, cols, 1, 12, max_slots=1) as ctx:
        ctx.set_background(0, zero, zero, 12)
        for rep in range(6):
            frames = np.zeros((12, rows, cols), np.uint8)
            for f in range(1, 12):
                img = np.zeros((rows, cols), np.uint8)
                kind = (f + rep) % 4
                if kind == 0:
                    img = ((rng.random((rows, cols)) < rng.uniform(0.02, 0.5)) * 200).astype(np.uint8)
                elif kind == 1:
                    for _ in range(rng.integers(1, 9)):
                        cv2.ellipse(img, (int(rng.integers(0, cols)), int(rng.integers(0, rows))),
                                    (int(rng.integers(1, 30)), int(rng.integers(1, 30))), float(rng.uniform(0, 180)), 0, 360,
                                    200, int(rng.choice([-1, 1, 2])))
                elif kind == 2:
                    for _ in range(rng.integers(1, 4)):
                        ctr = (int(rng.integers(0, cols)), int(rng.integers(0, rows)))
                        for r in range(int(rng.integers(4, 40)), 2, -int(rng.integers(3, 7))):
                            cv2.circle(img, ctr, r, 200, 1)
                    img |= ((rng.random((rows, cols)) < 0.02) * 200).astype(np.uint8)
                else:
                    g = cv2.GaussianBlur(rng.random((rows, cols)).astype(np.float32), (0, 0), 3)
                    img = ((g > np.quantile(g, 0.6)) * 200).astype(np.uint8)
                frames[f] = img
            ctx.event_upload(0, 0, frames)
            res = ctx.localize_batch([(0, 0, 1, 0, 3, 10)])[0]
            for k in range(1, 11):
                _check_blobs(o, res[k], o.post_trigger_binary(frames[1 + k], zero, zero))


def test_end_to_end_text_matches_oracle(capi, small_run, small_trained):
    """Training + trigger + localizer on the device, host bookkeeping as the reference, output
    text byte-identical to the oracle's (BASELINE.json: 'bit-exact abub3 recon output')."""
    import oracle.cv2_oracle as o
    from tests import glue
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    ref_text, _ = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks)
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        tss = []
        for c in range(4):
            ctx.train_begin(c)
            for ev in range(len(run.events)):
                ctx.train_add_event(c, run.events[ev][c][0], run.events[ev][c][1])
            _, _, n, st = ctx.train_finish(c)
            assert st == 0
            tss.append(n)
            ctx.set_background_from_training(c)
        text = []
        for evn, ev in zip(run.event_numbers, run.events):
            for c in range(4):
                ctx.event_upload(0, c, ev[c])
            cams = [glue.gpu_any_cam_analysis(ctx, 0, c, len(ev[c]), tss[c], run.cam_masks[c], run.bellows_masks[c], evn)
                    for c in range(4)]
            text.append(o.write_camera_output(run.run_id, cams, run.frame_offset))
    assert "".join(text) == ref_text


@pytest.mark.parametrize("shape", [(40, 16), (67, 248), (130, 496), (33, 1696)])
def test_post_trigger_binary_stream_kernel(capi, shape):
    """cols % 8 == 0 takes the streaming K5b kernel (and cols % 16 == 0 the vectorised candidate scan):
    binarised post-trigger frames and blob records against cv2, dense and sparse content, image borders."""
    import oracle.cv2_oracle as o
    rows, cols = shape
    rng = np.random.default_rng(rows + cols)
    with capi.Context(rows, cols, 1, 12, max_slots=1) as ctx:
        for trial in range(3):
            avg = rng.integers(0, 256, (rows, cols)).astype(np.uint8)
            sigma = rng.integers(0, 2 + 20 * trial, (rows, cols)).astype(np.uint8)
            ctx.set_background(0, avg, sigma, 12)
            noise = rng.normal(0, 3 + 25 * trial, (12, rows, cols))
            frames = np.clip(avg[None].astype(np.float64) + noise, 0, 255).astype(np.uint8)
            frames[3:, : min(rows, 9), :7] = 255 - avg[: min(rows, 9), :7]          # blobs touching the corners
            frames[5:, -6:, -9:] = 255 - avg[-6:, -9:]
            ctx.event_upload(0, 0, frames)
            res = ctx.localize_batch([(0, 0, 1, 0, 3, 10)], cap=1 << 17)[0]
            for k in range(1, 11):
                ref = o.post_trigger_binary(frames[1 + k], avg, sigma)
                _check_blobs(o, res[k], ref)
            for k in (1, 4, 10):
                assert np.array_equal(ctx.localize_binary(k), o.post_trigger_binary(frames[1 + k], avg, sigma)), (shape, trial, k)


# ---- K4 fast path (16 px per thread, reciprocal-and-correct division) and batched K3 gate ----------------
def test_fdiv_shortcut_is_ieee_division_exhaustively(capi):
    """+0 and every float a with 2^-50 <= |a| <= 256 (the values sample - mean can take) against every frame count the
    one-step kernel may see; the two-step form (larger counts) on three windows of divisors up to 2^24."""
    fmax = capi.load().abub_k4_fdiv1_max()
    with capi.Context(32, 32, 1, 4, max_slots=1) as ctx:
        for k0 in range(1, fmax + 1, 512):
            bad, ex = ctx.selftest_fdiv(1, k0, min(k0 + 511, fmax))
            assert bad == 0, f"one-step quotient differs from IEEE division, e.g. a bits {ex[0]:#x} / {ex[1]}"
        for k0 in (fmax + 1, 1_000_000, (1 << 24) - 64):
            bad, ex = ctx.selftest_fdiv(2, k0, k0 + 63)
            assert bad == 0, f"two-step quotient differs from IEEE division, e.g. a bits {ex[0]:#x} / {ex[1]}"


@pytest.mark.parametrize("shape,n", [((48, 64), 2), ((48, 64), 6), ((48, 64), 36), ((64, 80), 300), ((50, 68), 36)])
def test_welford_fast_path_vs_oracle(capi, shape, n):
    """Noisy frames (every sample differs) over enough frames to exercise many divisors; (50,68) is the generic kernel."""
    import oracle.cv2_oracle as o
    rows, cols = shape
    rng = np.random.default_rng(1000 + n)
    base = rng.integers(0, 256, (rows, cols))
    frames = np.clip(base[None] + rng.normal(0, 9, (n, rows, cols)).round(), 0, 255).astype(np.uint8)
    pairs = np.stack([frames[0::2], frames[0::2]], axis=1)            # identical pairs pass the gate
    kept = [f for p in pairs for f in p]
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        ctx.train_begin(0)
        e, k = ctx.train_add_events(0, pairs)
        assert k.all() and np.abs(e).max() <= 1e-6
        mean, sigma, cnt, st = ctx.train_finish(0)
    m_ref, s_ref = o.welford_mean_sigma(kept)
    assert st == 0 and cnt == len(kept)
    assert np.array_equal(mean, m_ref) and np.array_equal(sigma, s_ref)


def test_welford_fast_equals_generic_kernel(capi, monkeypatch):
    """Same frames through k4_welford16_kernel and the __fdiv_rn kernel (ABUB_K4_GENERIC): identical planes."""
    rows, cols, n = 96, 128, 500
    rng = np.random.default_rng(77)
    frames = rng.integers(0, 256, (n // 2, 1, rows, cols), dtype=np.uint8)
    pairs = np.concatenate([frames, frames], axis=1)
    out = []
    for generic in (False, True):
        if generic:
            monkeypatch.setenv("ABUB_K4_GENERIC", "1")
        else:
            monkeypatch.delenv("ABUB_K4_GENERIC", raising=False)
        with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
            ctx.train_begin(0)
            ctx.train_add_events(0, pairs)
            out.append(ctx.train_finish(0))
    assert out[0][2] == out[1][2] == n
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])


def test_train_batched_gate_equals_single_calls(capi, small_run, small_trained):
    """abub_train_add_events / abub_train_add_slots vs one abub_train_add_event per event and vs the oracle."""
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    with capi.Context(rows, cols, 4, 40, max_slots=nev) as ctx:
        for s in range(nev):
            for c in range(4):
                ctx.event_upload(s, c, np.stack(run.events[s][c]))
        for c in range(4):
            ctx.train_begin(c)
            single = [ctx.train_add_event(c, run.events[ev][c][0], run.events[ev][c][1]) for ev in range(nev)]
            ref = ctx.train_finish(c)
            ctx.train_begin(c)
            pairs = np.stack([np.stack(run.events[ev][c][:2]) for ev in range(nev)])
            e1, k1 = ctx.train_add_events(c, pairs)
            got1 = ctx.train_finish(c)
            ctx.train_begin(c)
            # two batches, to cross a batch boundary
            e2a, k2a = ctx.train_add_slots(c, list(range(0, nev // 2)))
            e2b, k2b = ctx.train_add_slots(c, list(range(nev // 2, nev)))
            got2 = ctx.train_finish(c)
            e2, k2 = np.concatenate([e2a, e2b]), np.concatenate([k2a, k2b])
            for ev in range(nev):
                assert single[ev][0] == e1[ev] == e2[ev]
                assert single[ev][1] == bool(k1[ev]) == bool(k2[ev]) == (ev in trained[c].kept_events)
            for got in (got1, got2):
                assert got[2:] == ref[2:] == (trained[c].training_set_size, trained[c].status)
                assert np.array_equal(got[0], trained[c].avg) and np.array_equal(got[1], trained[c].sigma)


def test_train_add_events_with_rejected_pairs(capi):
    """Rejected pairs in the middle of a batch leave no trace in the statistics (compaction of the parked frames)."""
    import oracle.cv2_oracle as o
    rows, cols = 48, 64
    rng = np.random.default_rng(5)
    quiet = [np.clip(120 + rng.normal(0, 2, (rows, cols)).round(), 0, 255).astype(np.uint8) for _ in range(5)]
    loud = rng.integers(0, 256, (rows, cols), dtype=np.uint8)
    dark = rng.integers(0, 30, (rows, cols), dtype=np.uint8)
    pairs = np.stack([np.stack([quiet[0], quiet[0]]), np.stack([dark, loud]), np.stack([quiet[1], quiet[1]]),
                      np.stack([dark, loud]), np.stack([dark, loud]), np.stack([quiet[2], quiet[2]])])
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        ctx.train_begin(0)
        e, k = ctx.train_add_events(0, pairs)
        assert k.tolist() == [True, False, True, False, False, True]
        e2, k2 = ctx.train_add_events(0, pairs[1:3])
        assert k2.tolist() == [False, True]
        mean, sigma, cnt, st = ctx.train_finish(0)
    kept = [quiet[0]] * 2 + [quiet[1]] * 2 + [quiet[2]] * 2 + [quiet[1]] * 2
    m_ref, s_ref = o.welford_mean_sigma(kept)
    assert (cnt, st) == (8, 0)
    assert np.array_equal(mean, m_ref) and np.array_equal(sigma, s_ref)


def _adversarial_stack(kind, n, rows, cols, rng):
    """n/2 distinct frames (each is pushed twice: an identical pair always passes the gate, Trainer.cpp:287)."""
    h = n // 2
    base = rng.integers(0, 250, (rows, cols))
    if kind == "noise":
        fr = base[None] + (rng.normal(0, 1, (h, rows, cols)) * rng.uniform(1, 6, (rows, cols))[None]).round()
    elif kind == "spikes":              # sparse 255 spikes on a dark floor: sigma >> mean, the case the old threshold missed
        fr = np.where(rng.random((h, rows, cols)) < 1.0 / 16, 255, (base % 8)[None] + rng.integers(0, 2, (h, rows, cols)))
    elif kind == "two_level":           # exact .5 / integer means, sigma on exact values
        lv = rng.integers(0, 128, (2, rows, cols))
        fr = np.where((np.arange(h) % 2 == 0)[:, None, None], lv[0][None], (lv[0] + 2 * lv[1])[None])
    elif kind == "drift":               # slow ramp: the recurrence's mean lags, order matters most
        fr = base[None] * 0.5 + np.linspace(0, 120, h)[:, None, None] + rng.normal(0, 1.5, (h, rows, cols))
    elif kind == "near_integer":        # mean = v + 2 j / n, -6 <= j <= 6: just above / just below an integer
        fr = np.repeat(base[None] + 1, h, 0)
        j = rng.integers(-6, 7, (rows, cols))
        for t in range(6):
            fr[t] += (j > t).astype(fr.dtype) - (-j > t).astype(fr.dtype)
        fr = fr[rng.permutation(h)]
    else:
        raise ValueError(kind)
    return np.clip(fr, 0, 255).astype(np.uint8)


@pytest.mark.parametrize("n,kind", [(40, "noise"), (40, "spikes"), (400, "two_level"), (2000, "near_integer"), (2000, "noise"),
                                    (2000, "spikes"), (2000, "drift")])
def test_partials_flag_bound_covers_every_float32_difference(capi, monkeypatch, n, kind):
    """Multi-GPU Trainer: planes from the exact integer sums (fp64) are truncated like the float32 recurrence's except
    where rounding decides; those pixels must ALL be flagged (k4_finish_partials_kernel's error bound), because only
    flagged pixels are repaired.  Single GPU, adversarial data: (1) K4 planes; (2) integer partials WITHOUT the repair
    (test hook ABUB_TRAIN_NO_FIXUP): the pixels that differ from K4 are counted -- the data must produce some, or the
    test has no teeth; (3) integer partials with the repair == K4 everywhere, i.e. every differing pixel was flagged."""
    rows, cols = 48, 64
    rng = np.random.default_rng(n * 7 + len(kind))
    fr = _adversarial_stack(kind, n, rows, cols, rng)
    pairs = np.stack([fr, fr], axis=1)                            # [n/2, 2, rows, cols]
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        def feed():
            ctx.train_begin(0)
            for i in range(0, len(pairs), 250):
                _, k = ctx.train_add_events(0, pairs[i:i + 250])
                assert k.all()
        feed()
        m4, s4, n4, st4 = ctx.train_finish(0)
        assert (n4, st4) == (n, 0)
        monkeypatch.setenv("ABUB_TRAIN_NO_FIXUP", "1")
        feed()
        ctx.train_partials(0)
        m0, s0, n0, st0, nflag = ctx.train_finish_partials(0)
        monkeypatch.delenv("ABUB_TRAIN_NO_FIXUP")
        feed()
        ctx.train_partials(0)
        m1, s1, n1, st1, nflag1 = ctx.train_finish_partials(0)
    assert (n0, st0, n1, st1) == (n, 0, n, 0) and nflag1 == nflag
    assert np.array_equal(m1, m4) and np.array_equal(s1, s4), (kind, n)
    ndiff = int(((m0 != m4) | (s0 != s4)).sum())
    assert ndiff <= nflag
    print(f"\n[flag bound] n={n} {kind}: {ndiff} pixels differ before the repair, {nflag} of {rows * cols} flagged")
    if kind in ("two_level", "near_integer"):
        assert ndiff > 0 or nflag > 0                              # means / sigmas sit on the boundaries by construction
    if kind == "noise":
        # generic data: exactly-integer means (a share ~ 1/n of the pixels at small n) plus the bound's width,
        # 2 e_mean + e_up + e_dn ~ 0.5e-4 n at sigma ~ 3
        assert nflag <= (0.06 + 1e-4 * n) * rows * cols, (nflag, n)
