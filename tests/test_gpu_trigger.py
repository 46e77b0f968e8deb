"""GPU parity: K1 (ProcessFrame + 256-bin histogram) and K2 (significance trigger) through
the C ABI against the cv2 oracle.  Bit-exact for histograms / trigger index / status /
loc_thres; significance doubles compared exactly too (same IEEE operations, no FMA)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def capi():
    from autobub3hs_b200 import capi
    capi.load()
    return capi


def _oracle_hists(o, frames, sigma, tss):
    n = len(frames)
    H = np.zeros((n, 256), np.uint32)
    d = 2 if tss >= 6 else 1
    for j in range(1, n):
        jp = max(j - d, 0)
        H[j] = o.hist256(o.process_frame(frames[j], frames[jp], sigma)).astype(np.uint32)
    return H


@pytest.mark.parametrize("shape", [(264, 420), (77, 131), (40, 244), (35, 120), (5, 5), (36, 121)])
def test_process_frame_single(capi, shape):
    import oracle.cv2_oracle as o
    rows, cols = shape
    rng = np.random.default_rng(rows * 1000 + cols)
    with capi.Context(rows, cols, ncams=1, max_frames=4, max_slots=1) as ctx:
        for trial in range(4):
            base = rng.integers(0, 256, (rows, cols)).astype(np.float32)
            a = np.clip(base + rng.normal(0, 4 + 20 * trial, (rows, cols)), 0, 255).astype(np.uint8)
            b = np.clip(base + rng.normal(0, 4 + 20 * trial, (rows, cols)), 0, 255).astype(np.uint8)
            sigma = rng.integers(0, 3 + trial * 20, (rows, cols)).astype(np.uint8)
            mean = np.zeros_like(sigma)
            ctx.set_background(0, mean, sigma, 12)
            diff, hist = ctx.process_frame(0, a, b)
            ref = o.process_frame(a, b, sigma)
            assert np.array_equal(diff, ref)
            assert np.array_equal(hist, o.hist256(ref).astype(np.uint32))
            assert int(hist.sum()) == rows * cols


def test_process_frame_extremes(capi):
    import oracle.cv2_oracle as o
    rows, cols = 64, 248
    with capi.Context(rows, cols, ncams=1, max_frames=4, max_slots=1) as ctx:
        cases = [(255, 0, 0), (0, 255, 0), (255, 0, 42), (255, 0, 255), (7, 7, 0)]
        for av, bv, sv in cases:
            a = np.full((rows, cols), av, np.uint8)
            b = np.full((rows, cols), bv, np.uint8)
            s = np.full((rows, cols), sv, np.uint8)
            ctx.set_background(0, s, s, 12)
            diff, hist = ctx.process_frame(0, a, b)
            ref = o.process_frame(a, b, s)
            assert np.array_equal(diff, ref), (av, bv, sv)
            assert np.array_equal(hist, o.hist256(ref).astype(np.uint32))
        # checkerboard of saturated values: exercises both planes and every rounding carry
        yy, xx = np.mgrid[0:rows, 0:cols]
        a = (((yy + xx) & 1) * 255).astype(np.uint8)
        b = (255 - a).astype(np.uint8)
        s = np.zeros((rows, cols), np.uint8)
        ctx.set_background(0, s, s, 12)
        diff, hist = ctx.process_frame(0, a, b)
        assert np.array_equal(diff, o.process_frame(a, b, s))


def test_trigger_batch_vs_oracle(capi, small_run, small_trained):
    import oracle.cv2_oracle as o
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=nev) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
        for ev in range(nev):
            for c in range(4):
                ctx.event_upload(ev, c, run.events[ev][c])
        ctx.trigger_hist(0, nev)
        ctx.trigger_scan(0, nev)
        trig, status, lt, sig = ctx.trigger_results(0, nev)
        for ev in range(nev):
            for c in range(4):
                fr = run.events[ev][c]
                H = ctx.trigger_hist_read(ev, c)
                Href = _oracle_hists(o, fr, trained[c].sigma, trained[c].training_set_size)
                assert np.array_equal(H[:len(fr)], Href), (ev, c)
                st = o.SignificanceState()
                r = o.find_trigger_frame(o.FrameSource(fr), trained[c].sigma, trained[c].training_set_size, 1, st, 0)
                assert status[ev, c] == r.status, (ev, c, run.specs[ev].klass)
                if r.status == 0:
                    assert trig[ev, c] == r.trig
                if r.status != -9:
                    assert lt[ev, c] == r.loc_thres
                for (i, s) in r.sig:
                    g = sig[ev, c, i]
                    assert (np.isnan(s) and np.isnan(g)) or s == g, (ev, c, i, s, g)


def test_trigger_retrigger_and_small_training_set(capi, small_run, small_trained):
    """start_frame > 1 resumes with history intact (AnalyzerUnit.cpp:165); TrainingSetSize < 6
    switches to the one-frame offset and the 5.0 threshold (:182-188)."""
    import oracle.cv2_oracle as o
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=1) as ctx:
        for tss in (12, 4):
            for c in range(4):
                ctx.set_background(c, trained[c].avg, trained[c].sigma, tss)
            for ev in (0, 1, 7):
                for c in range(4):
                    ctx.event_upload(0, c, run.events[ev][c])
                for c in range(4):
                    fr = run.events[ev][c]
                    st = o.SignificanceState()
                    hp = o.HistProvider(o.FrameSource(fr), trained[c].sigma)
                    mat = 0
                    for attempt in range(4):
                        r = o.find_trigger_frame(o.FrameSource(fr), trained[c].sigma, tss, mat + 1, st, mat, hp)
                        t, s, l, sig = ctx.trigger(0, c, mat + 1)
                        assert s == r.status, (tss, ev, c, attempt)
                        if r.status != 0:
                            break
                        assert t == r.trig and l == r.loc_thres
                        for (i, sv) in r.sig:
                            assert sig[i] == sv or (np.isnan(sv) and np.isnan(sig[i]))
                        mat = r.trig


def test_significance_corner_cases(capi):
    """Hand-built frames that drive bins >= 2 above 46340 counts (int32 wrap of val*val,
    AnalyzerUnit.cpp:529), zero-variance histories and NaN/inf propagation."""
    import oracle.cv2_oracle as o
    rows, cols = 300, 400      # 120000 px > 46340
    n = 12
    rng = np.random.default_rng(5)
    frames = np.zeros((n, rows, cols), np.uint8)
    for f in range(n):
        frames[f] = 100 + (f % 2) * 0
    # big uniform step at frame 6 and 7 -> whole image in one non-zero bin
    frames[6:] += 40
    frames[9:] += 3
    frames[10, :50] += 90
    sigma = np.zeros((rows, cols), np.uint8)
    with capi.Context(rows, cols, ncams=1, max_frames=n, max_slots=1) as ctx:
        for tss in (12, 2):
            ctx.set_background(0, sigma, sigma, tss)
            ctx.event_upload(0, 0, frames)
            t, s, l, sig = ctx.trigger(0, 0, 1)
            st = o.SignificanceState()
            r = o.find_trigger_frame(o.FrameSource(frames), sigma, tss, 1, st, 0)
            assert s == r.status
            if r.status == 0:
                assert t == r.trig and l == r.loc_thres
            for (i, sv) in r.sig:
                assert sig[i] == sv or (np.isnan(sv) and np.isnan(sig[i])), (tss, i, sv, sig[i])


@pytest.mark.parametrize("shape", [(64, 16), (37, 256), (151, 240), (300, 496), (163, 1696), (9, 32)])
def test_trigger_hist_tma_path(capi, shape):
    """cols % 16 == 0 takes the TMA-fed trigger kernel (k1_trigger_tma.cu): strips that start
    and end inside the reflected border rows, partially filled last warp, image-edge lanes,
    ragged frame counts, saturated sigma, both frame offsets."""
    import oracle.cv2_oracle as o
    rows, cols = shape
    rng = np.random.default_rng(rows * 7 + cols)
    ncams, nfr = 2, 7
    with capi.Context(rows, cols, ncams=ncams, max_frames=nfr, max_slots=2) as ctx:
        for tss in (12, 3):
            sig = []
            for c in range(ncams):
                s = rng.integers(0, 3 + 45 * c, (rows, cols)).astype(np.uint8)
                sig.append(s)
                ctx.set_background(c, np.zeros_like(s), s, tss)
            stacks = {}
            for slot in range(2):
                for c in range(ncams):
                    n = nfr - (slot + c) % 3
                    base = rng.integers(0, 256, (rows, cols)).astype(np.float32)
                    fr = np.clip(base[None] + rng.normal(0, 3 + 25 * slot, (n, rows, cols)), 0, 255).astype(np.uint8)
                    fr[n // 2:, rows // 3: rows // 3 + 6, : min(cols, 9)] = 255      # blob on the left border
                    fr[n // 2:, -3:, -5:] = 0                                          # and in the bottom-right corner
                    stacks[(slot, c)] = fr
                    ctx.event_upload(slot, c, fr)
            ctx.trigger_hist(0, 2)
            for (slot, c), fr in stacks.items():
                H = ctx.trigger_hist_read(slot, c)
                Href = _oracle_hists(o, fr, sig[c], tss)
                assert np.array_equal(H[:len(fr)], Href), (shape, tss, slot, c)


@pytest.mark.parametrize("chunk", [1, 4, 7, 40])
def test_trigger_stream_equals_batch(capi, small_run, small_trained, chunk):
    """abub_trigger_stream (frames sent chunk by chunk from pinned memory, nothing sent to a camera after its
    search is decided) returns exactly what the resident batch path returns."""
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=nev) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
        nf = np.zeros((nev, 4), np.int32)
        for ev in range(nev):
            pin = ctx.pinned(ev)
            for c in range(4):
                fr = run.events[ev][c]
                nf[ev, c] = len(fr)
                pin[c, :len(fr)] = fr
                ctx.event_upload(ev, c, fr)
        ctx.trigger_hist(0, nev)
        ctx.trigger_scan(0, nev)
        trig, status, lt, _ = ctx.trigger_results(0, nev)
        t2, s2, l2, sent = ctx.trigger_stream(0, nev, nf, chunk)
        assert np.array_equal(s2, status)
        ok = status == 0
        assert np.array_equal(t2[ok], trig[ok])
        live = status != -9
        assert np.array_equal(l2[live], lt[live])
        assert np.all(sent <= nf) and np.all(sent[status == -9] == 0) and np.all(sent[status == -3] == nf[status == -3])
        # a triggered camera needs frames up to trig+2 and gets at most one chunk beyond that
        need = np.minimum(trig + 3, nf)
        assert np.all(sent[ok] >= need[ok]) and np.all(sent[ok] < need[ok] + chunk)
        # the result table is left as abub_trigger_results expects it
        t3, s3, l3, _ = ctx.trigger_results(0, nev, want_sig=False)
        assert np.array_equal(s3, status) and np.array_equal(t3[ok], trig[ok])


def test_set_nframes_range_equals_per_slot_calls(capi, small_run, small_trained):
    """abub_event_set_nframes_range (one launch for a range of slots) leaves the device with the same frame counts as one
    abub_event_set_nframes per slot: frames are placed through the device pointer, the counts -- shortened for some
    cameras, which changes status and trigger frame -- go in both ways, and the trigger results must agree."""
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    rng = np.random.default_rng(3)
    counts = np.array([[len(run.events[ev][c]) for c in range(4)] for ev in range(nev)], np.int32)
    cut = counts.copy()
    for ev in range(nev):
        c = int(rng.integers(0, 4))
        cut[ev, c] = max(0, int(cut[ev, c]) - int(rng.integers(1, 30)))
    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=nev) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
        for ev in range(nev):
            for c in range(4):
                ctx.event_upload(ev, c, run.events[ev][c])
        res = []
        for form in ("per_slot", "range", "range_tail"):
            if form == "per_slot":
                for ev in range(nev):
                    ctx.event_set_nframes(ev, cut[ev])
            elif form == "range":
                ctx.event_set_nframes_range(0, cut)
            else:                                            # a range that starts in the middle, after resetting everything
                ctx.event_set_nframes_range(0, counts)
                ctx.event_set_nframes_range(0, cut[:2])
                ctx.event_set_nframes_range(2, cut[2:])
            ctx.trigger_hist(0, nev)
            ctx.trigger_scan(0, nev)
            res.append(ctx.trigger_results(0, nev, want_sig=False)[:3])
        for other in res[1:]:
            for a, b in zip(res[0], other):
                assert np.array_equal(a, b)
        ctx.event_set_nframes_range(0, counts)
        ctx.trigger_hist(0, nev)
        ctx.trigger_scan(0, nev)
        full = ctx.trigger_results(0, nev, want_sig=False)[:3]
        assert not all(np.array_equal(a, b) for a, b in zip(res[0], full))      # the shortened counts did matter
        with pytest.raises(capi.AbubError):
            ctx.event_set_nframes_range(nev - 1, cut[:2])                       # runs past the last slot


@pytest.mark.parametrize("start_now", [False, True])
def test_trigger_early_deferred_equals_direct(capi, small_run, small_trained, start_now):
    """abub_trigger_early_deferred: the search of a SECOND slot range is registered, launched by the localizer call of the
    first range between its own stages (optionally its first round at once, abub_trigger_deferred_start), and gives what
    abub_trigger_early gives when called directly; so does a registration that is only flushed (no localizer call), and
    the localizer's own results are not disturbed by the launches in its middle."""
    from tests import glue
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    half = nev // 2

    def same(got, want, lo, hi):
        t, s, l, sg = got
        tw, sw, lw, sgw = (a[lo:hi] for a in want)
        assert np.array_equal(s, sw)
        ok = sw == 0
        assert np.array_equal(t[ok], tw[ok])
        live = sw != -9
        assert np.array_equal(l[live], lw[live])
        for e in range(hi - lo):
            for c in range(4):
                upto = tw[e, c] + 1 if sw[e, c] == 0 else (len(run.events[lo + e][c]) if sw[e, c] == -3 else 0)
                assert np.array_equal(sg[e, c, :upto], sgw[e, c, :upto], equal_nan=True), (e, c)

    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=nev) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
        for ev in range(nev):
            for c in range(4):
                ctx.event_upload(ev, c, run.events[ev][c])
        ctx.trigger_early(0, nev, 5)
        want = ctx.trigger_results(0, nev)
        tw, sw, lw, _ = want
        items = [(ev, c, int(tw[ev, c]), max(0, int(tw[ev, c]) - 2), int(lw[ev, c]), glue.n_post_frames(int(tw[ev, c]), len(run.events[ev][c])))
                 for ev in range(half) for c in range(4) if sw[ev, c] == 0]
        assert items
        ref_blobs = ctx.localize_batch(items)
        # first range searched directly, second range registered and launched inside the first range's localizer call
        ctx.trigger_early(0, half, 5)
        same(ctx.trigger_results(0, half), want, 0, half)
        ctx.trigger_early_deferred(half, nev - half, 5)
        if start_now:
            ctx.trigger_deferred_start()
        blobs = ctx.localize_batch(items)
        ctx.trigger_deferred_flush()                                   # nothing left to launch
        same(ctx.trigger_results(half, nev - half), want, half, nev)
        assert [[[(b.x, b.y, b.w, b.h, b.npts, b.area) for b in fr] for fr in it] for it in blobs] == \
               [[[(b.x, b.y, b.w, b.h, b.npts, b.area) for b in fr] for fr in it] for it in ref_blobs]
        # a registration that never meets a localizer call: the flush launches all of it
        ctx.trigger_early_deferred(0, nev, 6)
        if start_now:
            ctx.trigger_deferred_start()
        ctx.trigger_deferred_flush()
        same(ctx.trigger_results(0, nev), want, 0, nev)


@pytest.mark.parametrize("chunk", [1, 4, 6, 40])
def test_trigger_early_equals_batch(capi, small_run, small_trained, chunk):
    """abub_trigger_early (resident frames, K1 in rounds of `chunk` frame pairs, decided cameras drop out of the later
    rounds: the reference's return at the trigger, AnalyzerUnit.cpp:297-300) returns exactly what the full-table path
    returns, leaves the significance of every evaluated frame, and a later abub_trigger rebuilds the table."""
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    nev = len(run.events)
    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=nev) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
        for ev in range(nev):
            for c in range(4):
                ctx.event_upload(ev, c, run.events[ev][c])
        ctx.trigger_hist(0, nev)
        ctx.trigger_scan(0, nev)
        trig, status, lt, sig = ctx.trigger_results(0, nev)
        ctx.trigger_early(0, nev, chunk)
        t2, s2, l2, sig2 = ctx.trigger_results(0, nev)
        assert np.array_equal(s2, status)
        ok = status == 0
        assert ok.any() and (status == -3).any() and (status == -9).any()
        assert np.array_equal(t2[ok], trig[ok])
        live = status != -9
        assert np.array_equal(l2[live], lt[live])
        for ev in range(nev):
            for c in range(4):
                upto = trig[ev, c] + 1 if status[ev, c] == 0 else (len(run.events[ev][c]) if status[ev, c] == -3 else 0)
                assert np.array_equal(sig2[ev, c, :upto], sig[ev, c, :upto], equal_nan=True), (ev, c)
        # the partial tables are not mistaken for full ones: a re-trigger after the first trigger sees every frame
        ev, c = [int(v[0]) for v in np.nonzero(ok)]
        again = ctx.trigger(ev, c, int(trig[ev, c]) + 1)
        ctx.trigger_hist(0, nev)
        st = np.ones((nev, 4), np.int32)
        st[ev, c] = trig[ev, c] + 1
        ctx.trigger_scan(0, nev, st)
        t3, s3, l3, _ = ctx.trigger_results(0, nev, want_sig=False)
        assert again[1] == s3[ev, c] and (again[1] != 0 or again[0] == t3[ev, c])


@pytest.mark.parametrize("metric", ["entropy", "entropy_significance"])
def test_trigger_entropy_variants(capi, small_run, small_trained, metric):
    """the two trigger quantities the reference keeps as commented alternatives (AnalyzerUnit.cpp:226-227,
    :290-291 -> calculateEntropyFrame :386-423, calculateEntropySignificance :425-433), selected with
    abub_set_trigger_metric on the same K1 table.  Status, trigger frame and loc_thres (3, never updated on these paths)
    must equal the oracle's, including re-triggers from a later start frame; the float entropy may differ from the
    host's log2f by an ulp per bin: 1e-6 absolute on the entropy, 1e-3 relative on the z-score built from it."""
    import oracle.cv2_oracle as o
    run, trained = small_run, small_trained
    rows, cols = run.cam_masks[0].shape
    exact = total = 0
    with capi.Context(rows, cols, ncams=4, max_frames=40, max_slots=1) as ctx:
        for thr in (None, 0.02 if metric == "entropy" else 3.0):
            ctx.set_trigger_metric(metric, 0.0 if thr is None else thr)
            for tss in (12, 4):
                for c in range(4):
                    ctx.set_background(c, trained[c].avg, trained[c].sigma, tss)
                for ev in range(len(run.events)):
                    for c in range(4):
                        ctx.event_upload(0, c, run.events[ev][c])
                    for c in range(4):
                        fr = run.events[ev][c]
                        st = o.SignificanceState()
                        hp = o.HistProvider(o.FrameSource(fr), trained[c].sigma)
                        mat = 0
                        for attempt in range(3):
                            r = o.find_trigger_frame(o.FrameSource(fr), trained[c].sigma, tss, mat + 1, st, mat, hp,
                                                     metric=metric, threshold=thr)
                            t, s, l, sig = ctx.trigger(0, c, mat + 1)
                            assert s == r.status, (metric, thr, tss, ev, c, attempt)
                            for (i, sv) in r.sig:
                                g = sig[i]
                                total += 1
                                exact += (sv == g) or (np.isnan(sv) and np.isnan(g))
                                if np.isnan(sv) or np.isinf(sv):
                                    assert np.isnan(g) or g == sv or (np.isinf(g) and np.isinf(sv)), (i, sv, g)
                                elif metric == "entropy":
                                    assert abs(g - sv) <= 1e-6, (i, sv, g)
                                else:
                                    assert abs(g - sv) <= 1e-3 * max(1.0, abs(sv)), (i, sv, g)
                            if r.status != -9:
                                assert l == 3
                            if r.status != 0:
                                break
                            assert t == r.trig
                            mat = r.trig
        ctx.set_trigger_metric("significance")
    # an entropy that differs in its last bit shifts every later z-score of that camera a little: only the plain entropy is
    # expected to be bit-equal most of the time
    # Every value is held to the tolerance above (1e-6 on the float32 entropy); on top of that most of them must be bit-equal.
    # The ones that are not differ in the last float32 bit only: the entropy is a float sum of up to 128 terms p*log2(p), and
    # the device's log2f (<= 1 ulp) and glibc's log2f (correctly rounded in nearly all cases) disagree on a few percent of
    # the arguments.  The decisions (status, trigger frame) are compared exactly either way.
    assert total > 0 and (metric != "entropy" or exact >= 0.9 * total), (exact, total)
