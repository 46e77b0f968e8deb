"""GPU parity of the bellows veto (L3Localizer.cpp:292-390, :462-543) through the C ABI.

TrackAFeature is floating point: the device forms the correlation numerators as exact integers, OpenCV with a
float32 DFT.  Tolerances (stated here, used below): CCORR_NORMED map 2e-5 absolute; sub-pixel position 2e-3 px;
the integer peak and the truncated positions must be equal whenever the oracle's position is at least 5e-4 px away
from an integer.  Everything downstream of the two positions is integer arithmetic and compared exactly."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAP_TOL = 2e-5
POS_TOL = 2e-3
INT_MARGIN = 5e-4


@pytest.fixture(scope="module")
def capi():
    from autobub3hs_b200 import capi
    capi.load()
    return capi


@pytest.fixture(scope="module")
def bellows_run():
    from autobub3hs_b200 import synth
    return synth.make_run(6, ["normal", "bellows", "bellows_bubble", "notrigger", "bellows", "normal"], rows=264, cols=420,
                          seed=5, bellows_object=True)


@pytest.fixture(scope="module")
def bellows_trained(bellows_run):
    import oracle.cv2_oracle as o
    return o.train_all(bellows_run.events, 4)


def _safe(v):
    return abs(float(v) - round(float(v))) >= INT_MARGIN


@pytest.mark.parametrize("shape,tshape", [((90, 131), (37, 23)), ((64, 96), (16, 16)), ((120, 333), (50, 141)),
                                          ((72, 200), (69, 5)), ((48, 64), (3, 61))])
def test_ccorr_normed_map_random(capi, shape, tshape):
    """Random frame with the template pasted in: map vs cv2.matchTemplate, peak, sub-pixel position."""
    import oracle.cv2_oracle as o
    rng = np.random.default_rng(shape[1] * 7 + tshape[1])
    rows, cols = shape
    th, tw = tshape
    templ = rng.integers(0, 256, (th, tw), dtype=np.uint8)
    frame = rng.integers(0, 120, (rows, cols), dtype=np.uint8)
    py, px = (rows - th) // 2 + 1, (cols - tw) // 3 + 1
    py, px = max(1, min(py, rows - th - 1)), max(1, min(px, cols - tw - 1))
    frame[py:py + th, px:px + tw] = templ
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        ctx.set_bellows_template(0, templ)
        ctx.event_upload(0, 0, frame[None])
        (x, y), peak = ctx.bellows_track(0, 0, 0)
        got = ctx.bellows_map()
    want = o.ccorr_normed_map(frame, templ)
    assert got.shape == want.shape
    assert np.abs(got - want).max() <= MAP_TOL
    ox, oy, opeak = o.subpixel_from_map(want)
    assert peak == opeak == (px, py)
    assert abs(float(x) - float(ox)) <= POS_TOL and abs(float(y) - float(oy)) <= POS_TOL


def test_track_positions_on_bellows_frames(capi, bellows_run):
    import oracle.cv2_oracle as o
    run = bellows_run
    rows, cols = run.cam_masks[0].shape
    checked = 0
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        for cam in (1, 3):
            ctx.set_bellows_template(cam, run.bellows_templates[cam])
            for ev in (1, 2):
                ctx.event_upload(0, cam, run.events[ev][cam])
                mv = run.specs[ev].bellows_move[cam][0]
                for f in (0, mv - 1, mv, mv + 1, mv + 3):
                    (x, y), peak = ctx.bellows_track(0, cam, f)
                    ox, oy, opeak = o.subpixel_from_map(o.ccorr_normed_map(run.events[ev][cam][f], run.bellows_templates[cam]))
                    assert peak == opeak
                    assert abs(float(x) - float(ox)) <= POS_TOL and abs(float(y) - float(oy)) <= POS_TOL
                    if _safe(ox) and _safe(oy):
                        assert (int(x), int(y)) == (int(ox), int(oy))
                        checked += 1
    assert checked >= 10


def test_veto_given_positions_is_exact(capi, bellows_run, bellows_trained):
    """The integer part of the branch: oracle positions in, binary image / ROI / contours out, bit-exact."""
    import oracle.cv2_oracle as o
    from tests import glue
    run, trained = bellows_run, bellows_trained
    rows, cols = run.cam_masks[0].shape
    n = 0
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        for cam in (1, 3):
            ctx.set_background(cam, trained[cam].avg, trained[cam].sigma, trained[cam].training_set_size)
            ctx.set_bellows_template(cam, run.bellows_templates[cam])
            templ = run.bellows_templates[cam]
            for ev in (1, 2, 4):
                fr = run.events[ev][cam]
                ctx.event_upload(0, cam, fr)
                mv = run.specs[ev].bellows_move[cam][0]
                for trig, pre, lt in ((mv, mv - 2, 3), (mv + 1, mv - 1, 2), (mv + 2, mv, 3)):
                    pt = o.track_a_feature(fr[trig], templ)
                    pp = o.track_a_feature(fr[pre], templ)
                    over = o.genesis_over_the_sigma(fr[trig], fr[pre], trained[cam].sigma)
                    over2, (npt, npp, roi) = o.bellows_veto(over, fr[trig], fr[pre], trained[cam].sigma, templ, (pt, pp))
                    thr = o.binarise_over_the_sigma(over2, lt)
                    blobs, st, groi = ctx.localize_bellows(0, cam, trig, pre, lt, pt, pp)
                    assert st == 0 and groi == tuple(int(v) for v in roi)
                    assert np.array_equal(ctx.localize_binary(0), thr)
                    want = [glue_frame for glue_frame in (o._measure(c, True) for c in o.find_external_contours(thr))]
                    got = [glue.blob_to_frame(b, True) for b in blobs]
                    assert len(got) == len(want)
                    for g, w in zip(got, want):
                        assert (g.x, g.y, g.w, g.h) == (w.x, w.y, w.w, w.h)
                        assert g.cont_area == w.cont_area and (g.m00, g.m10, g.m01) == (w.m00, w.m10, w.m01)
                    n += 1
    assert n == 18


def test_veto_rect_outside_frame_is_minus6(capi, bellows_run, bellows_trained):
    run, trained = bellows_run, bellows_trained
    rows, cols = run.cam_masks[0].shape
    th, tw = run.bellows_templates[1].shape
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        ctx.set_background(1, trained[1].avg, trained[1].sigma, trained[1].training_set_size)
        ctx.set_bellows_template(1, run.bellows_templates[1])
        ctx.event_upload(0, 1, run.events[1][1])
        # x + 1 + tw > cols after the nudge
        _, st, _ = ctx.localize_bellows(0, 1, 10, 8, 3, (cols - tw + 0.5, 5.0), (cols - tw - 3.0, 5.0))
        assert st == -6
        _, st, _ = ctx.localize_bellows(0, 1, 10, 8, 3, (-0.2, 5.0), (3.0, 5.0))     # int(x - 1) = -1
        assert st == -6
        _, st, _ = ctx.localize_bellows(0, 1, 10, 8, 3, (20.0, 5.0), (22.0, 5.0))
        assert st == 0


def test_bellows_run_text_matches_oracle(capi, bellows_run, bellows_trained):
    """Whole detect loop on a run whose cameras 1 and 3 show a moving bellows: text identical to the oracle's."""
    import oracle.cv2_oracle as o
    from tests import glue
    run, trained = bellows_run, bellows_trained
    rows, cols = run.cam_masks[0].shape
    want, recs = o.run_detect(run.run_id, run.event_numbers, run.events, trained, run.cam_masks, run.bellows_masks, 30,
                              run.bellows_templates)
    taken = sum(1 for rr in recs for a in rr for info in a.bellows if info.get("bellows_branch"))
    assert taken >= 6
    # the comparison below is only meaningful where truncation of the oracle's positions is stable
    for rr in recs:
        for a in rr:
            for info in a.bellows:
                if "bellows_positions" in info:
                    pt, pp, _ = info["bellows_positions"]
                    assert all(_safe(v) for v in (*pt, *pp))
    text = []
    log = []
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
            if run.bellows_templates[c] is not None:
                ctx.set_bellows_template(c, run.bellows_templates[c])
        for evn, ev in zip(run.event_numbers, run.events):
            for c in range(4):
                ctx.event_upload(0, c, ev[c])
            ctx.trigger_hist(0, 1)
            cams = [glue.gpu_any_cam_analysis(ctx, 0, c, len(ev[c]), trained[c].training_set_size, run.cam_masks[c],
                                              run.bellows_masks[c], evn, run.bellows_templates[c] is not None, log)
                    for c in range(4)]
            text.append(o.write_camera_output(run.run_id, cams, 30))
    assert len(log) == taken
    assert "".join(text) == want
    import os
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bellows_run.txt")
    assert o.write_header() + "".join(text) == open(gold).read()              # the committed vector of the same run


def test_branch_without_template_keeps_all_contours(capi, bellows_run, bellows_trained):
    """Template not loadable (:297-301): the contours found before are used without the bellows filter."""
    import oracle.cv2_oracle as o
    from tests import glue
    run, trained = bellows_run, bellows_trained
    rows, cols = run.cam_masks[0].shape
    ev, cam = 1, 1
    a = o.any_cam_analysis(o.FrameSource(run.events[ev][cam]), trained[cam], run.cam_masks[cam], run.bellows_masks[cam], ev, True, None)
    assert any(i.get("bellows_branch") for i in a.bellows)
    with capi.Context(rows, cols, 4, 40, max_slots=1) as ctx:
        for c in range(4):
            ctx.set_background(c, trained[c].avg, trained[c].sigma, trained[c].training_set_size)
            ctx.event_upload(0, c, run.events[ev][c])
        ctx.trigger_hist(0, 1)
        got = glue.gpu_any_cam_analysis(ctx, 0, cam, len(run.events[ev][cam]), trained[cam].training_set_size, run.cam_masks[cam],
                                        run.bellows_masks[cam], ev, False)
    assert o.write_camera_output(run.run_id, [got], 30) == o.write_camera_output(run.run_id, [a.out], 30)


def test_track_full_size_frame(capi):
    """BASELINE.json frame size (1050x1680) with a template of the reference's size (557x178, cam1_bellows_template.png)."""
    import oracle.cv2_oracle as o
    from autobub3hs_b200 import synth
    rows, cols = synth.ROWS, synth.COLS
    pat, (bx, by) = synth.make_bellows_object(1, rows, cols)
    assert pat.shape == (556, 178) or pat.shape[1] == 178
    rng = np.random.default_rng(3)
    bg = synth.make_background(1, rows, cols).astype(np.float32)
    frames = []
    for dx, dy in ((0, 0), (1, 3)):
        scene = bg.copy()
        scene[by + dy:by + dy + pat.shape[0], bx + dx:bx + dx + pat.shape[1]] = pat
        frames.append(np.clip(np.rint(scene + 2.0 * rng.standard_normal((rows, cols), dtype=np.float32)), 0, 255).astype(np.uint8))
    with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
        ctx.set_bellows_template(0, pat)
        ctx.event_upload(0, 0, np.stack(frames))
        for f, (dx, dy) in enumerate(((0, 0), (1, 3))):
            (x, y), peak = ctx.bellows_track(0, 0, f)
            got = ctx.bellows_map()
            want = o.ccorr_normed_map(frames[f], pat)
            assert np.abs(got - want).max() <= MAP_TOL
            ox, oy, opeak = o.subpixel_from_map(want)
            assert peak == opeak == (bx + dx, by + dy)
            assert abs(float(x) - float(ox)) <= POS_TOL and abs(float(y) - float(oy)) <= POS_TOL
