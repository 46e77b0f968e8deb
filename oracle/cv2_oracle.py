"""cv2-backed oracle: restatement of the AutoBub3hs hot path (TEST INFRASTRUCTURE).

Every function cites the reference file:line it follows (paths relative to the
reference checkout).  Arithmetic that the reference delegates to OpenCV is
delegated to the same OpenCV entry point via cv2; first-party arithmetic
(significance, Welford, entropy, tracking, text formatting) is restated with
explicit float32 / float64 / int32 types so that it rounds like the C++.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import cv2
import numpy as np

cv2.setNumThreads(1)

NUM_FRAMES_BUBBLE_TRACK = 10  # common/CommonParameters.h:5
F32 = np.float32
F64 = np.float64


# ----------------------------------------------------------------------------
# a1  AnalyzerUnit::ProcessFrame  (AnalyzerUnit.cpp:341-377)
# ----------------------------------------------------------------------------
def sub6sigma(a: np.ndarray, b: np.ndarray, sigma: np.ndarray) -> np.ndarray:
    """`a - b - 6*sigma` as the MatExpr evaluates it (AnalyzerUnit.cpp:351-352):
    saturating u8 subtract materialised first, then addWeighted(.,1,sigma,-6)."""
    return cv2.addWeighted(cv2.subtract(a, b), 1.0, sigma, -6.0, 0.0)


def process_frame(work: np.ndarray, prev: np.ndarray, sigma: np.ndarray,
                  blur_diam: int = 5) -> np.ndarray:
    """AnalyzerUnit.cpp:346-377 (full-frame ROI)."""
    pos = sub6sigma(work, prev, sigma)
    neg = sub6sigma(prev, work, sigma)
    pos = cv2.GaussianBlur(pos, (blur_diam, blur_diam), 0)   # :359
    neg = cv2.GaussianBlur(neg, (blur_diam, blur_diam), 0)   # :360
    return cv2.absdiff(pos, neg)                             # :370


def process_frame_roi(work: np.ndarray, prev: np.ndarray, sigma: np.ndarray, roi, blur_diam: int = 5) -> np.ndarray:
    """AnalyzerUnit.cpp:346-377 with a ROI (x, y, w, h): the MatExpr results are fresh ROI-sized matrices, so the
    blur reflects at the ROI's own edges; outside the ROI the result is zero (:349, :370).  A ROI that leaves the
    frame makes cv::Mat::operator() throw (caught in AnyCamAnalysis -> -6)."""
    x, y, w, h = roi
    rows, cols = work.shape
    if not (0 <= x and 0 <= w and x + w <= cols and 0 <= y and 0 <= h and y + h <= rows):
        raise ValueError("ROI outside the frame (cv::Mat::operator() assertion)")
    if w == 0 or h == 0:
        raise ValueError("empty ROI (cv::GaussianBlur asserts on an empty matrix)")
    out = np.zeros_like(work)                                # :349
    cw = np.ascontiguousarray(work[y:y + h, x:x + w])
    cp = np.ascontiguousarray(prev[y:y + h, x:x + w])
    cs = np.ascontiguousarray(sigma[y:y + h, x:x + w])
    out[y:y + h, x:x + w] = process_frame(cw, cp, cs, blur_diam)
    return out


# ----------------------------------------------------------------------------
# a2  calcHist, 256 bins, no mask  (AnalyzerUnit.cpp:441-456)
# ----------------------------------------------------------------------------
def hist256(img: np.ndarray) -> np.ndarray:
    h = cv2.calcHist([img], [0], None, [256], [0, 256])
    return h.reshape(256)            # float32 counts


# ----------------------------------------------------------------------------
# a3  calculateSignificanceFrame + CalcMean/CalcStdDev
#     (AnalyzerUnit.cpp:435-504, 514-532)
# ----------------------------------------------------------------------------
def _wrap_i32(v: int) -> int:
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


class SignificanceState:
    """pix_counts (vector<vector<int>>, AnalyzerUnit.hpp:71) + loc_thres."""

    def __init__(self) -> None:
        self.pix_counts: List[List[int]] = [[] for _ in range(256)]
        self.entropies: List[np.float64] = []    # AnalyzerUnit.hpp:72
        self.loc_thres = 3          # AnalyzerUnit.cpp:29

    def clear(self) -> None:        # AnalyzerUnit.cpp:165-169
        self.pix_counts = [[] for _ in range(256)]
        self.entropies = []


def calc_mean(vec: Sequence[int], size: int) -> np.float64:
    """AnalyzerUnit.cpp:514-521.  sum in double; sum/size (int -> double)."""
    s = F64(0.0)
    for v in vec:
        s = s + F64(v)
    with np.errstate(all="ignore"):
        return s / F64(size)


def calc_stddev(vec: Sequence[int], mean: np.float64, size: int) -> np.float64:
    """AnalyzerUnit.cpp:523-532.  `val*val` is int*int (wraps mod 2^32)."""
    s = F64(0.0)
    for v in vec:
        s = s + F64(_wrap_i32(int(v) * int(v)))
    with np.errstate(all="ignore"):
        return np.sqrt(s / F64(size) - mean * mean)


def calculate_significance(hist: np.ndarray, total: int, store: bool,
                           st: SignificanceState, training_set_size: int,
                           loc_thres_max: int = 3) -> float:
    """AnalyzerUnit.cpp:435-504 on a precomputed 256-bin histogram."""
    significance = F64(0.0)
    pix_rem = int(total)
    first_over = -1
    max_adc = 0
    with np.errstate(all="ignore"):
        for i in range(256):
            bin_entry = F32(hist[i])
            if pix_rem > 0:
                if store:
                    st.pix_counts[i].append(int(bin_entry))          # :461
                if i > 1:
                    n = len(st.pix_counts[0])
                    mean = calc_mean(st.pix_counts[i], n)            # :474
                    sigma = calc_stddev(st.pix_counts[i], mean, n)   # :475
                    if (F64(bin_entry) != mean) or (sigma > 0):      # :478
                        significance = significance + (F64(bin_entry) - mean) / sigma
                    if significance < 0:                              # :479
                        significance = F64(0.0)
                if significance > 3.5 and first_over < 0:             # :483
                    first_over = i
                if i > max_adc:
                    max_adc = i
                if store:                                             # :487-491
                    lt = max(first_over - 1, max_adc - 1)
                    if lt < 2:
                        lt = 2
                    if lt > loc_thres_max or training_set_size < 6:
                        lt = loc_thres_max
                    st.loc_thres = lt
                # `pix_rem -= binEntry` : int -= float -> (int)((float)pix_rem - binEntry)
                pix_rem = int(F32(pix_rem) - bin_entry)               # :493
            else:
                break
    return float(significance)


# ----------------------------------------------------------------------------
# a4  AnalyzerUnit::FindTriggerFrame  (AnalyzerUnit.cpp:119-324)
# ----------------------------------------------------------------------------
@dataclass
class TriggerResult:
    status: int                 # TriggerFrameIdentificationStatus: 0, -3, -9
    trig: int                   # MatTrigFrame (unchanged when status != 0)
    ok: bool                    # okToProceed
    loc_thres: int
    sig: List[Tuple[int, float]] = field(default_factory=list)  # (frame, significance as double, before the float cast)


class FrameSource:
    """Minimal Parser::GetImage stand-in over an in-memory [N,rows,cols] u8 stack.
    `bad` = frame indices for which GetImage returns -1 (ZipParser.cpp:198,227)."""

    def __init__(self, frames: np.ndarray, bad: Sequence[int] = ()):
        self.frames = frames
        self.bad = set(bad)

    def __len__(self) -> int:
        return len(self.frames)

    def get(self, i: int) -> Tuple[int, np.ndarray]:
        if i in self.bad:
            return -1, None
        return 1, self.frames[i]


class HistProvider:
    """H[j] = hist256(ProcessFrame(frame[j], frame[j-d])) with memoisation, so the
    oracle pays for a frame pair once even across look-ahead and retriggers."""

    def __init__(self, src: FrameSource, sigma: np.ndarray):
        self.src, self.sigma, self.cache = src, sigma, {}

    def hist(self, j: int, jprev: int) -> np.ndarray:
        key = (j, jprev)
        if key not in self.cache:
            d = process_frame(self.src.frames[j], self.src.frames[jprev], self.sigma)
            self.cache[key] = hist256(d)
        return self.cache[key]


class TableHistProvider:
    """Histogram rows given up front (golden vectors / corner cases): H[j] for every j."""

    def __init__(self, table: np.ndarray):
        self.table = np.asarray(table)

    def hist(self, j: int, jprev: int) -> np.ndarray:
        return self.table[j].astype(np.float32)


class TableSource(FrameSource):
    """A frame source that only knows how many frames there are and the pixel count."""

    def __init__(self, nframes: int, total: int):
        self.frames = [np.zeros(total, np.uint8)] * nframes
        self.bad = set()


# The two alternatives the reference keeps beside the significance (AnalyzerUnit.cpp:226-227, :290-291; one of the three
# lines is compiled) and the entropyThreshold each one goes with (:140-142).
TRIGGER_THRESHOLDS = {"significance": 3.5, "entropy": 8e-5, "entropy_significance": 1.9}


def calculate_entropy_frame128(h256: np.ndarray, total: int) -> np.float32:
    """AnalyzerUnit::calculateEntropyFrame (AnalyzerUnit.cpp:386-423) from the 256-bin histogram of the frame: calcHist
    with 128 bins over [0,256) pairs adjacent grey levels (:392-394, :410); then the arithmetic of the Trainer's copy."""
    h128 = np.asarray(h256, dtype=np.int64).reshape(128, 2).sum(axis=1)
    return entropy_from_hist16(h128, total)


def calculate_entropy_significance(h256: np.ndarray, total: int, store: bool, st: SignificanceState) -> np.float64:
    """AnalyzerUnit::calculateEntropySignificance (AnalyzerUnit.cpp:425-433) with CalcMean/CalcStdDev on doubles."""
    e = F64(calculate_entropy_frame128(h256, total))              # double entropy = (float) :426
    if store:
        st.entropies.append(e)                                    # :427-429
    s = F64(0.0)
    for v in st.entropies:
        s = s + v
    s2 = F64(0.0)
    for v in st.entropies:
        s2 = s2 + v * v
    with np.errstate(all="ignore"):
        n = F64(len(st.entropies))
        mean = s / n                                              # :430
        sigma = np.sqrt(s2 / n - mean * mean)                     # :431
        return (e - mean) / sigma                                 # :432


def find_trigger_frame(src: FrameSource, sigma: np.ndarray, training_set_size: int,
                       startframe: int, st: SignificanceState, mat_trig_frame: int,
                       hp: Optional[HistProvider] = None,
                       min_eval_frame: int = 2, metric: str = "significance",
                       threshold: Optional[float] = None) -> TriggerResult:
    """AnalyzerUnit.cpp:119-324.  gammaCorrection is a no-op (:329).  `metric` picks which of the lines :226-228 (and
    :290-292) is the compiled one."""
    def measure(h, store):
        if metric == "significance":
            return calculate_significance(h, total, store, st, training_set_size)
        if metric == "entropy":
            return F64(calculate_entropy_frame128(h, total))
        return calculate_entropy_significance(h, total, store, st)

    n = len(src)
    if n < 5:                                                     # :122-126
        return TriggerResult(-9, mat_trig_frame, False, st.loc_thres)
    if hp is None:
        hp = HistProvider(src, sigma)
    total = int(src.frames[0].size)
    thr = F32(TRIGGER_THRESHOLDS[metric] if threshold is None else threshold)   # :140-142
    if startframe < 1:
        startframe = 1
    if startframe == 1:                                           # :165-169
        st.clear()
    prev = startframe - 1                                         # :175
    prevprev = prev if startframe < 2 else startframe - 2         # :176-177
    status = -3                                                   # :180
    two = True
    if training_set_size < 6:                                     # :185-188
        two = False
        thr = F32(F64(thr) * (5 / 3.5))
    res_sig: List[Tuple[int, float]] = []
    trig = mat_trig_frame
    for i in range(startframe, n):                                # :191
        err, _ = src.get(i)
        if err == -1:                                             # :208-213
            return TriggerResult(-9, trig, False, st.loc_thres, res_sig)
        h = hp.hist(i, prevprev if two else prev)                 # :218
        sg = measure(h, True)                                     # :226-228
        single = F32(sg)                                          # float singleEntropy (:146)
        res_sig.append((i, float(sg)))
        if single > thr and i >= min_eval_frame:                  # :254
            if i != n - 1:                                        # :261
                tpp, tp = prev, i                                 # :277-278
                max_so_far = F64(single)                          # :280
                ii = 1
                while ii <= 2 and ii + i < n:                     # :282
                    h2 = hp.hist(i + ii, tpp if two else tp)      # :286
                    single = F32(measure(h2, False))              # :290-292
                    with np.errstate(all="ignore"):
                        lhs = F64(single) / (F64(thr) / 3.5 * 5) + F64(single) / max_so_far
                    if lhs <= 3:                                  # :296
                        break
                    elif ii == 2:                                 # :297-300
                        status = 0
                        trig = i
                    if F64(single) > max_so_far:                  # :302
                        max_so_far = F64(single)
                    tpp, tp = tp, i + ii                          # :304-305
                    ii += 1
                if status == 0:
                    break
        prevprev, prev = prev, i                                  # :312-313
    ok = status != -3                                             # :317-321
    return TriggerResult(status, trig, ok, st.loc_thres, res_sig)


# ----------------------------------------------------------------------------
# a5  Trainer::calculateEntropyFrame  (Trainer.cpp:341-376) -- the training gate
# ----------------------------------------------------------------------------
def hist16(img: np.ndarray) -> np.ndarray:
    return cv2.calcHist([img], [0], None, [16], [0, 256]).reshape(16)   # :365


def entropy_from_hist16(h: np.ndarray, npix: int) -> np.float32:
    """Trainer.cpp:367-373.  `hist/(rows*cols)` is a MatExpr scale by 1/N applied by
    convertTo on a CV_32F Mat; `log2` resolves to the float overload."""
    # Mat/int -> MatExpr scale alpha=1./N, applied by convertTo as src*(float)alpha in float32
    p = np.asarray(h, dtype=F32) * F32(1.0 / float(npix))
    e = F32(0.0)
    for v in p:
        v = F32(v)
        if v != 0:
            e = F32(e - F32(v * F32(np.log2(v))))
    return e


def training_gate_entropy(frame0: np.ndarray, frame1: np.ndarray) -> np.float32:
    """Trainer.cpp:279-280: entropy of the saturating difference frame1 - frame0."""
    d = cv2.subtract(frame1, frame0)
    return entropy_from_hist16(hist16(d), d.size)


# ----------------------------------------------------------------------------
# a6  Trainer::CalculateMeanSigmaImageVector  (Trainer.cpp:144-216)
# ----------------------------------------------------------------------------
def welford_mean_sigma(frames: Sequence[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """Sequential float32 Welford per pixel, frames in push order (:180-189);
    var = m2/(N-1) (:191); sigma=(int)sqrt(var) -> u8 (:195); mean truncated (:196).
    Vectorised over pixels only: each pixel sees exactly the scalar recurrence."""
    n = len(frames)
    shape = frames[0].shape
    mean = np.zeros(shape, F32)
    m2 = np.zeros(shape, F32)
    for f, img in enumerate(frames):
        x = img.astype(F32)
        delta = x - mean
        mean = mean + delta / F32(f + 1)
        m2 = m2 + delta * (x - mean)
    with np.errstate(all="ignore"):
        var = m2 / F32(n - 1)              # n==1 -> 0/0 = NaN
        sig = np.sqrt(var)
    # (int)NaN on x86 = INT_MIN -> (uchar) = 0 ; finite values truncate
    sig_i = np.where(np.isfinite(sig), sig, 0).astype(np.int64)
    sigma_u8 = (sig_i & 0xFF).astype(np.uint8)
    mean_u8 = mean.astype(np.int64).astype(np.uint8)       # float -> uchar truncation
    return mean_u8, sigma_u8


@dataclass
class TrainedData:
    """Public result fields of Trainer (Trainer.hpp:35-54)."""
    avg: Optional[np.ndarray]
    sigma: Optional[np.ndarray]
    training_set_size: int
    status: int                      # 0 ok, -7 nothing kept (Trainer.cpp:308-313)
    kept_events: List[int] = field(default_factory=list)
    entropies: List[float] = field(default_factory=list)


def make_avg_sigma_image(event_frames01: Sequence[Optional[Tuple[np.ndarray, np.ndarray]]]) -> TrainedData:
    """Trainer::MakeAvgSigmaImage(false)  (Trainer.cpp:221-331).
    `event_frames01[i]` = (frame0, frame1) of event i for this camera, or None when the
    event has no frames / a corrupt training frame (isThisAGoodEvent=false, :253-276)."""
    kept: List[np.ndarray] = []
    kept_ev: List[int] = []
    ents: List[float] = []
    for i, pair in enumerate(event_frames01):
        if pair is None:
            ents.append(0.0)
            continue
        f0, f1 = pair
        e = training_gate_entropy(f0, f1)                      # :279-280
        ents.append(float(e))
        if e <= F32(0.0005):                                   # :287 (float <= double literal)
            kept.append(f0)
            kept.append(f1)
            kept_ev.append(i)
    if not kept:
        return TrainedData(None, None, 0, -7, kept_ev, ents)
    avg, sig = welford_mean_sigma(kept)
    return TrainedData(avg, sig, len(kept), 0, kept_ev, ents)


# ----------------------------------------------------------------------------
# a13 types: BubbleImageFrame (bubble.hpp:22-31), bubble (bubble.hpp:35-82)
# ----------------------------------------------------------------------------
@dataclass
class BubbleImageFrame:
    x: int
    y: int
    w: int
    h: int
    cont_area: float            # double
    cont_radius: float          # double
    m00: float
    m10: float
    m01: float
    cx: np.float32              # MassCentres (Point2f)
    cy: np.float32


class Bubble:
    """bubble.cpp:34-118."""

    def __init__(self, g: BubbleImageFrame):            # :34-48
        self.known = [g]
        self.last_x = F32(g.cx)
        self.last_y = F32(g.cy)
        self.genesis = g
        self.lock = True
        self.dz: List[np.float32] = []

    def push(self, f: BubbleImageFrame) -> None:        # operator<<  :54-72
        if not self.lock:
            self.known.append(f)
            self.dz.append(F32(self.last_x - F32(f.x)))
            self.last_x = F32(f.cx)
            self.last_y = F32(f.cy)
            self.lock = True

    def dzdt(self) -> np.float32:                       # :101-108
        n = len(self.known)
        total_z = F32(self.known[0].x - self.known[n - 1].x)
        with np.errstate(all="ignore"):
            return F32(F64(total_z) / (F64(F32(n)) - 1.0))

    def drdt(self) -> np.float32:                       # :110-118
        n = len(self.known)
        dx = F32(self.known[0].w - self.known[n - 1].w)
        dy = F32(self.known[0].h - self.known[n - 1].h)
        dr = F32(np.sqrt(F64(F32(F32(dx * dx) + F32(dy * dy)))))
        with np.errstate(all="ignore"):
            return F32(F64(dr) / (F64(F32(n)) - 1.0))


# ----------------------------------------------------------------------------
# a10  L3Localizer::isInMask  (L3Localizer.cpp:971-1012)
# ----------------------------------------------------------------------------
def is_in_mask(x: int, y: int, w: int, h: int, mask: Optional[np.ndarray], bellows: bool,
               mask_dir_set: bool = True) -> bool:
    if not mask_dir_set:
        return not bellows                               # :989
    if mask is None:                                     # unreadable mask :990-995
        return not bellows
    xpix = int(x + w / 2.0)                              # :974  double -> int
    ypix = y + h // 2                                    # :975  int division (h >= 0)
    return int(mask[ypix, xpix]) > 0                     # :1003


# ----------------------------------------------------------------------------
# contour measurement helpers (L3Localizer.cpp:264-278, 401-419, 818-823)
# ----------------------------------------------------------------------------
def find_external_contours(binary: np.ndarray):
    cs, _ = cv2.findContours(binary, cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_TC89_L1)
    return list(cs)


def _measure(c: np.ndarray, genesis: bool) -> BubbleImageFrame:
    x, y, w, h = cv2.boundingRect(c)
    area = float(cv2.contourArea(c))
    m = cv2.moments(c, False)
    with np.errstate(all="ignore"):
        radius = float(np.sqrt(F64(area) / F64(3.14159)))            # :404 / :819
        if (not genesis) or m["m00"] > 0:                            # :405 ; no guard at :822
            cx = F32(F64(m["m10"]) / F64(m["m00"]))
            cy = F32(F64(m["m01"]) / F64(m["m00"]))
        else:                                                        # :409-419
            pts = c.reshape(-1, 2)
            sx = F64(0.0)
            sy = F64(0.0)
            n = F64(0.0)
            for px, py in pts:
                sx = sx + F64(int(px))
                sy = sy + F64(int(py))
                n = n + 1
            cx = F32(sx / n)
            cy = F32(sy / n)
    return BubbleImageFrame(x, y, w, h, area, radius, m["m00"], m["m10"], m["m01"], cx, cy)


# ----------------------------------------------------------------------------
# a8  L3Localizer::CalculateInitialBubbleParams  (L3Localizer.cpp:215-460)
# ----------------------------------------------------------------------------
class BellowsBranchTaken(Exception):
    """Kept for callers that want to know: never raised since the bellows veto (a14) is restated below."""


def genesis_over_the_sigma(trig_frame, pre_frame, sigma) -> np.ndarray:
    return process_frame(trig_frame, pre_frame, sigma, 5)                        # :232


def binarise_over_the_sigma(over, loc_thres) -> np.ndarray:
    _, thr = cv2.threshold(over, loc_thres, 255, cv2.THRESH_TOZERO)              # :252 / :365
    _, thr = cv2.threshold(thr, 0, 255, cv2.THRESH_BINARY | cv2.THRESH_OTSU)     # :254 / :367
    return thr


def genesis_binary(trig_frame, pre_frame, sigma, loc_thres) -> np.ndarray:
    return binarise_over_the_sigma(genesis_over_the_sigma(trig_frame, pre_frame, sigma), loc_thres)


# ---- a14  bellows veto: TrackAFeature / GetDiffROI / block :292-390 -------------------------------------
def ccorr_normed_map(frame: np.ndarray, templ: np.ndarray) -> np.ndarray:
    """cv::matchTemplate(CV_TM_CCORR_NORMED) (L3Localizer.cpp:500), float32 [rows-th+1, cols-tw+1]."""
    return cv2.matchTemplate(frame, templ, cv2.TM_CCORR_NORMED)


def subpixel_from_map(result: np.ndarray):
    """L3Localizer.cpp:501-540 on a correlation map: min-max normalisation to [0,1] (:501), first maximum in raster
    order (:507), 3x3 value-weighted centroid accumulated in float (:522-533).  Returns (x, y) as np.float32 and the
    integer peak.  The reference reads the 3x3 neighbourhood without a bounds check; a peak on the map's edge is
    undefined behaviour there and an IndexError here."""
    res = cv2.normalize(result, None, 0, 1, cv2.NORM_MINMAX, -1)
    _, _, _, max_loc = cv2.minMaxLoc(res)
    mx, my = int(max_loc[0]), int(max_loc[1])
    if not (1 <= mx < res.shape[1] - 1 and 1 <= my < res.shape[0] - 1):
        raise IndexError("correlation peak on the edge of the map")
    sx = F32(0.0)
    sy = F32(0.0)
    total = F32(0.0)
    for i in (-1, 0, 1):                                   # :522 i = x offset (outer), j = y offset (inner)
        for j in (-1, 0, 1):
            v = F32(res[my + j, mx + i])
            sx = F32(sx + F32(F32(mx + i) * v))            # Point2f * float, then +=
            sy = F32(sy + F32(F32(my + j) * v))
            total = F32(total + v)
    with np.errstate(all="ignore"):
        inv = F32(F64(1.0) / F64(total))                   # :531  double division stored to float
    x = F32(F32(sx * inv) + F32(0))                        # :532, :538 (+ TemplateSearchZone.x = 0)
    y = F32(F32(sy * inv) + F32(0))
    return x, y, (mx, my)


def track_a_feature(frame: np.ndarray, templ: np.ndarray):
    """L3Localizer::TrackAFeature (L3Localizer.cpp:473-543): (x, y) float32 of the best template match."""
    x, y, _ = subpixel_from_map(ccorr_normed_map(frame, templ))
    return x, y


def nudge_apart(pos_trig, pos_pre):
    """:316-323  the two x positions are pushed one pixel apart (float ++ / --)."""
    tx, ty = pos_trig
    px, py = pos_pre
    if tx >= px:
        tx = F32(tx + F32(1))
        px = F32(px - F32(1))
    else:
        tx = F32(tx - F32(1))
        px = F32(px + F32(1))
    return (tx, ty), (px, py)


def get_diff_roi(p1, p2, templ_shape):
    """L3Localizer::GetDiffROI (:462-470): float arithmetic, truncated into ints."""
    th, tw = templ_shape
    start_x = int(max(F32(p1[0]), F32(p2[0])))
    start_y = int(max(F32(p1[1]), F32(p2[1])))
    delta_x = int(F32(F32(min(F32(p1[0]), F32(p2[0])) + F32(tw)) - F32(start_x)))
    delta_y = int(F32(F32(min(F32(p1[1]), F32(p2[1])) + F32(th)) - F32(start_y)))
    return start_x, start_y, delta_x, delta_y


def _paste(shape, templ, pos):
    """:347-349  Mat::zeros + TemplateImage.copyTo(frame(Rect(x, y, tw, th))) with Rect built from floats (truncation)."""
    rows, cols = shape
    th, tw = templ.shape
    x, y = int(pos[0]), int(pos[1])
    if not (0 <= x and x + tw <= cols and 0 <= y and y + th <= rows):
        raise ValueError("template pasted outside the frame (cv::Mat::operator() assertion)")
    out = np.zeros(shape, np.uint8)
    out[y:y + th, x:x + tw] = templ
    return out


def bellows_veto(over, trig_frame, pre_frame, sigma, templ, positions=None):
    """Block :303-364: returns overTheSigma with the bellows-motion model subtracted, and the positions used.
    `positions` = ((tx,ty),(px,py)) float32 from TrackAFeature; computed with cv2 when None."""
    if positions is None:
        positions = (track_a_feature(trig_frame, templ), track_a_feature(pre_frame, templ))   # :313-314
    pos_trig, pos_pre = nudge_apart(*positions)
    trig_copy = _paste(trig_frame.shape, templ, pos_trig)                        # :347
    pre_copy = _paste(pre_frame.shape, templ, pos_pre)                           # :349
    roi = get_diff_roi(pos_trig, pos_pre, templ.shape)                           # :355
    diff = process_frame_roi(trig_copy, pre_copy, sigma, roi, 5)                 # :356
    return cv2.subtract(over, diff), (pos_trig, pos_pre, roi)                    # :363 saturating -=


def calculate_initial_bubble_params(trig_frame, pre_frame, sigma, loc_thres,
                                    cam_mask, bellows_mask, mask_dir_set=True, bellows_template=None,
                                    info: Optional[dict] = None):
    """Returns the list of genesis BubbleImageFrame that become bubbles.  `bellows_template`: the camera's
    cam<c>_bellows_template.png (None = not loadable, :297-301).  `info` (optional dict) receives what the bellows
    branch did."""
    over = genesis_over_the_sigma(trig_frame, pre_frame, sigma)
    thr = binarise_over_the_sigma(over, loc_thres)
    contours = find_external_contours(thr)                                       # :264
    rects = []
    kept = []
    largest = 0
    all_in_bellows = len(contours) > 0                                           # :273
    for c in contours:                                                           # :275-290
        r = cv2.boundingRect(c)
        if not is_in_mask(r[0], r[1], r[2], r[3], bellows_mask, True, mask_dir_set):
            all_in_bellows = False
            if largest < r[2] * r[3]:
                largest = r[2] * r[3]
            kept.append(c)
            rects.append(r)
    if info is not None:
        info["bellows_branch"] = bool(all_in_bellows)
    if all_in_bellows:                                                           # :292-390
        if bellows_template is not None:
            over, used = bellows_veto(over, trig_frame, pre_frame, sigma, bellows_template)
            thr = binarise_over_the_sigma(over, loc_thres)                       # :365-367
            if info is not None:
                info["bellows_positions"] = used
        kept = find_external_contours(thr)                                       # :374 (all of them, no bellows filter)
        rects = [cv2.boundingRect(c) for c in kept]
        largest = 0
        for r in rects:                                                          # :383-389
            if largest < r[2] * r[3]:
                largest = r[2] * r[3]
    out = []
    for c, r in zip(kept, rects):                                                # :392-444
        box_area = r[2] * r[3]
        if box_area > 10 or box_area >= largest:
            f = _measure(c, True)
            if not is_in_mask(f.x, f.y, f.w, f.h, cam_mask, False, mask_dir_set):  # :432
                continue
            out.append(f)
    return out


# ----------------------------------------------------------------------------
# a9  L3Localizer::CalculatePostTriggerFrameParams  (L3Localizer.cpp:764-869)
# ----------------------------------------------------------------------------
def post_trigger_binary(frame, avg, sigma) -> np.ndarray:
    d = cv2.absdiff(frame, avg)                                                  # :779
    over = cv2.addWeighted(d, 1.0, sigma, -6.0, 0.0)                             # :782
    over = cv2.blur(over, (3, 3))                                                # :785
    _, over = cv2.threshold(over, 3, 255, cv2.THRESH_TOZERO)                     # :786
    _, over = cv2.threshold(over, 0, 255, cv2.THRESH_BINARY | cv2.THRESH_OTSU)   # :787
    return over


def post_trigger_candidates(frame, avg, sigma, cam_mask, mask_dir_set=True):
    over = post_trigger_binary(frame, avg, sigma)
    out = []
    for c in find_external_contours(over):                                       # :793
        r = cv2.boundingRect(c)
        if r[2] * r[3] > 10:                                                     # :812
            f = _measure(c, False)
            if not is_in_mask(f.x, f.y, f.w, f.h, cam_mask, False, mask_dir_set):  # :826
                continue
            out.append(f)
    return out


def track_candidates(bubbles: List[Bubble], cands: List[BubbleImageFrame]) -> None:
    """L3Localizer.cpp:842-863 (float compares; NaN never matches; no abs on x)."""
    for b in bubbles:
        b.lock = False
    for c in cands:
        tx, ty = F32(c.cx), F32(c.cy)
        for b in bubbles:
            with np.errstate(all="ignore"):
                if (F32(b.last_x - tx) < 5) and (abs(F64(F32(b.last_y - ty))) < 5):
                    b.push(c)
                    break


# ----------------------------------------------------------------------------
# a11  L3Localizer::LocalizeOMatic  (L3Localizer.cpp:881-968)
# ----------------------------------------------------------------------------
def localize_o_matic(src: FrameSource, trained: TrainedData, trig: int, loc_thres: int,
                     cam_mask, bellows_mask, bubbles: List[Bubble], mask_dir_set=True, bellows_template=None,
                     info: Optional[dict] = None) -> bool:
    """Appends to `bubbles` (AnalyzerUnit::BubbleList).  Returns okToProceed."""
    n = len(src)
    if n <= 5:                                                                   # :889
        return False
    prev_offset = 1 if trained.training_set_size < 6 else 2                      # :919-926
    pre = max(0, trig - prev_offset)                                             # :932-933
    for g in calculate_initial_bubble_params(src.frames[trig], src.frames[pre], trained.sigma,
                                             loc_thres, cam_mask, bellows_mask, mask_dir_set, bellows_template, info):
        bubbles.append(Bubble(g))
    kmax = NUM_FRAMES_BUBBLE_TRACK if trig < 29 else 39 - trig                   # :944-954
    for k in range(1, kmax + 1):
        if trig + k >= n:
            break
        cands = post_trigger_candidates(src.frames[trig + k], trained.avg, trained.sigma,
                                        cam_mask, mask_dir_set)
        track_candidates(bubbles, cands)
    return True


# ----------------------------------------------------------------------------
# PICOFormatWriterV4  (PICOFormatWriter/PICOFormatWriterV4.cpp:54-302)
# ----------------------------------------------------------------------------
def _ff(v) -> str:
    """ostream << fixed << setprecision(2) of a float/double (libstdc++: printf %.2f;
    NaN prints as "nan"/"-nan" by sign bit)."""
    v = float(v)
    if math.isnan(v):
        return "-nan" if math.copysign(1.0, v) < 0 else "nan"
    return "%.2f" % v


def write_header(nt: int = NUM_FRAMES_BUBBLE_TRACK) -> str:
    """PICOFormatWriterV4.cpp:54-88."""
    s = "Output of AutoBub v3 - the automatic unified bubble finder code by Pitam, using OpenCV.\n"
    s += "run  ev  ibubimage  TotalBub4CamImg  camera  frame0  hori  vert  GenesisW  GenesisH  dZdt  dRdt  "
    s += f"TrkFrame({nt})  TrkHori({nt})  TrkVert({nt})  "
    s += f"TrkBubW({nt})  TrkBubH({nt})  TrkBubRadius({nt})  FakeValue\n"
    s += "%12s  %5d  %d  %d  %d  %d  %.02f  %.02f  %d  %d  %.02f  %.02f  "
    s += "%d  " * nt
    s += "%.02f  " * (5 * nt)
    s += "%d"
    s += "\n8\n\n\n"
    return s


@dataclass
class CameraOutput:
    """OutputWriter::BubbleData (PICOFormatWriterV4.hpp:35-43)."""
    status: int = 0
    frame0: int = 0
    event: int = 0
    bubbles: List[Bubble] = field(default_factory=list)


def stage_camera_output(bubbles: List[Bubble], frame0: int, event: int) -> CameraOutput:
    """:99-110 -- empty list stages -1."""
    return CameraOutput(-1 if len(bubbles) == 0 else 0, frame0, event, list(bubbles))


def form_each_bubble_output(run: str, cam: int, wd: CameraOutput, ibub_start: int, nbub_total: int,
                            frame_offset: int, nt: int = NUM_FRAMES_BUBBLE_TRACK) -> Tuple[str, int]:
    """:132-283."""
    o = []
    if wd.status != 0:
        o.append(f"{run}  {wd.event}  0  0  {cam}  {wd.status}  0.00  0.00  0  0")
        o.append("  0.00  0.00  ")
        o.append("0  " * nt)
        o.append("0.00  " * (5 * nt))
        o.append("1  \n")
        return "".join(o), ibub_start
    for i, b in enumerate(wd.bubbles):
        o.append(f"{run}  {wd.event}  {ibub_start + i}  {nbub_total}  {cam}  ")
        f0 = wd.frame0 + frame_offset
        o.append(f"{f0}  ")
        g = b.genesis
        width, height = F32(g.w), F32(g.h)
        o.append(f"{_ff(F32(g.cx))}  {_ff(F32(g.cy))}  {int(width)}  {int(height)}  {_ff(b.dzdt())}  {_ff(b.drdt())}  ")
        npt = len(b.known) - 1
        nex = nt - npt if nt > npt else 0
        for j in range(1, npt + 1):
            o.append(f"{f0 + j}  ")
        for j in range(nex):
            o.append(f"{f0 + npt + j}  ")
        for getter in (lambda d: _ff(d.cx), lambda d: _ff(d.cy), lambda d: str(d.w),
                       lambda d: str(d.h), lambda d: _ff(d.cont_radius)):
            for j in range(1, npt + 1):
                o.append(getter(b.known[j]) + "  ")
            for j in range(nex):
                o.append("-1  ")
        o.append("1  \n")
    return "".join(o), ibub_start + len(wd.bubbles)


def write_camera_output(run: str, cams: List[CameraOutput], frame_offset: int) -> str:
    """:287-302."""
    nbt = sum(len(c.bubbles) if c.status == 0 else 0 for c in cams)
    ib = 1
    out = []
    for icam, c in enumerate(cams):
        s, ib = form_each_bubble_output(run, icam, c, ib, nbt, frame_offset)
        out.append(s)
    return "".join(out)


# ----------------------------------------------------------------------------
# AnyCamAnalysis (AutoBubStart3.cpp:67-125) and the detect loop (:342-388)
# ----------------------------------------------------------------------------
@dataclass
class CamAnalysis:
    out: CameraOutput
    triggers: List[TriggerResult] = field(default_factory=list)
    bellows: List[dict] = field(default_factory=list)        # one per LocalizeOMatic call: what the a14 branch did


def any_cam_analysis(src: FrameSource, trained: TrainedData, cam_mask, bellows_mask,
                     event_number: int, mask_dir_set: bool = True, bellows_template=None,
                     metric: str = "significance", threshold: Optional[float] = None) -> CamAnalysis:
    st = SignificanceState()
    hp = HistProvider(src, trained.sigma)
    bubbles: List[Bubble] = []
    mat_trig = 0                                                 # AnalyzerUnit.cpp:28
    ok = True
    res = CamAnalysis(CameraOutput())
    try:
        while True:
            tr = find_trigger_frame(src, trained.sigma, trained.training_set_size,
                                    mat_trig + 1, st, mat_trig, hp, metric=metric, threshold=threshold)   # :88
            res.triggers.append(tr)
            mat_trig = tr.trig
            ok = ok and tr.ok
            if ok:
                info: dict = {}
                ok = localize_o_matic(src, trained, mat_trig, tr.loc_thres, cam_mask,
                                      bellows_mask, bubbles, mask_dir_set, bellows_template, info)  # :94
                res.bellows.append(info)
                if ok:
                    res.out = stage_camera_output(bubbles, mat_trig, event_number)  # :97
                else:
                    res.out = CameraOutput(-8, 0, event_number)                      # :100
                    break
            else:
                res.out = CameraOutput(tr.status, 0, event_number)                   # :106
                break
            if len(bubbles) != 0:                                                     # :110
                break
    except Exception:                                                                 # :114-118
        res.out = CameraOutput(-6, 0, event_number)
    return res


def train_all(run_events: Sequence[Sequence[Optional[np.ndarray]]], ncams: int) -> List[TrainedData]:
    """run_events[ev][cam] = [N,rows,cols] u8 stack (or None).  AutoBubStart3.cpp:295-307."""
    out = []
    for cam in range(ncams):
        pairs = []
        for ev in run_events:
            fr = ev[cam]
            pairs.append(None if fr is None or len(fr) == 0 else (fr[0], fr[1]))
        out.append(make_avg_sigma_image(pairs))
    return out


def run_detect(run: str, event_numbers: Sequence[int], run_events, trained: List[TrainedData],
               cam_masks, bellows_masks, frame_offset: int = 30, bellows_templates=None,
               metric: str = "significance", threshold: Optional[float] = None) -> Tuple[str, list]:
    """Detect loop of main (AutoBubStart3.cpp:342-388): returns the text appended after
    the header and the per-(event,cam) analysis records."""
    text = []
    recs = []
    ncams = len(trained)
    for evn, ev in zip(event_numbers, run_events):
        cams = []
        rr = []
        for cam in range(ncams):
            a = any_cam_analysis(FrameSource(ev[cam]), trained[cam], cam_masks[cam],
                                 bellows_masks[cam], evn, True,
                                 None if bellows_templates is None else bellows_templates[cam], metric, threshold)
            cams.append(a.out)
            rr.append(a)
        text.append(write_camera_output(run, cams, frame_offset))
        recs.append(rr)
    return "".join(text), recs
