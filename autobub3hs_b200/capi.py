"""ctypes binding of libabub_cuda.so (include/abub_cuda.h) -- harness plumbing only.

The product is the shared library; this module exists so that tests and bench.py can
call the C ABI with numpy buffers.  There is no CPU fallback: a missing library or a
failing call raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ABUB_LIB") or os.path.join(_HERE, "libabub_cuda.so")   # ABUB_LIB: kernel-variant sweeps (tools/k1_sweep.sh)

ABUB_MAX_TRACK = 10

EXPORTS = [
    "abub_ctx_create", "abub_ctx_destroy", "abub_last_error", "abub_sync", "abub_stream",
    "abub_set_background", "abub_set_background_from_training",
    "abub_pinned_acquire", "abub_event_submit", "abub_event_upload", "abub_event_submit_cam", "abub_event_upload_frame", "abub_set_trigger_metric", "abub_event_decode_png", "abub_event_flip_rows", "abub_event_device_ptr",
    "abub_event_set_nframes",
    "abub_trigger_hist", "abub_trigger_hist_read", "abub_trigger_hist_write", "abub_trigger_scan", "abub_trigger_results", "abub_trigger", "abub_trigger_stream", "abub_trigger_early",
    "abub_train_begin", "abub_train_add_event", "abub_train_add_slot", "abub_train_add_slots",
    "abub_train_add_events", "abub_train_finish", "abub_selftest_fdiv", "abub_k4_fdiv1_max",
    "abub_train_partials", "abub_train_allreduce", "abub_train_finish_partials", "abub_train_finish_allreduced",
    "abub_nccl_comm_create", "abub_nccl_comm_destroy", "abub_nccl_barrier",
    "abub_set_bellows_template", "abub_bellows_track", "abub_bellows_map", "abub_localize_bellows",
    "abub_localize", "abub_localize_batch", "abub_localize_binary", "abub_process_frame",
    "abub_event_submit_range", "abub_launch_count", "abub_profile", "abub_profile_read",
    "abub_gate_entropy_static", "abub_event_set_nframes_range", "abub_trigger_early_deferred", "abub_trigger_deferred_flush", "abub_trigger_deferred_start",
]


class Blob(C.Structure):
    """struct abub_blob."""
    _fields_ = [("k", C.c_int32), ("x", C.c_int32), ("y", C.c_int32), ("w", C.c_int32), ("h", C.c_int32),
                ("npts", C.c_int32), ("first_px", C.c_int32), ("reserved", C.c_int32),
                ("sum_x", C.c_int64), ("sum_y", C.c_int64),
                ("area", C.c_double), ("m00", C.c_double), ("m10", C.c_double), ("m01", C.c_double)]


_cudart = None


class PngJob(C.Structure):
    """struct abub_png_job."""
    _fields_ = [("slot", C.c_int32), ("cam", C.c_int32), ("frame", C.c_int32), ("zlen", C.c_uint32)]


class LocItem(C.Structure):
    """struct abub_loc_item."""
    _fields_ = [("slot", C.c_int32), ("cam", C.c_int32), ("trig", C.c_int32), ("pre_trig", C.c_int32),
                ("loc_thres", C.c_int32), ("n_post", C.c_int32)]


class AbubError(RuntimeError):
    pass


_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AbubError(f"{LIB_PATH} is missing: run `python -m autobub3hs_b200.build` (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    vp, ip, u8p = C.c_void_p, C.POINTER(C.c_int), C.c_void_p
    lib.abub_ctx_create.argtypes = [C.c_int] * 6 + [C.POINTER(vp)]
    lib.abub_ctx_destroy.argtypes = [vp]
    lib.abub_ctx_destroy.restype = None
    lib.abub_last_error.argtypes = [vp]
    lib.abub_last_error.restype = C.c_char_p
    lib.abub_sync.argtypes = [vp]
    lib.abub_stream.argtypes = [vp]
    lib.abub_stream.restype = vp
    lib.abub_set_background.argtypes = [vp, C.c_int, u8p, u8p, C.c_int]
    lib.abub_set_background_from_training.argtypes = [vp, C.c_int]
    lib.abub_pinned_acquire.argtypes = [vp, C.c_int, C.POINTER(vp)]
    lib.abub_event_submit.argtypes = [vp, C.c_int, ip]
    lib.abub_event_upload.argtypes = [vp, C.c_int, C.c_int, u8p, C.c_int]
    lib.abub_event_device_ptr.argtypes = [vp, C.c_int, C.POINTER(vp)]
    lib.abub_event_set_nframes.argtypes = [vp, C.c_int, ip]
    lib.abub_trigger_hist.argtypes = [vp, C.c_int, C.c_int]
    lib.abub_trigger_hist_read.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.abub_trigger_hist_write.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int]
    lib.abub_trigger_scan.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.abub_trigger_results.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp]
    lib.abub_trigger.argtypes = [vp, C.c_int, C.c_int, C.c_int, ip, ip, ip, vp]
    lib.abub_trigger_stream.argtypes = [vp, C.c_int, C.c_int, ip, C.c_int, vp, vp, vp, vp]
    lib.abub_trigger_early.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    lib.abub_train_begin.argtypes = [vp, C.c_int]
    lib.abub_train_add_event.argtypes = [vp, C.c_int, u8p, C.POINTER(C.c_float), ip]
    lib.abub_train_add_slot.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_float), ip]
    lib.abub_event_submit_cam.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    lib.abub_event_upload_frame.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    lib.abub_set_trigger_metric.argtypes = [vp, C.c_int, C.c_float]
    lib.abub_event_decode_png.argtypes = [vp, C.POINTER(PngJob), C.c_int, ip]
    lib.abub_event_flip_rows.argtypes = [vp, vp, C.c_int]
    lib.abub_train_add_slots.argtypes = [vp, C.c_int, ip, C.c_int, C.POINTER(C.c_float), ip]
    lib.abub_train_add_events.argtypes = [vp, C.c_int, u8p, C.c_int, C.POINTER(C.c_float), ip]
    lib.abub_selftest_fdiv.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
    lib.abub_k4_fdiv1_max.argtypes = []
    lib.abub_train_finish.argtypes = [vp, C.c_int, u8p, u8p, ip, ip]
    lib.abub_train_partials.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.abub_train_allreduce.argtypes = [vp, C.c_int, vp]
    lib.abub_train_finish_partials.argtypes = [vp, C.c_int, u8p, u8p, ip, ip, ip]
    lib.abub_train_finish_allreduced.argtypes = [vp, C.c_int, vp, u8p, u8p, ip, ip, ip]
    lib.abub_nccl_comm_create.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, C.POINTER(vp)]
    lib.abub_nccl_comm_destroy.argtypes = [vp]
    lib.abub_nccl_barrier.argtypes = [vp, vp]
    lib.abub_localize.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(Blob), C.c_int, ip]
    lib.abub_localize_batch.argtypes = [vp, C.POINTER(LocItem), C.c_int, C.POINTER(Blob), C.c_int, ip, ip]
    lib.abub_localize_binary.argtypes = [vp, C.c_int, u8p]
    fp = C.POINTER(C.c_float)
    lib.abub_set_bellows_template.argtypes = [vp, C.c_int, u8p, C.c_int, C.c_int]
    lib.abub_bellows_track.argtypes = [vp, C.c_int, C.c_int, C.c_int, fp, fp, ip]
    lib.abub_bellows_map.argtypes = [vp, fp, C.c_int, ip, ip]
    lib.abub_localize_bellows.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, fp, fp, C.POINTER(Blob), C.c_int,
                                          ip, ip, ip]
    lib.abub_process_frame.argtypes = [vp, C.c_int, u8p, u8p, u8p, vp]
    lib.abub_event_submit_range.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.abub_launch_count.argtypes = [vp]
    lib.abub_profile.argtypes = [vp, C.c_int]
    lib.abub_profile_read.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    lib.abub_event_set_nframes_range.argtypes = [vp, C.c_int, C.c_int, ip]
    lib.abub_trigger_early_deferred.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    lib.abub_trigger_deferred_flush.argtypes = [vp]
    lib.abub_trigger_deferred_start.argtypes = [vp]
    lib.abub_gate_entropy_static.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, C.c_char_p, C.c_int]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name not in ("abub_ctx_destroy", "abub_last_error", "abub_stream", "abub_launch_count"):
            fn.restype = C.c_int
    lib.abub_launch_count.restype = C.c_longlong
    _lib = lib
    return lib


ABUB_GATE_STATIC_MAX_PAIRS = 1024


def gate_entropy_static(d_pairs: int, n_pairs: int, rows: int, cols: int, d_entropy: int, d_kept: int = 0, d_hist16: int = 0,
                        stream: int = 0) -> Tuple[int, str]:
    """abub_gate_entropy_static: every pointer is a raw DEVICE address (int); returns (return code, error text)."""
    lib = load()
    err = C.create_string_buffer(256)
    rc = lib.abub_gate_entropy_static(C.c_void_p(d_pairs), n_pairs, rows, cols, C.c_void_p(d_entropy), C.c_void_p(d_kept or None),
                                      C.c_void_p(d_hist16 or None), C.c_void_p(stream or None), err, len(err))
    return rc, err.value.decode()


def _p(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u8(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


class Context:
    """Thin RAII wrapper around abub_ctx*."""

    def __init__(self, rows: int, cols: int, ncams: int = 4, max_frames: int = 40, max_slots: int = 1,
                 device: int = 0):
        self.lib = load()
        self.rows, self.cols, self.ncams, self.max_frames, self.max_slots = rows, cols, ncams, max_frames, max_slots
        self.npix = rows * cols
        h = C.c_void_p()
        rc = self.lib.abub_ctx_create(device, rows, cols, ncams, max_frames, max_slots, C.byref(h))
        self.h = h
        if rc != 0:
            msg = self.lib.abub_last_error(h).decode() if h else "abub_ctx_create failed"
            if h:
                self.lib.abub_ctx_destroy(h)
                self.h = None
            raise AbubError(f"abub_ctx_create rc={rc}: {msg}")

    @classmethod
    def adopt(cls, handle: int, rows: int, cols: int, ncams: int, max_frames: int, max_slots: int) -> "Context":
        """Wraps an abub_ctx* owned by someone else (the host library's AbubDevice): close() leaves it alone."""
        self = cls.__new__(cls)
        self.lib = load()
        self.rows, self.cols, self.ncams, self.max_frames, self.max_slots = rows, cols, ncams, max_frames, max_slots
        self.npix = rows * cols
        self.h = C.c_void_p(handle)
        self._borrowed = True
        return self

    def close(self) -> None:
        if getattr(self, "h", None):
            if not getattr(self, "_borrowed", False):
                self.lib.abub_ctx_destroy(self.h)
            self.h = None

    def launch_count(self) -> int:
        return int(self.lib.abub_launch_count(self.h))

    def profile(self, enable: bool) -> None:
        self._ck(self.lib.abub_profile(self.h, 1 if enable else 0), "abub_profile")

    def profile_read(self):
        """[(name, ms, calls)] per kernel group since profile(True)."""
        cap = 32
        names = (C.c_char_p * cap)()
        ms = (C.c_double * cap)()
        calls = (C.c_longlong * cap)()
        n = self.lib.abub_profile_read(self.h, cap, names, ms, calls)
        return [(names[i].decode(), float(ms[i]), int(calls[i])) for i in range(min(n, cap))]

    def event_submit_range(self, slot: int, cam: int, frame_first: int, nframes: int) -> None:
        self._ck(self.lib.abub_event_submit_range(self.h, slot, cam, frame_first, nframes), "abub_event_submit_range")

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int, what: str) -> None:
        if rc != 0:
            raise AbubError(f"{what} rc={rc}: {self.lib.abub_last_error(self.h).decode()}")

    # ---- plumbing ----
    def sync(self) -> None:
        self._ck(self.lib.abub_sync(self.h), "abub_sync")

    def stream(self) -> int:
        return int(self.lib.abub_stream(self.h) or 0)

    def set_background(self, cam: int, mean: np.ndarray, sigma: np.ndarray, training_set_size: int) -> None:
        m, s = _u8(mean), _u8(sigma)
        assert m.size == self.npix and s.size == self.npix
        self._ck(self.lib.abub_set_background(self.h, cam, _p(m), _p(s), training_set_size), "abub_set_background")

    def set_background_from_training(self, cam: int) -> None:
        self._ck(self.lib.abub_set_background_from_training(self.h, cam), "abub_set_background_from_training")

    def pinned(self, slot: int) -> np.ndarray:
        """numpy view [ncams, max_frames, rows, cols] of the slot's pinned staging buffer."""
        p = C.c_void_p()
        self._ck(self.lib.abub_pinned_acquire(self.h, slot, C.byref(p)), "abub_pinned_acquire")
        n = self.ncams * self.max_frames * self.npix
        buf = (C.c_uint8 * n).from_address(p.value)
        return np.frombuffer(buf, dtype=np.uint8).reshape(self.ncams, self.max_frames, self.rows, self.cols)

    def event_submit(self, slot: int, nframes: Sequence[int]) -> None:
        nf = (C.c_int * self.ncams)(*nframes)
        self._ck(self.lib.abub_event_submit(self.h, slot, nf), "abub_event_submit")

    def event_upload(self, slot: int, cam: int, frames: np.ndarray) -> None:
        f = _u8(frames)
        n = 0 if f.size == 0 else f.shape[0]
        self._ck(self.lib.abub_event_upload(self.h, slot, cam, _p(f) if n else None, n), "abub_event_upload")

    def event_device_ptr(self, slot: int) -> int:
        p = C.c_void_p()
        self._ck(self.lib.abub_event_device_ptr(self.h, slot, C.byref(p)), "abub_event_device_ptr")
        return int(p.value)

    def download_frames(self, slot: int) -> np.ndarray:
        """[ncams, max_frames, rows, cols] copy of the slot's device memory (test helper; cudaMemcpy through libcudart)."""
        global _cudart
        if _cudart is None:
            for name in ("libcudart.so.12", "libcudart.so"):
                try:
                    _cudart = C.CDLL(name)
                    break
                except OSError:
                    continue
            if _cudart is None:
                raise AbubError("libcudart not found")
            _cudart.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        self.sync()
        out = np.empty((self.ncams, self.max_frames, self.rows, self.cols), np.uint8)
        rc = _cudart.cudaMemcpy(out.ctypes.data, self.event_device_ptr(slot), out.size, 2)
        if rc != 0:
            raise AbubError(f"cudaMemcpy D2H failed ({rc})")
        return out

    def event_set_nframes(self, slot: int, nframes: Sequence[int]) -> None:
        nf = (C.c_int * self.ncams)(*nframes)
        self._ck(self.lib.abub_event_set_nframes(self.h, slot, nf), "abub_event_set_nframes")

    def event_set_nframes_range(self, slot_first: int, nframes) -> None:
        """nframes: [nslots, ncams] frame counts of slots slot_first ... (frames placed in HBM by a producer)."""
        nf = np.ascontiguousarray(nframes, dtype=np.int32).reshape(-1, self.ncams)
        self._ck(self.lib.abub_event_set_nframes_range(self.h, slot_first, len(nf), nf.ctypes.data_as(C.POINTER(C.c_int))),
                 "abub_event_set_nframes_range")

    # ---- trigger ----
    def trigger_hist(self, slot_first: int = 0, nslots: int = 1) -> None:
        self._ck(self.lib.abub_trigger_hist(self.h, slot_first, nslots), "abub_trigger_hist")

    def trigger_hist_read(self, slot: int, cam: int) -> np.ndarray:
        out = np.zeros((self.max_frames, 256), np.uint32)
        self._ck(self.lib.abub_trigger_hist_read(self.h, slot, cam, _p(out)), "abub_trigger_hist_read")
        return out

    def trigger_hist_write(self, slot: int, cam: int, hist: np.ndarray, nframes: int) -> None:
        h = np.zeros((self.max_frames, 256), np.uint32)
        h[:hist.shape[0]] = hist
        self._ck(self.lib.abub_trigger_hist_write(self.h, slot, cam, _p(h), nframes), "abub_trigger_hist_write")

    METRICS = {"significance": 0, "entropy": 1, "entropy_significance": 2}

    def set_trigger_metric(self, metric: str = "significance", threshold: float = 0.0) -> None:
        """Which quantity the trigger thresholds (AnalyzerUnit.cpp:226-228); threshold <= 0: the reference's value."""
        self._ck(self.lib.abub_set_trigger_metric(self.h, self.METRICS[metric], float(threshold)), "abub_set_trigger_metric")

    def trigger_scan(self, slot_first: int = 0, nslots: int = 1, start_frames: Optional[np.ndarray] = None) -> None:
        st = None if start_frames is None else np.ascontiguousarray(start_frames, dtype=np.int32)
        self._ck(self.lib.abub_trigger_scan(self.h, slot_first, nslots, _p(st)), "abub_trigger_scan")

    def trigger_results(self, slot_first: int = 0, nslots: int = 1, want_sig: bool = True):
        n = nslots * self.ncams
        trig, status, lt = (np.zeros(n, np.int32) for _ in range(3))
        sig = np.zeros((n, self.max_frames), np.float64) if want_sig else None
        self._ck(self.lib.abub_trigger_results(self.h, slot_first, nslots, _p(trig), _p(status), _p(lt), _p(sig)),
                 "abub_trigger_results")
        sh = (nslots, self.ncams)
        return trig.reshape(sh), status.reshape(sh), lt.reshape(sh), (None if sig is None else sig.reshape(sh + (-1,)))

    def trigger(self, slot: int, cam: int, start_frame: int = 1):
        t, s, l = C.c_int(), C.c_int(), C.c_int()
        sig = np.zeros(self.max_frames, np.float64)
        self._ck(self.lib.abub_trigger(self.h, slot, cam, start_frame, C.byref(t), C.byref(s), C.byref(l), _p(sig)),
                 "abub_trigger")
        return t.value, s.value, l.value, sig

    # ---- training ----
    def train_begin(self, cam: int) -> None:
        self._ck(self.lib.abub_train_begin(self.h, cam), "abub_train_begin")

    def train_add_event(self, cam: int, frame0: np.ndarray, frame1: np.ndarray) -> Tuple[float, bool]:
        f = np.ascontiguousarray(np.stack([_u8(frame0), _u8(frame1)]))
        e, k = C.c_float(), C.c_int()
        self._ck(self.lib.abub_train_add_event(self.h, cam, _p(f), C.byref(e), C.byref(k)), "abub_train_add_event")
        return e.value, bool(k.value)

    def train_add_slot(self, cam: int, slot: int) -> Tuple[float, bool]:
        e, k = C.c_float(), C.c_int()
        self._ck(self.lib.abub_train_add_slot(self.h, cam, slot, C.byref(e), C.byref(k)), "abub_train_add_slot")
        return e.value, bool(k.value)

    def train_add_slots(self, cam: int, slots) -> Tuple[np.ndarray, np.ndarray]:
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        e = np.zeros(len(sl), np.float32)
        k = np.zeros(len(sl), np.int32)
        self._ck(self.lib.abub_train_add_slots(self.h, cam, sl.ctypes.data_as(C.POINTER(C.c_int)), len(sl),
                                               e.ctypes.data_as(C.POINTER(C.c_float)), k.ctypes.data_as(C.POINTER(C.c_int))),
                 "abub_train_add_slots")
        return e, k.astype(bool)

    def train_add_events(self, cam: int, frames01: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """frames01: uint8 [n, 2, rows, cols]."""
        f = np.ascontiguousarray(frames01, dtype=np.uint8)
        assert f.ndim == 4 and f.shape[1:] == (2, self.rows, self.cols)
        n = f.shape[0]
        e = np.zeros(n, np.float32)
        k = np.zeros(n, np.int32)
        self._ck(self.lib.abub_train_add_events(self.h, cam, _p(f), n, e.ctypes.data_as(C.POINTER(C.c_float)),
                                                k.ctypes.data_as(C.POINTER(C.c_int))), "abub_train_add_events")
        return e, k.astype(bool)

    def selftest_fdiv(self, steps: int, k_first: int, k_last: int) -> Tuple[int, Tuple[int, int]]:
        n = C.c_ulonglong()
        ex = (C.c_ulonglong * 2)()
        self._ck(self.lib.abub_selftest_fdiv(self.h, steps, k_first, k_last, C.byref(n), ex), "abub_selftest_fdiv")
        return n.value, (ex[0], ex[1])

    def train_finish(self, cam: int, read: bool = True):
        """read=False leaves the planes on the device (abub_set_background_from_training installs them from there)."""
        mean = np.zeros((self.rows, self.cols), np.uint8) if read else None
        sigma = np.zeros((self.rows, self.cols), np.uint8) if read else None
        n, st = C.c_int(), C.c_int()
        self._ck(self.lib.abub_train_finish(self.h, cam, _p(mean), _p(sigma), C.byref(n), C.byref(st)),
                 "abub_train_finish")
        return mean, sigma, n.value, st.value

    def train_partials(self, cam: int) -> Tuple[int, int, int]:
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._ck(self.lib.abub_train_partials(self.h, cam, C.byref(a), C.byref(b), C.byref(c)), "abub_train_partials")
        return int(a.value), int(b.value), int(c.value)

    def train_finish_partials(self, cam: int):
        mean = np.zeros((self.rows, self.cols), np.uint8)
        sigma = np.zeros((self.rows, self.cols), np.uint8)
        n, st, nf = C.c_int(), C.c_int(), C.c_int()
        self._ck(self.lib.abub_train_finish_partials(self.h, cam, _p(mean), _p(sigma), C.byref(n), C.byref(st),
                                                     C.byref(nf)), "abub_train_finish_partials")
        return mean, sigma, n.value, st.value, nf.value

    def trigger_stream(self, slot_first: int, nslots: int, nframes, chunk_frames: int = 4):
        """Streamed FindTriggerFrame from the pinned staging buffers; returns trig, status, loc_thres, frames_sent
        as [nslots, ncams] arrays."""
        n = nslots * self.ncams
        nf = (C.c_int * n)(*[int(v) for v in np.asarray(nframes).ravel()])
        out = [np.zeros((nslots, self.ncams), np.int32) for _ in range(4)]
        self._ck(self.lib.abub_trigger_stream(self.h, slot_first, nslots, nf, chunk_frames, *[_p(o) for o in out]),
                 "abub_trigger_stream")
        return tuple(out)

    def trigger_early_deferred(self, slot_first: int, nslots: int, chunk_frames: int = 6) -> None:
        """Registers the same search for another slot range; the next localize_batch launches it between its stages."""
        self._ck(self.lib.abub_trigger_early_deferred(self.h, slot_first, nslots, chunk_frames), "abub_trigger_early_deferred")

    def trigger_deferred_start(self) -> None:
        self._ck(self.lib.abub_trigger_deferred_start(self.h), "abub_trigger_deferred_start")

    def trigger_deferred_flush(self) -> None:
        self._ck(self.lib.abub_trigger_deferred_flush(self.h), "abub_trigger_deferred_flush")

    def trigger_early(self, slot_first: int, nslots: int, chunk_frames: int = 6) -> None:
        """Resident FindTriggerFrame with the reference's early exit (K1 in rounds of chunk_frames frame pairs, decided
        cameras drop out); asynchronous, results through trigger_results."""
        self._ck(self.lib.abub_trigger_early(self.h, slot_first, nslots, chunk_frames), "abub_trigger_early")

    # ---- multi-GPU (one process per GPU) ----
    def nccl_comm_create(self, id_file: str, rank: int, world: int) -> int:
        c = C.c_void_p()
        self._ck(self.lib.abub_nccl_comm_create(self.h, id_file.encode(), rank, world, C.byref(c)), "abub_nccl_comm_create")
        return int(c.value)

    def nccl_comm_destroy(self, comm: int) -> None:
        self.lib.abub_nccl_comm_destroy(C.c_void_p(comm))

    def nccl_barrier(self, comm: int) -> None:
        self._ck(self.lib.abub_nccl_barrier(self.h, C.c_void_p(comm)), "abub_nccl_barrier")

    def train_allreduce(self, cam: int, comm: int) -> None:
        self._ck(self.lib.abub_train_allreduce(self.h, cam, C.c_void_p(comm)), "abub_train_allreduce")

    def train_finish_allreduced(self, cam: int, comm: int):
        mean = np.zeros((self.rows, self.cols), np.uint8)
        sigma = np.zeros((self.rows, self.cols), np.uint8)
        n, st, nf = C.c_int(), C.c_int(), C.c_int()
        self._ck(self.lib.abub_train_finish_allreduced(self.h, cam, C.c_void_p(comm), _p(mean), _p(sigma), C.byref(n),
                                                       C.byref(st), C.byref(nf)), "abub_train_finish_allreduced")
        return mean, sigma, n.value, st.value, nf.value

    # ---- localizer ----
    def localize(self, slot: int, cam: int, trig: int, pre_trig: int, loc_thres: int, n_post: int, cap: int = 4096):
        blobs = (Blob * cap)()
        npf = (C.c_int * (ABUB_MAX_TRACK + 1))()
        self._ck(self.lib.abub_localize(self.h, slot, cam, trig, pre_trig, loc_thres, n_post, blobs, cap, npf),
                 "abub_localize")
        counts = [npf[k] for k in range(n_post + 1)]
        return [blobs[i] for i in range(sum(counts))], counts

    def localize_batch_raw(self, items: Sequence[Tuple[int, int, int, int, int, int]], cap: int = 65536, buffers=None):
        """The bare C call (no per-blob Python objects): returns (blobs ctypes array, npf ctypes array, total, buffers);
        pass `buffers` back in to reuse the allocations."""
        n = len(items)
        if buffers is None or buffers[3] != n or buffers[4] != cap:
            buffers = ((LocItem * n)(*[LocItem(*it) for it in items]), (Blob * cap)(), (C.c_int * (n * (ABUB_MAX_TRACK + 1)))(), n, cap)
        arr, blobs, npf = buffers[0], buffers[1], buffers[2]
        tot = C.c_int()
        self._ck(self.lib.abub_localize_batch(self.h, arr, n, blobs, cap, npf, C.byref(tot)), "abub_localize_batch")
        return blobs, npf, tot.value, buffers

    def localize_batch(self, items: Sequence[Tuple[int, int, int, int, int, int]], cap: int = 65536):
        """items: (slot, cam, trig, pre_trig, loc_thres, n_post).  Returns per item a list (k=0..n_post) of blob lists."""
        n = len(items)
        arr = (LocItem * n)(*[LocItem(*it) for it in items])
        blobs = (Blob * cap)()
        npf = (C.c_int * (n * (ABUB_MAX_TRACK + 1)))()
        tot = C.c_int()
        self._ck(self.lib.abub_localize_batch(self.h, arr, n, blobs, cap, npf, C.byref(tot)), "abub_localize_batch")
        out, off = [], 0
        for i, it in enumerate(items):
            frames = []
            for k in range(it[5] + 1):
                c = npf[i * (ABUB_MAX_TRACK + 1) + k]
                frames.append([blobs[off + j] for j in range(c)])
                off += c
            out.append(frames)
        assert off == tot.value
        return out

    def decode_png(self, jobs) -> np.ndarray:
        """jobs: (slot, cam, frame, zlib_bytes).  The streams are copied into the slots' pinned staging regions and
        decoded on the device; returns ok[n] (bool)."""
        n = len(jobs)
        arr = (PngJob * n)()
        pins = {}
        for i, (slot, cam, frame, z) in enumerate(jobs):
            if slot not in pins:
                pins[slot] = self.pinned(slot)
            zb = np.frombuffer(z, dtype=np.uint8)
            pins[slot][cam, frame].reshape(-1)[:len(zb)] = zb
            arr[i] = PngJob(slot, cam, frame, len(zb))
        ok = (C.c_int * n)()
        self._ck(self.lib.abub_event_decode_png(self.h, arr, n, ok), "abub_event_decode_png")
        return np.array([ok[i] for i in range(n)], dtype=bool)

    def event_flip_rows(self, refs) -> None:
        """refs: (slot, cam, frame) of resident frames whose rows are stored bottom-up (BMP pixel arrays read as
        they lie in the file); turned over in place on the device."""
        a = np.ascontiguousarray(np.asarray(refs, dtype=np.int32).reshape(-1, 3))
        self._ck(self.lib.abub_event_flip_rows(self.h, _p(a), len(a)), "abub_event_flip_rows")

    # ---- bellows veto ----
    def set_bellows_template(self, cam: int, templ: Optional[np.ndarray]) -> None:
        if templ is None:
            self._ck(self.lib.abub_set_bellows_template(self.h, cam, None, 0, 0), "abub_set_bellows_template")
            return
        t = _u8(templ)
        self._ck(self.lib.abub_set_bellows_template(self.h, cam, _p(t), t.shape[0], t.shape[1]), "abub_set_bellows_template")

    def bellows_track(self, slot: int, cam: int, frame: int):
        """-> (x, y) np.float32, (peak_x, peak_y)"""
        x, y = C.c_float(), C.c_float()
        pk = (C.c_int * 2)()
        self._ck(self.lib.abub_bellows_track(self.h, slot, cam, frame, C.byref(x), C.byref(y), pk), "abub_bellows_track")
        return (np.float32(x.value), np.float32(y.value)), (pk[0], pk[1])

    def bellows_map(self) -> np.ndarray:
        r, c = C.c_int(), C.c_int()
        self._ck(self.lib.abub_bellows_map(self.h, None, 0, C.byref(r), C.byref(c)), "abub_bellows_map")
        out = np.zeros((r.value, c.value), np.float32)
        self._ck(self.lib.abub_bellows_map(self.h, out.ctypes.data_as(C.POINTER(C.c_float)), out.size, None, None), "abub_bellows_map")
        return out

    def localize_bellows(self, slot: int, cam: int, trig: int, pre_trig: int, loc_thres: int, pos_trig, pos_pre, cap: int = 65536):
        """-> (blobs, status, roi)"""
        pt = (C.c_float * 2)(float(pos_trig[0]), float(pos_trig[1]))
        pp = (C.c_float * 2)(float(pos_pre[0]), float(pos_pre[1]))
        out = (Blob * cap)()
        n, st = C.c_int(), C.c_int()
        roi = (C.c_int * 4)()
        self._ck(self.lib.abub_localize_bellows(self.h, slot, cam, trig, pre_trig, loc_thres, pt, pp, out, cap, C.byref(n),
                                                C.byref(st), roi), "abub_localize_bellows")
        return [out[i] for i in range(n.value)], st.value, tuple(roi)

    def localize_binary(self, k: int) -> np.ndarray:
        out = np.zeros((self.rows, self.cols), np.uint8)
        self._ck(self.lib.abub_localize_binary(self.h, k, _p(out)), "abub_localize_binary")
        return out

    # ---- parity taps ----
    def process_frame(self, cam: int, work: np.ndarray, prev: np.ndarray):
        w, p = _u8(work), _u8(prev)
        diff = np.zeros((self.rows, self.cols), np.uint8)
        hist = np.zeros(256, np.uint32)
        self._ck(self.lib.abub_process_frame(self.h, cam, _p(w), _p(p), _p(diff), _p(hist)), "abub_process_frame")
        return diff, hist


# ---------------------------------------------------------------------------------------
# libabubhost.so: the host C++ classes (Trainer / L3Localizer / OutputWriter, AbubDevice) with the harness entry
# points of host/ReconDriver.cpp -- the full recon of events produced in memory, through the abub3hs detect loop.
# ---------------------------------------------------------------------------------------
HOST_LIB_PATH = os.path.join(_HERE, "host", "libabubhost.so")
HOST_EXPORTS = ["abubhost_open", "abubhost_begin_run", "abubhost_recon", "abubhost_end_run", "abubhost_close", "abubhost_last_error",
                "abubhost_drain", "abubhost_set_threads", "abubhost_stats", "abubhost_recon_hint"]
_hostlib = None


def load_host() -> C.CDLL:
    global _hostlib
    if _hostlib is not None:
        return _hostlib
    load()                                     # libabub_cuda.so first (same copy for both bindings)
    if not os.path.exists(HOST_LIB_PATH):
        raise AbubError(f"{HOST_LIB_PATH} is missing: run `make -C autobub3hs_b200/host`")
    lib = C.CDLL(HOST_LIB_PATH)
    ip = C.POINTER(C.c_int)
    lib.abubhost_open.argtypes = [C.c_int] * 7 + [C.POINTER(C.c_void_p)]
    lib.abubhost_begin_run.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, ip]
    lib.abubhost_recon.argtypes = [C.c_int, C.c_int, ip, ip, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.abubhost_recon_hint.argtypes = [C.c_int, C.c_int, ip]
    lib.abubhost_end_run.argtypes = []
    lib.abubhost_drain.argtypes = []
    lib.abubhost_set_threads.argtypes = [C.c_int, C.c_int]
    lib.abubhost_stats.argtypes = [C.POINTER(C.c_double)]
    lib.abubhost_close.argtypes = []
    lib.abubhost_close.restype = None
    lib.abubhost_last_error.argtypes = []
    lib.abubhost_last_error.restype = C.c_char_p
    _hostlib = lib
    return lib


class HostRecon:
    """The abub3hs detect loop over events whose frames live in device slots / pinned staging buffers.
    `ctx` is a borrowed Context on the host library's abub_ctx (training, slot pointers, pinned buffers)."""

    STAT_KEYS = ("trigger_s", "trigger_readback_s", "localizer_s", "host_analysis_s", "file_write_s", "frames_sent")

    def __init__(self, rows: int, cols: int, ncams: int = 4, max_frames: int = 40, max_slots: int = 1, device: int = 0,
                 early_chunk: int = 6):
        self.hl = load_host()
        h = C.c_void_p()
        rc = self.hl.abubhost_open(device, rows, cols, ncams, max_frames, max_slots, early_chunk, C.byref(h))
        if rc != 0:
            raise AbubError("abubhost_open: " + self.hl.abubhost_last_error().decode())
        self.ctx = Context.adopt(h.value, rows, cols, ncams, max_frames, max_slots)
        self.ncams = ncams
        self.stats = (C.c_double * 8)()

    def begin_run(self, run_id: str, out_dir: str, mask_dir: str, frame_offset: int, training_set_sizes: Sequence[int]) -> str:
        tss = (C.c_int * self.ncams)(*[int(v) for v in training_set_sizes])
        rc = self.hl.abubhost_begin_run(run_id.encode(), out_dir.encode(), (mask_dir or "").encode(), frame_offset, tss)
        if rc != 0:
            raise AbubError("abubhost_begin_run: " + self.hl.abubhost_last_error().decode())
        return os.path.join(out_dir, f"abub3hs_{run_id}.txt")

    def recon(self, slot_first: int, event_numbers: Sequence[int], nframes, pinned: bool = False, chunk_frames: int = 4) -> None:
        n = len(event_numbers)
        ev = (C.c_int * n)(*[int(v) for v in event_numbers])
        nf = np.ascontiguousarray(nframes, dtype=np.int32).reshape(n * self.ncams)
        rc = self.hl.abubhost_recon(slot_first, n, ev, nf.ctypes.data_as(C.POINTER(C.c_int)), 1 if pinned else 0, chunk_frames,
                                    self.stats)
        if rc != 0:
            raise AbubError("abubhost_recon: " + self.hl.abubhost_last_error().decode())

    def recon_hint(self, slot_first: int, nslots: int, nframes) -> None:
        """Announces the next recon call (resident form): its trigger search is queued inside the current call's localizer."""
        nf = np.ascontiguousarray(nframes, dtype=np.int32).reshape(nslots * self.ncams)
        if self.hl.abubhost_recon_hint(slot_first, nslots, nf.ctypes.data_as(C.POINTER(C.c_int))) != 0:
            raise AbubError("abubhost_recon_hint: bad arguments")

    def set_threads(self, threads: int = 0, sub_batch: int = 32) -> None:
        """Host threads of the event loop (0 = OpenMP default) and events per device->host hand-over; call before begin_run."""
        self.hl.abubhost_set_threads(threads, sub_batch)

    def drain(self) -> None:
        """Returns when every submitted event has been analysed and its rows are in the file."""
        if self.hl.abubhost_drain() != 0:
            raise AbubError("abubhost_drain: " + self.hl.abubhost_last_error().decode())

    def read_stats(self, reset: bool = True) -> dict:
        """Seconds per stage since the previous call (drains the pipeline first) + frames sent over PCIe."""
        t = (C.c_double * 5)()
        if self.hl.abubhost_stats(t) != 0:
            raise AbubError("abubhost_stats: " + self.hl.abubhost_last_error().decode())
        d = {k: float(t[i]) for i, k in enumerate(self.STAT_KEYS[:5])}
        d["frames_sent"] = float(self.stats[5])
        if reset:
            for i in range(8):
                self.stats[i] = 0.0
        return d

    def end_run(self) -> None:
        self.hl.abubhost_end_run()

    def close(self) -> None:
        if getattr(self, "ctx", None) is not None:
            self.ctx.close()
            self.ctx = None
            self.hl.abubhost_close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
