"""ctypes binding of the C-ABI in include/ebt.h (libebt_cuda.so).

This is the product path used by tests and bench.py: every computing call goes
into the CUDA library.  There is no CPU fallback — if the library is missing
``load()`` raises, and without a GPU ``Tracker()`` raises ``EbtError``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .records import EVENT, OPT_RECORD, PATCH_PIXELS, PROBLEM, SOLUTION, TRACK

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EBT_LIB", os.path.join(_HERE, "libebt_cuda.so"))
_vp = C.c_void_p


class EbtError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("max_patches", C.c_int32),
        ("init_num_events", C.c_int32), ("min_num_events", C.c_int32),
        ("max_num_events", C.c_int32), ("batch_stride", C.c_int32),
        ("max_iterations", C.c_int32), ("border_margin", C.c_int32),
        ("record_capacity", C.c_int32),
        ("association_distance", C.c_double), ("huber_loss", C.c_double),
        ("optimizer_threshold", C.c_double),
    ]


class FrontendParams(C.Structure):
    _fields_ = [("quality_level", C.c_double), ("min_distance", C.c_double), ("lk_levels", C.c_int32),
                ("lk_max_count", C.c_int32), ("lk_epsilon", C.c_double)]


class Camera(C.Structure):
    """common::CameraModelParams<double>"""
    _fields_ = [(k, C.c_double) for k in ("fx", "fy", "cx", "cy", "k1", "k2", "k3", "p1", "p2")]


EXPORTS = {
    "ebt_default_params": (C.c_int, [C.POINTER(Params), C.c_int, C.c_int]),
    "ebt_create": (C.c_int, [C.POINTER(Params), C.c_int, C.POINTER(_vp)]),
    "ebt_destroy": (C.c_int, [_vp]),
    "ebt_last_error": (C.c_char_p, [_vp]),
    "ebt_new_frame_begin": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int]),
    "ebt_new_frame_begin_device": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int]),
    "ebt_reset": (C.c_int, [_vp]),
    "ebt_measure_fma_peak": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "ebt_get_uninit": (C.c_int, [_vp, _vp, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_set_flow": (C.c_int, [_vp, C.c_int32, C.c_int, C.c_float, C.c_float]),
    "ebt_set_flows": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int]),
    "ebt_new_frame_end": (C.c_int, [_vp]),
    "ebt_new_frame_end_multi": (C.c_int, [_vp, C.c_int]),
    "ebt_default_frontend_params": (C.c_int, [C.POINTER(FrontendParams)]),
    "ebt_set_frontend_params": (C.c_int, [_vp, C.POINTER(FrontendParams)]),
    "ebt_detect_features": (C.c_int, [_vp, _vp, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_lk_flow": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, _vp, _vp]),
    "ebt_new_image": (C.c_int, [_vp, _vp, C.c_int64]),
    "ebt_get_features": (C.c_int, [_vp, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_push_events": (C.c_int, [_vp, _vp, C.c_int64, C.POINTER(C.c_int64)]),
    "ebt_push_events_device": (C.c_int, [_vp, _vp, C.c_int64, C.POINTER(C.c_int64)]),
    "ebt_push_events_multi": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int, _vp]),
    "ebt_set_shard": (C.c_int, [_vp, C.c_int, C.c_int]),
    "ebt_sizeof_patch_state": (C.c_size_t, []),
    "ebt_get_patch_state": (C.c_int, [_vp, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_merge_patch_state": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "ebt_parse_events_text": (C.c_int, [_vp, C.c_char_p, C.c_size_t, C.c_int, _vp, C.c_int64,
                                        C.POINTER(C.c_int64), C.POINTER(C.c_size_t)]),
    "ebt_parse_events_text_static": (C.c_int, [C.c_int, _vp, _vp, C.c_size_t, C.c_int, _vp, C.c_int64,
                                               C.POINTER(C.c_int64), C.POINTER(C.c_size_t), C.c_char_p, C.c_size_t]),
    "ebt_replay": (C.c_int, [_vp, _vp, C.c_int64, _vp, _vp, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "ebt_get_tracks": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "ebt_get_patch": (C.c_int, [_vp, C.c_int32, C.c_int, _vp]),
    "ebt_set_products": (C.c_int, [_vp, C.c_int]),
    "ebt_debug_feature_load": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp]),
    "ebt_set_stream_kernel": (C.c_int, [_vp, C.c_int]),
    "ebt_set_cost_map": (C.c_int, [_vp, C.c_int]),
    "ebt_get_cost_map": (C.c_int, [_vp, C.c_int32, _vp]),
    "ebt_get_batch_cost_maps": (C.c_int, [_vp, _vp, C.c_int64]),
    "ebt_get_landmarks": (C.c_int, [_vp, C.POINTER(Camera), _vp, _vp, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_unproject": (C.c_int, [_vp, C.POINTER(Camera), _vp, C.c_int, _vp]),
    "ebt_get_trajectory": (C.c_int, [_vp, C.c_int32, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_get_final_costs": (C.c_int, [_vp, C.c_int32, _vp, C.c_int, C.POINTER(C.c_int)]),
    "ebt_get_records": (C.c_int, [_vp, _vp, C.c_int64, C.POINTER(C.c_int64)]),
    "ebt_clear_records": (C.c_int, [_vp]),
    "ebt_get_gradients": (C.c_int, [_vp, _vp, _vp]),
    "ebt_frames_add": (C.c_int, [_vp, _vp]),
    "ebt_frames_get": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "ebt_optimize_batch": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int64, _vp, _vp]),
    "ebt_optimize_batch_device": (C.c_int, [_vp, _vp, C.c_int64, _vp, _vp, _vp]),
    "ebt_sync": (C.c_int, [_vp]),
    "ebt_timing_enable": (C.c_int, [_vp, C.c_int]),
    "ebt_timing_get": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "ebt_timing_reset": (C.c_int, [_vp]),
    "ebt_set_prefilter_margin": (C.c_int, [_vp, C.c_double]),
    "ebt_debug_phase_clocks": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "ebt_sizeof_event": (C.c_size_t, []),
    "ebt_sizeof_track": (C.c_size_t, []),
    "ebt_sizeof_opt_record": (C.c_size_t, []),
    "ebt_sizeof_problem": (C.c_size_t, []),
    "ebt_sizeof_solution": (C.c_size_t, []),
    "ebt_sizeof_params": (C.c_size_t, []),
    "ebt_sizeof_frontend_params": (C.c_size_t, []),
    "ebt_sizeof_camera": (C.c_size_t, []),
    "ebt_abi_version": (C.c_int, []),
}

_lib = None


def load():
    """Load libebt_cuda.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EbtError(f"{LIB_PATH} not built: run __graft_entry__.build() "
                           "(make -C event_based_odomety_b200/csrc)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in EXPORTS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return _vp(a)
    return a.ctypes.data_as(_vp)


K_TRACK, K_BATCH, K_GRAD, K_NUMEV, K_CORNERS, K_PYRAMID, K_LK, K_PARSE = range(8)


def push_events_multi(trackers, events, counts=None, device_resident=False):
    """ebt_push_events_multi: one batch per tracker in one launch.  events: list of
    numpy EVENT arrays (host) or of raw pointers with `counts` (host pinned or device).
    Returns the optimisation count per stream."""
    lib = load()
    S = len(trackers)
    ctxs = (_vp * S)(*[t.ctx for t in trackers])
    keep = []
    if counts is None:
        arrs = [np.ascontiguousarray(e) for e in events]
        for a in arrs:
            assert a.dtype == EVENT
        keep = arrs
        ptrs = (_vp * S)(*[a.ctypes.data if len(a) else None for a in arrs])
        counts = [len(a) for a in arrs]
    else:
        ptrs = (_vp * S)(*[int(p) if p else None for p in events])
    n = (C.c_int64 * S)(*[int(c) for c in counts])
    out = (C.c_int64 * S)()
    rc = lib.ebt_push_events_multi(ctxs, S, ptrs, n, int(bool(device_resident)), out)
    del keep
    if rc != 0:
        raise EbtError(f"ebt error {rc}: {lib.ebt_last_error(trackers[0].ctx).decode()}")
    return [int(v) for v in out]


def new_frame_end_multi(trackers):
    """ebt_new_frame_end_multi: the frame tails of several trackers, device work overlapped."""
    lib = load()
    S = len(trackers)
    ctxs = (_vp * S)(*[t.ctx for t in trackers])
    rc = lib.ebt_new_frame_end_multi(ctxs, S)
    if rc != 0:
        raise EbtError(f"ebt error {rc}: {lib.ebt_last_error(trackers[0].ctx).decode()}")


STATIC_PARSE_MAX_BYTES = 64 << 20  # EBT_STATIC_PARSE_MAX_BYTES (include/ebt.h)


def parse_events_text_static(device, d_text_ptr, n_bytes, d_out_ptr, cap, final=True, stream=0):
    """ebt_parse_events_text_static: text and output are device pointers the caller
    owns; nothing is allocated by the call.  Returns (lines parsed, bytes consumed)."""
    lib = load()
    n, used = C.c_int64(), C.c_size_t()
    msg = C.create_string_buffer(512)
    rc = lib.ebt_parse_events_text_static(int(device), _vp(stream), _vp(d_text_ptr), n_bytes, int(final),
                                          _vp(d_out_ptr), cap, C.byref(n), C.byref(used), msg, len(msg))
    if rc != 0:
        raise EbtError(f"ebt error {rc}: {msg.value.decode()}")
    return n.value, used.value


class Tracker:
    """One tracker context = one CUDA device + one stream (ebt_ctx)."""

    def __init__(self, width, height, max_patches=100, device=0, **overrides):
        self.lib = load()
        self.width, self.height = width, height
        self.params = Params()
        self.lib.ebt_default_params(C.byref(self.params), width, height)
        self.params.max_patches = max_patches
        for k, v in overrides.items():
            setattr(self.params, k, v)
        self.ctx = _vp()
        rc = self.lib.ebt_create(C.byref(self.params), device, C.byref(self.ctx))
        if rc != 0:
            msg = self.lib.ebt_last_error(None)
            self.ctx = None
            raise EbtError(f"ebt_create failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.ebt_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        self.close()

    def _ck(self, rc):
        if rc < 0:
            raise EbtError(f"ebt error {rc}: {self.lib.ebt_last_error(self.ctx).decode()}")
        return rc

    # ---- per-frame ------------------------------------------------------
    def new_frame_begin(self, gray, ts, corners):
        gray = np.ascontiguousarray(gray, np.uint8)
        assert gray.shape == (self.height, self.width)
        corners = np.ascontiguousarray(corners, np.float64).reshape(-1, 2)
        self._ck(self.lib.ebt_new_frame_begin(self.ctx, _ptr(gray), int(ts), _ptr(corners),
                                              len(corners)))

    def new_frame_begin_device(self, d_gray_ptr, ts, corners):
        corners = np.ascontiguousarray(corners, np.float64).reshape(-1, 2)
        self._ck(self.lib.ebt_new_frame_begin_device(self.ctx, _vp(d_gray_ptr), int(ts), _ptr(corners),
                                                     len(corners)))

    def reset(self):
        self._ck(self.lib.ebt_reset(self.ctx))

    def measure_fma_peak(self):
        a, b = C.c_double(), C.c_double()
        self._ck(self.lib.ebt_measure_fma_peak(self.ctx, C.byref(a), C.byref(b)))
        return a.value, b.value

    def get_uninit(self, cap=8192):
        ids = np.zeros(cap, np.int32)
        xy = np.zeros((cap, 2))
        n = C.c_int()
        self._ck(self.lib.ebt_get_uninit(self.ctx, _ptr(ids), _ptr(xy), cap, C.byref(n)))
        return ids[:n.value].copy(), xy[:n.value].copy()

    def set_flow(self, track_id, found, nx, ny):
        self._ck(self.lib.ebt_set_flow(self.ctx, int(track_id), int(found), float(nx), float(ny)))

    def set_flows(self, track_ids, found, next_xy):
        ids = np.ascontiguousarray(track_ids, np.int32)
        fnd = np.ascontiguousarray(found, np.int32)
        nxt = np.ascontiguousarray(next_xy, np.float32).reshape(-1, 2)
        assert len(ids) == len(fnd) == len(nxt)
        self._ck(self.lib.ebt_set_flows(self.ctx, ids.ctypes.data_as(_vp), fnd.ctypes.data_as(_vp),
                                        nxt.ctypes.data_as(_vp), len(ids)))

    def new_frame_end(self):
        self._ck(self.lib.ebt_new_frame_end(self.ctx))

    # ---- per-frame re-arm on the device (corner detector, LK) -------------
    def detect_features(self, gray, cap=8192):
        gray = np.ascontiguousarray(gray, np.uint8)
        assert gray.shape == (self.height, self.width)
        xy = np.zeros((cap, 2))
        n = C.c_int()
        self._ck(self.lib.ebt_detect_features(self.ctx, _ptr(gray), _ptr(xy), cap, C.byref(n)))
        return xy[:min(n.value, cap)].copy()

    def lk_flow(self, prev, cur, pts):
        prev = np.ascontiguousarray(prev, np.uint8)
        cur = np.ascontiguousarray(cur, np.uint8)
        assert prev.shape == cur.shape == (self.height, self.width)
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        nxt = np.zeros_like(pts)
        st = np.zeros(len(pts), np.uint8)
        self._ck(self.lib.ebt_lk_flow(self.ctx, _ptr(prev), _ptr(cur), _ptr(pts), len(pts), _ptr(nxt), _ptr(st)))
        return nxt, st

    def new_image(self, gray, ts):
        gray = np.ascontiguousarray(gray, np.uint8)
        assert gray.shape == (self.height, self.width)
        self._ck(self.lib.ebt_new_image(self.ctx, _ptr(gray), int(ts)))

    def features(self, cap=8192):
        xy = np.zeros((cap, 2))
        n = C.c_int()
        self._ck(self.lib.ebt_get_features(self.ctx, _ptr(xy), cap, C.byref(n)))
        return xy[:min(n.value, cap)].copy()

    def set_frontend_params(self, **kw):
        p = FrontendParams()
        self.lib.ebt_default_frontend_params(C.byref(p))
        for k, v in kw.items():
            setattr(p, k, v)
        self._ck(self.lib.ebt_set_frontend_params(self.ctx, C.byref(p)))

    # ---- events ---------------------------------------------------------
    def push_events(self, events):
        events = np.ascontiguousarray(events)
        assert events.dtype == EVENT
        n = C.c_int64()
        self._ck(self.lib.ebt_push_events(self.ctx, _ptr(events), len(events), C.byref(n)))
        return n.value

    def push_events_ptr(self, host_ptr, n_events):
        n = C.c_int64()
        self._ck(self.lib.ebt_push_events(self.ctx, _vp(host_ptr), n_events, C.byref(n)))
        return n.value

    def push_events_device(self, dev_ptr, n_events):
        n = C.c_int64()
        self._ck(self.lib.ebt_push_events_device(self.ctx, _vp(dev_ptr), n_events, C.byref(n)))
        return n.value

    # ---- ingestion ------------------------------------------------------
    # ---- feature-batch sharding ------------------------------------------
    def set_shard(self, index, count):
        self._ck(self.lib.ebt_set_shard(self.ctx, int(index), int(count)))

    def patch_state(self):
        """Opaque per-patch state of this shard's list, (n, ebt_sizeof_patch_state()) bytes."""
        sz = self.lib.ebt_sizeof_patch_state()
        n = C.c_int()
        self._ck(self.lib.ebt_get_patch_state(self.ctx, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), sz), np.uint8)
        self._ck(self.lib.ebt_get_patch_state(self.ctx, _ptr(out), n.value, C.byref(n)))
        return out[:n.value]

    def merge_patch_state(self, states):
        """states: one patch_state() array per shard, in shard order."""
        states = [np.ascontiguousarray(s, np.uint8) for s in states]
        ptrs = (_vp * len(states))(*[s.ctypes.data if len(s) else None for s in states])
        self._ck(self.lib.ebt_merge_patch_state(self.ctx, ptrs, len(states), len(states[0])))

    def parse_events_text(self, text: bytes, final=True, cap=None):
        """(events, bytes consumed) of a buffer of "t_sec x y p" lines."""
        if cap is None:
            cap = text.count(b"\n") + 1
        out = np.zeros(max(cap, 1), EVENT)
        n = C.c_int64()
        used = C.c_size_t()
        self._ck(self.lib.ebt_parse_events_text(self.ctx, text, len(text), int(final), _ptr(out), cap,
                                                C.byref(n), C.byref(used)))
        return out[:n.value].copy(), used.value

    def replay(self, events, frames, frame_ts):
        """Replayer order over in-memory samples; returns (events delivered, optimisations)."""
        events = np.ascontiguousarray(events)
        assert events.dtype == EVENT
        frames = np.ascontiguousarray(np.stack(frames), np.uint8)
        ts = np.ascontiguousarray(frame_ts, np.int64)
        assert frames.shape == (len(ts), self.height, self.width)
        nd, no = C.c_int64(), C.c_int64()
        self._ck(self.lib.ebt_replay(self.ctx, _ptr(events), len(events), _ptr(frames), _ptr(ts), len(ts),
                                     C.byref(nd), C.byref(no)))
        return nd.value, no.value

    # ---- getters --------------------------------------------------------
    def tracks(self, archived=False):
        n = C.c_int()
        self._ck(self.lib.ebt_get_tracks(self.ctx, None, 0, int(archived), C.byref(n)))
        out = np.zeros(max(n.value, 1), TRACK)
        self._ck(self.lib.ebt_get_tracks(self.ctx, _ptr(out), n.value, int(archived), C.byref(n)))
        return out[:n.value]

    def records(self):
        n = C.c_int64()
        self._ck(self.lib.ebt_get_records(self.ctx, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), OPT_RECORD)
        self._ck(self.lib.ebt_get_records(self.ctx, _ptr(out), n.value, C.byref(n)))
        return out[:n.value]

    def clear_records(self):
        self._ck(self.lib.ebt_clear_records(self.ctx))

    def patch(self, track_id, which=0):
        out = np.zeros(PATCH_PIXELS)
        self._ck(self.lib.ebt_get_patch(self.ctx, int(track_id), which, _ptr(out)))
        return out

    # ---- tracker -> VO hand-off ---------------------------------------------
    def landmarks(self, camera=None, cap=8192):
        """(track ids, corners (n,2), bearings (n,3) or None) of the live patches."""
        ids = np.zeros(cap, np.int32)
        xy = np.zeros((cap, 2))
        b = np.zeros((cap, 3)) if camera is not None else None
        n = C.c_int()
        self._ck(self.lib.ebt_get_landmarks(self.ctx, C.byref(camera) if camera is not None else None,
                                            _ptr(ids), _ptr(xy), _ptr(b), cap, C.byref(n)))
        m = min(n.value, cap)
        return ids[:m].copy(), xy[:m].copy(), (b[:m].copy() if b is not None else None)

    def unproject(self, camera, pts):
        pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 2)
        out = np.zeros((len(pts), 3))
        self._ck(self.lib.ebt_unproject(self.ctx, C.byref(camera), _ptr(pts), len(pts), _ptr(out)))
        return out

    def set_products(self, mode):
        self._ck(self.lib.ebt_set_products(self.ctx, int(mode)))

    def feature_load(self, cap=8192):
        """(track ids, optimisations, t0 ns, t1 ns, sm) of every live patch's CTA in the previous push."""
        ids, nopt, sm = (np.zeros(cap, np.int32) for _ in range(3))
        t0, t1 = np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
        n = C.c_int(0)
        self._ck(self.lib.ebt_debug_feature_load(self.ctx, _ptr(ids), _ptr(nopt), _ptr(t0), _ptr(t1), _ptr(sm),
                                                 cap, C.byref(n)))
        k = min(n.value, cap)
        return ids[:k], nopt[:k], t0[:k], t1[:k], sm[:k]

    def set_stream_kernel(self, mode):
        """0: one CTA per feature; 2: persistent CTAs of two leader warps sharing four pixel warps; 1: automatic."""
        self._ck(self.lib.ebt_set_stream_kernel(self.ctx, int(mode)))

    def set_cost_map(self, on=True):
        """OptimizerParams::drawCostMap."""
        self._ck(self.lib.ebt_set_cost_map(self.ctx, int(on)))

    def cost_map(self, track_id):
        """Patch::getCostMap(): 11x11."""
        out = np.zeros((11, 11), np.float64)
        self._ck(self.lib.ebt_get_cost_map(self.ctx, int(track_id), _ptr(out)))
        return out

    def batch_cost_maps(self, n):
        out = np.zeros((int(n), 11, 11), np.float64)
        self._ck(self.lib.ebt_get_batch_cost_maps(self.ctx, _ptr(out), int(n)))
        return out

    def trajectory(self, track_id):
        n = C.c_int()
        self._ck(self.lib.ebt_get_trajectory(self.ctx, int(track_id), None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 3))
        self._ck(self.lib.ebt_get_trajectory(self.ctx, int(track_id), _ptr(out), n.value, C.byref(n)))
        return out[:n.value]

    def final_costs(self, track_id):
        n = C.c_int()
        self._ck(self.lib.ebt_get_final_costs(self.ctx, int(track_id), None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1))
        self._ck(self.lib.ebt_get_final_costs(self.ctx, int(track_id), _ptr(out), n.value, C.byref(n)))
        return out[:n.value]

    def gradients(self):
        gx = np.zeros((self.height, self.width))
        gy = np.zeros((self.height, self.width))
        self._ck(self.lib.ebt_get_gradients(self.ctx, _ptr(gx), _ptr(gy)))
        return gx, gy

    # ---- batched --------------------------------------------------------
    def add_frame(self, gray):
        gray = np.ascontiguousarray(gray, np.uint8)
        assert gray.shape == (self.height, self.width)
        return self._ck(self.lib.ebt_frames_add(self.ctx, _ptr(gray)))

    def frame_gradients(self, slot):
        gx = np.zeros((self.height, self.width))
        gy = np.zeros((self.height, self.width))
        self._ck(self.lib.ebt_frames_get(self.ctx, slot, _ptr(gx), _ptr(gy)))
        return gx, gy

    def optimize_batch(self, problems, events, want_nabla=True):
        problems = np.ascontiguousarray(problems)
        events = np.ascontiguousarray(events)
        assert problems.dtype == PROBLEM and events.dtype == EVENT
        sol = np.zeros(len(problems), SOLUTION)
        nabla = np.zeros((len(problems), PATCH_PIXELS), np.int32) if want_nabla else None
        self._ck(self.lib.ebt_optimize_batch(self.ctx, _ptr(problems), len(problems), _ptr(events),
                                             len(events), _ptr(sol), _ptr(nabla)))
        return sol, nabla

    def optimize_batch_device(self, d_problems, n, d_events, d_solutions, d_nabla=0):
        self._ck(self.lib.ebt_optimize_batch_device(self.ctx, _vp(d_problems), n, _vp(d_events),
                                                    _vp(d_solutions), _vp(d_nabla) if d_nabla else None))

    def sync(self):
        self._ck(self.lib.ebt_sync(self.ctx))

    def phase_clocks(self):
        out = (C.c_uint64 * 8)()
        self._ck(self.lib.ebt_debug_phase_clocks(self.ctx, out))
        return [int(v) for v in out]

    def set_prefilter_margin(self, pixels):
        self._ck(self.lib.ebt_set_prefilter_margin(self.ctx, float(pixels)))

    def timing(self, on=True):
        self._ck(self.lib.ebt_timing_enable(self.ctx, int(on)))
        self._ck(self.lib.ebt_timing_reset(self.ctx))

    def timing_get(self, kernel):
        ms = C.c_double()
        n = C.c_int64()
        self._ck(self.lib.ebt_timing_get(self.ctx, kernel, C.byref(ms), C.byref(n)))
        return ms.value, n.value
