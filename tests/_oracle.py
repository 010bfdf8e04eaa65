"""ctypes loaders for the two CPU checkers (TEST INFRASTRUCTURE ONLY).

* ``load_oracle()``  -> oracle/libebt_oracle.so  (plain-C restatement; built on
  demand with gcc, works anywhere).
* ``load_ref()``     -> oracle/_ref/libebt_ref.so (the reference's own
  Patch/Optimizer + vendored Ceres; prebuilt in the authoring container, may be
  absent elsewhere -> returns None).

Both expose the same call shapes, wrapped by :class:`CpuTracker` and
:class:`CpuBatch`.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from event_based_odomety_b200.records import (EVENT, OPT_RECORD, PATCH_PIXELS, PROBLEM,
                                              SOLUTION, TRACK)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

_vp = C.c_void_p


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def load_oracle():
    so = os.path.join(ORACLE_DIR, "libebt_oracle.so")
    src = os.path.join(ORACLE_DIR, "ebt_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
    return _declare(C.CDLL(so), "oracle_")


def load_ref():
    so = os.path.join(ORACLE_DIR, "_ref", "libebt_ref.so")
    if not os.path.exists(so):
        return None
    try:
        return _declare(C.CDLL(so), "ref_")
    except OSError:
        return None


class _Lib:
    pass


def _declare(dll, prefix):
    L = _Lib()
    L.prefix = prefix
    L.dll = dll

    def fn(name, restype, *argtypes):
        f = getattr(dll, prefix + name)
        f.restype = restype
        f.argtypes = list(argtypes)
        setattr(L, name, f)

    if prefix == "ref_":
        fn("create", _vp, C.c_int, C.c_int, C.c_int, C.c_int)
    else:
        fn("create", _vp, C.c_int, C.c_int, C.c_int)
    fn("destroy", None, _vp)
    fn("set_draw_images", None, _vp, C.c_int)
    fn("unproject", None, _vp, _vp, C.c_int, _vp)
    fn("project", None, _vp, _vp, C.c_int, _vp)
    fn("new_frame_begin", C.c_int if prefix != "ref_" else None, _vp, _vp, C.c_int64, _vp, C.c_int)
    fn("get_uninit", C.c_int, _vp, _vp, _vp, C.c_int)
    fn("set_flow", None, _vp, C.c_int32, C.c_int, C.c_float, C.c_float)
    fn("new_frame_end", None, _vp)
    fn("push_events", C.c_int64, _vp, _vp, C.c_int64)
    fn("num_patches", C.c_int, _vp)
    fn("num_archived", C.c_int, _vp)
    fn("get_tracks", C.c_int, _vp, _vp, C.c_int, C.c_int)
    fn("num_records", C.c_int64, _vp)
    fn("get_records", C.c_int64, _vp, _vp, C.c_int64)
    fn("clear_records", None, _vp)
    fn("get_patch", C.c_int, _vp, C.c_int32, C.c_int, _vp)
    fn("get_trajectory", C.c_int, _vp, C.c_int32, _vp, C.c_int)
    fn("get_gradients", C.c_int, _vp, _vp, _vp)
    fn("frames_create", _vp, C.c_int, C.c_int)
    fn("frames_destroy", None, _vp)
    fn("frames_add", C.c_int, _vp, _vp)
    fn("frames_get", C.c_int, _vp, C.c_int, _vp, _vp)
    fn("set_draw_cost_map", None, C.c_int)
    fn("optimize_batch_cm", C.c_double, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp)
    if prefix == "ref_":
        fn("optimize_batch", C.c_double, _vp, _vp, C.c_int64, _vp, _vp, _vp, C.c_int)
        fn("eval_functor", C.c_int, _vp, C.c_int, _vp, C.c_double, C.c_double, _vp, C.c_double,
           _vp, _vp, _vp, _vp)
        fn("patch_integrate", C.c_int, C.c_double, C.c_double, C.c_int, _vp, C.c_int, _vp, _vp)
        fn("warp_index_map", None, _vp, C.c_int, C.c_int, _vp, _vp)
    else:
        fn("optimize_batch", C.c_double, _vp, _vp, C.c_int64, _vp, _vp, _vp, _vp)
        fn("eval_functor", C.c_int, _vp, C.c_int, _vp, C.c_double, C.c_double, _vp, C.c_double,
           _vp, _vp)
        fn("patch_integrate", C.c_int, C.c_double, C.c_double, _vp, C.c_int, _vp, _vp)
        fn("warp_index_map", None, _vp, C.c_int, C.c_int, _vp, _vp)
        fn("cosf_sinf", None, _vp, C.c_int, _vp, _vp)
    return L


class CpuTracker:
    """Stream-mode CPU tracker (reference harness or C restatement)."""

    def __init__(self, lib, width, height, max_patches=100, warp_image=False, draw_images=False):
        self.L = lib
        self.w, self.h = width, height
        if lib.prefix == "ref_":
            # warp_image: also run Patch::warpImage after every optimisation as the reference
            # does (two full-frame warps; off by default because it only feeds a getter)
            self.h_ = lib.create(width, height, max_patches, int(warp_image or draw_images))
        else:
            self.h_ = lib.create(width, height, max_patches)
        if draw_images:
            lib.set_draw_images(self.h_, 1)

    def close(self):
        if self.h_:
            self.L.destroy(self.h_)
            self.h_ = None

    def __del__(self):
        self.close()

    def new_frame_begin(self, gray, ts, corners):
        gray = np.ascontiguousarray(gray, np.uint8)
        corners = np.ascontiguousarray(corners, np.float64).reshape(-1, 2)
        self.L.new_frame_begin(self.h_, _ptr(gray), int(ts), _ptr(corners), len(corners))

    def get_uninit(self, cap=8192):
        ids = np.zeros(cap, np.int32)
        xy = np.zeros((cap, 2), np.float64)
        n = self.L.get_uninit(self.h_, _ptr(ids), _ptr(xy), cap)
        return ids[:n].copy(), xy[:n].copy()

    def set_flow(self, track_id, found, nx, ny):
        self.L.set_flow(self.h_, int(track_id), int(found), float(nx), float(ny))

    def new_frame_end(self):
        self.L.new_frame_end(self.h_)

    def push_events(self, events):
        events = np.ascontiguousarray(events)
        assert events.dtype == EVENT
        return self.L.push_events(self.h_, _ptr(events), len(events))

    def tracks(self, archived=False):
        n = self.L.num_archived(self.h_) if archived else self.L.num_patches(self.h_)
        out = np.zeros(max(n, 1), TRACK)
        self.L.get_tracks(self.h_, _ptr(out), n, int(archived))
        return out[:n]

    def records(self):
        n = self.L.num_records(self.h_)
        out = np.zeros(max(n, 1), OPT_RECORD)
        self.L.get_records(self.h_, _ptr(out), n)
        return out[:n]

    def patch(self, track_id, which=0):
        out = np.zeros(PATCH_PIXELS, np.float64)
        n = self.L.get_patch(self.h_, int(track_id), which, _ptr(out))
        return out if n == PATCH_PIXELS else None

    def cost_map(self, track_id):
        """Patch::getCostMap(): 11x11 after Optimizer::drawCostMap ran, else zeros."""
        out = np.zeros(PATCH_PIXELS, np.float64)
        n = self.L.get_patch(self.h_, int(track_id), 3, _ptr(out))
        return out[:121].reshape(11, 11).copy() if n == 121 else np.zeros((11, 11))

    def trajectory(self, track_id, cap=1 << 16):
        out = np.zeros((cap, 3), np.float64)
        n = self.L.get_trajectory(self.h_, int(track_id), _ptr(out), cap)
        return out[:max(n, 0)].copy()

    def gradients(self):
        gx = np.zeros((self.h, self.w))
        gy = np.zeros((self.h, self.w))
        self.L.get_gradients(self.h_, _ptr(gx), _ptr(gy))
        return gx, gy


class CpuBatch:
    """Batched independent feature-patch problems on the CPU checker."""

    def __init__(self, lib, width, height):
        self.L = lib
        self.w, self.h = width, height
        self.f_ = lib.frames_create(width, height)

    def close(self):
        if self.f_:
            self.L.frames_destroy(self.f_)
            self.f_ = None

    def __del__(self):
        self.close()

    def add_frame(self, gray):
        gray = np.ascontiguousarray(gray, np.uint8)
        return self.L.frames_add(self.f_, _ptr(gray))

    def gradients(self, slot):
        gx = np.zeros((self.h, self.w))
        gy = np.zeros((self.h, self.w))
        self.L.frames_get(self.f_, slot, _ptr(gx), _ptr(gy))
        return gx, gy

    def optimize(self, problems, events, want_nabla=True, threads=1):
        problems = np.ascontiguousarray(problems)
        events = np.ascontiguousarray(events)
        assert problems.dtype == PROBLEM and events.dtype == EVENT
        n = len(problems)
        sol = np.zeros(n, SOLUTION)
        nabla = np.zeros((n, PATCH_PIXELS)) if want_nabla else None
        if self.L.prefix == "ref_":
            secs = self.L.optimize_batch(self.f_, _ptr(problems), n, _ptr(events), _ptr(sol),
                                         _ptr(nabla), int(threads))
            evals = None
        else:
            evals = np.zeros((n, 2), np.int32)
            secs = self.L.optimize_batch(self.f_, _ptr(problems), n, _ptr(events), _ptr(sol),
                                         _ptr(nabla), _ptr(evals))
        return sol, nabla, evals, secs

    def optimize_cost_maps(self, problems, events):
        """Optimizer::optimize with OptimizerParams::drawCostMap on -> (solutions, n x 11 x 11 cost maps)."""
        problems = np.ascontiguousarray(problems)
        events = np.ascontiguousarray(events)
        n = len(problems)
        sol = np.zeros(n, SOLUTION)
        cm = np.zeros((n, 11, 11))
        self.L.optimize_batch_cm(self.f_, _ptr(problems), n, _ptr(events), _ptr(sol), None, _ptr(cm))
        return sol, cm


def cpu_unproject(lib, camera, pts):
    """CameraModel::unproject on the CPU checker; camera = cabi.Camera."""
    pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 2)
    out = np.zeros((len(pts), 3))
    lib.unproject(C.byref(camera), _ptr(pts), len(pts), _ptr(out))
    return out


def cpu_project(lib, camera, xyz):
    xyz = np.ascontiguousarray(xyz, np.float64).reshape(-1, 3)
    out = np.zeros((len(xyz), 2))
    lib.project(C.byref(camera), _ptr(xyz), len(xyz), _ptr(out))
    return out


def load_ingest():
    """oracle/libebt_ingest.so: C++ restatement of the reference's text event reader."""
    so = os.path.join(ORACLE_DIR, "libebt_ingest.so")
    src = os.path.join(ORACLE_DIR, "ingest_baseline.cpp")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "oracle"], stdout=subprocess.DEVNULL)
    dll = C.CDLL(so)
    dll.ingest_parse_text.restype = C.c_int64
    dll.ingest_parse_text.argtypes = [C.c_char_p, C.c_size_t, _vp, C.c_int64]
    return dll


def cpu_parse_text(dll, text: bytes):
    out = np.zeros(text.count(b"\n") + 1, EVENT)
    n = dll.ingest_parse_text(text, len(text), _ptr(out), len(out))
    if n < 0:
        raise ValueError("malformed event text")
    return out[:n]
