"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE (the CPU restatement of the
reference).  Imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from raysnail_b200._abi import RsCamera, RsRenderParams, RsSceneDesc, RsStats

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_LIB = os.path.join(ROOT, "oracle", "liboracle.so")

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(ORACLE_LIB):
        subprocess.check_call(["make", "-C", ROOT, "oracle"])
    lib = C.CDLL(ORACLE_LIB)
    vp, i32 = C.c_void_p, C.c_int32
    P = C.POINTER
    lib.rso_world_create.restype = vp
    lib.rso_world_create.argtypes = [P(RsSceneDesc), C.c_int]
    lib.rso_world_free.restype = None
    lib.rso_world_free.argtypes = [vp]
    lib.rso_root_bbox.argtypes = [vp, C.c_double, C.c_double, vp]
    lib.rso_philox4x32_10.argtypes = [vp, vp, vp]
    lib.rso_set_marcher_mode.argtypes = [C.c_int]
    lib.rso_render.argtypes = [vp, P(RsCamera), P(RsRenderParams), vp, vp, C.c_int, P(RsStats)]
    lib.rso_debug_primary.argtypes = [vp, P(RsCamera), P(RsRenderParams), i32, vp, vp, vp, vp, vp, vp, vp]
    lib.rso_debug_scatter.argtypes = [vp, P(RsCamera), P(RsRenderParams), i32, vp, vp, vp, vp]
    lib.rso_combine_passes.argtypes = [vp, vp, vp, C.c_int64, C.c_float]
    lib.rso_noise_map.argtypes = [vp, vp, i32, i32, i32]
    lib.rso_quantize_rgb8.argtypes = [vp, vp, C.c_int64]
    lib.rso_render_passes.argtypes = [vp, P(RsCamera), P(RsRenderParams), i32, i32, C.c_float, C.c_int, vp, vp]
    lib.rso_texture_color.argtypes = [vp, i32, vp, C.c_int64, vp]
    _lib = lib
    return lib


class OracleWorld:
    def __init__(self, scene, use_desc_bbox=True):
        self._lib = load()
        self._scene = scene
        self._w = self._lib.rso_world_create(scene.desc, 1 if use_desc_bbox else 0)
        if not self._w:
            raise RuntimeError("oracle could not build the world from the scene description")
        self.n_roots = scene.desc.contents.n_roots

    def __del__(self):
        try:
            if self._w:
                self._lib.rso_world_free(self._w)
                self._w = None
        except Exception:
            pass

    def root_bbox(self, time_range=(0.0, 0.0)):
        """`Hittable::bbox(time_limit)` of every world object (the range only matters for moving spheres)."""
        out = np.zeros((self.n_roots, 6))
        self._lib.rso_root_bbox(self._w, float(time_range[0]), float(time_range[1]), out.ctypes.data)
        return out

    def render(self, cam, params, mask=None, threads=0):
        out = np.empty((cam.height, cam.width, 4), dtype=np.float32)
        stats = RsStats()
        mptr = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            mptr = mask.ctypes.data
        rc = self._lib.rso_render(self._w, C.byref(cam), C.byref(params), mptr, out.ctypes.data, threads, C.byref(stats))
        if rc != 0:
            raise RuntimeError(f"rso_render failed: {rc}")
        return out, stats

    def render_passes(self, cam, params, n_passes, redo_mode=0, threshold=0.01, threads=0):
        """The CLI's passes loop (src/bin/raysnail.rs:379-427): (final old_pixels, redo count per pass)."""
        out = np.empty((cam.height, cam.width, 4), dtype=np.float32)
        counts = np.zeros(max(n_passes, 1), dtype=np.int64)
        rc = self._lib.rso_render_passes(self._w, C.byref(cam), C.byref(params), n_passes, redo_mode, threshold, threads, out.ctypes.data,
                                         counts.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"rso_render_passes failed: {rc}")
        return out, counts[:n_passes]

    def texture_color(self, tex, uvp):
        """Texture::color of texture index `tex` at rows (u, v, x, y, z) -> (n, 4) f32."""
        uvp = np.ascontiguousarray(uvp, dtype=np.float64).reshape(-1, 5)
        out = np.empty((uvp.shape[0], 4), dtype=np.float32)
        rc = self._lib.rso_texture_color(self._w, int(tex), uvp.ctypes.data, uvp.shape[0], out.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"rso_texture_color failed: {rc}")
        return out

    def debug_primary(self, cam, params, sample_index=0):
        n = cam.width * cam.height
        o = dict(obj_id=np.empty(n, np.int32), t=np.empty(n), normal=np.empty((n, 3)), point=np.empty((n, 3)),
                 front=np.empty(n, np.int32), material=np.empty(n, np.int32), ray=np.empty((n, 6)))
        self._lib.rso_debug_primary(self._w, C.byref(cam), C.byref(params), sample_index, o["obj_id"].ctypes.data, o["t"].ctypes.data,
                                    o["normal"].ctypes.data, o["point"].ctypes.data, o["front"].ctypes.data, o["material"].ctypes.data,
                                    o["ray"].ctypes.data)
        return o

    def debug_scatter(self, cam, params, sample_index=0):
        n = cam.width * cam.height
        o = dict(kind=np.empty(n, np.int32), next=np.empty((n, 6)), weight=np.empty((n, 3)), emitted=np.empty((n, 3)))
        rc = self._lib.rso_debug_scatter(self._w, C.byref(cam), C.byref(params), sample_index, o["kind"].ctypes.data, o["next"].ctypes.data,
                                         o["weight"].ctypes.data, o["emitted"].ctypes.data)
        if rc != 0:
            raise RuntimeError(f"rso_debug_scatter failed: {rc}")
        return o


def philox(ctr, key):
    lib = load()
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib.rso_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
    return out


def set_marcher_mode(mode):
    """0 = canonical (IEEE-exact) Mandelbulb arithmetic, 1 = literal libm restatement (oracle.cpp, namespace marcher)"""
    load().rso_set_marcher_mode(int(mode))
