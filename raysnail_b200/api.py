"""Host-side Python mirror of the reference's render interface, over the C ABI.

Names follow the reference (src/camera.rs, src/painter.rs, src/sdl_parser.rs):
`CameraBuilder(...).build()`, `camera.take_photo().samples(n).depth(d).shot(world)`,
`SdlParser.parse(path)`.  Everything that computes pixels goes through
libraysnail_b200.so's `rs_*` entry points (CUDA, sm_100a); nothing here falls back to the
CPU — without a GPU `Renderer()` raises.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional, Sequence

import numpy as np

from . import _abi
from ._abi import RsCamera, RsRenderParams, RsStats

DEFAULT_SEED = 0x5EED0007


class RaysnailError(RuntimeError):
    pass


def _vec3(v: Sequence[float]):
    return (C.c_double * 3)(float(v[0]), float(v[1]), float(v[2]))


def painter_samples(samples_requested: int) -> int:
    """`Painter::samples` (src/painter.rs:110-118): sqrt_spp = floor(sqrt(n))."""
    return int(math.floor(math.sqrt(float(samples_requested))))


class Scene:
    """Owner of an `rs_host_scene` (the C++ mirror of the reference scene graph)."""

    def __init__(self, handle=None):
        self._lib = _abi.load()
        self._h = C.c_void_p(handle) if handle is not None else C.c_void_p(self._lib.rs_host_scene_new())
        self._desc = None

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self._lib.rs_host_scene_free(self._h)
                self._h = C.c_void_p(None)
        except Exception:
            pass

    # ---- whole scenes ---------------------------------------------------------------------
    @classmethod
    def _from_call(cls, fn, *args) -> "Scene":
        lib = _abi.load()
        out = C.c_void_p()
        rc = getattr(lib, fn)(*args, C.byref(out))
        if rc != 0:
            raise RaysnailError(f"{fn} failed: {lib.rs_host_last_error().decode()}")
        return cls(out.value)

    @classmethod
    def from_sdl(cls, path: str) -> "Scene":
        return cls._from_call("rs_host_scene_from_sdl_file", path.encode())

    @classmethod
    def from_sdl_text(cls, text: str) -> "Scene":
        return cls._from_call("rs_host_scene_from_sdl_text", text.encode())

    @classmethod
    def builtin(cls, name: str, seed: int = 7, iparam: int = 0) -> "Scene":
        return cls._from_call("rs_host_scene_builtin", name.encode(), C.c_uint64(seed), C.c_int32(iparam))

    # ---- incremental construction -----------------------------------------------------------
    def _ck(self, rc: int, what: str) -> int:
        if rc < 0:
            raise RaysnailError(f"{what}: {self._lib.rs_host_last_error().decode()}")
        self._desc = None
        return rc

    def solid(self, r, g, b, a=1.0):
        return self._ck(self._lib.rs_host_tex_solid(self._h, r, g, b, a), "tex_solid")

    def checker(self, odd, even, scale):
        return self._ck(self._lib.rs_host_tex_checker(self._h, odd, even, scale), "tex_checker")

    def image(self, rgba: np.ndarray):
        rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
        h, w, c = rgba.shape
        assert c == 4
        return self._ck(self._lib.rs_host_tex_image(self._h, w, h, rgba.ctypes.data), "tex_image")

    def image_file(self, path: str):
        return self._ck(self._lib.rs_host_tex_image_file(self._h, path.encode()), "tex_image_file")

    def perlin(self, point_count=256, vector=True, seed=1, kind=0, depth=0, smooth=2, scale=1.0):
        return self._ck(self._lib.rs_host_tex_perlin(self._h, point_count, int(vector), seed, kind, depth, smooth, scale), "tex_perlin")

    def lambertian(self, tex):
        return self._ck(self._lib.rs_host_mat_lambertian(self._h, tex), "mat_lambertian")

    def metal(self, tex):
        return self._ck(self._lib.rs_host_mat_metal(self._h, tex), "mat_metal")

    def diffuse_metal(self, exponent, tex):
        return self._ck(self._lib.rs_host_mat_diffuse_metal(self._h, exponent, tex), "mat_diffuse_metal")

    def dielectric(self, color, refractive, glass_curve=True):
        return self._ck(self._lib.rs_host_mat_dielectric(self._h, color[0], color[1], color[2], refractive, int(glass_curve)), "mat_dielectric")

    def diffuse_light(self, tex, multiplier=1.0):
        return self._ck(self._lib.rs_host_mat_diffuse_light(self._h, tex, multiplier), "mat_diffuse_light")

    def isotropic(self, color):
        return self._ck(self._lib.rs_host_mat_isotropic(self._h, color[0], color[1], color[2]), "mat_isotropic")

    def blinn_phong(self, k_specular, exponent, tex):
        return self._ck(self._lib.rs_host_mat_blinn_phong(self._h, k_specular, exponent, tex), "mat_blinn_phong")

    def mixed(self, mat1, mat2, probability_1):
        return self._ck(self._lib.rs_host_mat_mixed(self._h, mat1, mat2, probability_1), "mat_mixed")

    def set_phong(self, mat, phong_factor, phong_exponent):
        return self._ck(self._lib.rs_host_mat_set_phong(self._h, mat, phong_factor, phong_exponent), "mat_set_phong")

    def sphere(self, center, radius, mat=-1, speed=None):
        sp = _vec3(speed) if speed is not None else None
        return self._ck(self._lib.rs_host_obj_sphere(self._h, _vec3(center), radius, sp, mat), "obj_sphere")

    def box(self, p0, p1, mat=-1):
        return self._ck(self._lib.rs_host_obj_box(self._h, _vec3(p0), _vec3(p1), mat), "obj_box")

    def rect(self, axis: str, k, a, b, mat=-1):
        kind = {"xy": 0, "xz": 1, "yz": 2}[axis]
        return self._ck(self._lib.rs_host_obj_rect(self._h, kind, k, a[0], a[1], b[0], b[1], mat), "obj_rect")

    def raymarcher(self, mat=-1):
        return self._ck(self._lib.rs_host_obj_raymarcher(self._h, mat), "obj_raymarcher")

    def quadric(self, q, mat=-1):
        return self._ck(self._lib.rs_host_obj_quadric(self._h, (C.c_double * 10)(*[float(x) for x in q]), mat), "obj_quadric")

    def difference(self, plus, minus, mat=-1):
        return self._ck(self._lib.rs_host_obj_difference(self._h, plus, minus, mat), "obj_difference")

    def intersection(self, o1, o2, mat=-1):
        return self._ck(self._lib.rs_host_obj_intersection(self._h, o1, o2, mat), "obj_intersection")

    def bvh(self, members):
        """`BVH::new(list)` used as ONE object (bvh.rs:173-192): a world member, under a transform, a medium boundary."""
        n = len(members)
        return self._ck(self._lib.rs_host_obj_bvh(self._h, n, (C.c_int32 * max(n, 1))(*[int(m) for m in members])), "obj_bvh")

    def transform(self, child, ops):
        """ops: list of ("translate", v3) | ("rotate_x", rad) | ("rotate_y", rad) | ("rotate_z", rad) | ("scale", v3)"""
        codes = {"translate": 0, "rotate_x": 1, "rotate_y": 2, "rotate_z": 3, "scale": 4}
        n = len(ops)
        kinds = (C.c_int32 * n)(*[codes[o[0]] for o in ops])
        args = []
        for name, val in ops:
            if name in ("translate", "scale"):
                args += [float(val[0]), float(val[1]), float(val[2])]
            else:
                args += [float(val), 0.0, 0.0]
        return self._ck(self._lib.rs_host_obj_transform(self._h, child, n, kinds, (C.c_double * (3 * n))(*args)), "obj_transform")

    def medium(self, boundary, color, density):
        return self._ck(self._lib.rs_host_obj_medium(self._h, boundary, color[0], color[1], color[2], density), "obj_medium")

    def add(self, obj):
        return self._ck(self._lib.rs_host_world_add(self._h, obj), "world_add")

    def add_light(self, obj):
        return self._ck(self._lib.rs_host_world_add_light(self._h, obj), "world_add_light")

    def add_mesh(self, positions, indices, normals=None, scale=1.0, offset=(0, 0, 0), rotation_angle=0.0, axis=1, mat=-1):
        pos = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1)
        idx = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        nrm = np.ascontiguousarray(normals, dtype=np.float32).reshape(-1) if normals is not None else None
        return self._ck(self._lib.rs_host_world_add_mesh(self._h, pos.ctypes.data, pos.size, nrm.ctypes.data if nrm is not None else None,
                                                         nrm.size if nrm is not None else 0, idx.ctypes.data, idx.size, scale, _vec3(offset),
                                                         rotation_angle, axis, mat), "world_add_mesh")

    def add_obj_file(self, path, scale=1.0, offset=(0, 0, 0), rotation_angle=0.0, axis=1, mat=-1):
        return self._ck(self._lib.rs_host_world_add_obj_file(self._h, path.encode(), scale, _vec3(offset), rotation_angle, axis, mat), "world_add_obj_file")

    def set_background(self, c0, c1):
        a = (C.c_float * 4)(*[float(x) for x in c0])
        b = (C.c_float * 4)(*[float(x) for x in c1])
        return self._ck(self._lib.rs_host_world_set_background(self._h, a, b), "set_background")

    def set_time_range(self, t0, t1):
        return self._ck(self._lib.rs_host_world_set_time_range(self._h, t0, t1), "set_time_range")

    def set_camera(self, look_from, look_at, vup=(0, 1, 0), fov=90.0, aperture=0.0, focus=1.0, shutter_speed=0.0):
        return self._ck(self._lib.rs_host_scene_set_camera(self._h, _vec3(look_from), _vec3(look_at), _vec3(vup), fov, aperture, focus, shutter_speed), "set_camera")

    # ---- flat description -------------------------------------------------------------------
    @property
    def desc(self):
        """POINTER(RsSceneDesc), valid while this Scene is alive and unmodified."""
        if self._desc is None:
            d = self._lib.rs_host_scene_desc(self._h)
            if not d:
                raise RaysnailError(f"flatten failed: {self._lib.rs_host_last_error().decode()}")
            self._desc = d
        return self._desc

    def camera(self, width: int, height: int) -> RsCamera:
        cam = RsCamera()
        rc = self._lib.rs_host_scene_camera(self._h, width, height, C.byref(cam))
        if rc != 0:
            raise RaysnailError(self._lib.rs_host_last_error().decode())
        return cam

    def root_bbox(self) -> np.ndarray:
        d = self.desc.contents
        return np.ctypeslib.as_array(d.root_bbox, shape=(d.n_roots, 6)).copy() if d.n_roots else np.zeros((0, 6))


class SdlParser:
    """`SdlParser::parse` (src/sdl_parser.rs:187-205) + the CLI's scene assembly."""

    @staticmethod
    def parse(filename: str) -> Scene:
        return Scene.from_sdl(filename)


class CameraBuilder:
    """`CameraBuilder` (src/camera.rs:299-414)."""

    def __init__(self):
        self._from, self._at, self._vup = (0.0, 0.0, 0.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0)
        self._fov, self._aperture, self._focus, self._shutter = 90.0, 0.0, 1.0, 0.0
        self._w, self._h = 400, 200

    def look_from(self, p): self._from = tuple(p); return self
    def look_at(self, p): self._at = tuple(p); return self
    def vup(self, v): self._vup = tuple(v); return self
    def fov(self, f): self._fov = float(f); return self
    def aperture(self, a): self._aperture = float(a); return self
    def focus(self, d): self._focus = float(d); return self
    def shutter_speed(self, d): self._shutter = float(d); return self
    def width(self, w): self._w = int(w); return self
    def height(self, h): self._h = int(h); return self

    def build(self) -> RsCamera:
        lib = _abi.load()
        cam = RsCamera()
        rc = lib.rs_host_camera_build(_vec3(self._from), _vec3(self._at), _vec3(self._vup), self._fov, self._aperture, self._focus,
                                      self._shutter, self._w, self._h, C.byref(cam))
        if rc != 0:
            raise RaysnailError(lib.rs_host_last_error().decode())
        return cam


def make_params(samples: int = 50, depth: int = 8, gamma: bool = True, seed: int = DEFAULT_SEED, flags: int = 0,
                rows: Optional[Sequence[int]] = None, max_wave_paths: int = 0, sqrt_spp: Optional[int] = None) -> RsRenderParams:
    q = sqrt_spp if sqrt_spp is not None else painter_samples(samples)
    p = RsRenderParams()
    p.sqrt_spp = q
    p.depth = depth
    p.gamma = 1 if gamma else 0
    p.stratum_row_begin, p.stratum_row_end = (rows if rows is not None else (0, q))
    p.max_wave_paths = max_wave_paths
    p.seed = seed
    p.flags = flags
    return p


class Renderer:
    """One CUDA device context (`rs_ctx`).  Raises when no sm_100 GPU is present."""

    def __init__(self, device: int = 0):
        self._lib = _abi.load()
        self._ctx = C.c_void_p()
        rc = self._lib.rs_create(C.byref(self._ctx), device)
        if rc != 0:
            raise RaysnailError(f"rs_create failed ({rc}): {self._lib.rs_last_error(None).decode()}")
        self._scene = None

    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            self._lib.rs_destroy(self._ctx)
            self._ctx = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise RaysnailError(f"{what} failed ({rc}): {self._lib.rs_last_error(self._ctx).decode()}")

    def set_bvh_width(self, width: int):
        """2 = binary world BVH (default), 4 = QBVH; takes effect at the next upload."""
        self._ck(self._lib.rs_set_option(self._ctx, 1, int(width)), "rs_set_option")
        if self._scene is not None:
            self.upload(self._scene)

    def set_list_max(self, n: int):
        """RS_OPT_LIST_MAX: worlds of at most n objects are searched as a list, not a tree (0 = always the tree)."""
        self._ck(self._lib.rs_set_option(self._ctx, 3, int(n)), "rs_set_option")
        if self._scene is not None:
            self.upload(self._scene)

    def set_wave_streams(self, n: int):
        """RS_OPT_WAVE_STREAMS: n in 1..4 (default 4) = consecutive waves rotate over n streams, 1 = one stream."""
        self._ck(self._lib.rs_set_option(self._ctx, 2, int(n)), "rs_set_option")

    def set_wait_budget_ms(self, ms: int):
        """RS_OPT_WAIT_BUDGET_MS: how long rs_wait_counter's kernel polls before it gives up (0 = for ever; default 10 min)."""
        self._ck(self._lib.rs_set_option(self._ctx, 4, int(ms)), "rs_set_option")

    def upload(self, scene: Scene):
        self._ck(self._lib.rs_upload_scene(self._ctx, scene.desc), "rs_upload_scene")
        self._scene = scene  # keep the host buffers alive

    def render(self, cam: RsCamera, params: RsRenderParams, mask: Optional[np.ndarray] = None, out: Optional[np.ndarray] = None) -> np.ndarray:
        """`shot_to_target`: returns the (H, W, 4) f32 image."""
        h, w = cam.height, cam.width
        if out is None:
            out = np.empty((h, w, 4), dtype=np.float32)
        mptr = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            mptr = mask.ctypes.data
        self._ck(self._lib.rs_render(self._ctx, C.byref(cam), C.byref(params), mptr, out.ctypes.data), "rs_render")
        return out

    def render_to_ptr(self, cam, params, mask_ptr, out_ptr):
        self._ck(self._lib.rs_render(self._ctx, C.byref(cam), C.byref(params), mask_ptr, out_ptr), "rs_render")

    def render_accumulate(self, cam, params, d_mask: Optional[int], d_accum: int, stream: int = 0):
        self._ck(self._lib.rs_render_accumulate(self._ctx, C.byref(cam), C.byref(params), d_mask, d_accum, stream), "rs_render_accumulate")

    def finalize(self, d_accum: int, d_mask: Optional[int], d_out: int, width: int, height: int, n_samples: int, gamma: bool, stream: int = 0):
        self._ck(self._lib.rs_finalize(self._ctx, d_accum, d_mask, d_out, width, height, n_samples, int(gamma), stream), "rs_finalize")

    # ---- multi-GPU: peer-memory buffers and the fused reduce + finalize -----------------------
    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._ck(self._lib.rs_device_alloc(self._ctx, int(nbytes), C.byref(p)), "rs_device_alloc")
        return int(p.value)

    def device_free(self, d_ptr: int):
        self._ck(self._lib.rs_device_free(self._ctx, d_ptr), "rs_device_free")

    def ipc_export(self, d_ptr: int) -> bytes:
        h = (C.c_uint8 * 64)()
        self._ck(self._lib.rs_ipc_export(self._ctx, d_ptr, h), "rs_ipc_export")
        return bytes(h)

    def ipc_open(self, handle: bytes) -> int:
        h = (C.c_uint8 * 64).from_buffer_copy(handle)
        p = C.c_void_p()
        self._ck(self._lib.rs_ipc_open(self._ctx, h, C.byref(p)), "rs_ipc_open")
        return int(p.value)

    def ipc_close(self, d_ptr: int):
        self._ck(self._lib.rs_ipc_close(self._ctx, d_ptr), "rs_ipc_close")

    def finalize_peers(self, d_accums, d_mask: Optional[int], d_out: int, pixel_begin: int, pixel_end: int, n_samples: int, gamma: bool,
                       stream: int = 0):
        arr = (C.c_void_p * len(d_accums))(*[int(a) for a in d_accums])
        self._ck(self._lib.rs_finalize_peers(self._ctx, arr, len(d_accums), d_mask, d_out, int(pixel_begin), int(pixel_end), int(n_samples),
                                             int(gamma), stream), "rs_finalize_peers")

    def signal(self, d_counter: int, stream: int = 0):
        """rs_signal: counter += 1 (release, system scope) once the work queued on `stream` so far is done."""
        self._ck(self._lib.rs_signal(self._ctx, int(d_counter), stream), "rs_signal")

    def wait_counter(self, d_counter: int, expect: int, stream: int = 0):
        """rs_wait_counter: work queued on `stream` after this starts once the counter has reached `expect`."""
        self._ck(self._lib.rs_wait_counter(self._ctx, int(d_counter), int(expect) & 0xffffffff, stream), "rs_wait_counter")

    def wait_check(self, stream: int = 0):
        """rs_wait_check: synchronise `stream`; raises if a wait of this context gave up (a peer never signalled)."""
        self._ck(self._lib.rs_wait_check(self._ctx, stream), "rs_wait_check")

    def render_passes(self, cam: RsCamera, params: RsRenderParams, n_passes: int, redo_mode: int = 0, threshold: float = 0.01,
                      want_rgb8: bool = False):
        """The CLI's passes loop (src/bin/raysnail.rs:379-427) in one call: (image, redo count per pass[, rgb8])."""
        h, w = cam.height, cam.width
        out = np.empty((h, w, 4), dtype=np.float32)
        rgb = np.empty((h, w, 3), dtype=np.uint8) if want_rgb8 else None
        counts = np.zeros(max(n_passes, 1), dtype=np.int64)
        self._ck(self._lib.rs_render_passes(self._ctx, C.byref(cam), C.byref(params), int(n_passes), int(redo_mode), float(threshold),
                                            out.ctypes.data, rgb.ctypes.data if want_rgb8 else None, counts.ctypes.data), "rs_render_passes")
        return (out, counts[:n_passes], rgb) if want_rgb8 else (out, counts[:n_passes])

    def combine_passes(self, d_old, d_new, d_out, n_pixels, pass_index: float, stream: int = 0):
        self._ck(self._lib.rs_combine_passes(self._ctx, d_old, d_new, d_out, n_pixels, pass_index, stream), "rs_combine_passes")

    def noise_map(self, d_pixels, d_noise, width, height, quirk=True, stream: int = 0):
        self._ck(self._lib.rs_noise_map(self._ctx, d_pixels, d_noise, width, height, int(quirk), stream), "rs_noise_map")

    def redo_map_static(self, d_pixels, d_redo_mask, d_redo_count, width, height, quirk=True, threshold: float = 0.01, stream: int = 0):
        """rs_redo_map_static: noise map -> redo mask + count, for hosts that cannot let the library allocate on the
        device (their own pool owns the card, or a cudaMalloc would stall or dead-lock what is running).  The three
        arguments are device addresses the caller owns; the noise map in between sits in a fixed 16 MiB array inside
        the library (at most RS_STATIC_MAX_PIXELS = 4 Mi pixels; larger frames raise).  Calls that overlap on one GPU
        share that array and corrupt each other: keep them on one stream."""
        self._ck(self._lib.rs_redo_map_static(self._ctx, d_pixels, d_redo_mask, d_redo_count, int(width), int(height), int(quirk),
                                              float(threshold), stream), "rs_redo_map_static")

    def quantize_rgb8(self, d_pixels, d_rgb, n_pixels, stream: int = 0):
        self._ck(self._lib.rs_quantize_rgb8(self._ctx, d_pixels, d_rgb, n_pixels, stream), "rs_quantize_rgb8")

    def debug_primary(self, cam, params, sample_index=0):
        n = cam.width * cam.height
        o = dict(obj_id=np.empty(n, np.int32), t=np.empty(n), normal=np.empty((n, 3)), point=np.empty((n, 3)),
                 front=np.empty(n, np.int32), material=np.empty(n, np.int32), ray=np.empty((n, 6)))
        self._ck(self._lib.rs_debug_primary(self._ctx, C.byref(cam), C.byref(params), sample_index, o["obj_id"].ctypes.data, o["t"].ctypes.data,
                                            o["normal"].ctypes.data, o["point"].ctypes.data, o["front"].ctypes.data, o["material"].ctypes.data,
                                            o["ray"].ctypes.data), "rs_debug_primary")
        return o

    def debug_scatter(self, cam, params, sample_index=0):
        n = cam.width * cam.height
        o = dict(kind=np.empty(n, np.int32), next=np.empty((n, 6)), weight=np.empty((n, 3)), emitted=np.empty((n, 3)))
        self._ck(self._lib.rs_debug_scatter(self._ctx, C.byref(cam), C.byref(params), sample_index, o["kind"].ctypes.data, o["next"].ctypes.data,
                                            o["weight"].ctypes.data, o["emitted"].ctypes.data), "rs_debug_scatter")
        return o

    def stats(self) -> RsStats:
        s = RsStats()
        self._ck(self._lib.rs_get_stats(self._ctx, C.byref(s)), "rs_get_stats")
        return s



class MultiRenderer:
    """Several CUDA devices behind one call (`rs_multi`): the scene is replicated, the strata rows are
    split over the devices, the image is combined on device 0 over NVLink peer memory."""

    def __init__(self, devices: Sequence[int]):
        self._lib = _abi.load()
        self._m = C.c_void_p()
        ids = (C.c_int32 * len(devices))(*[int(d) for d in devices])
        rc = self._lib.rs_create_multi(C.byref(self._m), ids, len(devices))
        if rc != 0:
            raise RaysnailError(f"rs_create_multi failed ({rc}): {self._lib.rs_multi_last_error(None).decode()}")
        self._scene = None
        self.stats = RsStats()

    def close(self):
        if getattr(self, "_m", None) and self._m.value:
            self._lib.rs_destroy_multi(self._m)
            self._m = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise RaysnailError(f"{what} failed ({rc}): {self._lib.rs_multi_last_error(self._m).decode()}")

    @property
    def device_count(self) -> int:
        return int(self._lib.rs_multi_device_count(self._m))

    def set_option(self, option: int, value: int):
        self._ck(self._lib.rs_multi_set_option(self._m, int(option), int(value)), "rs_multi_set_option")

    def upload(self, scene: Scene):
        self._ck(self._lib.rs_multi_upload_scene(self._m, scene.desc), "rs_multi_upload_scene")
        self._scene = scene

    def render(self, cam: RsCamera, params: RsRenderParams, mask: Optional[np.ndarray] = None, out: Optional[np.ndarray] = None) -> np.ndarray:
        h, w = cam.height, cam.width
        if out is None:
            out = np.empty((h, w, 4), dtype=np.float32)
        mptr = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.uint8)
            mptr = mask.ctypes.data
        self._ck(self._lib.rs_multi_render(self._m, C.byref(cam), C.byref(params), mptr, out.ctypes.data, C.byref(self.stats)), "rs_multi_render")
        return out

    def render_to_ptr(self, cam, params, mask_ptr, out_ptr):
        self._ck(self._lib.rs_multi_render(self._m, C.byref(cam), C.byref(params), mask_ptr, out_ptr, C.byref(self.stats)), "rs_multi_render")

class TakePhotoSettings:
    """`TakePhotoSettings` (src/camera.rs:103-154): camera.take_photo().samples(n).depth(d).shot(world)."""

    def __init__(self, cam: RsCamera, renderer: Renderer):
        self._cam, self._r = cam, renderer
        self._depth, self._gamma, self._samples, self._seed = 8, True, 50, DEFAULT_SEED

    def depth(self, d): self._depth = int(d); return self
    def gamma(self, g): self._gamma = bool(g); return self
    def samples(self, s): self._samples = int(s); return self
    def seed(self, s): self._seed = int(s); return self
    def threads(self, _n): return self    # CPU knobs of the reference: accepted, meaningless on the GPU
    def parallel(self, _p): return self

    def shot_to_target(self, world: Optional[Scene] = None, target=None, pixel_map: Optional[np.ndarray] = None) -> np.ndarray:
        """camera.rs:261-287.  `target.register_pixels(row, pixels)` is called for rows 0..H-1 and then
        once with (height, []) as the reference's end-of-pass marker (painter.rs:214,332)."""
        if world is not None:
            self._r.upload(world)
        params = make_params(self._samples, self._depth, self._gamma, self._seed)
        img = self._r.render(self._cam, params, mask=pixel_map)
        if target is not None:
            for row in range(self._cam.height):
                target.register_pixels(row, img[row])
            target.register_pixels(self._cam.height, img[0:0])
        return img

    def shot(self, world: Optional[Scene] = None) -> np.ndarray:
        return self.shot_to_target(world)


def take_photo(cam: RsCamera, renderer: Renderer) -> TakePhotoSettings:
    return TakePhotoSettings(cam, renderer)


def pass_seed(seed: int, pass_index: int) -> int:
    """Philox key of pass k of the passes loop (`rs_pass_seed`)."""
    return int(_abi.load().rs_pass_seed(C.c_uint64(seed & 0xFFFFFFFFFFFFFFFF), int(pass_index)))


def philox4x32_10(counter: Sequence[int], key: Sequence[int]) -> np.ndarray:
    """`rs_philox4x32_10`: the kernels' counter-based RNG, evaluated on the host (for harnesses)."""
    c = np.asarray(counter, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    rc = _abi.load().rs_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
    if rc != 0:
        raise RaysnailError(f"rs_philox4x32_10 failed ({rc})")
    return out


def philox_uniform(seed: int, pixel: int, sample: int, depth: int, purpose: int, call: int = 0) -> np.ndarray:
    """`rs_philox_uniform`: the four uniforms of one draw site of a path."""
    out = np.zeros(4, dtype=np.float64)
    rc = _abi.load().rs_philox_uniform(C.c_uint64(seed), pixel, sample, depth, purpose, call, out.ctypes.data)
    if rc != 0:
        raise RaysnailError(f"rs_philox_uniform failed ({rc})")
    return out
