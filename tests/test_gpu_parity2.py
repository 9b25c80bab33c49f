"""GPU parity, round 2: the branches and configurations the first round's scenes never reached
(VERDICT r01 "untested configs"): several lights of mixed kind (list.rs:49-52), every Perlin variant
(noise.rs), Dielectric without the Glass curve, nested MixedMaterial, a DiffuseLight under a
MixedMaterial, degenerate PDF values (camera.rs:236-243), the opt-in MIS integrator, the passes
driver, BASELINE config 4 at its full 1 M triangles, and the reference's own render of
sdl/quadric.sdl.  Same bars as test_gpu_parity.py.
"""
import ctypes as C
import math
import os

import numpy as np
import pytest

from conftest import load_scene
from oracle_binding import OracleWorld
from raysnail_b200 import (RS_FLAG_MIS, RS_FLAG_PIXEL_CENTRE, RS_REDO_NOISE, RS_REDO_NOISE_FIXED, RS_REDO_REFERENCE, CameraBuilder, RsCamera,
                           Scene, make_params, pass_seed)

pytestmark = pytest.mark.gpu

W, H = 192, 120
REL = 1e-5


def _rel_err(a, b, floor=1e-12):
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)


def _luminance(img):
    return 0.2126 * img[..., 0] + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]


def feature_scene(tier):
    """tier "spheres": static spheres only, three sphere lights (lean kernels); "flat": + an xz-rect light, boxes
    (flat tier); "generic": + a sphere light under a TfFacade and a CSG object (expression trees everywhere)."""
    s = Scene()
    grey = s.lambertian(s.solid(0.55, 0.55, 0.55))
    s.add(s.sphere((0, -1000, 0), 1000.0, s.lambertian(s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.9, 0.9), 3.0))))
    lights = [s.sphere((-6, 7, 3), 1.5, s.diffuse_light(s.solid(1.0, 0.8, 0.6), 6.0)),
              s.sphere((7, 5, -2), 1.0, s.diffuse_light(s.solid(0.6, 0.8, 1.0), 9.0))]
    if tier == "spheres":
        lights.append(s.sphere((0, 9, -6), 2.0, s.diffuse_light(s.solid(1, 1, 1), 3.0)))
    else:
        lights.append(s.rect("xz", 8.0, (-2.0, 2.0), (-1.0, 3.0), s.diffuse_light(s.solid(1, 1, 1), 4.0)))
    if tier == "generic":
        lights.append(s.transform(s.sphere((0, 0, 0), 1.0, s.diffuse_light(s.solid(0.9, 1.0, 0.7), 5.0)),
                                  [("scale", (1.0, 0.5, 2.0)), ("rotate_y", 0.6), ("translate", (3.0, 6.0, 5.0))]))
    for l in lights:
        s.add(l)
        s.add_light(l)
    # Perlin: Normal / turbulence(7) / marble(7), the three smooth types, both lattices (noise.rs:100-210)
    variants = [(True, 1, 7, 2, 1.0), (True, 2, 7, 2, 4.0), (True, 0, 0, 0, 2.0), (False, 0, 0, 1, 3.0), (False, 0, 0, 2, 3.0),
                (False, 1, 7, 1, 1.0), (False, 2, 7, 0, 2.0), (True, 0, 0, 1, 2.5)]
    for k, (vector, kind, depth, smooth, scale) in enumerate(variants):
        tex = s.perlin(64 if k % 2 else 256, vector, 100 + k, kind, depth, smooth, scale)
        s.add(s.sphere((-5.25 + 1.5 * k, 0.6, 0.8 * math.sin(k)), 0.6, s.lambertian(tex)))
    # Dielectric without / with the Glass curve (dielectric.rs:17-25, 68-74), nested mixes (mixed_material.rs:43-58)
    s.add(s.sphere((-2.0, 0.7, 2.6), 0.7, s.dielectric((1.0, 1.0, 1.0), 1.5, glass_curve=False)))
    s.add(s.sphere((-0.4, 0.7, 2.9), 0.7, s.dielectric((0.9, 1.0, 0.9), 1.3, glass_curve=True)))
    a = s.lambertian(s.solid(0.8, 0.2, 0.2))
    s.set_phong(a, 2.0, 12)
    metal = s.metal(s.solid(0.9, 0.9, 0.9))
    nested = s.mixed(s.mixed(a, metal, 0.6), s.diffuse_metal(40.0, s.solid(0.3, 0.4, 0.9)), 0.5)
    s.add(s.sphere((1.3, 0.7, 2.7), 0.7, nested))
    glow = s.mixed(s.diffuse_light(s.solid(1, 1, 1), 20.0), grey, 0.5)  # emits nothing (mixed_material.rs:53-55)
    s.add(s.sphere((3.0, 0.7, 2.4), 0.7, glow))
    s.add(s.sphere((4.6, 0.7, 2.0), 0.7, s.blinn_phong(0.3, 12.0, s.solid(0.2, 0.7, 0.4))))
    if tier != "spheres":
        s.add(s.box((-4.5, 0.0, -3.5), (-3.0, 1.5, -2.0), a))
        s.add(s.transform(s.box((-0.5, 0.0, -0.5), (0.5, 2.0, 0.5), s.mixed(metal, grey, 0.5)), [("rotate_y", 0.5), ("translate", (3.5, 0.0, -3.0))]))
    if tier == "generic":
        s.add(s.transform(s.difference(s.box((-1, -1, -1), (1, 1, 1), nested), s.sphere((0, 0, 0), 1.25, grey)),
                          [("rotate_x", 0.4), ("translate", (0.0, 1.2, -3.0))]))
    s.set_background((0.3, 0.4, 0.5), (0.7, 0.89, 1.0))
    s.set_camera((0.5, 4.5, 13.0), (0.0, 0.8, 0.0), (0, 1, 0), fov=42.0, aperture=0.05, focus=12.0)
    return s


def check_primary(g, o, name):
    mism = np.flatnonzero(g["obj_id"] != o["obj_id"])
    assert mism.size == 0, f"{name}: {mism.size} primary-hit ids differ, first {mism[:5]}: gpu {g['obj_id'][mism[:5]]} oracle {o['obj_id'][mism[:5]]}"
    hit = o["obj_id"] >= 0
    assert hit.any()
    assert _rel_err(g["t"][hit], o["t"][hit]).max() < REL
    assert np.abs(g["normal"][hit] - o["normal"][hit]).max() < REL * max(1.0, np.abs(o["normal"][hit]).max())
    assert np.abs(g["point"][hit] - o["point"][hit]).max() < REL * max(1.0, np.abs(o["point"][hit]).max())
    for key in ("front", "material"):
        assert np.array_equal(g[key][hit], o[key][hit]), f"{name}: {key} differs"


def check_scatter(g, o, name, w_tol=1e-4):
    mism = np.flatnonzero(g["kind"] != o["kind"])
    assert mism.size == 0, f"{name}: scatter kind differs at {mism[:5]}: gpu {g['kind'][mism[:5]]} oracle {o['kind'][mism[:5]]}"
    sc = o["kind"] >= 2
    if sc.any():
        assert np.abs(g["next"][sc][:, 3:] - o["next"][sc][:, 3:]).max() < REL, f"{name}: scatter direction"
        assert np.abs(g["next"][sc][:, :3] - o["next"][sc][:, :3]).max() < REL * max(1.0, np.abs(o["next"][sc][:, :3]).max())
        gw, ow = g["weight"][sc], o["weight"][sc]
        assert np.array_equal(np.isnan(gw), np.isnan(ow)), f"{name}: NaN weights differ"
        ok = ~np.isnan(ow)
        assert _rel_err(gw[ok], ow[ok], floor=1e-6).max() < w_tol, f"{name}: weight"
    assert np.abs(g["emitted"] - o["emitted"]).max() < 1e-5 * max(1.0, np.abs(o["emitted"]).max())


def check_image(g, o, name):
    rmse = float(np.sqrt(np.mean((g[..., :3] - o[..., :3]) ** 2))) / max(float(np.sqrt(np.mean(o[..., :3] ** 2))), 1e-3)
    lum_g, lum_o = _luminance(g).mean(), _luminance(o).mean()
    assert rmse < 0.01, f"{name}: relative RMSE {rmse}"
    assert abs(lum_g - lum_o) <= 0.005 * lum_o, f"{name}: mean luminance {lum_g} vs {lum_o}"


@pytest.mark.parametrize("tier", ["spheres", "flat", "generic"])
def test_feature_scene_parity(renderer, tier):
    scene = feature_scene(tier)
    renderer.upload(scene)
    oracle = OracleWorld(scene)
    cam = scene.camera(W, H)
    n_lights = scene.desc.contents.n_lights
    assert n_lights == (4 if tier == "generic" else 3)
    for flags in (RS_FLAG_PIXEL_CENTRE, 0):
        params = make_params(samples=9, flags=flags)
        for sample in (0, 7):
            check_primary(renderer.debug_primary(cam, params, sample), oracle.debug_primary(cam, params, sample), tier)
    params = make_params(samples=9)
    picked = set()
    for sample in (0, 3, 5):
        g, o = renderer.debug_scatter(cam, params, sample), oracle.debug_scatter(cam, params, sample)
        check_scatter(g, o, tier)
        # which light did the light-directed rays aim at?  (list.rs:49-52: next_u32 % n) — every light must be used
        light_rays = o["next"][o["kind"] == 3]
        picked.add(len(light_rays))
    assert min(picked) > 100
    # the light pick: rays of kind 3 must point at every one of the lights somewhere in the image
    d = scene.desc.contents
    o = oracle.debug_scatter(cam, params, 0)
    rays = o["next"][o["kind"] == 3]
    centres = [(-6, 7, 3), (7, 5, -2), (0, 9, -6) if tier == "spheres" else (0, 8, 1)] + ([(3, 6, 5)] if tier == "generic" else [])
    for c in centres:
        to_c = np.asarray(c, dtype=np.float64)[None, :] - rays[:, :3]
        cosang = (to_c * rays[:, 3:]).sum(1) / np.linalg.norm(to_c, axis=1)
        assert (np.abs(cosang) > 0.97).sum() > 20, f"no light-directed ray aims at (or, AARect::random: away from) the light at {c}"
    gi = renderer.render(scene.camera(72, 45), make_params(samples=1024))
    oi, _ = oracle.render(scene.camera(72, 45), make_params(samples=1024))
    check_image(gi, oi, tier)
    assert np.isfinite(gi).all()


def test_degenerate_pdf_values_parity(renderer):
    """camera.rs:236-243 on the BRDF branch: value 0 -> weight 0; value < 0 -> value / 1e-5; NaN -> NaN
    (see tests/test_oracle_kat2.py::test_degenerate_pdf_values_follow_camera_rs_236_243 for the arithmetic)."""
    s = Scene()
    lamp = s.sphere((0, 50, 0), 5.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
    s.add(lamp)
    s.add_light(lamp)
    for k, ks in enumerate((1.0, 3.0, float("nan"))):
        s.add(s.sphere((-2.5 + 2.5 * k, 0, 0), 1.0, s.blinn_phong(ks, -1.0, s.solid(0.5, 0.6, 0.7))))
    s.add(s.sphere((0, -1001, 0), 1000.0, s.diffuse_metal(0.0, s.solid(0.8, 0.8, 0.8))))  # ReflectionPdf with exponent 0
    s.set_background((0.3, 0.4, 0.5), (0.7, 0.89, 1.0))
    s.set_camera((0, 2, 9), (0, 0, 0), (0, 1, 0), fov=40.0, aperture=0.0, focus=9.0)
    renderer.upload(s)
    oracle = OracleWorld(s)
    cam = s.camera(W, H)
    params = make_params(samples=9)
    seen = {"zero": 0, "negative": 0, "nan": 0}
    for sample in (0, 4, 8):
        g, o = renderer.debug_scatter(cam, params, sample), oracle.debug_scatter(cam, params, sample)
        check_scatter(g, o, "degenerate")
        brdf = o["kind"] == 4
        seen["zero"] += int((brdf & (o["weight"] == 0.0).all(axis=1)).sum())
        seen["negative"] += int((brdf & (o["weight"] < -1.0).all(axis=1)).sum())
        seen["nan"] += int((brdf & np.isnan(o["weight"]).all(axis=1)).sum())
    assert min(seen.values()) > 100, seen


@pytest.mark.parametrize("name", ["cornell", "rtow_final", "mesh_blinn", "csg"])
def test_mis_integrator_parity(renderer, name):
    """RS_FLAG_MIS: the MixturePdf(HittablePdf, srec.pdf) integrator the reference keeps commented out
    (camera.rs:188-192, 222-233; pdf.rs:237-307), GPU vs oracle on all three shade tiers."""
    scene = load_scene(name)
    renderer.upload(scene)
    oracle = OracleWorld(scene)
    cam = scene.camera(W, H)
    params = make_params(samples=9, flags=RS_FLAG_MIS)
    for sample in (0, 5):
        check_scatter(renderer.debug_scatter(cam, params, sample), oracle.debug_scatter(cam, params, sample), name + "/mis")
    size, spp = ((48, 48), 1024) if name != "mesh_blinn" else ((64, 64), 256)
    cam = scene.camera(*size)
    g = renderer.render(cam, make_params(samples=spp, flags=RS_FLAG_MIS))
    o, _ = oracle.render(cam, make_params(samples=spp, flags=RS_FLAG_MIS))
    fin = np.isfinite(o).all(axis=-1) & np.isfinite(g).all(axis=-1)
    assert fin.mean() > 0.999
    check_image(g[fin], o[fin], name + "/mis")
    plain = renderer.render(cam, make_params(samples=spp))
    assert not np.array_equal(plain, g)


def test_render_passes_driver(renderer):
    """rs_render_passes == the chain rs_render -> rs_combine_passes -> rs_noise_map -> threshold, pass by pass
    (bit for bit), in all three redo modes; and it tracks the oracle's passes loop (raysnail.rs:379-427)."""
    import torch
    from oracle_binding import load as load_oracle
    ol = load_oracle()
    scene = load_scene("example")
    renderer.upload(scene)
    oracle = OracleWorld(scene)
    w, h = 80, 50
    cam = scene.camera(w, h)
    params = make_params(samples=4)
    n_passes = 3
    for mode, thr in ((RS_REDO_REFERENCE, 0.01), (RS_REDO_NOISE, 0.05), (RS_REDO_NOISE_FIXED, 0.05)):
        img, counts, rgb = renderer.render_passes(cam, params, n_passes, mode, thr, want_rgb8=True)
        # the chain, with the bit-exact CPU twins of the combine / noise kernels between the GPU renders
        old = np.zeros((h, w, 4), np.float32)
        old[..., 3] = 1.0
        mask, exp_counts, per_pass = None, [], []
        for k in range(n_passes):
            new = renderer.render(cam, make_params(samples=4, seed=pass_seed(params.seed, k)), mask=mask)
            per_pass.append(new)
            out = np.empty_like(old)
            ol.rso_combine_passes(old.ctypes.data, new.ctypes.data, out.ctypes.data, w * h, C.c_float(float(k)))
            noise = np.empty((h, w), np.float32)
            ol.rso_noise_map(out.ctypes.data, noise.ctypes.data, w, h, 0 if mode == RS_REDO_NOISE_FIXED else 1)
            redo = (noise >= np.float32(thr)).astype(np.uint8)
            exp_counts.append(int(redo.sum()))
            mask = redo if mode != RS_REDO_REFERENCE else None
            old = out
        assert np.array_equal(img, old), f"mode {mode}: driver image differs from the chained calls"
        assert list(counts) == exp_counts
        exp_rgb = np.empty((h, w, 3), np.uint8)
        ol.rso_quantize_rgb8(old.ctypes.data, exp_rgb.ctypes.data, w * h)
        assert np.array_equal(rgb, exp_rgb)
        assert not np.array_equal(per_pass[0], per_pass[1])  # a fresh Philox key per pass
        # oracle twin: same loop on the CPU; redo maps may differ in a few borderline pixels, images agree closely
        oi, ocounts = oracle.render_passes(cam, params, n_passes, mode, thr)
        assert np.abs(img - oi).mean() < 2e-4
        assert np.abs(np.asarray(counts) - ocounts).max() <= max(4, 0.01 * w * h)
        st = renderer.stats()
        assert st.n_samples >= w * h * 4 and st.n_samples <= n_passes * w * h * 4
    # more passes, less noise (the point of the loop): error against a converged render falls
    conv = renderer.render(cam, make_params(samples=1024, seed=99))
    one, _ = renderer.render_passes(cam, params, 1)
    six, _ = renderer.render_passes(cam, params, 6)
    e1 = float(np.sqrt(np.mean((one[..., :3] - conv[..., :3]) ** 2)))
    e6 = float(np.sqrt(np.mean((six[..., :3] - conv[..., :3]) ** 2)))
    assert e6 < 0.6 * e1, (e1, e6)
    del torch


def test_gpu_render_of_quadric_sdl_matches_the_references_own_image(renderer):
    """The same image-level pin as tests/test_oracle_kat2.py, for the CUDA path at the reference's 800x500."""
    from test_oracle_kat2 import check_object_layout, compare_with_reference_jpeg
    scene = load_scene("quadric")
    renderer.upload(scene)
    img = renderer.render(scene.camera(800, 500), make_params(samples=121))  # the CLI's default --samples 122 -> 11 x 11
    corr, rmse, mean, a, b = compare_with_reference_jpeg(img)
    assert min(corr) > 0.999, corr
    assert rmse < 0.015, rmse
    assert mean < 0.01, mean
    check_object_layout(a, b)


def crop_camera(cam, x0, y0, w, h):
    """The camera whose w x h image is the pixel window [x0, x0+w) x [y0, y0+h) of `cam`'s image: same u, v per pixel
    (painter.rs:133-139: u = x / W, v = (H - 1 - y) / H)."""
    out = RsCamera()
    C.memmove(C.byref(out), C.byref(cam), C.sizeof(RsCamera))
    W_, H_ = cam.width, cam.height
    u0 = x0 / W_
    # v of row y in the crop: (h - 1 - y) / h * V' + v0 must equal (H - 1 - (y0 + y)) / H of the full frame
    v0 = (H_ - y0 - h) / H_
    for k in range(3):
        out.lb[k] = cam.lb[k] + u0 * cam.horizontal_full[k] + v0 * cam.vertical_full[k]
        out.horizontal_full[k] = cam.horizontal_full[k] * (w / W_)
        out.vertical_full[k] = cam.vertical_full[k] * (h / H_)
    out.width, out.height = w, h
    return out


_C4 = {}


def _config4_oracle():
    """Scene + oracle answers of the 1 M-triangle test, computed once for both BVH widths (the CPU restatement
    needs ~20 s per 64 k-ray probe on this mesh)."""
    if not _C4:
        scene = Scene.builtin("mesh_blinn", 7, 708)
        oracle = OracleWorld(scene)
        full = scene.camera(2048, 2048)
        win = crop_camera(full, 896, 1000, 256, 256)
        _C4.update(scene=scene, full=full, win=win)
        for flags in (RS_FLAG_PIXEL_CENTRE, 0):
            _C4["primary", flags] = oracle.debug_primary(win, make_params(samples=4, flags=flags), 1)
        _C4["scatter"] = oracle.debug_scatter(win, make_params(samples=4), 2)
        rng = np.random.default_rng(13)
        mask = np.zeros((2048, 2048), np.uint8)
        mask[rng.integers(0, 2048, 3000), rng.integers(0, 2048, 3000)] = 1
        _C4["mask"] = mask
        _C4["pixels"], _ = oracle.render(full, make_params(sqrt_spp=1), mask=mask)
    return _C4


@pytest.mark.parametrize("bvh_width", [2, 4])
def test_config4_one_million_triangles(renderer, bvh_width):
    """BASELINE.json config 4 at its quoted size: the 708-grid mesh (1 001 112 triangles), Blinn-Phong, 2048x2048.
    Primary ids/t/normals on a 256x256 window of the full frame, single-bounce scatter on the same window, and
    3000 oracle-rendered pixels of the full frame."""
    c4 = _config4_oracle()
    scene, full, win = c4["scene"], c4["full"], c4["win"]
    assert scene.desc.contents.n_tris > 1_000_000
    try:
        renderer.set_bvh_width(bvh_width)
        renderer.upload(scene)
        for flags in (RS_FLAG_PIXEL_CENTRE, 0):
            g, o = renderer.debug_primary(win, make_params(samples=4, flags=flags), 1), c4["primary", flags]
            mism = np.flatnonzero(g["obj_id"] != o["obj_id"])
            assert mism.size == 0, f"{mism.size} ids differ: gpu {g['obj_id'][mism[:5]]} oracle {o['obj_id'][mism[:5]]} t {g['t'][mism[:5]]} / {o['t'][mism[:5]]}"
            hit = o["obj_id"] >= 0
            assert hit.mean() > 0.5 and len(np.unique(o["obj_id"][hit])) > 5000
            assert _rel_err(g["t"][hit], o["t"][hit]).max() < REL
            assert np.abs(g["normal"][hit] - o["normal"][hit]).max() < REL * max(1.0, np.abs(o["normal"][hit]).max())
            assert np.array_equal(g["material"][hit], o["material"][hit]) and np.array_equal(g["front"][hit], o["front"][hit])
        check_scatter(renderer.debug_scatter(win, make_params(samples=4), 2), c4["scatter"], "mesh_blinn:708")
        img = renderer.render(full, make_params(sqrt_spp=1))
        st = renderer.stats()
        assert st.n_samples == 2048 * 2048 and 2048 * 2048 <= st.n_segments <= 8 * 2048 * 2048
        sel = c4["mask"] == 1
        ref = c4["pixels"]
        diff = np.abs(img[sel][:, :3] - ref[sel][:, :3])
        scale = max(float(ref[sel][:, :3].mean()), 1e-3)
        assert np.median(diff) <= 1e-5 * scale
        assert (diff.max(axis=1) > 1e-3 * scale).mean() <= 0.002
        assert float(diff.mean()) <= 1e-3 * scale
    finally:
        renderer.set_bvh_width(2)


def test_surfaces_outside_their_reference_bbox_are_found_by_list_and_trees(renderer):
    """csg.sdl's rotated Intersection reaches outside the bbox the reference computes for it.  The canonical
    rule (oracle header) visits an object whenever the ray passes that bbox over [tmin, inf), wherever the
    hit lies, so neither tree may drop it behind a closer hit found first: the trees are built over
    (reference bbox ∪ bound of the surface), the object list and both trees must return the oracle's hit."""
    scene = load_scene("csg")
    oracle = OracleWorld(scene)
    o3 = np.array([-0.67466441, -0.99993399, -2.40198036])
    d3 = np.array([0.42554444, 0.63877392, 0.64099907])
    probe = CameraBuilder().look_from(o3).look_at(o3 + d3).vup((0.123, 1, 0.0456)).fov(1.0).aperture(0.0).focus(1.0).width(1).height(1).build()
    pp = make_params(sqrt_spp=1, depth=1, flags=RS_FLAG_PIXEL_CENTRE)
    cam = scene.camera(96, 60)
    params = make_params(samples=9)
    expect = oracle.debug_primary(probe, pp, 0)
    assert expect["obj_id"][0] == 1  # the Intersection, in front of the box (object 2) the ray reaches later
    imgs = []
    try:
        for list_max, width in ((12, 2), (0, 2), (0, 4)):
            renderer.set_list_max(list_max)
            renderer.set_bvh_width(width)
            renderer.upload(scene)
            g = renderer.debug_primary(probe, pp, 0)
            assert g["obj_id"][0] == 1 and math.isclose(g["t"][0], expect["t"][0], rel_tol=1e-13), (list_max, width)
            imgs.append((g["t"][0], renderer.render(cam, params)))
    finally:
        renderer.set_list_max(12)
        renderer.set_bvh_width(2)
    assert imgs[0][0] == imgs[1][0] == imgs[2][0]
    assert np.array_equal(imgs[0][1], imgs[1][1]) and np.array_equal(imgs[0][1], imgs[2][1])


def _device_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("devices", [(0,), (0, 0), (0, 0, 0), (0, 1), (0, 1, 2, 3)])
def test_rs_multi_render_equals_the_one_device_image(renderer, devices):
    """rs_create_multi / rs_multi_render: strata rows split over the contexts, arrival counters instead of host
    barriers, fused reduce + finalize of pixel slices into device 0's image.  The union of the shards is the
    one-device sample set; every pixel is a sum of the same per-sample values, grouped per shard, so the
    image may differ from rs_render's by the rounding of that grouping only.  A device listed twice gives
    separate contexts on one GPU, which exercises all of it on a one-GPU box."""
    from raysnail_b200 import MultiRenderer
    if max(devices) >= _device_count():
        pytest.skip(f"needs {max(devices) + 1} CUDA devices")
    scene = load_scene("cornell")
    cam = scene.camera(160, 120)
    params = make_params(sqrt_spp=6, depth=8)
    renderer.upload(scene)
    ref = renderer.render(cam, params)
    mask = np.ones((120, 160), np.uint8)
    mask[10:40, 20:90] = 0
    ref_masked = renderer.render(cam, params, mask=mask)
    m = MultiRenderer(devices)
    try:
        assert m.device_count == len(devices)
        m.upload(scene)
        for rep in range(3):  # the counters are epochs: calls may follow each other without a reset
            img = m.render(cam, params)
            np.testing.assert_allclose(img, ref, rtol=2e-6, atol=1e-7)
            assert m.stats.n_samples == 160 * 120 * 36
        if len(devices) == 1:
            assert np.array_equal(img, ref)
        img_m = m.render(cam, params, mask=mask)
        np.testing.assert_allclose(img_m, ref_masked, rtol=2e-6, atol=1e-7)
        assert not img_m[10:40, 20:90].any()
        # more devices than strata rows: the extra ones get an empty range
        p1 = make_params(sqrt_spp=1, depth=4)
        np.testing.assert_allclose(m.render(cam, p1), renderer.render(cam, p1), rtol=2e-6, atol=1e-7)
    finally:
        m.close()


def test_rs_multi_render_recovers_after_a_failed_frame(renderer):
    """A frame that fails part-way (here: on purpose, RS_OPT_INJECT_FAULT) leaves some devices' arrival signals out.
    The call has to come back with an error — the peers' wait kernels give up after RS_OPT_WAIT_BUDGET_MS instead of
    spinning for ever — and re-base its counters, so that the NEXT frame is complete and correct again.  Two contexts
    on device 0, so that a one-GPU box runs it."""
    from raysnail_b200 import MultiRenderer, RaysnailError
    RS_OPT_WAIT_BUDGET_MS, RS_OPT_INJECT_FAULT = 4, 100
    scene = load_scene("cornell")
    cam = scene.camera(96, 72)
    params = make_params(sqrt_spp=4, depth=6)
    renderer.upload(scene)
    ref = renderer.render(cam, params)
    m = MultiRenderer((0, 0))
    try:
        m.upload(scene)
        m.set_option(RS_OPT_WAIT_BUDGET_MS, 1500)
        np.testing.assert_allclose(m.render(cam, params), ref, rtol=2e-6, atol=1e-7)
        for fault in (1, 2, 2, 1):  # 1 = before anything is signalled, 2 = between the two signals of one device
            m.set_option(RS_OPT_INJECT_FAULT, fault)
            with pytest.raises(RaysnailError, match="injected fault"):
                m.render(cam, params)
            for _ in range(2):
                np.testing.assert_allclose(m.render(cam, params), ref, rtol=2e-6, atol=1e-7)
                assert m.stats.n_samples == 96 * 72 * 16
    finally:
        m.close()


def test_rs_multi_first_frame_with_more_devices_than_strata_rows(renderer):
    """sqrt_spp = 1 on two FRESH contexts: the second one has no strata row, never sets up its wave state, and its
    counters must read zero (not fail, not repeat an earlier frame's)."""
    from raysnail_b200 import MultiRenderer
    scene = load_scene("cornell")
    cam = scene.camera(80, 60)
    p1 = make_params(sqrt_spp=1, depth=4)
    renderer.upload(scene)
    ref = renderer.render(cam, p1)
    m = MultiRenderer((0, 0))
    try:
        m.upload(scene)
        np.testing.assert_allclose(m.render(cam, p1), ref, rtol=2e-6, atol=1e-7)
        assert m.stats.n_samples == 80 * 60
        p4 = make_params(sqrt_spp=4, depth=4)
        m.render(cam, p4)
        assert m.stats.n_samples == 80 * 60 * 16
        m.render(cam, p1)  # after a full frame: the idle context must not report that frame's counters again
        assert m.stats.n_samples == 80 * 60
    finally:
        m.close()


def test_a_wait_without_a_signal_ends_after_its_budget():
    """rs_wait_counter on a counter nobody signals: with RS_OPT_WAIT_BUDGET_MS the kernel stops polling, the stream
    drains, and rs_wait_check reports it (once); a satisfied wait reports nothing."""
    import time
    import torch
    from raysnail_b200 import RaysnailError, Renderer
    r = Renderer(0)
    try:
        r.set_wait_budget_ms(300)
        c = r.device_alloc(256)
        r.signal(c)
        r.wait_counter(c, 1)
        r.wait_check()  # satisfied
        t0 = time.time()
        r.wait_counter(c, 2)  # nobody will bring the counter to 2
        with pytest.raises(RaysnailError, match="gave up"):
            r.wait_check()
        assert 0.25 < time.time() - t0 < 30.0
        r.wait_check()  # the flag was cleared
        r.signal(c)
        r.wait_counter(c, 2)
        r.wait_check()
        torch.cuda.synchronize()
        r.device_free(c)
    finally:
        r.close()


def test_reuploads_keep_or_rebuild_the_tree_correctly():
    """rs_upload_scene keeps the SAH tree (and its device copy) while the objects' bounds and ids are unchanged.
    One context sees: scene A, A again (tree kept), A with a sphere moved (same count: must rebuild), a scene with
    other kinds at the same bounds (tree kept, leaf references must be re-encoded), a far-away camera (boxes
    re-padded) and A once more — every image must equal the one a fresh context renders, bit for bit."""
    from raysnail_b200 import Renderer

    def spheres(shift, as_boxes=False):
        s = Scene()
        s.add(s.sphere((0, -1000, 0), 1000.0, s.lambertian(s.solid(0.5, 0.5, 0.5))))
        light = s.sphere((0, 12, 0), 3.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
        s.add(light)
        s.add_light(light)
        for k in range(40):
            c = (-6.0 + 0.62 * (k % 20), 0.3 + 0.7 * (k // 20), -2.0 + 0.21 * k + (shift if k == 17 else 0.0))
            m = s.lambertian(s.solid(0.2 + 0.02 * k, 0.9 - 0.02 * k, 0.4))
            # a box with the bounds of the sphere: same tree members, another leaf kind
            s.add(s.box((c[0] - 0.3, c[1] - 0.3, c[2] - 0.3), (c[0] + 0.3, c[1] + 0.3, c[2] + 0.3), m) if as_boxes else s.sphere(c, 0.3, m))
        s.set_camera((0, 3, 12), (0, 0.5, 0), fov=40.0)
        return s

    params = make_params(samples=16)

    def fresh(scene, cam):
        r = Renderer(0)
        try:
            r.upload(scene)
            return r.render(cam, params)
        finally:
            r.close()

    a, moved, boxes = spheres(0.0), spheres(0.37), spheres(0.0, as_boxes=True)
    r = Renderer(0)
    try:
        for label, scene in (("A", a), ("A again", a), ("moved", moved), ("A back", a), ("boxes at the same bounds", boxes), ("A after boxes", a)):
            cam = scene.camera(96, 64)
            r.upload(scene)
            assert np.array_equal(r.render(cam, params), fresh(scene, cam)), label
        # a camera far outside the scene extent: the f32 boxes are re-padded for its origin, then A is uploaded again
        far = CameraBuilder().look_from((0.0, 900.0, 4000.0)).look_at((0.0, 0.5, 0.0)).fov(0.5).width(96).height(64).build()
        img_far = r.render(far, params)
        r.upload(a)
        assert np.array_equal(r.render(far, params), img_far)
        rf = Renderer(0)
        try:
            rf.upload(a)
            assert np.array_equal(rf.render(far, params), img_far)
        finally:
            rf.close()
        assert np.array_equal(r.render(a.camera(96, 64), params), fresh(a, a.camera(96, 64)))
    finally:
        r.close()


def bvh_scene():
    """A `BVH` as ONE object (bvh.rs:173-192): bare, under a TfFacade, nested in another BVH with a CSG member, and as the
    boundary of a ConstantMedium."""
    s = Scene()
    s.add(s.sphere((0, -1000, 0), 1000.0, s.lambertian(s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.9, 0.9), 2.0))))
    light = s.sphere((2, 14, 5), 3.0, s.diffuse_light(s.solid(1, 1, 1), 6.0))
    s.add(light)
    s.add_light(light)

    def cluster(x0, z0, seed):
        rng = np.random.default_rng(seed)
        out = []
        for k in range(7):
            c = (x0 + float(rng.uniform(-1.5, 1.5)), float(rng.uniform(0.3, 2.2)), z0 + float(rng.uniform(-1.0, 1.0)))
            m = [s.lambertian(s.solid(0.2 + 0.1 * k, 0.7 - 0.05 * k, 0.5)), s.metal(s.solid(0.8, 0.7, 0.6)), s.dielectric((1, 1, 1), 1.5)][k % 3]
            out.append(s.sphere(c, 0.35, m) if k % 2 else s.box((c[0] - 0.3, c[1] - 0.3, c[2] - 0.3), (c[0] + 0.3, c[1] + 0.3, c[2] + 0.3), m))
        return out

    s.add(s.bvh(cluster(-4.0, 0.0, 1)))                                                                  # bare
    s.add(s.transform(s.bvh(cluster(0.0, 0.0, 2)), [("rotate_y", 0.5), ("scale", (1.0, 1.2, 1.0)), ("translate", (0.0, 0.0, -1.0))]))
    csg = s.difference(s.box((3.2, 0.2, -0.6), (4.4, 1.4, 0.6), s.lambertian(s.solid(0.7, 0.3, 0.3))), s.sphere((3.8, 0.8, 0.6), 0.5), -1)
    s.add(s.bvh([s.bvh(cluster(4.0, 2.0, 3)[:3]), csg, s.sphere((5.2, 0.5, 0.0), 0.5, s.blinn_phong(0.5, 20.0, s.solid(0.3, 0.4, 0.8)))]))  # nested + CSG member
    s.add(s.transform(s.medium(s.bvh([s.sphere((0, 0, 0), 0.8), s.box((0.5, -0.5, -0.5), (1.6, 0.5, 0.5))]), (0.9, 0.9, 0.95), 1.5),
                      [("translate", (-1.0, 0.9, 3.0))]))                                               # medium boundary
    s.add(s.bvh([s.sphere((-6.5, 0.6, 2.0), 0.6, s.lambertian(s.solid(0.9, 0.8, 0.2)))]))                # one member: left child only
    s.set_camera((0.5, 4.0, 11.0), (0.0, 0.8, 0.0), fov=42.0)
    return s


def test_a_bvh_of_400_members_as_one_world_object(renderer):
    """The reference's all_feature scene adds `BVH::new` over 400 boxes as ONE world object
    (examples/common/scene.rs:361, bvh.rs:173-192).  A balanced tree over 400 members is 9-10 levels deep; the object
    program keeps one running result per BVH object, so its depth does not count against RS_MAX_TREE_DEPTH.  The same
    400 boxes (a few spheres and one CSG member among them) added to the world one by one must give the same hits."""
    rng = np.random.default_rng(400)

    def build(as_one_object):
        s = Scene()
        s.add(s.sphere((0, -1000, 0), 1000.0, s.lambertian(s.solid(0.5, 0.5, 0.5))))
        light = s.sphere((0, 30, 10), 8.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
        s.add(light)
        s.add_light(light)
        r = np.random.default_rng(7)
        members = []
        for i in range(20):
            for j in range(20):
                x0, z0, hgt = -10.0 + i, -10.0 + j, float(r.uniform(0.1, 1.6))
                m = s.lambertian(s.solid(float(r.uniform(0.2, 0.9)), float(r.uniform(0.2, 0.9)), float(r.uniform(0.2, 0.9))))
                if (i * 20 + j) % 37 == 5:
                    members.append(s.sphere((x0 + 0.5, hgt, z0 + 0.5), 0.45, s.metal(s.solid(0.8, 0.8, 0.7))))
                elif (i, j) == (9, 12):
                    members.append(s.difference(s.box((x0, 0.0, z0), (x0 + 0.9, 1.5, z0 + 0.9), m), s.sphere((x0 + 0.45, 1.5, z0 + 0.45), 0.4), -1))
                else:
                    members.append(s.box((x0, 0.0, z0), (x0 + 0.9, hgt, z0 + 0.9), m))
        if as_one_object:
            s.add(s.bvh(members))
        else:
            for m in members:
                s.add(m)
        s.set_camera((3.0, 9.0, 16.0), (0.0, 0.5, 0.0), fov=48.0)
        return s

    one, flat = build(True), build(False)
    cam = one.camera(W, H)
    params = make_params(samples=9, flags=RS_FLAG_PIXEL_CENTRE)
    renderer.upload(flat)
    f = renderer.debug_primary(cam, params, 0)
    renderer.upload(one)  # refused before round 3: "object tree deeper than RS_MAX_TREE_DEPTH"
    g = renderer.debug_primary(cam, params, 0)
    o = OracleWorld(one).debug_primary(cam, params, 0)
    assert np.array_equal(g["obj_id"], o["obj_id"])
    hit = o["obj_id"] >= 0
    assert (o["obj_id"] == 2).mean() > 0.2  # the BVH object fills a good part of the frame
    assert _rel_err(g["t"][hit], o["t"][hit]).max() < REL
    assert np.abs(g["normal"][hit] - o["normal"][hit]).max() < REL
    for key in ("front", "material"):
        assert np.array_equal(g[key][hit], o[key][hit]), key
    # one object or 400: the same surfaces are hit at the same distances (ids differ by construction)
    assert np.array_equal(f["obj_id"] >= 2, g["obj_id"] == 2)
    assert _rel_err(g["t"][hit], f["t"][hit]).max() < REL
    assert np.abs(g["normal"][hit] - f["normal"][hit]).max() < REL  # (material indices are per scene and not comparable)
    check_scatter(renderer.debug_scatter(cam, make_params(samples=9), 4), OracleWorld(one).debug_scatter(cam, make_params(samples=9), 4), "bvh400")
    small = one.camera(64, 40)
    gi = renderer.render(small, make_params(samples=64))
    oi, _ = OracleWorld(one).render(small, make_params(samples=64))
    rmse = float(np.sqrt(np.mean((gi[..., :3] - oi[..., :3]) ** 2))) / float(np.sqrt(np.mean(oi[..., :3] ** 2)))
    assert rmse < 0.01, rmse


def test_a_bvh_as_one_world_object(renderer):
    scene = bvh_scene()
    renderer.upload(scene)
    oracle = OracleWorld(scene)
    cam = scene.camera(W, H)
    for flags, sample in ((RS_FLAG_PIXEL_CENTRE, 0), (0, 5)):
        params = make_params(samples=9, flags=flags)
        g, o = renderer.debug_primary(cam, params, sample), oracle.debug_primary(cam, params, sample)
        assert np.array_equal(g["obj_id"], o["obj_id"])
        hit = o["obj_id"] >= 0
        assert set(np.unique(o["obj_id"][hit])) >= {0, 2, 3, 4, 5, 6}
        assert _rel_err(g["t"][hit], o["t"][hit]).max() < REL
        assert np.abs(g["normal"][hit] - o["normal"][hit]).max() < REL
        for key in ("front", "material"):
            assert np.array_equal(g[key][hit], o[key][hit]), key
    check_scatter(renderer.debug_scatter(cam, make_params(samples=9), 2), oracle.debug_scatter(cam, make_params(samples=9), 2), "bvh")
    small = scene.camera(64, 40)
    gi = renderer.render(small, make_params(samples=256))
    oi, _ = oracle.render(small, make_params(samples=256))
    rmse = float(np.sqrt(np.mean((gi[..., :3] - oi[..., :3]) ** 2))) / float(np.sqrt(np.mean(oi[..., :3] ** 2)))
    assert rmse < 0.01, rmse
    # BVH::contains is unimplemented!() in the reference (bvh.rs:194-200): a BVH as a CSG operand is refused
    bad = Scene()
    light = bad.sphere((0, 10, 0), 2.0, bad.diffuse_light(bad.solid(1, 1, 1), 5.0))
    bad.add(light)
    bad.add_light(light)
    bad.add(bad.intersection(bad.box((-1, -1, -1), (1, 1, 1)), bad.bvh([bad.sphere((0, 0, 0), 1.2)]), bad.lambertian(bad.solid(0.5, 0.5, 0.5))))
    with pytest.raises(Exception, match="BVH"):
        renderer.upload(bad)
    renderer.upload(scene)
