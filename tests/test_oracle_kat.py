"""Known-answer tests that pin the CPU oracle.

The reference's own tests hold exactly one usable known answer — the (stale) unit test
`test_y_rotation` in src/hittable/transform/transform.rs:187-206: rotate_by_y_axis(pi/2) maps
(0,0,1) to (1,0,0) within 1e-10 and round-trips.  Everything else here is a self-made KAT with
an analytically known answer, plus the published Philox4x32-10 vectors (Random123) and the
committed golden vectors the oracle itself produced (tests/golden/oracle_vectors.npz).
"""
import ctypes as C
import math
import os

import numpy as np
import pytest

from conftest import load_scene
from oracle_binding import OracleWorld, philox, set_marcher_mode
from raysnail_b200 import RS_FLAG_PIXEL_CENTRE, Scene, make_params

HERE = os.path.dirname(os.path.abspath(__file__))


def centre_probe(scene, look_from, look_at, vup=(0, 1, 0), what="primary"):
    """Shoots the ray look_from -> look_at: with a 3x3 image the pixel (row 0, col 1) at its centre
    has u = 0.5 and v = (H-1-0.5)/H = 0.5 (painter.rs:133-139), i.e. the view axis exactly."""
    scene.set_camera(look_from, look_at, vup, fov=40.0, aperture=0.0, focus=1.0)
    cam = scene.camera(3, 3)
    params = make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE)
    w = OracleWorld(scene)
    o = w.debug_primary(cam, params, 0) if what == "primary" else w.debug_scatter(cam, params, 0)
    return {k: v[1] for k, v in o.items()}


def mat_color(scene, flat_index):
    """solid colour behind a flat material index (handles are not flat indices)"""
    d = scene.desc.contents
    m = d.materials[int(flat_index)]
    return tuple(round(float(c), 6) for c in d.textures[m.texture].color[:3])


def simple_scene():
    s = Scene()
    white = s.lambertian(s.solid(0.8, 0.8, 0.8))
    light = s.sphere((0, 1000, 0), 1.0, s.diffuse_light(s.solid(1, 1, 1), 2.0))
    s.add(light)
    s.add_light(light)
    return s, white


def test_philox4x32_10_random123_known_answers():
    assert [hex(x) for x in philox([0, 0, 0, 0], [0, 0])] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in philox([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0])] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_reference_unit_test_y_rotation():
    """transform.rs:187-206 — the reference's only #[test]."""
    s, white = simple_scene()
    sp = s.sphere((0, 0, 1), 0.25, white)
    s.add(s.transform(sp, [("rotate_y", math.pi / 2)]))
    d = s.desc.contents
    m = np.array(d.xforms[0].m[:]).reshape(4, 4)
    inv = np.array(d.xforms[0].inv[:]).reshape(4, 4)
    r = m @ np.array([0.0, 0.0, 1.0, 1.0])
    assert np.abs(r[:3] - [1.0, 0.0, 0.0]).max() < 1e-10
    r2 = inv @ r
    assert np.abs(r2[:3] - [0.0, 0.0, 1.0]).max() < 1e-10
    # and through the oracle's TfFacade: the rotated sphere sits at (1,0,0)
    h = centre_probe(s, (5, 0, 0), (0, 0, 0))
    assert h["obj_id"] == 1
    assert abs(h["t"] - 3.75) < 1e-9
    assert np.abs(h["point"] - [1.25, 0, 0]).max() < 1e-9


def test_axis_ray_vs_unit_sphere():
    s, white = simple_scene()
    s.add(s.sphere((0, 0, 0), 1.0, white))
    h = centre_probe(s, (0, 0, 5), (0, 0, 0))
    assert h["obj_id"] == 1 and h["front"] == 1
    assert abs(h["t"] - 4.0) < 1e-12          # |o| - 1
    assert np.abs(h["normal"] - [0, 0, 1]).max() < 1e-12
    # from inside: far root, normal flipped, outside = false (hit.rs:36-40)
    h = centre_probe(s, (0, 0, 0.5), (0, 0, -1))
    assert abs(h["t"] - 1.5) < 1e-12 and h["front"] == 0
    assert np.abs(h["normal"] - [0, 0, 1]).max() < 1e-12


def test_ray_vs_box():
    s, white = simple_scene()
    s.add(s.box((-1, -1, -1), (1, 1, 1), white))
    h = centre_probe(s, (0, 0, 5), (0, 0, 0))
    assert h["obj_id"] == 1 and abs(h["t"] - 4.0) < 1e-12
    assert np.abs(h["normal"] - [0, 0, 1]).max() == 0.0 and h["front"] == 1
    # one face only (ray starts inside): HitRecord::new semantics -> flipped normal, outside false
    # (the face normal is +axis, face-forwarded against the ray: rect.rs:85-89, hit.rs:36-40)
    h = centre_probe(s, (0, 0, 0), (0, 0, 3))
    assert abs(h["t"] - 1.0) < 1e-12 and h["front"] == 0
    assert np.abs(h["normal"] - [0, 0, -1]).max() == 0.0
    h = centre_probe(s, (0, 0, 0), (0, 0, -3))
    assert abs(h["t"] - 1.0) < 1e-12 and h["front"] == 1
    assert np.abs(h["normal"] - [0, 0, 1]).max() == 0.0


@pytest.mark.parametrize("q,axis_from,expect_t", [
    # sdl/quadric.sdl shapes (coefficients after the parser's reordering, sdl_parser.rs:659-661)
    ((1, 0, 0, 0, 0, 0, 0, 1, 0, -1), (5, 0.3, 0), 4.0),            # cylinder x^2+z^2-1, along -x
    ((1, 0, 0, 0, -1, 0, 0, 1, 0, -1), (5, 0, 0), 4.0),             # hyperboloid x^2-y^2+z^2-1 at y=0
    ((1, 0, 0, 0, 0, 0, -1, 1, 0, 0), (0, 5, 0), 5.0),              # paraboloid x^2+z^2-y, along -y: y=0
])
def test_quadrics_along_an_axis(q, axis_from, expect_t):
    s, white = simple_scene()
    s.add(s.quadric(q, white))
    target = (0, axis_from[1] if axis_from[0] else 0, 0)
    h = centre_probe(s, axis_from, target, vup=(0, 0, 1) if axis_from[0] == 0 else (0, 1, 0))
    assert h["obj_id"] == 1
    assert abs(h["t"] - expect_t) < 1e-9


def test_difference_synthetic_inner_hit():
    """sdl/csg.sdl object 0 without its rotation: Box(-1,-1,-1 .. 1,0,1) minus Sphere((0,0.1,0),0.9),
    looked at straight down: the ray enters the sphere first (minus.t1 < plus.t1), leaves it inside
    the box -> synthetic hit at minus.t2 with normal -minus.normal and outside = true
    (difference.rs:83-99)."""
    s, white = simple_scene()
    blue = s.lambertian(s.solid(0.0, 0.1, 0.3))
    plus = s.box((-1, -1, -1), (1, 0, 1), blue)
    minus = s.sphere((0, 0.1, 0), 0.9, white)
    s.add(s.difference(plus, minus))
    h = centre_probe(s, (0, 5, 0), (0, 0, 0), vup=(0, 0, 1))
    assert h["obj_id"] == 1
    assert abs(h["t"] - (5 - (0.1 - 0.9))) < 1e-12   # sphere bottom at y = -0.8
    assert np.abs(h["normal"] - [0, 1, 0]).max() < 1e-12
    assert h["front"] == 1
    assert mat_color(s, h["material"]) == (0.8, 0.8, 0.8)   # minus.material()


def test_intersection_picks_the_inner_surface():
    s, white = simple_scene()
    sph = s.sphere((0, 0, 0), 1.0)
    bx = s.box((-2, -0.5, -2), (2, 0.5, 2))
    s.add(s.intersection(sph, bx, white))
    h = centre_probe(s, (0, 5, 0), (0, 0, 0), vup=(0, 0, 1))
    assert h["obj_id"] == 1 and abs(h["t"] - 4.5) < 1e-12 and mat_color(s, h["material"]) == (0.8, 0.8, 0.8)
    h = centre_probe(s, (5, 0, 0), (0, 0, 0))
    assert abs(h["t"] - 4.0) < 1e-12                  # the sphere surface inside the slab


def test_glass_schlick_and_branch_kinds():
    """Dielectric is a skip_pdf material (kind 2); at normal incidence (cos = 1) reflect_prob = r0 =
    ((1-n)/(1+n))^2 = 0.04 for n = 1.5, so the scattered ray is (almost always) the straight
    refraction; Lambertian yields kinds 3/4; a light yields kind 1 with emitted = colour x multiplier."""
    s, white = simple_scene()
    glass = s.dielectric((1, 1, 1), 1.5, True)
    s.add(s.sphere((0, 0, 0), 1.0, glass))
    sc = centre_probe(s, (0, 0, 5), (0, 0, 0), what="scatter")
    assert sc["kind"] == 2
    assert np.abs(sc["weight"] - 1.0).max() == 0.0
    d = sc["next"][3:]
    assert abs(abs(d[2]) - 1.0) < 1e-12               # straight through or straight back
    s2, white2 = simple_scene()
    s2.add(s2.sphere((0, 0, 0), 1.0, white2))
    sc = centre_probe(s2, (0, 0, 5), (0, 0, 0), what="scatter")
    assert sc["kind"] in (3, 4)
    if sc["kind"] == 4:                                # pdf branch: weight = albedo * value/value
        assert np.abs(sc["weight"] - np.float32(0.8)).max() < 1e-6
    sc = centre_probe(s2, (0, 995, 0), (0, 1000, 0), vup=(0, 0, 1), what="scatter")
    assert sc["kind"] == 1 and np.abs(sc["emitted"] - 2.0).max() < 1e-12


def test_cosine_pdf_light_branch_weight_is_cosine():
    """Light branch: pdf_val = 1/pi, scattering pdf = cos/pi -> factor = albedo * cos(theta)
    (camera.rs:198-245); CosinePdf.value(n) = 1/pi."""
    s, white = simple_scene()
    s.add(s.sphere((0, 0, 0), 1.0, white))
    scene_cam = dict(look_from=(0, 5, 0), look_at=(0, 0, 0), vup=(0, 0, 1))
    s.set_camera(fov=40.0, aperture=0.0, focus=1.0, **scene_cam)
    cam = s.camera(3, 3)
    w = OracleWorld(s)
    found = 0
    for seed in range(40):
        params = make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE, seed=seed)
        sc = w.debug_scatter(cam, params, 0)
        if sc["kind"][1] == 3:
            d = sc["next"][1][3:]
            cos = d[1]                                  # normal at the top of the sphere is +y
            assert abs(np.linalg.norm(d) - 1.0) < 1e-12
            assert np.abs(sc["weight"][1] - np.float32(0.8) * max(cos, 0.0)).max() < 1e-6
            found += 1
    assert found > 5


def test_checker_sign_table_and_image_lookup():
    s = Scene()
    odd, even = s.solid(1, 0, 0), s.solid(0, 1, 0)
    chk = s.lambertian(s.checker(odd, even, 2.0))
    lt = s.sphere((0, 1000, 0), 1.0, s.diffuse_light(s.solid(1, 1, 1), 1.0))
    s.add(lt)
    s.add_light(lt)
    s.add(s.rect("xz", 0.5, (-10, 10), (-10, 10), chk))
    for (x, z) in [(0.3, 0.3), (-0.3, 0.3), (2.0, 0.3), (0.3, -2.0)]:
        value = math.sin(2 * x) * math.sin(2 * 0.5) * math.sin(2 * z)
        found = False
        for seed in range(64):                          # find a pdf-branch sample: weight = albedo
            s.set_camera((x, 5, z), (x, 0.5, z), (0, 0, 1), fov=40.0, aperture=0.0, focus=1.0)
            cam = s.camera(3, 3)
            sc = OracleWorld(s).debug_scatter(cam, make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE, seed=seed), 0)
            if sc["kind"][1] == 4:
                expect = [1, 0, 0] if value < 0 else [0, 1, 0]
                assert np.abs(sc["weight"][1] - expect).max() < 1e-6
                found = True
                break
        assert found
    # image texture: nearest texel, v flipped (image.rs:36-49), through an emissive rect
    img = np.zeros((4, 8, 4), np.uint8)
    img[..., 3] = 255
    img[1, 5] = [255, 128, 0, 255]                     # row 1 from the top, column 5
    s2 = Scene()
    tex = s2.image(img)
    em = s2.diffuse_light(tex, 1.0)
    r = s2.rect("xz", 0.0, (0, 8), (0, 4), em)
    s2.add(r)
    s2.add_light(r)
    # rect uv: u = (x - 0)/8, v = (z - 0)/4 ; texel px = floor(u*8) = 5, py = floor((1-v)*4) = 1 -> v in (0.5, 0.75]
    s2.set_camera((5.5, 3, 2.6), (5.5, 0, 2.6), (0, 0, 1), fov=40.0, aperture=0.0, focus=1.0)
    sc = OracleWorld(s2).debug_scatter(s2.camera(3, 3), make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE), 0)
    assert sc["kind"][1] == 1
    assert np.abs(sc["emitted"][1] - [1.0, np.float32(128) / np.float32(255), 0.0]).max() < 1e-7


def test_sphere_uv_and_earth_texture_lookup():
    """Image texel lookup on the reference's earth-map.png (1920x960), read by the host PNG reader."""
    path = os.path.join(HERE, "golden", "earth-map.png")  # the reference's examples/earth-map.png (tests/golden/make_fixtures.py)
    if not os.path.exists(path):
        pytest.skip("reference assets are not on this machine")
    from PIL import Image
    ref = np.asarray(Image.open(path).convert("RGB"))
    s = Scene()
    tex = s.image_file(path)
    d_scene = s
    em = d_scene.diffuse_light(tex, 1.0)
    sp = d_scene.sphere((0, 0, 0), 1.0, em)
    d_scene.add(sp)
    d_scene.add_light(sp)
    for look_from in [(0, 0, 5), (5, 0, 0), (3, 2, -4)]:
        d_scene.set_camera(look_from, (0, 0, 0), (0, 1, 0), fov=40.0, aperture=0.0, focus=1.0)
        cam = d_scene.camera(3, 3)
        w = OracleWorld(d_scene)
        params = make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE)
        h = w.debug_primary(cam, params, 0)
        p = h["point"][1] / np.linalg.norm(h["point"][1])
        u = math.atan2(-p[2], p[0]) / 2 / math.pi + 0.5      # sphere.rs:64-71
        v = math.asin(p[1]) / math.pi + 0.5
        px, py = min(int(u * 1920), 1919), min(int((1 - v) * 960), 959)
        sc = w.debug_scatter(cam, params, 0)
        assert np.abs(sc["emitted"][1] - ref[py, px].astype(np.float32) / np.float32(255)).max() < 1e-6


def test_painter_v_offset_quirk_and_sample_rounding():
    from raysnail_b200 import painter_samples
    assert [painter_samples(n) for n in (65, 122, 256, 1024, 4096, 1, 3)] == [8, 11, 16, 32, 64, 1, 1]
    s, white = simple_scene()
    s.add(s.sphere((0, 0, 0), 1.0, white))
    s.set_camera((0, 0, 5), (0, 0, 0), (0, 1, 0), fov=90.0, aperture=0.0, focus=1.0)
    cam = s.camera(4, 2)
    o = OracleWorld(s).debug_primary(cam, make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE), 0)
    # v = (H-1-yo)/H (painter.rs:137): row 0 centre -> v = 0.25, row 1 -> v = -0.25: below the frame
    dirs = o["ray"][:, 3:]
    vh = 2.0 * math.tan(math.radians(45.0))              # viewport height at focus 1
    for row, v in ((0, 0.25), (1, -0.25)):
        dy = (v - 0.5) * vh
        d = dirs[row * 4 + 1]
        assert abs(d[1] / -d[2] - dy) < 1e-12


def test_oracle_matches_committed_golden_vectors():
    g = np.load(os.path.join(HERE, "golden", "oracle_vectors.npz"))
    for name in ["example", "quadric", "csg", "declares", "transforms", "rtow_final", "cornell", "mesh"]:
        s = load_scene(name)
        w = OracleWorld(s)
        cam = s.camera(48, 30)
        o = w.debug_primary(cam, make_params(samples=4, flags=RS_FLAG_PIXEL_CENTRE), 0)
        assert np.array_equal(o["obj_id"], g[name + "_id"]), name
        assert np.allclose(o["t"], g[name + "_t"], rtol=1e-12, atol=1e-12), name
        assert np.allclose(o["normal"], g[name + "_n"], rtol=1e-9, atol=1e-9), name
        sc = w.debug_scatter(cam, make_params(samples=4), 1)
        assert np.array_equal(sc["kind"], g[name + "_kind"]), name
        assert np.allclose(sc["next"], g[name + "_next"], rtol=1e-9, atol=1e-9), name


def test_oracle_image_is_deterministic_and_thread_count_invariant():
    s = load_scene("example")
    w = OracleWorld(s)
    cam = s.camera(40, 24)
    p = make_params(samples=9)
    a, sa = w.render(cam, p, threads=1)
    b, sb = w.render(cam, p, threads=5)
    assert np.array_equal(a, b) and sa.n_segments == sb.n_segments and sa.n_samples == 40 * 24 * 9
    mask = np.zeros((24, 40), np.uint8)
    mask[3:9, 5:20] = 1
    m, _ = w.render(cam, p, mask=mask)
    assert (m[mask == 0] == 0).all() and np.array_equal(m[mask == 1], a[mask == 1])


# ---------------------------------------------------------------------------------------------
# RayMarcher (src/hittable/geometry/raymarching.rs): the power-8 Mandelbulb
# ---------------------------------------------------------------------------------------------
def test_raymarcher_along_the_z_axis():
    """On the z axis the bulb's iteration is the 1-D map z -> |z|^8 + c (theta = 0 or pi, 8*pi wraps).
    For c > 0 it stays bounded iff z^8 + c = z has a root: c <= c* = 7/8 * 8^(-1/7) = 0.650146.  A ray
    from (0,0,5) straight down the axis starts outside the bailout sphere (|p|^2 > 8: the distance
    estimate is ln(0)*0 = NaN -> 0.1 -> search_surface, raymarching.rs:127-134), walks in 0.05 steps
    and bisects 8 times: t = 5 - c_num with c_num in (c*, c* + a few 1e-3] (100 iterations let orbits
    just above the tangent bifurcation survive) and t a multiple of 0.05/256."""
    s, white = simple_scene()
    s.add(s.raymarcher(white))
    h = centre_probe(s, (0, 0, 5), (0, 0, 0))
    c_star = 7.0 / 8.0 * 8.0 ** (-1.0 / 7.0)
    assert h["obj_id"] == 1 and h["front"] == 1
    assert 5 - c_star - 6e-3 < h["t"] < 5 - c_star + 0.05 / 256
    assert abs(h["t"] / (0.05 / 256) - round(h["t"] / (0.05 / 256))) < 1e-6
    # the bulb has 7-fold (power - 1) symmetry about z, so the surface normal on the axis is +z; the
    # reference's normal is a central difference of the distance estimate over 0.01 (raymarching.rs:83-96):
    # noisy on a fractal, but it must point mostly along +z
    assert h["normal"][2] > 0.9
    # from farther than 200 * 0.05 = 10 units the search gives up: no hit (raymarching.rs:64)
    s2, white2 = simple_scene()
    s2.add(s2.raymarcher(white2))
    assert centre_probe(s2, (0, 0, 12), (0, 0, 0))["obj_id"] == -1


def test_raymarcher_contains_through_difference():
    """`contains` = is_inside (raymarching.rs:167-170).  Box(-2..2) minus the bulb, ray down the z axis:
    the plus hit (box face z = 2, t = 3) is outside the bulb -> it is the hit (difference.rs:64-72).
    Box(-0.3..0.3) minus the bulb: its face z = 0.3 < c* lies inside the bulb and so does its far
    face -> no hit from the box; with the bulb as `plus` and the box as `minus` the surface hit stays."""
    s, white = simple_scene()
    s.add(s.difference(s.box((-2, -2, -2), (2, 2, 2), white), s.raymarcher(white), white))
    h = centre_probe(s, (0, 0, 5), (0, 0, 0))
    assert h["obj_id"] == 1 and abs(h["t"] - 3.0) < 1e-12
    s, white = simple_scene()
    s.add(s.difference(s.raymarcher(white), s.box((-0.3, -0.3, -0.3), (0.3, 0.3, 0.3), white), white))
    h = centre_probe(s, (0, 0, 5), (0, 0, 0))
    c_star = 7.0 / 8.0 * 8.0 ** (-1.0 / 7.0)
    assert h["obj_id"] == 1 and abs(h["t"] - (5 - c_star)) < 6e-3


def test_raymarching_scene_golden_vectors():
    """Oracle vectors of the `render_raymarching_test` recipe (examples/preview_sdl2.rs:338-407),
    made by tests/golden/make_fixtures.py.  The canonical Mandelbulb arithmetic uses IEEE-exact
    operations only, so the vectors are reproducible on any host."""
    g = np.load(os.path.join(HERE, "golden", "raymarching_vectors.npz"))
    s = load_scene("raymarching")
    w = OracleWorld(s)
    o = w.debug_primary(s.camera(48, 30), make_params(samples=4, flags=RS_FLAG_PIXEL_CENTRE), 0)
    assert (o["obj_id"] == 1).sum() > 150
    assert np.array_equal(o["obj_id"], g["id"])
    assert np.allclose(o["t"], g["t"], rtol=1e-12, atol=1e-12)
    bulb = o["obj_id"] == 1
    assert np.allclose(o["normal"][~bulb], g["n"][~bulb], rtol=1e-9, atol=1e-9)
    assert np.allclose(o["normal"][bulb], g["n"][bulb], rtol=1e-6, atol=1e-6)


def test_raymarcher_canonical_vs_libm():
    """Pins the canonical (IEEE-exact, GPU-identical) form of the Mandelbulb map to the literal
    restatement of raymarching.rs:202-241 that calls sqrt/atan2/powf/sin/cos/ln of the platform libm.
    Both evaluate the same map to a few ulp per call; orbits near the surface amplify that, so a
    small share of the rays lands on another point of the 0.05/256 search grid — the same spread
    two libms show between themselves (glibc vs CUDA, measured on B200: 40 of 3881 rays).  Bar:
    same object on >= 98.5 % of the rays that meet the bulb's box, same t (1e-9) on >= 97 % of the
    common hits, equal silhouettes elsewhere; on the z axis both give the same grid point."""
    s = load_scene("raymarching")
    w = OracleWorld(s)
    cam = s.camera(96, 60)
    p = make_params(samples=4, flags=RS_FLAG_PIXEL_CENTRE)
    try:
        set_marcher_mode(1)
        lit = w.debug_primary(cam, p, 0)
        sa, white = simple_scene()
        sa.add(sa.raymarcher(white))
        axis_lit = centre_probe(sa, (0, 0, 5), (0, 0, 0))
    finally:
        set_marcher_mode(0)
    can = w.debug_primary(cam, p, 0)
    sa, white = simple_scene()
    sa.add(sa.raymarcher(white))
    axis_can = centre_probe(sa, (0, 0, 5), (0, 0, 0))
    assert axis_can["t"] == axis_lit["t"]
    bulb = (lit["obj_id"] == 1) | (can["obj_id"] == 1)
    assert bulb.sum() > 600
    assert np.array_equal(lit["obj_id"][~bulb], can["obj_id"][~bulb])
    assert np.mean(lit["obj_id"][bulb] == can["obj_id"][bulb]) >= 0.985
    both = (lit["obj_id"] == 1) & (can["obj_id"] == 1)
    same_t = np.abs(lit["t"][both] - can["t"][both]) <= 1e-9 * lit["t"][both]
    print("bulb rays", int(bulb.sum()), "id mismatches", int((lit["obj_id"][bulb] != can["obj_id"][bulb]).sum()),
          "t mismatches", int((~same_t).sum()), "of", int(both.sum()))
    assert same_t.mean() >= 0.97
