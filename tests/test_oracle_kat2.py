"""More known-answer tests for the CPU oracle (round 2): the Perlin variants of
src/texture/noise.rs against an independent numpy restatement and hand-computed lattice values,
the Philox entry points exported for host harnesses, MixedMaterial::emitted, the opt-in MIS
integrator's weights, the passes loop — and the ONE artefact the reference's authors produced
with their own program: examples/sdl_quadrics.jpg, an 800x500 render of sdl/quadric.sdl, which
pins oracle + host layer at image level (camera, handedness, quadric coefficient order, CSG,
checker, light model, sky, gamma).
"""
import math
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, load_scene
from oracle_binding import OracleWorld, philox
from raysnail_b200 import (RS_FLAG_MIS, RS_FLAG_PIXEL_CENTRE, RS_REDO_NOISE, RS_REDO_NOISE_FIXED, RS_REDO_REFERENCE, Scene, make_params,
                           pass_seed, philox4x32_10, philox_uniform)


# ---------------------------------------------------------------------------------------------
# Perlin (src/texture/noise.rs), restated with numpy scalars straight from the Rust text
# ---------------------------------------------------------------------------------------------
def perlin_tables(scene, tex_index):
    d = scene.desc.contents
    t = d.textures[tex_index]
    pc = t.point_count
    perm = np.ctypeslib.as_array(d.perlin_perm, shape=(d.n_perlin_perm,))[t.perm_offset:t.perm_offset + 3 * pc].copy()
    n = pc * (3 if t.perlin_vector else 1)
    val = np.ctypeslib.as_array(d.perlin_values, shape=(d.n_perlin_values,))[t.value_offset:t.value_offset + n].copy()
    return t, perm.reshape(3, pc), (val.reshape(pc, 3) if t.perlin_vector else val)


def np_noise(t, perm, val, p):
    pc = t.point_count
    if t.perlin_smooth == 0:  # noise.rs:102-112: `(4.0 * x) as isize & (n - 1)`
        i, j, k = (int(math.trunc(4.0 * c)) & (pc - 1) for c in p)
        v = val[perm[0][i] ^ perm[1][j] ^ perm[2][k]]
        return float(v[0]) if t.perlin_vector else float(v)
    i, j, k = (int(math.floor(c)) for c in p)  # :114-119
    u, v, w = p[0] - i, p[1] - j, p[2] - k
    uu, vv, ww = u, v, w
    if t.perlin_smooth == 2:  # :160-164
        uu, vv, ww = u * u * (3.0 - 2.0 * u), v * v * (3.0 - 2.0 * v), w * w * (3.0 - 2.0 * w)
    total = 0.0
    for di in range(2):
        sj = 0.0
        for dj in range(2):
            sk = 0.0
            for dk in range(2):
                idx = perm[0][(i + di) & (pc - 1)] ^ perm[1][(j + dj) & (pc - 1)] ^ perm[2][(k + dk) & (pc - 1)]
                lat = float(np.dot(val[idx], [u - di, v - dj, w - dk])) if t.perlin_vector else float(val[idx])
                sk += (di * uu + (1 - di) * (1.0 - uu)) * (dj * vv + (1 - dj) * (1.0 - vv)) * (dk * ww + (1 - dk) * (1.0 - ww)) * lat
            sj += sk
        total += sj
    return total


def np_perlin_color(t, perm, val, p):
    if t.perlin_kind == 0:  # noise.rs:196-204
        f = np_noise(t, perm, val, [t.scale * c for c in p])
        if t.perlin_vector:
            f = 0.5 * (f + 1.0)
        return f
    q, weight, s = list(p), 1.0, 0.0  # :139-152
    for _ in range(t.perlin_depth):
        s += weight * np_noise(t, perm, val, q)
        weight *= 0.5
        q = [2.0 * c for c in q]
    f = abs(s)
    if t.perlin_kind == 2:  # :206-209
        f = (math.sin(t.scale * p[2] + 10.0 * f) + 1.0) * 0.5
    return f


PERLIN_VARIANTS = [(vector, kind, smooth) for vector in (False, True) for kind in (0, 1, 2) for smooth in (0, 1, 2)]


def perlin_scene(vector, kind, smooth, scale=1.7, point_count=64, seed=11):
    s = Scene()
    tex = s.perlin(point_count, vector, seed, kind, 7 if kind else 0, smooth, scale)
    light = s.sphere((0, 1000, 0), 1.0, s.diffuse_light(s.solid(1, 1, 1), 2.0))
    s.add(s.sphere((0, 0, 0), 1.0, s.lambertian(tex)))
    s.add(light)
    s.add_light(light)
    d = s.desc.contents
    idx = [i for i in range(d.n_textures) if d.textures[i].type == 3]
    assert len(idx) == 1
    return s, idx[0]


@pytest.mark.parametrize("vector,kind,smooth", PERLIN_VARIANTS)
def test_perlin_variants_against_numpy_restatement(vector, kind, smooth):
    s, ti = perlin_scene(vector, kind, smooth)
    t, perm, val = perlin_tables(s, ti)
    assert (t.perlin_vector, t.perlin_kind, t.perlin_smooth) == (int(vector), kind, smooth)
    rng = np.random.default_rng(5)
    pts = np.concatenate([rng.uniform(-9.0, 9.0, (200, 3)), rng.uniform(-0.3, 0.3, (50, 3)), [[0.0, 0.0, 0.0], [-1.0, 2.0, -3.0], [63.5, -64.25, 0.75]]])
    got = OracleWorld(s).texture_color(ti, np.concatenate([np.zeros((len(pts), 2)), pts], axis=1))
    exp = np.array([np_perlin_color(t, perm, val, p) for p in pts])
    assert np.abs(got[:, 0] - exp.astype(np.float32)).max() <= 4e-7 * max(1.0, np.abs(exp).max())
    assert np.array_equal(got[:, 0], got[:, 1]) and np.array_equal(got[:, 0], got[:, 2]) and (got[:, 3] == 1.0).all()  # Color(1,1,1,1) * f


def test_perlin_hand_computed_lattice_values():
    """On lattice points the interpolation weights are exactly 0 or 1 (noise.rs:156-189), so the float
    lattice returns the table entry behind perm_x[i] ^ perm_y[j] ^ perm_z[k]; `SmoothType::None` reads
    the cell of (4x, 4y, 4z) truncated towards zero; the vector lattice at a lattice point is 0 -> 0.5."""
    for smooth in (1, 2):
        s, ti = perlin_scene(False, 0, smooth, scale=1.0)
        t, perm, val = perlin_tables(s, ti)
        w = OracleWorld(s)
        for (i, j, k) in [(0, 0, 0), (3, 5, 7), (-1, -2, 70), (63, 64, -64)]:
            exp = val[perm[0][i & 63] ^ perm[1][j & 63] ^ perm[2][k & 63]]
            assert w.texture_color(ti, [[0, 0, i, j, k]])[0, 0] == np.float32(exp)
        # midway along x between two lattice points: the mean of the two entries (linear), also for Hermite (uu(0.5) = 0.5)
        a, b = val[perm[0][2] ^ perm[1][1] ^ perm[2][4]], val[perm[0][3] ^ perm[1][1] ^ perm[2][4]]
        assert abs(w.texture_color(ti, [[0, 0, 2.5, 1.0, 4.0]])[0, 0] - np.float32(0.5 * a + 0.5 * b)) < 1e-7
    s, ti = perlin_scene(False, 0, 0, scale=1.0)
    t, perm, val = perlin_tables(s, ti)
    w = OracleWorld(s)
    for p, cell in [((0.26, 0.0, 0.74), (1, 0, 2)), ((-0.26, 3.0, 15.99), (-1 & 63, 12, 63)), ((-0.2, 0.2, 16.0), (0, 0, 0))]:
        exp = val[perm[0][cell[0]] ^ perm[1][cell[1]] ^ perm[2][cell[2]]]
        assert w.texture_color(ti, [[0, 0, *p]])[0, 0] == np.float32(exp)
    s, ti = perlin_scene(True, 0, 2, scale=1.0)
    assert OracleWorld(s).texture_color(ti, [[0, 0, 4.0, -2.0, 9.0]])[0, 0] == np.float32(0.5)


# ---------------------------------------------------------------------------------------------
# Philox entry points for host harnesses (include/raysnail_b200.h "the RNG, for host harnesses")
# ---------------------------------------------------------------------------------------------
def test_exported_philox_matches_random123_and_the_oracle():
    assert [hex(x) for x in philox4x32_10([0, 0, 0, 0], [0, 0])] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in philox4x32_10([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in philox4x32_10([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0])] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]
    rng = np.random.default_rng(1)
    for _ in range(200):
        c, k = rng.integers(0, 2 ** 32, 4, dtype=np.uint64).astype(np.uint32), rng.integers(0, 2 ** 32, 2, dtype=np.uint64).astype(np.uint32)
        assert np.array_equal(philox4x32_10(c, k), philox(c, k))
    seed = 0x0123456789ABCDEF
    u = philox_uniform(seed, 77, 5, 3, 5, 2)
    w = philox([77, 5, 3, 5 | (2 << 8)], [seed & 0xFFFFFFFF, seed >> 32])
    assert np.array_equal(u, (w >> 8).astype(np.float64) / 2.0 ** 24)
    assert ((u >= 0.0) & (u < 1.0)).all()
    assert pass_seed(seed, 0) == seed and pass_seed(seed, 3) == (seed + 3 * 0x9E3779B97F4A7C15) % 2 ** 64


def test_oracle_draws_the_exported_stream():
    """The jitter of pixel p / sample s is (word0, word1) of (p, s, depth 0, RS_RNG_PIXEL): the camera ray
    the oracle generates must be the one computed here from rs_philox_uniform (painter.rs:167-172)."""
    s = load_scene("example")
    w, h, q = 12, 8, 3
    cam = s.camera(w, h)
    params = make_params(sqrt_spp=q)
    o = OracleWorld(s).debug_primary(cam, params, 4)  # s_j = 1, s_i = 1
    for pixel in (0, 17, w * h - 1):
        row, col = divmod(pixel, w)
        u = philox_uniform(params.seed, pixel, 4, 0, 0, 0)
        xo, yo = col + (1 + u[0]) / q, row + (1 + u[1]) / q
        uu, vv = xo / w, (h - 1.0 - yo) / h
        # aperture 0.01: the lens offset moves the origin by <= 0.005; compare the point on the focus plane
        target = np.array(cam.lb) + uu * np.array(cam.horizontal_full) + vv * np.array(cam.vertical_full)
        ray_o, ray_d = o["ray"][pixel, :3], o["ray"][pixel, 3:]
        t = np.linalg.norm(target - ray_o)
        assert np.abs(ray_o + t * ray_d - target).max() < 1e-9


# ---------------------------------------------------------------------------------------------
# materials / integrator
# ---------------------------------------------------------------------------------------------
def probe(scene, look_from, look_at, flags=RS_FLAG_PIXEL_CENTRE, what="scatter", seed=None):
    scene.set_camera(look_from, look_at, (0, 1, 0), fov=40.0, aperture=0.0, focus=1.0)
    cam = scene.camera(3, 3)
    params = make_params(samples=1, flags=flags) if seed is None else make_params(samples=1, flags=flags, seed=seed)
    w = OracleWorld(scene)
    o = w.debug_primary(cam, params, 0) if what == "primary" else w.debug_scatter(cam, params, 0)
    return {k: v[1] for k, v in o.items()}


def test_light_under_a_mixed_material_emits_nothing():
    """mixed_material.rs:53-55: MixedMaterial::emitted is None, and the DiffuseLight it may select does not
    scatter (light.rs) => the hit contributes nothing, whichever branch the draw takes."""
    for prob in (0.0, 1.0, 0.5):
        s = Scene()
        lamp = s.diffuse_light(s.solid(1, 1, 1), 5.0)
        mixed = s.mixed(lamp, s.lambertian(s.solid(0.5, 0.5, 0.5)), prob)
        far = s.sphere((0, 1000, 0), 1.0, lamp)
        s.add(s.sphere((0, 0, 0), 1.0, mixed))
        s.add(far)
        s.add_light(far)
        for seed in range(6):
            r = probe(s, (0, 0, 5), (0, 0, 0), seed=1000 + seed)
            assert np.array_equal(r["emitted"], [0.0, 0.0, 0.0])
            assert r["kind"] in (1, 3, 4)
            if prob == 1.0:
                assert r["kind"] == 1  # the light was selected: no scatter, no emission
            if prob == 0.0:
                assert r["kind"] >= 3


def test_mis_flag_weights_a_lambertian_sample_by_the_mixture_density():
    """RS_FLAG_MIS: weight = albedo * CosinePdf(n).value(d) / (0.5 * CosinePdf(d).value(d) + 0.5 * CosinePdf(n).value(d))
    = albedo * 2 cos / (1 + cos) (pdf.rs:237-307, camera.rs:222-233); the ray starts at the hit point."""
    s = Scene()
    lamp = s.sphere((0, 50, 0), 5.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
    s.add(s.sphere((0, 0, 0), 1.0, s.lambertian(s.solid(0.5, 0.25, 0.75))))
    s.add(lamp)
    s.add_light(lamp)
    kinds = set()
    for seed in range(24):
        hit = probe(s, (0, 3, 4), (0, 0, 0), what="primary", seed=seed)
        r = probe(s, (0, 3, 4), (0, 0, 0), flags=RS_FLAG_PIXEL_CENTRE | RS_FLAG_MIS, seed=seed)
        kinds.add(int(r["kind"]))
        d, n = r["next"][3:], hit["normal"]
        assert abs(np.linalg.norm(d) - 1.0) < 1e-12 and np.abs(r["next"][:3] - hit["point"]).max() < 1e-12
        cos = max(float(np.dot(d, n)), 0.0)
        exp = np.array([0.5, 0.25, 0.75]) * (2.0 * cos / (1.0 + cos))
        assert np.abs(r["weight"] - exp).max() < 1e-9
        plain = probe(s, (0, 3, 4), (0, 0, 0), seed=seed)
        assert plain["kind"] == r["kind"]
        if plain["kind"] == 4:  # the live integrator: BRDF sample, weight = albedo
            assert np.abs(plain["weight"] - [0.5, 0.25, 0.75]).max() < 1e-12 and np.abs(plain["next"][3:] - d).max() < 1e-12
    assert kinds == {3, 4}


def test_degenerate_pdf_values_follow_camera_rs_236_243():
    """camera.rs:236-243: a pdf value <= 0 or NaN becomes 1e-5 and the scattering value is divided by it.
    BlinnPhong(k_specular 1, exponent -1): normal_pdf = 0/(2 pi) * ... = 0, cosine term * (1 - 1) = 0 => value 0 => weight 0.
    BlinnPhong(k_specular 3, exponent -1): value = cos/pi * (1 - 3) < 0 => weight = albedo * value / 1e-5 (negative).
    BlinnPhong(k_specular NaN): value NaN => weight NaN."""
    for ks, expect in ((1.0, "zero"), (3.0, "negative"), (float("nan"), "nan")):
        s = Scene()
        lamp = s.sphere((0, 50, 0), 5.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
        s.add(s.sphere((0, 0, 0), 1.0, s.blinn_phong(ks, -1.0, s.solid(0.5, 0.5, 0.5))))
        s.add(lamp)
        s.add_light(lamp)
        seen = 0
        at = (0.45, 0.3, 0.0)  # off-centre, so that the reflected direction is not the normal
        for seed in range(16):
            r = probe(s, (0, 3, 4), at, seed=seed)
            if r["kind"] != 4:
                continue
            seen += 1
            hit = probe(s, (0, 3, 4), at, what="primary", seed=seed)
            if expect == "zero":
                assert np.array_equal(r["weight"], [0.0, 0.0, 0.0])
            elif expect == "nan":
                assert np.isnan(r["weight"]).all()
            else:
                cos = float(np.dot(r["next"][3:], hit["normal"]))
                assert cos > 1e-3 and np.abs(r["weight"] / (0.5 * (cos / math.pi) * (1.0 - 3.0) / 1e-5) - 1.0).max() < 1e-9
        assert seen >= 3


# ---------------------------------------------------------------------------------------------
# passes loop (src/bin/raysnail.rs:379-427)
# ---------------------------------------------------------------------------------------------
def test_passes_loop_of_the_oracle():
    import ctypes as C
    from oracle_binding import load as load_oracle
    ol = load_oracle()
    s = load_scene("example")
    w = OracleWorld(s)
    cam = s.camera(40, 24)
    params = make_params(samples=4)
    one, c1 = w.render_passes(cam, params, 1)
    plain, _ = w.render(cam, params)
    assert np.array_equal(one, plain)  # pass 0: (old * 0 + new) / 1 with the base seed
    ref3, counts = w.render_passes(cam, params, 3, RS_REDO_REFERENCE)
    # by hand: three full renders with the pass seeds, running mean in gamma space (combine_pixel)
    old = np.zeros((24, 40, 4), np.float32)
    old[..., 3] = 1.0
    imgs = []
    for k in range(3):
        img, _ = w.render(cam, make_params(samples=4, seed=pass_seed(params.seed, k)))
        imgs.append(img)
        out = np.empty_like(old)
        ol.rso_combine_passes(old.ctypes.data, img.ctypes.data, out.ctypes.data, 40 * 24, C.c_float(float(k)))
        old = out
    assert np.array_equal(ref3, old)
    assert not np.array_equal(imgs[0], imgs[1])  # fresh random numbers every pass
    # averaging passes reduces the noise: pass-0 image vs the 3-pass image against a converged render
    conv, _ = w.render(cam, make_params(samples=400, seed=99))
    e1 = np.sqrt(np.mean((imgs[0][..., :3] - conv[..., :3]) ** 2))
    e3 = np.sqrt(np.mean((ref3[..., :3] - conv[..., :3]) ** 2))
    assert e3 < 0.8 * e1
    assert len(counts) == 3 and (counts >= 0).all() and (counts <= 40 * 24).all()
    # opt-in redo maps: pixels outside the map keep their previous value exactly
    for mode in (RS_REDO_NOISE, RS_REDO_NOISE_FIXED):
        two, c2 = w.render_passes(cam, params, 2, mode, threshold=0.05)
        noise = np.empty((24, 40), np.float32)
        ol.rso_noise_map(one.ctypes.data, noise.ctypes.data, 40, 24, 0 if mode == RS_REDO_NOISE_FIXED else 1)
        keep = noise < 0.05
        assert c2[0] == int((~keep).sum())
        assert np.array_equal(two[keep], one[keep])
        if (~keep).any():
            assert not np.array_equal(two[~keep], one[~keep])


# ---------------------------------------------------------------------------------------------
# the reference's own render
# ---------------------------------------------------------------------------------------------
def box_down(a, k=8):
    h, w, c = a.shape
    return a[:h // k * k, :w // k * k].reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


def compare_with_reference_jpeg(img):
    """img: (500, 800, >=3) f32 render of sdl/quadric.sdl.  Returns (per-channel correlation, relative RMSE,
    relative mean-colour difference) after the CLI's PNG quantisation (raysnail.rs:437-441) and 8x box down-sampling."""
    from PIL import Image
    ref = np.asarray(Image.open(os.path.join(GOLDEN_DIR, "sdl_quadrics.jpg")).convert("RGB")).astype(np.float64) / 255.0
    ours = np.floor(np.clip(img[..., :3].astype(np.float64), 0.0, 1.0) * 255.5) / 255.0
    a, b = box_down(ref), box_down(ours)
    corr = [float(np.corrcoef(a[..., c].ravel(), b[..., c].ravel())[0, 1]) for c in range(3)]
    rmse = float(np.sqrt(np.mean((a - b) ** 2)) / np.sqrt(np.mean(a ** 2)))
    mean = float(np.abs(a.mean(axis=(0, 1)) - b.mean(axis=(0, 1))).max() / a.mean())
    return corr, rmse, mean, a, b


# 8x8-pixel cells (row, column) inside each solid of examples/sdl_quadrics.jpg, left to right, with the channel order
# their material shows under the yellowish light (sdl/quadric.sdl:22-84: sand, yellow-green, mint, blue)
QUADRIC_PROBES = [("paraboloid", (30, 25), (0, 1, 2)), ("cone", (31, 40), (1, 0, 2)), ("cylinder", (30, 60), (1, 2, 0)), ("hyperboloid", (24, 80), (2, 1, 0))]


def check_object_layout(a, b):
    """Same solids at the same places, left to right, in the reference's image `a` and ours `b`."""
    for name, (r, c), order in QUADRIC_PROBES:
        for img in (a, b):
            px = img[r, c]
            assert px[order[0]] > px[order[1]] > px[order[2]], (name, px)
        assert np.abs(a[r, c] - b[r, c]).max() < 0.04, (name, a[r, c], b[r, c])
    # and the sky / ground split: top rows are the sky gradient, the bottom-left corner is the dark ground
    assert np.abs(a[2, 50] - b[2, 50]).max() < 0.03 and a[2, 50][2] > 0.8
    assert np.abs(a[60, 2] - b[60, 2]).max() < 0.03 and a[60, 2].max() < 0.3


def test_oracle_render_of_quadric_sdl_matches_the_references_own_image():
    """examples/sdl_quadrics.jpg is the only output of the Rust program in its repository.  It is a lossy JPEG
    at an unknown sample count, so this is not a bit-level parity target — but agreement at this level (measured:
    correlation 0.9998, 8x-down-sampled RMSE 1.0 %, mean colour 0.3 %) cannot be reached with a wrong camera,
    handedness, coefficient order, CSG rule, checker phase, light or gamma."""
    s = load_scene("quadric")
    img, _ = OracleWorld(s).render(s.camera(800, 500), make_params(samples=16))
    corr, rmse, mean, a, b = compare_with_reference_jpeg(img)
    assert min(corr) > 0.999, corr
    assert rmse < 0.02, rmse
    assert mean < 0.01, mean
    check_object_layout(a, b)


def test_mixed_settings_forward_to_material_1_recursively():
    """mixed_material.rs:56-58: settings() = material_1.settings(), which recurses through nested mixes."""
    s = Scene()
    a = s.lambertian(s.solid(0.5, 0.5, 0.5))
    s.set_phong(a, 2.0, 7)
    b = s.metal(s.solid(0.9, 0.9, 0.9))
    inner = s.mixed(a, b, 0.5)
    outer = s.mixed(inner, b, 0.5)
    lamp = s.sphere((0, 50, 0), 5.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
    s.add(s.sphere((0, 0, 0), 1.0, outer))
    s.add(lamp)
    s.add_light(lamp)
    d = s.desc.contents
    mixed = [d.materials[i] for i in range(d.n_materials) if d.materials[i].type == 7]
    assert len(mixed) == 2
    for m in mixed:
        assert (m.phong_factor, m.phong_exponent) == (2.0, 7)


# ---------------------------------------------------------------------------------------------
# a BVH used as one object (bvh.rs:173-192)
# ---------------------------------------------------------------------------------------------
def _bvh_members(s, mats):
    """five solids scattered over x / y so that both sort axes of BVH::new matter"""
    return [s.sphere((-2.0, 0.5, 0.0), 0.5, mats[0]), s.box((-1.2, 0.0, -0.4), (-0.4, 0.8, 0.4), mats[1]), s.sphere((0.3, 1.4, 0.2), 0.45, mats[2]),
            s.box((1.0, 0.0, -0.3), (1.6, 1.6, 0.3), mats[3]), s.sphere((2.3, 0.4, -0.1), 0.4, mats[4])]


def _bvh_world(grouped, transform=None):
    s = Scene()
    s.add(s.sphere((0, -1000, 0), 1000.0, s.lambertian(s.solid(0.5, 0.5, 0.5))))
    light = s.sphere((0, 12, 3), 3.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
    s.add(light)
    s.add_light(light)
    mats = [s.lambertian(s.solid(0.1 + 0.15 * k, 0.8 - 0.1 * k, 0.3)) for k in range(5)]
    members = _bvh_members(s, mats)
    if grouped:
        obj = s.bvh(members)
        s.add(s.transform(obj, transform) if transform else obj)
    else:
        for m in members:
            s.add(s.transform(m, transform) if transform else m)
    s.set_camera((0.5, 2.0, 7.0), (0.0, 0.6, 0.0), fov=45.0)
    return s


@pytest.mark.parametrize("transform", [None, [("rotate_y", 0.7), ("scale", (1.0, 1.3, 0.8)), ("translate", (0.2, 0.1, -0.5))]], ids=["bare", "tf_facade"])
def test_a_bvh_as_one_object_finds_the_hits_of_its_members(transform):
    """BVH::hit returns the closest member hit (bvh.rs:173-192): the same surfaces as the members
    added to the world one by one — same t, normal, side and material wherever either is hit; one object id instead of five."""
    cam_params = make_params(samples=1, flags=RS_FLAG_PIXEL_CENTRE)
    g, f = _bvh_world(True, transform), _bvh_world(False, transform)
    a = OracleWorld(g).debug_primary(g.camera(160, 100), cam_params, 0)
    b = OracleWorld(f).debug_primary(f.camera(160, 100), cam_params, 0)
    assert np.array_equal(a["obj_id"] >= 2, b["obj_id"] >= 2)
    grp = a["obj_id"] == 2
    assert grp.sum() > 800 and len(np.unique(b["obj_id"][grp])) == 5
    assert np.array_equal(a["t"][grp], b["t"][grp])
    assert np.array_equal(a["normal"][grp], b["normal"][grp])
    assert np.array_equal(a["front"][grp], b["front"][grp])
    # (material ids are handed out in flatten order, which differs between the two worlds: the pairing must be a bijection)
    pairs = set(zip(a["material"][grp].tolist(), b["material"][grp].tolist()))
    assert len(pairs) == 5 and len({p[0] for p in pairs}) == 5 and len({p[1] for p in pairs}) == 5


def test_bvh_node_layout_follows_bvh_new():
    """bvh.rs:47-112: one object -> left only; two -> both; more -> sorted by the bbox minimum, split in the middle; every
    node's bbox is the union of its children's."""
    s = Scene()
    mats = [s.lambertian(s.solid(0.5, 0.5, 0.5))] * 5
    s.add(s.bvh(_bvh_members(s, mats)))
    light = s.sphere((0, 12, 3), 3.0, s.diffuse_light(s.solid(1, 1, 1), 5.0))
    s.add(light)
    s.add_light(light)
    d = s.desc.contents
    nodes = [d.nodes[i] for i in range(d.n_nodes)]
    bvh = [n for n in nodes if n.type == 9]
    assert len(bvh) == 5  # 5 -> (2, 3), 3 -> (1, 2): the root, its pair, its triple, the triple's single and the triple's pair
    root = nodes[d.roots[0]]
    assert root.type == 9 and root.p[6] == 1.0

    def leaves(n):
        if n.type != 9:
            return 1
        return sum(leaves(nodes[c]) for c in (n.child0, n.child1) if c >= 0)

    def box_of(n):
        if n.type == 9:
            return np.array(n.p[:6])
        if n.type == 0:
            c, r = np.array(n.p[:3]), n.p[3]
            return np.concatenate([c - r, c + r])
        return np.array(n.p[:6])

    assert leaves(root) == 5 and leaves(nodes[root.child0]) == 2 and leaves(nodes[root.child1]) == 3
    for n in bvh:
        kids = [box_of(nodes[c]) for c in (n.child0, n.child1) if c >= 0]
        lo = np.min([k[:3] for k in kids], axis=0)
        hi = np.max([k[3:] for k in kids], axis=0)
        assert np.allclose(n.p[:3], lo, atol=2e-4) and np.allclose(n.p[3:6], hi, atol=2e-4)  # (Box::bbox pads by 0.0001, box.rs)
    single = [n for n in bvh if n.child1 < 0]
    assert len(single) == 1 and single[0].child0 >= 0
