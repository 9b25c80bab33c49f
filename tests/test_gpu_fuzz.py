"""Randomised parity: seeded random worlds built from every node type (spheres static and moving, boxes, rects, quadrics,
Difference / Intersection of random operands up to three levels deep, TfFacade stacks, ConstantMedium) and every material,
GPU against the CPU oracle on primary hits (ids bit-exact) and one shaded bounce — the same bars as test_gpu_parity.py, on
scenes nobody hand-picked.  Small worlds exercise the object list (and its skip-ahead walk when the trees happen to share a
shape), large ones the tree; both BVH widths run.
"""
import math

import numpy as np
import pytest

from oracle_binding import OracleWorld
from raysnail_b200 import RS_FLAG_PIXEL_CENTRE, Scene, make_params

pytestmark = pytest.mark.gpu

W, H = 128, 80
REL = 1e-5


def _rel_err(a, b, floor=1e-12):
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)


def random_world(seed, n_objects, uniform_trees=False):
    rng = np.random.default_rng(seed)
    s = Scene()
    u = lambda a, b: float(rng.uniform(a, b))  # noqa: E731

    def colour():
        return (u(0.1, 0.9), u(0.1, 0.9), u(0.1, 0.9))

    def material():
        k = int(rng.integers(0, 7))
        if k == 0:
            return s.lambertian(s.solid(*colour()))
        if k == 1:
            return s.lambertian(s.checker(s.solid(*colour()), s.solid(*colour()), u(1.0, 6.0)))
        if k == 2:
            return s.metal(s.solid(*colour()))
        if k == 3:
            return s.diffuse_metal(u(5.0, 200.0), s.solid(*colour()))
        if k == 4:
            return s.dielectric(colour(), u(1.2, 1.8), glass_curve=bool(rng.integers(0, 2)))
        if k == 5:
            return s.blinn_phong(u(0.1, 0.9), u(4.0, 60.0), s.solid(*colour()))
        return s.mixed(s.lambertian(s.solid(*colour())), s.metal(s.solid(*colour())), u(0.2, 0.8))

    def primitive(mat, allow_rect=True, centre=None, bounded=True):
        c = centre if centre is not None else (u(-1.0, 1.0), u(-1.0, 1.0), u(-1.0, 1.0))
        k = int(rng.integers(0, 4 if allow_rect else 3))
        if k == 0:
            return s.sphere(c, u(0.4, 1.2), mat)
        if k == 1:
            h = (u(0.3, 1.0), u(0.3, 1.0), u(0.3, 1.0))
            return s.box((c[0] - h[0], c[1] - h[1], c[2] - h[2]), (c[0] + h[0], c[1] + h[1], c[2] + h[2]), mat)
        if k == 2:
            # ellipsoid / cylinder / cone / hyperboloid about the origin (quadric.rs coefficient order: a b c d e f g h i j)
            kind = 0 if bounded else int(rng.integers(0, 4))  # (an unbounded quadric only as the second operand of a CSG node)
            q = {0: [u(0.5, 2.0), 0, 0, 0, u(0.5, 2.0), 0, 0, u(0.5, 2.0), 0, -1.0],
                 1: [1.0, 0, 0, 0, 0.0, 0, 0, 1.0, 0, -u(0.3, 1.0)],
                 2: [1.0, 0, 0, 0, -1.0, 0, 0, 1.0, 0, 0.0],
                 3: [1.0, 0, 0, 0, -1.0, 0, 0, 1.0, 0, -u(0.2, 0.8)]}[kind]
            return s.quadric(q, mat)
        axis = ["xy", "xz", "yz"][int(rng.integers(0, 3))]
        return s.rect(axis, u(-0.5, 0.5), (u(-1.5, -0.2), u(0.2, 1.5)), (u(-1.5, -0.2), u(0.2, 1.5)), mat)

    def solid(depth, mat):
        """a closed shape CSG may combine (no rects: they contain nothing)"""
        if depth == 0 or rng.random() < 0.4:
            return primitive(mat if rng.random() < 0.8 else -1, allow_rect=False)
        if rng.random() < 0.5:
            # the minus operand is a primitive: the cavity's surface takes `minus.normal(p)`, which only primitives
            # implement (hit.rs:82-93: `unimplemented!()` otherwise — the reference aborts there)
            node = s.difference(solid(depth - 1, mat), primitive(material() if rng.random() < 0.8 else -1, allow_rect=False, bounded=False), mat)
        else:
            second = primitive(material(), allow_rect=False, bounded=False) if rng.random() < 0.5 else solid(depth - 1, material())
            node = s.intersection(solid(depth - 1, mat), second, mat)
        if rng.random() < 0.4:
            node = s.transform(node, transform_ops(0.5))
        return node

    def transform_ops(spread):
        ops = []
        for _ in range(int(rng.integers(1, 4))):
            k = int(rng.integers(0, 5))
            if k == 0:
                ops.append(("translate", (u(-spread, spread), u(-spread, spread), u(-spread, spread))))
            elif k == 4:
                ops.append(("scale", (u(0.6, 1.6), u(0.6, 1.6), u(0.6, 1.6))))
            else:
                ops.append((["rotate_x", "rotate_y", "rotate_z"][k - 1], u(-math.pi, math.pi)))
        return ops

    s.add(s.sphere((0.0, -1003.0, 0.0), 1000.0, s.lambertian(s.checker(s.solid(0.2, 0.3, 0.1), s.solid(0.9, 0.9, 0.9), 2.0))))
    lights = [s.sphere((u(-8, 8), u(9, 14), u(-8, 8)), u(1.0, 2.5), s.diffuse_light(s.solid(1.0, 0.9, 0.8), u(3.0, 9.0)))]
    if rng.random() < 0.6:
        lights.append(s.rect("xz", u(10.0, 13.0), (-2.0, 2.0), (-2.0, 2.0), s.diffuse_light(s.solid(0.8, 0.9, 1.0), u(2.0, 6.0))))
    for l in lights:
        s.add(l)
        s.add_light(l)
    side = max(1, int(math.ceil(math.sqrt(n_objects))))
    for k in range(n_objects):
        cell = (-1.6 * side + 3.2 * (k % side) + 1.6, u(-1.0, 2.0), -1.6 * side + 3.2 * (k // side) + 1.6)
        place = [("translate", cell)]
        mat = material()
        if uniform_trees:
            # every object the same expression shape (the skip-ahead list walk): TfFacade(Intersection(quadric, box))
            q = s.quadric([1.0, 0, 0, 0, [0.0, -1.0][k % 2], 0, 0, 1.0, 0, -u(0.3, 1.0)])
            obj = s.transform(s.intersection(q, s.box((-1, -1, -1), (1, 1, 1)), mat), [("rotate_y", u(-3, 3))] + place)
        else:
            pick = rng.random()
            if pick < 0.30:
                obj = s.transform(primitive(mat), transform_ops(0.3) + place)
            elif pick < 0.45:
                obj = primitive(mat, centre=cell)
            elif pick < 0.55:
                obj = s.sphere(cell, u(0.4, 1.0), mat, speed=(u(-0.5, 0.5), u(0.0, 0.8), u(-0.5, 0.5)))
            elif pick < 0.65:
                obj = s.transform(s.medium(primitive(-1, allow_rect=False), colour(), u(0.3, 2.0)), place)
            elif pick < 0.75:
                # a BVH as one object (bvh.rs:173-192): a few primitives, sometimes a nested group, under the cell's transform
                members = [primitive(material(), centre=(u(-1.0, 1.0), u(-0.8, 0.8), u(-1.0, 1.0))) for _ in range(int(rng.integers(1, 6)))]
                if rng.random() < 0.4:
                    members.append(s.bvh([primitive(material(), centre=(u(-1.0, 1.0), u(-0.8, 0.8), u(-1.0, 1.0))) for _ in range(int(rng.integers(1, 4)))]))
                obj = s.transform(s.bvh(members), (transform_ops(0.3) if rng.random() < 0.5 else []) + place)
            else:
                obj = s.transform(solid(int(rng.integers(1, 4)), mat), transform_ops(0.3) + place)
        s.add(obj)
    reach = 3.2 * side
    s.set_time_range(0.0, 1.0)
    s.set_camera((u(-0.3, 0.3) * reach, u(2.0, 6.0), 1.2 * reach + 4.0), (0.0, 0.0, 0.0), fov=u(35.0, 60.0), aperture=[0.0, 0.1][seed % 2],
                 focus=1.2 * reach + 4.0, shutter_speed=1.0)
    return s


CASES = [(101, 5, False), (102, 9, False), (103, 11, True), (104, 20, False), (105, 40, False), (106, 6, True), (107, 90, False), (108, 23, False),
         (109, 3, False), (110, 12, False), (111, 13, False), (112, 24, False), (113, 25, False), (114, 160, False), (115, 22, True), (116, 30, True),
         (117, 7, False), (118, 16, False), (119, 60, False), (120, 10, False), (121, 33, False), (122, 4, True), (123, 18, False), (124, 250, False)]


@pytest.mark.parametrize("seed,n_objects,uniform", CASES)
@pytest.mark.parametrize("bvh_width", [2, 4])
def test_random_worlds_match_the_oracle(seed, n_objects, uniform, bvh_width):
    from raysnail_b200 import Renderer
    scene = random_world(seed, n_objects, uniform)
    oracle = OracleWorld(scene)
    cam = scene.camera(W, H)
    r = Renderer(0)
    try:
        r.set_bvh_width(bvh_width)
        r.upload(scene)
        for flags, sample in ((RS_FLAG_PIXEL_CENTRE, 0), (0, 3)):
            params = make_params(samples=4, flags=flags)
            g, o = r.debug_primary(cam, params, sample), oracle.debug_primary(cam, params, sample)
            mism = np.flatnonzero(g["obj_id"] != o["obj_id"])
            assert mism.size == 0, (f"seed {seed}: {mism.size} ids differ, pixels {mism[:5]}: gpu {g['obj_id'][mism[:5]]} t {g['t'][mism[:5]]} "
                                    f"oracle {o['obj_id'][mism[:5]]} t {o['t'][mism[:5]]}")
            hit = o["obj_id"] >= 0
            assert hit.sum() > W * H // 4
            assert _rel_err(g["t"][hit], o["t"][hit]).max() < REL
            assert np.abs(g["normal"][hit] - o["normal"][hit]).max() < REL * max(1.0, np.abs(o["normal"][hit]).max())
            for key in ("front", "material"):
                bad = np.flatnonzero(hit & (g[key] != o[key]))
                assert bad.size == 0, f"seed {seed}: {key} differs at {bad[:5]} (ids {o['obj_id'][bad[:5]]})"
        params = make_params(samples=4)
        g, o = r.debug_scatter(cam, params, 1), oracle.debug_scatter(cam, params, 1)
        mism = np.flatnonzero(g["kind"] != o["kind"])
        assert mism.size == 0, f"seed {seed}: scatter kind differs at {mism[:5]}: gpu {g['kind'][mism[:5]]} oracle {o['kind'][mism[:5]]}"
        sc = o["kind"] >= 2
        # (a degenerate sample is NaN where the reference's arithmetic makes it so — a zero-length scatter direction, a 0 / 0
        #  pdf ratio — and then on both sides)
        for key in ("next", "weight", "emitted"):
            assert np.array_equal(np.isnan(g[key]), np.isnan(o[key])), f"seed {seed}: NaN pattern of {key} differs"
        ok = sc & ~np.isnan(o["next"]).any(axis=1) & ~np.isnan(o["weight"]).any(axis=1)
        assert ok.sum() > 0.98 * sc.sum()
        assert np.abs(g["next"][ok] - o["next"][ok]).max() < REL * max(1.0, np.abs(o["next"][ok]).max())
        assert _rel_err(g["weight"][ok], o["weight"][ok], floor=1e-6).max() < 1e-4
        fe = ~np.isnan(o["emitted"]).any(axis=1)
        assert np.abs(g["emitted"][fe] - o["emitted"][fe]).max() < 1e-5 * max(1.0, np.abs(o["emitted"][fe]).max())
        # and a short render: finite, and the oracle's image within Monte-Carlo-free tolerance
        small = scene.camera(48, 30)
        gi = r.render(small, make_params(samples=16))
        oi, _ = oracle.render(small, make_params(samples=16))
        # (a pixel is NaN where the reference's arithmetic makes it so — a 0 / 0 pdf ratio, camera.rs:236-243 — on both sides)
        fin_o, fin_g = np.isfinite(oi).all(axis=-1), np.isfinite(gi).all(axis=-1)
        assert (fin_o != fin_g).mean() < 0.002, f"seed {seed}: finite masks differ on {(fin_o != fin_g).sum()} pixels"
        fin = fin_o & fin_g
        assert fin.mean() > 0.9
        rmse = float(np.sqrt(np.mean((gi[fin][:, :3] - oi[fin][:, :3]) ** 2))) / max(float(np.sqrt(np.mean(oi[fin][:, :3] ** 2))), 1e-3)
        assert rmse < 0.01, f"seed {seed}: relative RMSE {rmse}"
    finally:
        r.close()


def random_lean_world(seed, n_spheres, n_tris):
    """Static spheres and a random triangle soup only — the lean tier: batch extend on bounce 0, the streaming extend kernel
    from bounce 1 on when triangles are present, the staged shade kernels.  Some triangles are axis-parallel (a bbox with no
    thickness: never hit, triangle_mesh.rs:43-45 / aabb.rs:33) and some are slivers."""
    rng = np.random.default_rng(seed)
    s = Scene()
    u = lambda a, b: float(rng.uniform(a, b))  # noqa: E731
    mats = [s.lambertian(s.solid(u(0.1, 0.9), u(0.1, 0.9), u(0.1, 0.9))), s.metal(s.solid(0.8, 0.8, 0.7)), s.dielectric((1, 1, 1), 1.5),
            s.blinn_phong(0.4, 30.0, s.solid(0.3, 0.5, 0.8)), s.diffuse_metal(40.0, s.solid(0.9, 0.6, 0.3)),
            s.lambertian(s.checker(s.solid(0.1, 0.1, 0.1), s.solid(0.9, 0.9, 0.9), 4.0))]
    s.add(s.sphere((0.0, -1000.5, 0.0), 1000.0, mats[5]))
    lights = [s.sphere((u(-6, 6), u(8, 12), u(-6, 6)), u(1.0, 2.0), s.diffuse_light(s.solid(1, 1, 1), u(4.0, 9.0)))]
    if seed % 2:
        lights.append(s.rect("xz", 11.0, (-2.0, 2.0), (-2.0, 2.0), s.diffuse_light(s.solid(1.0, 0.9, 0.8), 5.0)))
    for l in lights:
        s.add(l)
        s.add_light(l)
    for _ in range(n_spheres):
        s.add(s.sphere((u(-5, 5), u(0.0, 3.0), u(-5, 5)), u(0.15, 0.7), mats[int(rng.integers(0, 5))]))
    if n_tris:
        centres = rng.uniform([-5, 0.0, -5], [5, 3.5, 5], size=(n_tris, 3))
        tri = centres[:, None, :] + rng.normal(0.0, 0.45, size=(n_tris, 3, 3))
        tri[::17, :, 1] = tri[::17, :1, 1]            # axis-parallel: all three vertices at one height
        tri[5::23, 2] = tri[5::23, 1] + 1e-3 * (tri[5::23, 2] - tri[5::23, 1])  # slivers
        pos = tri.reshape(-1, 3).astype(np.float32)
        idx = np.arange(3 * n_tris, dtype=np.uint32)
        for k in range(3):  # three materials: three meshes
            sel = np.arange(k, n_tris, 3)
            s.add_mesh(pos, idx.reshape(-1, 3)[sel].reshape(-1), mat=mats[[0, 1, 3][k]])
    s.set_camera((u(-2, 2), u(3.0, 5.0), 13.0), (0.0, 1.0, 0.0), fov=45.0, aperture=[0.0, 0.08][seed % 2], focus=13.0)
    return s


LEAN_CASES = [(201, 40, 0), (202, 10, 300), (203, 200, 2000), (204, 0, 5000), (205, 3, 40), (206, 600, 600)]


@pytest.mark.parametrize("seed,n_spheres,n_tris", LEAN_CASES)
@pytest.mark.parametrize("bvh_width", [2, 4])
def test_random_lean_worlds_match_the_oracle(seed, n_spheres, n_tris, bvh_width):
    from raysnail_b200 import Renderer
    scene = random_lean_world(seed, n_spheres, n_tris)
    oracle = OracleWorld(scene)
    cam = scene.camera(W, H)
    r = Renderer(0)
    try:
        r.set_bvh_width(bvh_width)
        r.upload(scene)
        for flags, sample in ((RS_FLAG_PIXEL_CENTRE, 0), (0, 2)):
            params = make_params(samples=4, flags=flags)
            g, o = r.debug_primary(cam, params, sample), oracle.debug_primary(cam, params, sample)
            mism = np.flatnonzero(g["obj_id"] != o["obj_id"])
            assert mism.size == 0, f"seed {seed}: {mism.size} ids differ at {mism[:5]}: gpu {g['obj_id'][mism[:5]]} oracle {o['obj_id'][mism[:5]]}"
            hit = o["obj_id"] >= 0
            assert _rel_err(g["t"][hit], o["t"][hit]).max() < REL
            assert np.abs(g["normal"][hit] - o["normal"][hit]).max() < REL * max(1.0, np.abs(o["normal"][hit]).max())
            for key in ("front", "material"):
                assert np.array_equal(g[key][hit], o[key][hit]), key
        params = make_params(samples=4)
        g, o = r.debug_scatter(cam, params, 1), oracle.debug_scatter(cam, params, 1)
        assert np.array_equal(g["kind"], o["kind"])
        sc = (o["kind"] >= 2) & ~np.isnan(o["next"]).any(axis=1)
        assert np.abs(g["next"][sc] - o["next"][sc]).max() < REL * max(1.0, np.abs(o["next"][sc]).max())
        assert _rel_err(g["weight"][sc], o["weight"][sc], floor=1e-6).max() < 1e-4
        # the whole wavefront, every bounce (the streaming extend kernel on the meshes): a converged image
        small = scene.camera(64, 40)
        gi = r.render(small, make_params(samples=64))
        oi, _ = oracle.render(small, make_params(samples=64))
        fin = np.isfinite(oi).all(axis=-1) & np.isfinite(gi).all(axis=-1)
        assert fin.mean() > 0.995
        rmse = float(np.sqrt(np.mean((gi[fin][:, :3] - oi[fin][:, :3]) ** 2))) / max(float(np.sqrt(np.mean(oi[fin][:, :3] ** 2))), 1e-3)
        assert rmse < 0.01, f"seed {seed}: relative RMSE {rmse}"
        st = r.stats()
        assert st.n_samples == 64 * 40 * 64
    finally:
        r.close()
