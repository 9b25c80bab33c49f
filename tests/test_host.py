"""CPU-only tests of the host layer and of the C-ABI boundary (no compute calls: there is no
GPU here, and the product has no CPU fallback)."""
import ctypes as C
import math
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ALL_SCENES, ROOT, load_scene
from oracle_binding import OracleWorld
from raysnail_b200 import CameraBuilder, Scene, _abi, make_params, painter_samples

HEADERS = [os.path.join(ROOT, "include", "raysnail_b200.h"), os.path.join(ROOT, "include", "raysnail_b200_host.h")]


def declared_functions(path):
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rs_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _abi.load()
    declared = declared_functions(HEADERS[0]) + declared_functions(HEADERS[1])
    assert len(declared) > 50
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/ but not exported"
    assert sorted(declared) == sorted(_abi.CORE_SYMBOLS + _abi.HOST_SYMBOLS)
    assert lib.rs_abi_version() == 3


def test_ctypes_layouts_match_the_compiled_structs():
    lib = _abi.load()
    for which, cls in enumerate([_abi.RsNode, _abi.RsXform, _abi.RsMaterial, _abi.RsTexture, _abi.RsSceneDesc, _abi.RsCamera,
                                 _abi.RsRenderParams, _abi.RsStats]):
        assert lib.rs_abi_struct_size(which) == C.sizeof(cls), cls.__name__
    assert lib.rs_abi_struct_size(99) == -1


def test_no_gpu_means_a_loud_error_not_a_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from raysnail_b200 import RaysnailError, Renderer
    with pytest.raises(RaysnailError, match="no CUDA device|no CPU fallback"):
        Renderer(0)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "raysnail_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in text and "oracle_binding" not in text and "oracle/" not in text, f


def test_sdl_parser_on_the_reference_scene_files():
    expect = {  # roots (objects + 1 light sphere), lights
        "example": (7, 1), "quadric": (7, 1), "csg": (6, 1), "declares": (13, 1), "transforms": (5, 1),
    }
    for name, (n_roots, n_lights) in expect.items():
        s = load_scene(name)
        d = s.desc.contents
        assert (d.n_roots, d.n_lights) == (n_roots, n_lights), name
        # the light is an r = 12 DiffuseLight x1.7 sphere appended last (raysnail.rs:351-362)
        ln = d.nodes[d.lights[0]]
        assert ln.type == 0 and ln.p[3] == 12.0
        lm = d.materials[ln.material]
        assert lm.type == 4 and lm.a == 1.7
        assert d.roots[d.n_roots - 1] == d.lights[0]
    # example.sdl: camera, first sphere, checker slab (Appendix A of SURVEY.md)
    s = load_scene("example")
    d = s.desc.contents
    n0 = d.nodes[d.roots[0]]
    assert (n0.type, tuple(n0.p[:4])) == (0, (1.0, 0.0, -0.7, 1.0))
    m0 = d.materials[n0.material]
    assert m0.type == 0 and np.allclose(d.textures[m0.texture].color[:3], [0.9, 0.5, 0.1])
    slab = d.nodes[d.roots[4]]
    assert slab.type == 1 and tuple(slab.p[:6]) == (-3.5, -1.2, -6.0, 3.5, -1.0, 4.0)
    chk = d.textures[d.materials[slab.material].texture]
    assert chk.type == 1 and chk.scale == 2.0
    assert np.allclose(d.textures[chk.odd].color[:3], [0.3] * 3) and np.allclose(d.textures[chk.even].color[:3], [0.01] * 3)
    cam = s.camera(800, 500)
    assert tuple(cam.origin) == (6.0, 1.0, 2.5) and cam.aperture == 0.01
    # csg.sdl: expressions (3.0*1.5), reflection -> MixedMaterial(Metal, Lambertian, r)
    s = load_scene("csg")
    d = s.desc.contents
    assert tuple(s.camera(8, 8).origin) == (4.5, 1.35, 1.875)
    top = d.nodes[d.roots[0]]
    assert top.type == 6 and d.nodes[top.child0].type == 4          # TfFacade over Difference
    plus = d.nodes[d.nodes[top.child0].child0]
    mm = d.materials[plus.material]
    assert mm.type == 7 and mm.a == 0.2 and d.materials[mm.mat1].type == 1 and d.materials[mm.mat2].type == 0
    # declares.sdl: #declare + #while instantiate the shared LEMON_CYLINDER 7 times (one node, 7 facades)
    s = load_scene("declares")
    d = s.desc.contents
    facades = [d.nodes[d.roots[i]] for i in range(7)]
    assert all(f.type == 6 for f in facades) and len({f.child0 for f in facades}) == 1
    assert [f.n_xform for f in facades] == [2] + [3] * 6           # rotate <0, 0*360/7, 0> pushes nothing for n = 0
    # quadric coefficient order <A,B,C>,<D,E,F>,<G,H,I>,J -> qa=A qb=D qc=E qd=G qe=B qf=F qg=H qh=C qi=I qj=J
    s = Scene.from_sdl_text("camera { location <0,0,5> look_at <0,0,0> angle 40 }\nlight { <0,9,0>, color rgb <1,1,1> }\n"
                            "quadric { <1,2,3>, <4,5,6>, <7,8,9>, 10 }\n")
    q = s.desc.contents.nodes[s.desc.contents.roots[0]]
    assert tuple(q.p[:10]) == (1.0, 4.0, 5.0, 7.0, 2.0, 6.0, 8.0, 3.0, 9.0, 10.0)


def test_sdl_finish_settings_and_errors():
    txt = ("camera { location <0,0,5> look_at <0,0,0> angle 40 }\nlight { <0,9,0>, color rgb <1,1,1> }\n"
           "sphere { <0,0,0>, 1 texture { pigment { color rgb <1,0,0> } finish { phong 0.2 phong_size 45 } } scale <1,2,3> translate <0,-1,0> }\n"
           "sphere { <0,0,0>, 1 texture { pigment { color rgb <1,0,0> } surface { metallic diffuse 300 } } }\n")
    s = Scene.from_sdl_text(txt)
    d = s.desc.contents
    f = d.nodes[d.roots[0]]
    assert f.type == 6 and f.n_xform == 2
    m = d.materials[d.nodes[f.child0].material]
    assert m.type == 0 and m.phong_factor == pytest.approx(0.8) and m.phong_exponent == 4   # 4p, (0.1*s) as i32
    m2 = d.materials[d.nodes[d.roots[1]].material]
    assert m2.type == 2 and m2.a == 300.0
    from raysnail_b200 import RaysnailError
    with pytest.raises(RaysnailError):
        Scene.from_sdl_text("camera { location <0,0,5> look_at <0,0,0> }\nbogus { }\n")
    with pytest.raises(RaysnailError):
        Scene.from_sdl_text("sphere { <0,0,0>, 1 }\n")     # no camera


@pytest.mark.parametrize("name", ALL_SCENES)
def test_host_bbox_rules_match_the_oracle(name):
    """Two independent restatements of every Hittable::bbox (host C++ layer vs oracle)."""
    s = load_scene(name)
    host = s.root_bbox()
    # the recipes' `time_limit`: 0..1 for all_feature_scene (scene.rs:337), 0..0 elsewhere
    oracle = OracleWorld(s, use_desc_bbox=False).root_bbox((0.0, 1.0) if name == "all_feature" else (0.0, 0.0))
    assert host.shape == oracle.shape and np.array_equal(host, oracle)


def test_intersection_bbox_quirk_is_reproduced():
    """intersection.rs:110: min.z = max(b1.min.y, b2.min.z)."""
    s = load_scene("csg")
    bb = s.root_bbox()
    d = s.desc.contents
    inter = d.nodes[d.nodes[d.roots[1]].child0]
    assert inter.type == 5
    s2 = Scene()
    a = s2.sphere((0, 1.0, 0), 0.9)
    b = s2.box((-1, 0.1, -1), (1, 0.9, 1))
    lt = s2.sphere((0, 50, 0), 1.0, s2.diffuse_light(s2.solid(1, 1, 1), 1.0))
    s2.add(s2.intersection(a, b))
    s2.add(lt)
    s2.add_light(lt)
    box = s2.root_bbox()[0]
    assert box[2] == pytest.approx(0.1)          # min.z took sphere.min.y = 0.1 instead of -0.9
    assert box[5] == pytest.approx(0.9) and box[0] == pytest.approx(-0.9)
    assert bb.shape[0] == d.n_roots


def test_camera_builder_matches_camera_new():
    cam = CameraBuilder().look_from((13, 2, 3)).look_at((0, 0, 0)).fov(20).aperture(0.02).focus(10).width(800).height(500).build()
    w = -np.array([13, 2, 3.0]) / math.sqrt(13 * 13 + 4 + 9)
    hu = np.cross(w, [0, 1, 0]); hu /= np.linalg.norm(hu)
    vu = np.cross(hu, w); vu /= np.linalg.norm(vu)
    vh = 2 * math.tan(math.radians(20) / 2) * 10
    vw = vh * 800 / 500
    assert np.allclose(cam.horizontal_unit[:], hu, atol=1e-15) and np.allclose(cam.vertical_unit[:], vu, atol=1e-15)
    assert np.allclose(cam.horizontal_full[:], hu * vw, atol=1e-13) and np.allclose(cam.vertical_full[:], vu * vh, atol=1e-13)
    lb = np.array([13, 2, 3.0]) - hu * vw / 2 - vu * vh / 2 + 10 * w
    assert np.allclose(cam.lb[:], lb, atol=1e-12)
    assert (cam.width, cam.height, cam.aperture) == (800, 500, 0.02)


def test_rtow_final_scene_recipe():
    s = Scene.builtin("rtow_final", 7)
    d = s.desc.contents
    assert 440 <= d.n_roots <= 490 and d.n_lights == 1
    ground = d.nodes[d.roots[0]]
    assert tuple(ground.p[:4]) == (0.0, -1000.0, 0.0, 1000.0)
    chk = d.textures[d.materials[ground.material].texture]
    assert chk.type == 1 and chk.scale == 10.0
    big = [d.nodes[d.roots[d.n_roots - 4 + i]] for i in range(3)]
    assert [tuple(b.p[:4]) for b in big] == [(0.0, 1.0, 0.0, 1.0), (-4.0, 1.0, 0.0, 1.0), (4.0, 1.0, 0.0, 1.0)]
    assert [d.materials[b.material].type for b in big] == [3, 0, 1]
    light = d.nodes[d.roots[d.n_roots - 1]]
    assert tuple(light.p[:4]) == (300.0, 400.0, 100.0, 12.0) and d.materials[light.material].a == 1.5
    # same seed -> same scene; other seed -> other scene
    d2 = Scene.builtin("rtow_final", 7).desc.contents
    assert d2.n_roots == d.n_roots and tuple(d2.nodes[5].p[:3]) == tuple(d.nodes[5].p[:3])
    assert tuple(Scene.builtin("rtow_final", 8).desc.contents.nodes[5].p[:3]) != tuple(d.nodes[5].p[:3])
    # small balls: r = 0.2, y in [0.2, 1.1], materials drawn 80/15/5
    small = [d.nodes[d.roots[i]] for i in range(1, d.n_roots - 4)]
    assert all(n.p[3] == 0.2 and 0.2 <= n.p[1] <= 1.1 for n in small)
    kinds = np.bincount([d.materials[n.material].type for n in small], minlength=4)
    assert kinds[0] > 300 and kinds[3] > 5 and kinds[1] + kinds[2] > 30


def test_all_feature_scene_recipe():
    """examples/common/scene.rs:336-468: 400 ground boxes (the nested BVH flattened), the ceiling light,
    a moving sphere, glass / metal spheres, two constant media, an image and a Perlin texture."""
    s = Scene.builtin("all_feature", 7)
    d = s.desc.contents
    assert d.n_roots == 409 and d.n_lights == 1
    boxes = [d.nodes[d.roots[i]] for i in range(400)]
    assert all(b.type == 1 for b in boxes)                                   # RS_NODE_BOX
    assert tuple(boxes[0].p[:3]) == (-1000.0, 0.0, -1000.0) and boxes[0].p[3] == -900.0 and boxes[0].p[5] == -900.0
    assert all(1.0 <= b.p[4] < 100.0 for b in boxes)                         # y1 = rng.range(1..100)
    light = d.nodes[d.roots[400]]
    assert light.type == 2 and d.materials[light.material].type == 4 and d.materials[light.material].a == 7.0
    moving = d.nodes[d.roots[401]]
    assert tuple(moving.p[:7]) == (400.0, 400.0, 200.0, 50.0, 30.0, 0.0, 0.0)
    assert [d.nodes[d.roots[i]].type for i in (405, 406)] == [7, 7]          # RS_NODE_MEDIUM
    assert d.nodes[d.roots[406]].p[0] == 0.0001 or 0.0001 in tuple(d.nodes[d.roots[406]].p[:8])
    tex_types = sorted(d.textures[i].type for i in range(d.n_textures))
    assert 2 in tex_types and 3 in tex_types                                 # image + Perlin
    assert d.n_texel_bytes == 1920 * 960 * 4 and d.n_perlin_perm == 3 * 256
    cam = s.camera(64, 64)
    assert cam.shutter_speed == 1.0


def test_chacha12_block_known_answer_and_stdrng_stream():
    """ChaCha12, 256-bit zero key, zero counter/nonce — the eSTREAM / draft-strombergson vector."""
    lib = _abi.load()
    key = (C.c_uint32 * 8)()
    out = (C.c_uint32 * 16)()
    lib.rs_host_chacha12_block(key, 0, out)
    got = np.frombuffer(bytes(out), dtype=np.uint8)
    expect = bytes.fromhex("9bf49a6a0755f953811fce125f2683d50429c3bb49e074147e0089a52eae155f"
                           "0564f879d27ae3c02ce82834acfa8c793a629f2ca0de6919610be82f411326be")
    assert bytes(got) == expect
    a = np.zeros(200, np.uint32)
    b = np.zeros(200, np.uint32)
    lib.rs_host_seed_random_u32(7, a.ctypes.data, 200)
    lib.rs_host_seed_random_u32(7, b.ctypes.data, 200)
    assert np.array_equal(a, b) and len(set(a.tolist())) > 190
    lib.rs_host_seed_random_u32(8, b.ctypes.data, 200)
    assert not np.array_equal(a, b)


def test_mesh_loader_semantics(tmp_path):
    """TriangleMesh::load: rotate(axis) -> scale -> offset, smooth normals synthesised when the file
    has no vn (triangle_mesh.rs:186-276); reads v / vn / f with fan triangulation."""
    obj = tmp_path / "quad.obj"
    obj.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nf 1 2 3 4\n")
    s = Scene()
    white = s.lambertian(s.solid(1, 1, 1))
    n = s.add_obj_file(str(obj), scale=2.0, offset=(1, 0, 0), rotation_angle=90.0, axis=1, mat=white)
    assert n == 2
    lt = s.sphere((0, 50, 0), 1.0, s.diffuse_light(s.solid(1, 1, 1), 1.0))
    s.add(lt)
    s.add_light(lt)
    d = s.desc.contents
    assert d.n_tris == 2 and d.roots[0] == -1 and d.roots[1] == -2
    v = np.ctypeslib.as_array(d.tri_verts, shape=(2, 3, 3))
    # rotate about y by 90 deg: (x,y,z) -> (x cos + z sin, y, -x sin + z cos); then *2, +(1,0,0)
    assert np.allclose(v[0], [[1, 0, 0], [1, 0, -2], [1, 2, -2]], atol=1e-12)
    nrm = np.ctypeslib.as_array(d.tri_normals, shape=(2, 3, 3))
    assert np.allclose(np.abs(nrm[0]), [[1, 0, 0]] * 3, atol=1e-12)      # synthesised, unit length
    # explicit normals are rotated but not normalised
    obj.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 2\nf 1//1 2//1 3//1\n")
    s = Scene()
    s.add_obj_file(str(obj), rotation_angle=0.0)
    nrm = np.ctypeslib.as_array(s._lib.rs_host_scene_desc(s._h).contents.tri_normals, shape=(1, 3, 3))
    assert np.allclose(nrm[0], [[0, 0, 2]] * 3)
    # degenerate-bbox (axis-aligned) triangles get a zero-thickness reference bbox
    bb = s.root_bbox() if False else None
    assert bb is None


def test_png_round_trip(tmp_path):
    lib = _abi.load()
    rgb = (np.arange(5 * 7 * 3) % 251).astype(np.uint8).reshape(5, 7, 3)
    path = str(tmp_path / "x.png")
    assert lib.rs_host_write_png_rgb8(path.encode(), rgb.ctypes.data, 7, 5) == 0
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(path)), rgb)
    s = Scene()
    t = s.image_file(path)
    lt = s.sphere((0, 0, 0), 1.0, s.diffuse_light(t, 1.0))
    s.add(lt)
    s.add_light(lt)
    d = s.desc.contents
    tex = d.textures[d.materials[d.nodes[d.roots[0]].material].texture]
    assert (tex.width, tex.height) == (7, 5)
    px = np.ctypeslib.as_array(d.texels, shape=(d.n_texel_bytes,))[tex.texel_offset:tex.texel_offset + 5 * 7 * 4].reshape(5, 7, 4)
    assert np.array_equal(px[..., :3], rgb) and (px[..., 3] == 255).all()


def test_stratum_rows_partition():
    from raysnail_b200.distributed import stratum_rows
    for q in (1, 5, 8, 32, 64):
        for world in (1, 2, 3, 4, 8):
            parts = [stratum_rows(r, world, q) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == q
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        stratum_rows(2, 2, 8)


def test_gloo_two_rank_sharded_render_equals_single_rank():
    """The N > 1 host path (strata-row sharding + sum-reduce + finalize) with world_size 2 on the
    gloo backend; the per-rank renderer here is the CPU oracle (test infrastructure)."""
    script = os.path.join(ROOT, "tests", "gloo_worker.py")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29533", script], capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "SHARDED_OK" in out.stdout


def test_pixel_slices_of_the_fused_reduce_cover_the_image_once():
    from raysnail_b200.distributed import pixel_slice
    for n_pixels in (1, 15, 16, 17, 1920 * 1080, 3840 * 2160, 800 * 500):
        for world in (1, 2, 3, 4, 8):
            cuts = [pixel_slice(r, world, n_pixels) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n_pixels
            for (b0, e0), (b1, e1) in zip(cuts, cuts[1:]):
                assert e0 == b1 and b0 <= e0
            assert all(b % 16 == 0 for b, _ in cuts if b < n_pixels)


def test_raymarching_recipe_contents():
    """`render_raymarching_test` (examples/preview_sdl2.rs:338-407): light sphere, RayMarcher with a
    Lambertian carrying Phong settings (4.0, 2), r = 1000 DiffuseMetal(800) checker ground; the bulb's
    bbox is the fixed +-1.3 cube (raymarching.rs:172-181)."""
    s = load_scene("raymarching")
    d = s.desc.contents
    assert (d.n_roots, d.n_lights) == (3, 1)
    types = [d.nodes[d.roots[i]].type for i in range(3)]
    assert types == [0, 8, 0]                                   # sphere, RS_NODE_RAYMARCHER, sphere
    bulb = d.nodes[d.roots[1]]
    assert bulb.p[0] == 100.0
    m = d.materials[bulb.material]
    assert m.type == 0 and m.phong_factor == 4.0 and m.phong_exponent == 2
    assert [d.root_bbox[6 + k] for k in range(6)] == [-1.3, -1.3, -1.3, 1.3, 1.3, 1.3]
    g = d.materials[d.nodes[d.roots[2]].material]
    assert g.type == 2 and g.a == 800.0
    cam = s.camera(800, 500)
    assert abs(cam.origin[0] - 9.1) < 1e-12 and abs(cam.origin[1] + 1.19) < 1e-12
