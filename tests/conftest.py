import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SDL_DIR = os.path.join(ROOT, "tests", "golden", "sdl")
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
# `all_feature_scene` reads examples/earth-map.png (examples/common/scene.rs:425-431); the committed copy
# under tests/golden/ travels to the GPU box, /root/reference does not
os.environ.setdefault("RAYSNAIL_ASSETS", GOLDEN_DIR)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (sm_100a); run with -m gpu on a B200")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def renderer():
    if not _have_gpu():
        pytest.skip("no CUDA device")
    from raysnail_b200 import Renderer
    r = Renderer(0)
    yield r
    r.close()


def load_scene(name):
    """Scenes by short name: the five reference .sdl files (copied verbatim as test inputs under
    tests/golden/sdl) and the example recipes."""
    from raysnail_b200 import Scene
    if name in ("example", "quadric", "csg", "declares", "transforms"):
        return Scene.from_sdl(os.path.join(SDL_DIR, name + ".sdl"))
    if name == "mesh":
        return Scene.builtin("mesh", 7, 48)
    if name == "mesh_blinn":
        return Scene.builtin("mesh_blinn", 7, 48)
    return Scene.builtin(name, 7, 0)


ALL_SCENES = ["example", "quadric", "csg", "declares", "transforms", "rtow_final", "cornell", "cornell_smoke", "mesh", "mesh_blinn", "all_feature", "raymarching"]
