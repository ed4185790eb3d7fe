import importlib
import os
import sys

import numpy as np
import pytest

# The fixture tests compare against the reference at the ~1e-7 level: they run K3 with the
# reference's roundings (the library's default is its fast build of K3; tests/test_gpu_fast.py
# and tests/test_gpu_converged.py cover that one against north_star's own bar).
os.environ.setdefault("SPTX_EXACT", "1")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLD = os.path.join(ROOT, "tests", "golden")
GEN = os.path.join(ROOT, "scenes", "_gen")
SCENES = ["cornell64", "cornell_lens48", "instanced", "furnace_point", "furnace_area", "killeroo128", "zoo",
          "zoo_gauss", "uber_box", "lobes_box", "cornell_sobol", "envbox_const", "envbox_map", "texbox", "texbox_lens", "texbox_real", "texbox_uber", "manylights"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def capi():
    return importlib.import_module("pbrt-v3-spectral_b200.capi")


def golden(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def blob_path(name):
    """Flattened scene for `name`: prebuilt in scenes/_gen (travels with the repo);
    procedural scenes are re-flattened with the product host binary if missing."""
    p = os.path.join(GEN, name + ".sptx")
    if os.path.exists(p):
        return p
    entry = importlib.import_module("__graft_entry__")
    return entry.make_scene_blob(name)


@pytest.fixture(scope="session")
def gpu_ctx(capi):
    ctx = capi.Context(0)
    yield ctx
    ctx.close()
