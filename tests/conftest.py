import ctypes
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_names():
    """Renderer fixtures of the fp32 path (the colouriser's fixture, the fp16-mode fixtures, the NaN case and the
    BASELINE-size "full_*" fixtures have their own tests)."""
    names = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    return [n for n in names if not n.startswith(("colorize", "half_", "registration_", "full_", "nan_", "disc_"))]


def full_golden_names():
    """BASELINE-size fixtures: the executed reference's outputs only; the clouds come back from the seeded recipe."""
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "full_*.npz")))


def load_full_golden(name):
    import hashlib
    from scream_b200 import synth
    g = load_golden(name)
    src, tgt = synth.cloud_from_recipe(str(g["recipe"]))
    sha = hashlib.sha1(np.ascontiguousarray(src).tobytes() + np.ascontiguousarray(tgt).tobytes()).hexdigest()
    assert sha == str(g["xyz_sha1"]), f"{name}: the seeded recipe no longer reproduces the fixture's clouds"
    g["src"], g["tgt"] = src, tgt
    return g


# winner-map budget per fixture: how many near-tie pixels (oracle top-2 gap within 2 ulp) may take the runner-up.
# 0 = the winner map must be bit-exact (regular grids: exact ties only; sparse clouds: no near-ties at all).
NEAR_TIE_BUDGET = {"cfg4_opengg_w64": 0, "cfg4_opengg_w128": 0, "sparse_far_pixels": 0,
                   "full_cfg4_w64": 0, "full_cfg4_w128": 0}


def half_golden_names():
    """Fixtures of the fp16 (autocast) input mode, SURVEY 8 f2: the reference executed under torch.autocast (CPU, fp16)."""
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "half_*.npz")))


def load_golden(name):
    with np.load(os.path.join(GOLDEN_DIR, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def c_oracle():
    """The plain-C oracle (oracle/render_oracle.c), built on demand.  Test infrastructure only."""
    so = os.path.join(ROOT, "oracle", "libscream_oracle.so")
    src = os.path.join(ROOT, "oracle", "render_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    lib = ctypes.CDLL(so)
    F, I, L = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
    lib.scream_oracle_forward.argtypes = [F, L, F, L, F, F, I, I, ctypes.c_double, F, F, F, F, F]
    lib.scream_oracle_forward.restype = I
    lib.scream_oracle_transform.argtypes = [F, L, F, L, F, F, I, F, F, F]
    lib.scream_oracle_transform.restype = I
    lib.scream_oracle_backward.argtypes = [F, L, F, L, F, F, I, I, ctypes.c_double, F, F, F, F, F]
    lib.scream_oracle_backward.restype = I
    return lib
