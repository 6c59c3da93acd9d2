import glob
import json
import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "cpu-raytracer_b200", "python"))
sys.path.insert(0, os.path.join(REPO, "oracle"))
sys.path.insert(0, REPO)

GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")) if not os.path.basename(p).startswith(("sbvh_", "switch_", "stock_")))


class Golden:
    """One fixture of tests/golden/: a ScenePack + what the real reference produced for it."""

    def __init__(self, name):
        from b200rt.pack import ScenePack
        with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
            arrays = {k: z[k] for k in z.files}
        self.meta = json.loads(bytes(arrays["meta"]).decode())
        self.pack = ScenePack({k: v for k, v in arrays.items() if k.startswith(("s_", "f_")) or k == "resolution"})
        self.ref_frame_buffer = arrays["ref_frame_buffer"]
        self.ref_ids = arrays["ref_ids"]
        self.ref_hit = arrays["ref_hit"]
        self.ref_rays = arrays["ref_rays"]
        self.ref_frame_inputs = arrays["ref_frame_inputs"]


@pytest.fixture(scope="session", params=golden_names())
def golden(request):
    return Golden(request.param)


def unpack_rgb(frame_buffer):
    fb = np.asarray(frame_buffer, dtype=np.uint32)
    return np.stack([(fb >> 16) & 255, (fb >> 8) & 255, fb & 255], axis=-1).astype(np.int32)


def psnr(a, b):
    mse = ((np.asarray(a, dtype=np.float64) - np.asarray(b, dtype=np.float64)) ** 2).mean()
    return 99.0 if mse == 0 else float(10.0 * np.log10(255.0 ** 2 / mse))
