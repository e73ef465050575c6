import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_SCENES = ["g1_deg0_128", "g2_deg3_100x70", "g3_treehill0_half", "g4_saturated_64x48"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    import oracle
    d = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    cam = oracle.Camera(int(d["image_height"]), int(d["image_width"]), float(d["tanfovx"]),
                        float(d["tanfovy"]), d["viewmatrix"], d["projmatrix"], d["campos"],
                        bg=d["bg"], sh_degree=int(d["sh_degree"]))
    return d, cam


@pytest.fixture(scope="session")
def oracle_lib():
    import oracle
    oracle.build()
    return oracle
