import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    arrays = {k: torch.from_numpy(z[k]) for k in z.files if k != "meta"}
    return meta, arrays


GOLDEN_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and not f.startswith(("next_", "front_", "full_", "batch_")))   # next_*: test_next_oracle.py, front_*: test_frontend_oracle.py, full_*: lean full-size fixtures
BATCH_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.startswith("batch_") and f.endswith(".npz"))   # 16-pair batches of BASELINE configs[1], configs[2]
FULL_CASES = sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.startswith("full_") and f.endswith(".npz"))


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture
def tuning_lib():
    """Routes the test's C-ABI calls to libpointdsc_b200_tuning.so, the -DPDSC_TUNING build that reads the PDSC_* knobs from
    the environment (the product library ignores them)."""
    from pointdsc_b200 import _cabi
    with _cabi.use_tuning_library() as lib:
        yield lib
