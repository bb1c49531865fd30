import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run by the driver with -m gpu)")
    # torch before anything loads a libnccl.so.2: a device group (mcsim_cuda_group_create, > 1 GPU) dlopen's the system NCCL,
    # and libtorch_cuda.so imported AFTER that would be bound to it instead of its own bundled, newer one (undefined
    # ncclDevCommCreate).  With torch first, the group finds torch's already-loaded library under the same soname.
    try:
        import torch  # noqa: F401
    except ImportError:
        pass


@pytest.fixture(scope="session")
def plan():
    import pcoll_b200
    return pcoll_b200.load_plan_json(os.path.join(GOLD, "plan_committed.json"))


@pytest.fixture(scope="session")
def scenes():
    out = {}
    for name in ("pr2test2", "pr2custom"):
        z = np.load(os.path.join(GOLD, "scene_%s.npz" % name))
        out[name] = dict(boxes=z["boxes"], hulls=z["hulls"], robot_pose=z["robot_pose"])
    return out


@pytest.fixture(scope="session")
def ctx():
    """One CUDA context for the whole GPU session (fails loudly without a device)."""
    import pcoll_b200
    c = pcoll_b200.cuda_abi.Context(0)
    yield c
    c.close()


def golden(name):
    return np.load(os.path.join(GOLD, name))
