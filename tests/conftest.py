import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
# GPT_SORT_MIN_TRIS = a size threshold for re-ordering the secondary rays before each extend; it REPLACES the library's default
# policy (a timed trial on scenes of >= 256 Ki packed triangles, gpt_trial_info) and is read once, when the library is loaded.
# Set here so that the 1.2 M-triangle scene of test_gpu_wide.py goes through the sorted pipeline while the 1.0 M-triangle scene of
# test_gpu_build.py stays on the unsorted one.  The default policy itself -- all nine trial calls, ray order included -- runs in
# the subprocesses of test_gpu_wide.py::test_traversal_trial_* and test_gpu_configs.py, which drop every GPT_* variable.
os.environ.setdefault("GPT_SORT_MIN_TRIS", "1100000")
# Calls of fewer than GPT_SMALL_FRAME_PATHS paths (default 393216) go to the megakernel.  The parity tests render small frames and
# mean to exercise the pipeline their flags name, so the switch is off here; tests/test_gpu_parity.py::test_small_calls_* covers it.
os.environ.setdefault("GPT_SMALL_FRAME_PATHS", "0")
# GPT_REBUILD=0: no second tree from the library's own builder over uploaded reference-layout arrays (gpt_rebuild_info).  The parity
# tests compare slots, visit order and frames across calls with the tree AS UPLOADED; the rebuilt tree (same distances and face
# indices, ties aside) and its trial run in tests/test_gpu_rebuild.py and test_gpu_configs.py, whose subprocesses drop the variable.
os.environ.setdefault("GPT_REBUILD", "0")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure_oracle():
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    src = os.path.join(ROOT, "oracle", "oracle.cpp")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


@pytest.fixture(scope="session", autouse=True)
def built_oracle():
    _ensure_oracle()


@pytest.fixture(scope="session")
def gpt():
    import gpt_b200
    gpt_b200.load_library()
    return gpt_b200


@pytest.fixture(scope="session")
def torch_cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device: the product has no CPU fallback")
    return torch
