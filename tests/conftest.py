import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")
    # the native artefacts are git-ignored: build them when a fresh checkout runs the tests
    native = os.path.join(ROOT, "distributed-pytorch-training_b200", "_native")
    if not (os.path.exists(os.path.join(native, "libdptb200.so"))
            and os.path.exists(os.path.join(native, "_dptb200_C.so"))):
        try:
            import build_native
            build_native.build_all(verbose=True)
        except Exception as e:  # noqa: BLE001
            print(f"[conftest] native build failed: {e}")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
