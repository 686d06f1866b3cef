import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "hazyresearch_flash-attention_b200")
# The drop-in packages (`flash_attn`, `flash_attn_cuda`) live in PKG; put it FIRST so that the upstream
# flash_attn 2.x wheel present in this image's site-packages can never be picked up instead.
for p in (PKG, ROOT):
    if p in sys.path:
        sys.path.remove(p)
    sys.path.insert(0, p)
for name in list(sys.modules):
    if name == "flash_attn" or name.startswith("flash_attn."):
        del sys.modules[name]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
