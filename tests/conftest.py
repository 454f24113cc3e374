"""Shared pytest setup: import paths, the `gpu` marker, golden-fixture loader."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "compactor-vllm_b200"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name: str) -> dict:
    """npz -> dict of torch tensors / python scalars (16-bit floats restored from bit patterns)."""
    raw = np.load(GOLDEN / f"{name}.npz")
    out = {}
    for key in raw.files:
        if key.endswith("__dtype"):
            continue
        arr = raw[key]
        tag = str(raw[key + "__dtype"]) if key + "__dtype" in raw.files else None
        if tag == "bf16":
            out[key] = torch.from_numpy(arr.view(np.int16).copy()).view(torch.bfloat16)
        elif tag == "f16":
            out[key] = torch.from_numpy(arr.view(np.int16).copy()).view(torch.float16)
        elif tag is not None:
            out[key] = torch.from_numpy(arr.copy())
        elif arr.ndim == 0:
            out[key] = arr.item()
        else:
            out[key] = arr
    return out


@pytest.fixture
def golden():
    return load_golden
