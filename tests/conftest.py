"""pytest configuration: markers and import paths.

`-m "not gpu"` : oracle vs golden vectors, host logic, C-ABI symbol checks (CPU only).
`-m gpu`       : parity tests proper; call the CUDA path through the C ABI.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "dau-convnet_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        have_gpu = False
    if have_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


def pytest_terminal_summary(terminalreporter):
    """How large the parameter-gradient tolerance concession actually is (tests/_tolerance.py)."""
    try:
        import json
        import _tolerance
    except Exception:  # pragma: no cover
        return
    lines = _tolerance.summary_lines()
    if not lines:
        return
    terminalreporter.section("parity tolerance bookkeeping")
    for ln in lines:
        terminalreporter.write_line(ln)
    try:
        out = os.path.join(ROOT, "gpurun_out")
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "strict_tolerance.json"), "w") as f:
            json.dump({k: dict(calls=v[0], elements=v[1], outside_unscaled_rule=v[2], worst_ratio=v[3],
                               max_abs_ref=v[4]) for k, v in _tolerance.REPORT.items()}, f, indent=1, sort_keys=True)
    except OSError:
        pass
