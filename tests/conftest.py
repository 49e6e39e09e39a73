import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a real B200 (run with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running CPU test")


def pytest_collection_modifyitems(config, items):
    # GPU tests are opt-in: they fail loudly (not skip) if selected on a box without CUDA,
    # but a plain `pytest tests/` on a CPU box should not try them.
    import torch
    if torch.cuda.is_available():
        return
    markexpr = config.getoption("-m") or ""
    if "gpu" in markexpr and "not gpu" not in markexpr:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


REF_ROOT = "/root/reference"
REF_VENDORED = os.path.join(ROOT, "baseline", "_ref")


def reference_src():
    for base in (REF_ROOT, REF_VENDORED):
        p = os.path.join(base, "src")
        if os.path.isdir(p):
            return p
    return None


@pytest.fixture(scope="session")
def ref_modules():
    """Import the reference's server module in-process (oracle). Skips if unavailable."""
    src = reference_src()
    if src is None:
        pytest.skip("reference sources not available")
    import importlib
    if src not in sys.path:
        sys.path.insert(0, src)
    ps_pb2 = importlib.import_module("communication.ps_pb2")
    ps_pb2_grpc = importlib.import_module("communication.ps_pb2_grpc")
    server = importlib.import_module("parameter_server.server")
    return {"ps_pb2": ps_pb2, "ps_pb2_grpc": ps_pb2_grpc, "server": server}
