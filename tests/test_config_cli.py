"""W17: one config object; the reference's flag and env-var names still work (reference:
src/parameter_server/server.py:405-433, src/workers/worker.py:455-482)."""
import pytest

from dps_b200 import config as CFG


def test_server_cli_reference_flags_and_env(monkeypatch):
    monkeypatch.setenv("SERVER_MODE", "async")
    monkeypatch.setenv("TOTAL_WORKERS_EXPECTED", "8")
    monkeypatch.setenv("SERVER_PORT", "9123")
    cfg = CFG.parse_server_cli([])
    assert (cfg.mode, cfg.workers, cfg.port, cfg.lr, cfg.staleness_bound) == ("async", 8, 9123, 0.1, 5)
    cfg = CFG.parse_server_cli(["--mode", "sync", "--workers", "2", "--lr", "0.05", "--port", "8001",
                                "--staleness-bound", "9"])
    assert (cfg.mode, cfg.workers, cfg.port, cfg.lr) == ("sync", 2, 8001, 0.05)
    assert cfg.staleness_bound == 9          # Q8: the reference parses this flag and then drops it


def test_worker_cli_reference_flags_and_env(monkeypatch):
    monkeypatch.setenv("PARAMETER_SERVER_ADDRESS", "ps.internal:8000")
    cfg = CFG.parse_worker_cli([])
    assert (cfg.server, cfg.worker_name, cfg.epochs, cfg.batch_size, cfg.sync_steps) == \
        ("ps.internal:8000", "worker-node", 3, 128, 1)
    cfg = CFG.parse_worker_cli(["--server", "h:1", "--worker-name", "w7", "--epochs", "1", "--batch-size", "64",
                                "--lr", "0.2", "--sync-steps", "4", "--compat-pickle"])
    assert (cfg.server, cfg.worker_name, cfg.epochs, cfg.batch_size, cfg.lr, cfg.sync_steps, cfg.compat_pickle) == \
        ("h:1", "w7", 1, 64, 0.2, 4, True)


@pytest.mark.parametrize("bad", [dict(workers=0), dict(workers=33), dict(mode="bsp"), dict(sync_steps=0),
                                 dict(bn_buffer_policy="x"), dict(compress="int8")])
def test_validation(bad):
    with pytest.raises(ValueError):
        CFG.Config(**bad).validate()


def test_config_roundtrips_into_metrics():
    d = CFG.Config(mode="async", workers=4).to_dict()
    assert d["mode"] == "async" and d["workers"] == 4 and "staleness_bound" in d


def test_proto_contract_script_in_sync():
    """scripts/check_proto.py: ps.proto (human-readable contract) == programmatic descriptors (role of the
    reference's regenerate_proto.sh:12-16)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "scripts", "check_proto.py")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "8 messages, 4 rpcs, OK" in r.stdout


def test_job_without_cuda_points_to_the_grpc_path():
    import pytest
    import torch
    from dps_b200 import Config
    from dps_b200.api import Job
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="gRPC roles"):
        Job(Config(workers=1), rank=0, world=1, device_index=0)


def test_nvls_flag_parses_and_validates():
    import argparse
    p = argparse.ArgumentParser()
    CFG.add_server_args(p)
    CFG.add_common_args(p)
    assert CFG.from_args(p.parse_args([])).nvls == "off"            # per-peer stores measured faster: off by default
    assert CFG.from_args(p.parse_args(["--nvls", "auto"])).nvls == "auto"
    with pytest.raises(ValueError):
        CFG.Config(nvls="maybe").validate()


def test_ctypes_mirror_of_the_pointer_tables_matches_the_library():
    """The kernels take DpsTables by value: a field added on one side only would shift every pointer after it."""
    import ctypes
    from dps_b200.ops import _lib as L
    if not L.available():
        pytest.skip("native library not built")
    assert L.lib().dps_tables_size() == ctypes.sizeof(L.DpsTables)
    names = [f[0] for f in L.DpsTables._fields_]
    assert names[-3:] == ["mc_wbf16", "mc_small", "mc_stage"] and names.index("flags") < names.index("step_ctr")
