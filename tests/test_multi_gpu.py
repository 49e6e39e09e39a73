"""Real multi-process / multi-GPU runs (CUDA IPC fabric, one process per GPU, NVLink peer stores).
Skipped unless >= 2 GPUs are visible.  SURVEY §4.5: a sync step over W workers == one large-batch step."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(n, script_args, port=29631, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port)] + script_args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode", ["sync", "async"])
def test_two_rank_job_over_ipc(mode, tmp_path):
    out = tmp_path / f"res_{mode}.json"
    r = _torchrun(2, [os.path.join(ROOT, "bench_tools", "mgpu_check.py"), mode, str(out)],
                  port=29631 + (mode == "async"))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads(out.read_text())
    assert res["error_bits"] == [0, 0]
    assert res["replicas_equal"] and res["finite"]
    if mode == "sync":
        assert res["versions"] == [res["steps"]] * 2
        assert res["max_abs_diff_vs_single_process_mean_step"] < 2e-3
    else:
        assert sum(res["accepted"]) + sum(res["rejected"]) == 2 * 2 * res["steps"]   # W shards x W workers x steps
        assert all(v > 0 for v in res["versions"])
