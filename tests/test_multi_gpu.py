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


def _torchrun(n, script_args, port=29631, timeout=600, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port)] + script_args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout,
                          env=dict(os.environ, **(env or {})))


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode", ["sync", "async"])
def test_multi_rank_job_over_ipc(mode, world, tmp_path):
    """W processes, one per GPU, CUDA-IPC fabric, the public API end to end.  sync: the first step equals the mean-
    gradient step computed with an NCCL all-reduce, replicas stay identical; async (>= 50 steps): shard versions only
    move forward, every push is accounted for (accepted + rejected + dropped == W*W*steps), the staleness histogram
    covers every arrival, accepted updates == version, no torn pulls."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    steps = 3 if mode == "sync" else 50
    out = tmp_path / f"res_{mode}_{world}.json"
    r = _torchrun(world, [os.path.join(ROOT, "bench_tools", "mgpu_check.py"), mode, str(out), str(steps)],
                  port=29631 + world + 10 * (mode == "async"))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads(out.read_text())
    assert res["error_bits"] == [0] * world
    assert res["replicas_equal"] and res["finite"]
    if mode == "sync":
        assert res["versions"] == [res["steps"]] * world
        assert res["max_abs_diff_vs_single_process_mean_step"] < 2e-3
    else:
        pushes = world * world * res["steps"]                       # W workers x W shards x steps
        assert sum(res["accepted"]) + sum(res["rejected"]) + sum(res["dropped"]) == pushes
        assert res["hist_total"] == [a + b for a, b in zip(res["accepted"], res["rejected"])]
        assert res["versions"] == res["accepted"] and all(v > 0 for v in res["versions"])
        assert res["versions_monotone"] == [1] * world and res["torn_pulls"] == [0] * world


@pytest.mark.parametrize("mode", ["sync", "async"])
def test_pull_through_the_nvls_multicast_window(mode, tmp_path):
    """Config.nvls: the weight arena + fp32 prefix live in a multicast-bound arena and the owners broadcast with one
    multimem.st per vector.  Same invariants as the per-peer path: every replica (bf16 arena AND fp32 prefix) equals
    the sharded master after the last pull, the first sync step equals the mean-gradient step.  Skips when the box
    has no multicast support (the job then falls back to per-peer stores and says so)."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    out = tmp_path / f"res_nvls_{mode}.json"
    r = _torchrun(2, [os.path.join(ROOT, "bench_tools", "mgpu_check.py"), mode, str(out), "20"],
                  port=29671 + (mode == "async"), env={"DPS_NVLS": "auto"})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads(out.read_text())
    if not res["nvls_pull"]:
        pytest.skip("no NVLS multicast on this box: " + r.stderr[-300:])
    assert res["error_bits"] == [0, 0] and res["replicas_equal"] and res["finite"]
    if mode == "sync":
        assert res["versions"] == [20, 20]
        assert res["max_abs_diff_vs_single_process_mean_step"] < 2e-3
    else:
        assert res["torn_pulls"] == [0, 0] and res["versions_monotone"] == [1, 1]


def test_sync_push_reduced_in_the_switch(tmp_path):
    """Config.nvls_reduce: sources stage their fp16 tiles into their OWN replica of the multicast-bound staging arena
    and the owners sum the replicas with multimem.ld_reduce.  The first step must still equal the mean-gradient step
    (the in-switch sum returns fp16: 2^-11 relative on the mean gradient) and the replicas must stay identical."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    out = tmp_path / "res_nvls_reduce.json"
    r = _torchrun(2, [os.path.join(ROOT, "bench_tools", "mgpu_check.py"), "sync", str(out), "10"], port=29681,
                  env={"DPS_NVLS": "auto", "DPS_NVLS_REDUCE": "1"})
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads(out.read_text())
    if not res["nvls_reduce"]:
        pytest.skip("no NVLS multicast on this box: " + r.stderr[-300:])
    assert res["error_bits"] == [0, 0] and res["replicas_equal"] and res["finite"]
    assert res["versions"] == [10, 10]
    assert res["max_abs_diff_vs_single_process_mean_step"] < 2e-3


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_injected_rank_death_is_recovered_from_the_last_checkpoint(tmp_path):
    """rank 1's process dies at step 12 of 30 (os._exit, nothing cleaned up).  The supervising launcher
    restarts both ranks from the periodic checkpoint of step 10; the job finishes all 30 steps.  The reference's
    target is recovery in < 30 s (baseline/results/baseline_summary.json:39)."""
    import re
    ck = tmp_path / "job.pt"
    cmd = [sys.executable, "-m", "dps_b200.launch", "--gpus", "2", "--supervise", "--max-restarts", "2",
           "--master-port", "29711", "--", "--mode", "sync", "--batch-size", "16", "--epochs", "100", "--max-steps", "30",
           "--train-samples", "2048", "--test-samples", "64", "--eval-batches", "1", "--checkpoint", str(ck),
           "--checkpoint-every", "5", "--inject-fault", "1:12", "--barrier-timeout-s", "10",
           "--metrics-dir", str(tmp_path / "m")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-4000:]
    assert "RECOVERY attempt 1" in out and "resumed from" in out
    t_fail = float(re.search(r"failure_detected_at=([0-9.]+)", out).group(1))
    t_died = float(re.search(r"FAULT_INJECTED_AT ([0-9.]+)", out).group(1))
    t_first = [float(x) for x in re.findall(r"FIRST_STEP_DONE_AT ([0-9.]+)", out)][-1]
    recovery = t_first - t_died
    print(f"recovery: process death -> detected {t_fail - t_died:.1f}s, -> first step of the resumed job {recovery:.1f}s")
    assert recovery < 30.0, recovery
    from dps_b200.utils import checkpoint as CK
    _, side = CK.load(str(ck))
    assert side["global_step"] == 30 and side["shard_versions"] == [30, 30]
