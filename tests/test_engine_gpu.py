"""the native engine (tcgen05 convs + fused BN/ReLU + fused CE head) vs the reference
architecture in plain PyTorch fp32 with the same (bf16-rounded) weights: loss and all 62 gradients."""
import os
import pytest
import torch

from dps_b200.models.resnet import build_model
from dps_b200.store.shardmap import build_layout

pytestmark = pytest.mark.gpu


def _make(N=16, seed=0):
    from dps_b200.models.engine import ResNetEngine
    dev = torch.device("cuda")
    model = build_model(100, seed=seed)
    layout = build_layout(model, 1)
    flat = torch.zeros(layout.total)
    layout.flatten_into(flat, dict(model.named_parameters()))
    wbf16 = flat.to(dev).to(torch.bfloat16)
    small = flat[: layout.n_small].to(dev).clone()
    grad = torch.zeros(layout.total, device=dev)
    eng = ResNetEngine(layout, wbf16, small, grad, batch=N, image_size=32, num_classes=100, device=dev)
    # torch reference with the SAME rounded weights (small params stay fp32)
    ref = build_model(100, seed=seed).to(dev)
    rounded = layout.unflatten(wbf16.float())
    sd = ref.state_dict()
    for n, p in ref.named_parameters():
        sd[n] = rounded[n].clone() if p.dim() > 1 else p.detach().clone()
    ref.load_state_dict(sd)
    return eng, ref, layout, grad


def test_forward_backward_matches_torch_fp32():
    N = 16
    eng, ref, layout, grad = _make(N)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(N, 3, 32, 32, generator=g)
    y = torch.randint(0, 100, (N,), generator=g)
    eng.load_batch(x, y)
    eng.forward_backward()
    torch.cuda.synchronize()
    ref.train()
    xb = x.cuda().to(torch.bfloat16).float()          # the engine stages the input in bf16
    loss = torch.nn.functional.cross_entropy(ref(xb), y.cuda())
    loss.backward()
    assert abs(float(eng.out[0]) - float(loss)) < 0.05, (float(eng.out[0]), float(loss))
    g32 = {n: p.grad.clone() for n, p in ref.named_parameters()}
    rsd = {k: v.clone() for k, v in ref.state_dict().items()}     # running stats after exactly ONE step
    # calibration: the SAME network through torch's bf16 autocast (cuDNN/cuBLAS).  At batch 16 with random
    # weights bf16 activation noise alone moves early-layer gradients by ~30% (ReLU mask flips), so the
    # engine is required to be as close to fp32 as the library bf16 path is, tensor by tensor.
    ref.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        l16 = torch.nn.functional.cross_entropy(ref(xb).float(), y.cuda())
    l16.backward()
    got = layout.unflatten(grad)
    for n, p in ref.named_parameters():
        a, b, c = got[n].float().flatten(), g32[n].flatten(), p.grad.flatten()
        rel = float((a - b).norm() / (b.norm() + 1e-12))
        rel_lib = float((c - b).norm() / (b.norm() + 1e-12))
        cos = float(torch.nn.functional.cosine_similarity(a, b, dim=0))
        assert rel <= 1.3 * rel_lib + 0.03, (n, rel, rel_lib)
        assert cos > 0.9, (n, cos)
    for n in ("fc.weight", "fc.bias", "layer4.1.bn2.bias"):          # close to the loss: tight
        a, b = got[n].float().flatten(), g32[n].flatten()
        assert float((a - b).norm() / b.norm()) < 0.05, n
    # BN running statistics were updated like torch's (momentum 0.1, unbiased variance)
    bufs = eng.bn_buffers()
    for k in ("bn1.running_mean", "layer2.0.shortcut.1.running_var", "layer4.1.bn2.running_mean"):
        assert torch.allclose(bufs[k], rsd[k], rtol=5e-2, atol=5e-3), k
    assert int(bufs["bn1.num_batches_tracked"]) == 1


def test_graph_replay_equals_eager_and_eval_mode():
    N = 16
    eng, ref, layout, grad = _make(N, seed=1)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(N, 3, 32, 32, generator=g)
    y = torch.randint(0, 100, (N,), generator=g)
    eng.load_batch(x, y)
    eng.forward_backward()
    torch.cuda.synchronize()
    eager_loss, eager_grad = float(eng.out[0]), grad.clone()
    grad.zero_()
    eng.capture()
    grad.zero_()
    eng.load_batch(x, y)
    eng.step()
    torch.cuda.synchronize()
    assert abs(float(eng.out[0]) - eager_loss) < 1e-3
    assert float((grad - eager_grad).norm() / eager_grad.norm()) < 2e-2     # fp32 atomics reorder
    # eval mode: running statistics, prediction count agrees with torch within bf16 noise
    ref.load_state_dict({**ref.state_dict(), **{k: v for k, v in eng.bn_buffers().items()}})
    ref.eval()
    with torch.no_grad():
        pred = ref(x.cuda().to(torch.bfloat16).float()).argmax(1)
    want = int((pred == y.cuda()).sum())
    got = eng.eval_batch(x, y)
    assert abs(got - want) <= 2


def test_training_reduces_loss_with_ps_on_one_gpu():
    """Minimum end-to-end slice on one GPU: Job -> Worker.train_step (graph: fwd/bwd/push/pull)."""
    from dps_b200 import Config
    from dps_b200.api import Job
    from dps_b200.utils import data as D
    cfg = Config(mode="sync", workers=1, batch_size=32, lr=0.05, train_samples=256, test_samples=64)
    job = Job(cfg, rank=0, world=1, device_index=0)
    tl, vl, _ = D.make_loaders(cfg, 0, 1)
    losses = []
    for epoch in range(3):
        for x, y in tl:
            losses.append(job.worker.train_step(x, y).item())
    assert losses[-1] < losses[0] - 0.3, (losses[0], losses[-1])
    assert job.server.global_step == len(losses)
    acc = job.worker.evaluate(vl)
    assert 0.0 <= acc <= 100.0
    sd = job.server.state_dict()
    assert len(sd) == 122 and float(sd["bn1.running_mean"].abs().sum()) > 0
    job.close()


def test_launcher_one_gpu_end_to_end(tmp_path):
    """W18: `python -m dps_b200.launch` trains, evaluates, checkpoints and aggregates metrics."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    md, ck = tmp_path / "metrics", tmp_path / "ck.pt"
    r = subprocess.run([sys.executable, "-m", "dps_b200.launch", "--gpus", "1", "--", "--mode", "sync", "--epochs", "2",
                        "--batch-size", "32", "--train-samples", "256", "--test-samples", "64", "--lr", "0.05",
                        "--metrics-dir", str(md), "--checkpoint", str(ck)], cwd=root, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.load(open(md / "sync_1workers.result.json"))
    assert res["server_metrics"]["global_steps_completed"] == 16
    agg = res["worker_metrics_aggregated"]
    assert agg["num_workers"] == 1 and agg["total_local_steps"] == 16 and len(agg["accuracy_by_epoch"]) == 2
    assert agg["images_per_second_sum"] > 0 and os.path.exists(ck)
    assert r.stdout.count("METRICS_JSON:") == 2           # one worker record, one server record (Q11)


@pytest.mark.gpu
def test_experiment_pipeline_one_gpu(tmp_path):
    """dps_b200.experiments: baseline + one sync run -> result file in the reference schema + charts."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = str(tmp_path / "exp")
    r = subprocess.run([sys.executable, "-m", "dps_b200.experiments", "--workers", "1", "--modes", "sync", "--epochs", "1",
                        "--train-samples", "1024", "--test-samples", "256", "--out", out, "--master-port", "29731"],
                       cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.load(open(os.path.join(out, "sync_1workers.json")))
    assert res["server_metrics"]["mode"] == "sync" and res["worker_metrics_aggregated"]["num_workers"] == 1
    assert res["worker_metrics_aggregated"]["total_local_steps"] == 8
    assert os.path.exists(os.path.join(out, "charts", "scaling_analysis.png"))
    assert os.path.exists(os.path.join(out, "baseline", "baseline_summary.json"))


@pytest.mark.gpu
def test_profile_phases_reports_device_time_per_phase():
    """Tracing: per-phase device time of real steps, before and after the CUDA graph exists."""
    from dps_b200 import Config
    from dps_b200.api import Job
    job = Job(Config(workers=1, batch_size=128), rank=0, world=1, device_index=0)
    try:
        x = torch.randn(128, 3, 32, 32).pin_memory()
        y = torch.randint(0, 100, (128,)).pin_memory()
        ph = job.worker.profile_phases(x, y, reps=3)
        assert set(ph) == {"stage_batch", "forward", "backward", "exchange", "total"}
        assert all(v > 0 for v in ph.values()) and ph["backward"] > ph["forward"] * 0.8
        assert 500 < ph["total"] < 50000                      # microseconds; eager launches, so slower than the graph
        steps0 = job.worker.local_step_counter
        l1 = job.worker.train_step(x, y).item()               # graph path still works after eager profiling steps
        ph2 = job.worker.profile_phases(x, y, reps=2)         # and eager steps still work after capture
        l2 = job.worker.train_step(x, y).item()
        assert job.worker.local_step_counter == steps0 + 2 + 3
        assert l1 == l1 and l2 == l2 and ph2["total"] > 0 and job.ps.error_bits() == 0
    finally:
        job.close()


@pytest.mark.gpu
def test_sync_steps_reference_policy_discards_and_accumulate_policy_averages():
    """--sync-steps k.  'reference': only the first batch of every k is pushed, the other k-1
    gradients are thrown away (worker.py:367-377).  'accumulate' (default): the mean of the k gradients is pushed."""
    from dps_b200 import Config
    from dps_b200.api import Job
    g = torch.Generator().manual_seed(0)
    xs = [torch.randn(32, 3, 32, 32, generator=g) for _ in range(3)]
    ys = [torch.randint(0, 100, (32,), generator=g) for _ in range(3)]

    def run(steps, policy, batches):
        cfg = Config(mode="sync", workers=1, batch_size=32, lr=0.05, sync_steps=steps, local_step_policy=policy, seed=3)
        job = Job(cfg, rank=0, world=1, device_index=0)
        try:
            for i in batches:
                job.worker.train_step(xs[i], ys[i]).item()
            torch.cuda.synchronize()
            return job.server.global_step, {k: v.clone() for k, v in job.server.state_dict().items()
                                            if v.dtype.is_floating_point and "running" not in k}
        finally:
            job.close()

    step_one, w_one = run(1, "accumulate", [0])                  # one plain step on batch 0
    step_ref, w_ref = run(3, "reference", [0, 1, 2])             # k = 3, reference policy: only batch 0 counts
    assert step_one == step_ref == 1
    for k in w_one:
        assert torch.allclose(w_one[k], w_ref[k], rtol=1e-4, atol=1e-5), k
    step_acc, w_acc = run(3, "accumulate", [0, 1, 2])            # k = 3, accumulate: a different (averaged) update
    assert step_acc == 1
    assert any(not torch.allclose(w_one[k], w_acc[k], rtol=1e-4, atol=1e-5) for k in w_one)
