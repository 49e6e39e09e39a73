"""Convergence parity: 200 sync steps on the learnable synthetic set, the native bf16 engine behind the
one-shard parameter server against plain fp32 PyTorch (same architecture -- reference: server.py:21-76 -- same initial
weights, same batches, same plain SGD p -= lr*g, server.py:133).  One-step gradient closeness does not prove that bf16
activations + fp16 wire train like fp32; loss curves that stay together do."""
import pytest
import torch

from dps_b200.config import Config
from dps_b200.models.resnet import build_model
from dps_b200.utils.data import synthetic_dataset

pytestmark = pytest.mark.gpu


def test_200_step_loss_curve_tracks_fp32_torch():
    from dps_b200.api import Job
    B, steps, lr = 128, 200, 0.05
    ds = synthetic_dataset(B * 40, num_classes=100, image_size=32, seed=7, learnable=True)
    mean = torch.tensor([0.5071, 0.4867, 0.4408]).view(1, 3, 1, 1)
    std = torch.tensor([0.2675, 0.2565, 0.2761]).view(1, 3, 1, 1)
    xs = (torch.from_numpy(ds.images).permute(0, 3, 1, 2).float() / 255.0 - mean) / std      # ToTensor + Normalize
    ys = torch.from_numpy(ds.labels).long()
    batches = [(xs[i * B:(i + 1) * B].contiguous(), ys[i * B:(i + 1) * B].contiguous()) for i in range(40)]

    cfg = Config(mode="sync", workers=1, batch_size=B, lr=lr, momentum=0.0, weight_decay=0.0, seed=3)
    job = Job(cfg, rank=0, world=1, device_index=0)
    ref = build_model(100, seed=3).cuda()
    ref.load_state_dict(job.model_cpu.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr=lr)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ours, theirs = [], []
    for s in range(steps):
        x, y = batches[s % len(batches)]
        ours.append(job.worker.train_step(x.pin_memory(), y.pin_memory()).item())
        ref.train()
        opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(ref(x.cuda()), y.cuda())
        loss.backward()
        opt.step()
        theirs.append(float(loss))
    assert job.ps.error_bits() == 0
    job.close()
    # Compare smoothed curves (mean over 20 steps) at every 20th step.  The fp32 torch arm is itself not repeatable to
    # better than ~1.5 % where the loss falls by 5x per 20 steps (cuDNN's atomics + chaotic amplification: 3.744 and
    # 3.791 at step 20 in two runs), so the criterion is: within 3 % of the oracle's window, allowing a lag of ONE step
    # (the oracle's window shifted by -1 / 0 / +1 steps), or within 0.01 absolute once the loss has collapsed.
    win = 20

    def wmean(v, k):
        k = max(win, min(len(v), k))
        return sum(v[k - win:k]) / win

    rows = []
    for k in range(win, steps + 1, win):
        a, b = wmean(ours, k), wmean(theirs, k)
        near = [wmean(theirs, k - 1), b, wmean(theirs, k + 1)]
        ok = 0.97 * min(near) <= a <= 1.03 * max(near) or abs(a - b) < 0.01
        rows.append((k, a, b, abs(a - b) / max(b, 1e-6), ok))
    print("\n".join(f"step {k:3d}: native {a:.4f}  fp32 torch {b:.4f}  rel {r:.4f}  {'ok' if ok else 'OFF'}"
                    for k, a, b, r, ok in rows))
    assert theirs[-1] < 0.7 * theirs[0], "the oracle itself must be learning"
    assert all(ok for *_, ok in rows), rows
