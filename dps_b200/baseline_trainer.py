"""Single-machine baseline trainer: the reference point every distributed run is compared with.

Same role as the reference's baseline/baseline_training.py (TrainingMetrics :97-147, train/test epoch :149-199,
main :201-269): an ordinary training loop *with* a local optimizer step -- SGD(momentum 0.9, weight decay 5e-4),
MultiStepLR(milestones [10, 15], gamma 0.1), batch 128, 3 epochs -- that reports per-epoch loss / train accuracy /
test accuracy / time, draws a 2x2 summary figure and writes a summary JSON.

B200 design: on a GPU the "local optimizer" is the one-shard parameter server of this framework (world size 1):
the fused push kernel applies momentum + weight decay + the step in the same launch that reduces the gradient,
the whole iteration is one CUDA-graph replay, and the learning rate lives in device memory so the schedule can
change it between epochs without re-capturing.  Without a GPU (or with ``--device cpu``) it runs the plain
PyTorch loop on the CPU, which is also what the CPU test exercises.  The figure is a headless PNG (no matplotlib,
no window): the reference's ``plt.show()`` has no counterpart.

    python -m dps_b200.baseline_trainer --epochs 3 --output-dir out/baseline
"""
from __future__ import annotations

import argparse
import json
import os
import time
from typing import List, Optional, Sequence

import torch

from .config import Config
from .utils import data as D
from .utils.png import Canvas


def multistep_lr(base_lr: float, milestones: Sequence[int], gamma: float, epoch: int) -> float:
    """Learning rate in effect DURING `epoch` (0-based) under torch's MultiStepLR stepped once per epoch."""
    return base_lr * (gamma ** sum(1 for m in milestones if epoch >= m))


class TrainingMetrics:
    """Per-epoch record + figure (reference: baseline_training.py:97-147)."""

    def __init__(self) -> None:
        self.train_losses: List[float] = []
        self.train_accuracies: List[float] = []
        self.test_accuracies: List[float] = []
        self.epoch_times: List[float] = []
        self.learning_rates: List[float] = []
        self.total_time = 0.0

    def add_epoch(self, train_loss: float, train_acc: float, test_acc: float, epoch_time: float,
                  lr: Optional[float] = None) -> None:
        self.train_losses.append(float(train_loss))
        self.train_accuracies.append(float(train_acc))
        self.test_accuracies.append(float(test_acc))
        self.epoch_times.append(float(epoch_time))
        self.learning_rates.append(float(lr) if lr is not None else float("nan"))
        self.total_time += float(epoch_time)

    def summary(self, **extra) -> dict:
        n = max(1, len(self.epoch_times))
        out = {
            "final_accuracy": self.test_accuracies[-1] if self.test_accuracies else None,
            "total_time_minutes": self.total_time / 60.0,
            "avg_epoch_time": sum(self.epoch_times) / n,
            "epochs": len(self.epoch_times),
            "performance_metrics": {
                f"epoch_{i + 1}": {"train_loss": self.train_losses[i], "train_accuracy": self.train_accuracies[i],
                                   "test_accuracy": self.test_accuracies[i],
                                   "epoch_time_seconds": self.epoch_times[i],
                                   "learning_rate": self.learning_rates[i]}
                for i in range(len(self.epoch_times))},
        }
        out.update(extra)
        return out

    def plot_results(self, path: str = "baseline_results.png") -> str:
        """2x2: loss, train/test accuracy, time per epoch, text summary."""
        from .utils.charts import _plot_series
        cv = Canvas(1000, 720)
        ep = list(range(1, len(self.epoch_times) + 1))
        _plot_series(cv, (70, 40, 480, 310), {"train": list(zip(ep, self.train_losses))},
                     "Training Loss", "Epoch", "Loss")
        _plot_series(cv, (570, 40, 980, 310), {"Train": list(zip(ep, self.train_accuracies)),
                                              "Test": list(zip(ep, self.test_accuracies))},
                     "Accuracy", "Epoch", "Accuracy (%)")
        _plot_series(cv, (70, 400, 480, 670), {"time": list(zip(ep, self.epoch_times))},
                     "Time per Epoch", "Epoch", "Time (seconds)")
        s = self.summary()
        cv.text(580, 400, "Training Summary", scale=2)
        lines = [f"Final Test Accuracy: {s['final_accuracy']:.2f}%" if s["final_accuracy"] is not None else "-",
                 f"Total Training Time: {s['total_time_minutes']:.3f} minutes",
                 f"Avg Time/Epoch: {s['avg_epoch_time']:.3f} seconds",
                 f"Epochs Completed: {s['epochs']}"]
        for i, ln in enumerate(lines):
            cv.text(580, 450 + 40 * i, ln, scale=1)
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        cv.save(path)
        return path


# ---------------------------------------------------------------------------------------------------- the loops
def _train_cpu(cfg: Config, metrics: TrainingMetrics, milestones, gamma, log, max_steps=None, eval_batches=None):
    """Plain PyTorch loop (reference: baseline_training.py:149-199)."""
    from .models.resnet import ResNet18
    torch.manual_seed(cfg.seed)
    model = ResNet18(num_classes=cfg.num_classes)
    opt = torch.optim.SGD(model.parameters(), lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay)
    sched = torch.optim.lr_scheduler.MultiStepLR(opt, milestones=list(milestones), gamma=gamma)
    train_loader, test_loader, _ = D.make_loaders(cfg, 0, 1)
    crit = torch.nn.CrossEntropyLoss()
    steps = 0
    for epoch in range(cfg.epochs):
        t0 = time.time()
        model.train()
        loss_sum = 0.0
        correct = seen = nb = 0
        for x, y in train_loader:
            opt.zero_grad()
            out = model(x)
            loss = crit(out, y)
            loss.backward()
            opt.step()
            loss_sum += float(loss)
            correct += int((out.argmax(1) == y).sum())
            seen += int(y.numel())
            nb += 1
            steps += 1
            if max_steps is not None and steps >= max_steps:
                break
        test_acc = D.evaluate(model, test_loader, max_batches=eval_batches)
        lr_now = opt.param_groups[0]["lr"]
        sched.step()
        metrics.add_epoch(loss_sum / max(1, nb), 100.0 * correct / max(1, seen), test_acc, time.time() - t0, lr_now)
        log(f"Epoch {epoch + 1}/{cfg.epochs}: loss {metrics.train_losses[-1]:.4f} train {metrics.train_accuracies[-1]:.2f}% "
            f"test {test_acc:.2f}% lr {lr_now:g} in {metrics.epoch_times[-1]:.2f}s")
        if max_steps is not None and steps >= max_steps:
            break
    return sum(p.numel() for p in model.parameters()), "cpu"


def _train_device(cfg: Config, metrics: TrainingMetrics, milestones, gamma, log, max_steps=None, eval_batches=None):
    """One-shard parameter server as the local optimizer; lr schedule through the device-resident lr word."""
    from .api import Job
    job = Job(cfg.replace(workers=1, mode="sync", sync_steps=1), rank=0, world=1, device_index=0)
    wk = job.worker
    train_loader, test_loader, _ = D.make_loaders(cfg, 0, 1, pin=True, device=job.device)
    steps = 0
    try:
        for epoch in range(cfg.epochs):
            lr_now = multistep_lr(cfg.lr, milestones, gamma, epoch)
            job.ps.set_lr(lr_now)
            t0 = time.time()
            loss_sum = 0.0
            correct = seen = nb = 0
            prev = None
            for x, y in train_loader:
                if int(y.numel()) != cfg.batch_size:
                    continue                                  # static shapes: the ragged tail batch is dropped
                h = wk.train_step(x, y)
                if prev is not None:                          # read one step late: never stalls the launch queue
                    loss_sum += prev.item()
                    correct += prev.correct()
                prev = h
                seen += cfg.batch_size
                nb += 1
                steps += 1
                if max_steps is not None and steps >= max_steps:
                    break
            if prev is not None:
                loss_sum += prev.item()
                correct += prev.correct()
            torch.cuda.synchronize(job.device)
            test_acc = wk.evaluate(test_loader, eval_batches)
            metrics.add_epoch(loss_sum / max(1, nb), 100.0 * correct / max(1, seen), test_acc, time.time() - t0, lr_now)
            log(f"Epoch {epoch + 1}/{cfg.epochs}: loss {metrics.train_losses[-1]:.4f} train "
                f"{metrics.train_accuracies[-1]:.2f}% test {test_acc:.2f}% lr {lr_now:g} in {metrics.epoch_times[-1]:.2f}s")
            if max_steps is not None and steps >= max_steps:
                break
        n_params = sum(e.numel for e in job.layout.entries.values())
        return n_params, f"cuda ({torch.cuda.get_device_name(0)}, sm_100a kernels)"
    finally:
        job.close()


def run(cfg: Config, milestones=(10, 15), gamma: float = 0.1, output_dir: Optional[str] = None, quiet: bool = False,
        max_steps: Optional[int] = None, eval_batches: Optional[int] = None) -> dict:
    log = (lambda s: None) if quiet else (lambda s: print(s, flush=True))
    metrics = TrainingMetrics()
    on_gpu = cfg.device != "cpu" and torch.cuda.is_available()
    log(f"Using device: {'cuda' if on_gpu else 'cpu'}")
    fn = _train_device if on_gpu else _train_cpu
    n_params, hw = fn(cfg, metrics, milestones, gamma, log, max_steps, eval_batches)
    log(f"Model has {n_params:,} parameters")
    summary = metrics.summary(model_parameters=n_params, batch_size=cfg.batch_size, learning_rate=cfg.lr,
                              optimizer=f"SGD momentum={cfg.momentum} weight_decay={cfg.weight_decay}",
                              lr_schedule={"type": "MultiStepLR", "milestones": list(milestones), "gamma": gamma},
                              hardware=hw, dataset=cfg.dataset)
    if output_dir:
        os.makedirs(output_dir, exist_ok=True)
        summary["figure"] = metrics.plot_results(os.path.join(output_dir, "baseline_results.png"))
        with open(os.path.join(output_dir, "baseline_summary.json"), "w") as f:
            json.dump(summary, f, indent=2)
    log("Baseline Results Summary:")
    for k in ("final_accuracy", "total_time_minutes", "avg_epoch_time", "model_parameters", "batch_size", "epochs"):
        log(f"{k}: {summary[k]}")
    return summary


def main(argv=None) -> int:
    p = argparse.ArgumentParser(description="single-machine baseline trainer (reference: baseline/baseline_training.py)")
    p.add_argument("--epochs", type=int, default=3)
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--lr", type=float, default=0.1)
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--weight-decay", type=float, default=5e-4)
    p.add_argument("--milestones", default="10,15")
    p.add_argument("--gamma", type=float, default=0.1)
    p.add_argument("--dataset", choices=["synthetic", "cifar100"], default="synthetic")
    p.add_argument("--data-root", default="./data")
    p.add_argument("--train-samples", type=int, default=50000)
    p.add_argument("--test-samples", type=int, default=10000)
    p.add_argument("--device", choices=["auto", "cpu", "cuda"], default="auto")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--max-steps", type=int, default=None)
    p.add_argument("--eval-batches", type=int, default=None)
    p.add_argument("--output-dir", default=".")
    a = p.parse_args(argv)
    cfg = Config(workers=1, epochs=a.epochs, batch_size=a.batch_size, lr=a.lr, momentum=a.momentum,
                 weight_decay=a.weight_decay, dataset=a.dataset, data_root=a.data_root,
                 train_samples=a.train_samples, test_samples=a.test_samples, device=a.device,
                 seed=a.seed).validate()
    ms = [int(t) for t in a.milestones.split(",") if t]
    run(cfg, ms, a.gamma, a.output_dir, max_steps=a.max_steps, eval_batches=a.eval_batches)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
