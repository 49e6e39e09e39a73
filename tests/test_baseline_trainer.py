"""Single-machine baseline trainer (reference: baseline/baseline_training.py:97-269)."""
import json
import os

import pytest
import torch

from dps_b200 import Config
from dps_b200 import baseline_trainer as BT
from dps_b200.utils.png import read_png_size


def test_multistep_schedule_matches_torch():
    m = torch.nn.Linear(2, 2)
    opt = torch.optim.SGD(m.parameters(), lr=0.1)
    sched = torch.optim.lr_scheduler.MultiStepLR(opt, milestones=[10, 15], gamma=0.1)
    for epoch in range(20):
        assert BT.multistep_lr(0.1, [10, 15], 0.1, epoch) == pytest.approx(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()


def test_metrics_summary_and_figure(tmp_path):
    tm = BT.TrainingMetrics()
    tm.add_epoch(4.0, 8.0, 12.0, 10.0, 0.1)
    tm.add_epoch(3.5, 15.0, 18.0, 9.0, 0.1)
    s = tm.summary(batch_size=128)
    assert s["final_accuracy"] == 18.0 and s["epochs"] == 2 and s["avg_epoch_time"] == pytest.approx(9.5)
    assert s["total_time_minutes"] == pytest.approx(19.0 / 60) and s["batch_size"] == 128
    assert s["performance_metrics"]["epoch_2"]["train_loss"] == 3.5
    path = tm.plot_results(str(tmp_path / "fig" / "baseline_results.png"))
    assert read_png_size(path) == (1000, 720)


def test_cpu_loop_runs_and_learns(tmp_path):
    cfg = Config(workers=1, epochs=2, batch_size=16, lr=0.05, momentum=0.9, weight_decay=5e-4, device="cpu",
                 train_samples=32, test_samples=16, num_classes=4)
    s = BT.run(cfg, milestones=[1], gamma=0.1, output_dir=str(tmp_path), quiet=True)
    assert s["epochs"] == 2 and s["model_parameters"] > 11_000_000
    pm = s["performance_metrics"]
    assert pm["epoch_1"]["learning_rate"] == pytest.approx(0.05) and pm["epoch_2"]["learning_rate"] == pytest.approx(0.005)
    assert all(pm[k]["train_loss"] == pm[k]["train_loss"] for k in pm)          # finite
    saved = json.load(open(os.path.join(tmp_path, "baseline_summary.json")))
    assert saved["lr_schedule"]["milestones"] == [1] and os.path.exists(saved["figure"])


@pytest.mark.gpu
def test_device_loop_momentum_schedule(tmp_path):
    cfg = Config(workers=1, epochs=3, batch_size=128, lr=0.05, momentum=0.9, weight_decay=5e-4,
                 train_samples=1024, test_samples=256)
    s = BT.run(cfg, milestones=[2], gamma=0.1, output_dir=str(tmp_path), quiet=True)
    pm = s["performance_metrics"]
    assert s["epochs"] == 3 and "cuda" in s["hardware"]
    assert pm["epoch_3"]["learning_rate"] == pytest.approx(0.005)
    # synthetic data is learnable: the loss must come down under momentum SGD
    assert pm["epoch_3"]["train_loss"] < pm["epoch_1"]["train_loss"]
    assert os.path.exists(os.path.join(tmp_path, "baseline_results.png"))
