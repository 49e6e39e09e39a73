"""Device-resident input pipeline (csrc/data.cu, dps_b200/utils/device_data.py) against a plain fp32 host restatement
of the reference's transform stack (reference: src/workers/worker.py:146-158)."""
import pytest
import torch

from dps_b200.utils import data as D


def test_splitmix_and_param_ranges():
    from dps_b200.utils.device_data import aug_params, splitmix64
    assert splitmix64(0) == 0xE220A8397B1DCDAF                      # first output of the published generator
    seen_dy, seen_dx, flips = set(), set(), 0
    for c in range(2000):
        dy, dx, f = aug_params(7, c)
        assert 0 <= dy <= 8 and 0 <= dx <= 8 and f in (0, 1)
        seen_dy.add(dy); seen_dx.add(dx); flips += f
    assert seen_dy == set(range(9)) and seen_dx == set(range(9)) and 800 < flips < 1200


def test_host_reference_matches_batchloader_semantics():
    """The restatement used as the GPU oracle applies pad -> crop -> flip -> normalise exactly like BatchLoader."""
    from dps_b200.utils.device_data import host_reference_batch
    ds = D.synthetic_dataset(8, 10, 32, 0)
    x, y = host_reference_batch(ds.images, ds.labels, [3, 1], seed=0, counter0=0, train=False)
    bl = D.BatchLoader(ds, 0, 8, 8, train=False)
    xb, yb = next(iter(bl))
    assert torch.allclose(x[0], xb[3], atol=1e-6) and torch.allclose(x[1], xb[1], atol=1e-6)
    assert int(y[0]) == int(yb[3])


@pytest.mark.gpu
@pytest.mark.parametrize("train", [True, False])
def test_augment_kernel_matches_host(train):
    from dps_b200.utils.device_data import DeviceLoader, host_reference_batch
    ds = D.synthetic_dataset(300, 10, 32, 1)
    dl = DeviceLoader(ds, 20, 300, 64, train=train, seed=11, device="cuda:0")
    counter = 0
    batches = 0
    for x, y in dl:
        order = dl.last_order[batches * 64: batches * 64 + x.shape[0]].tolist()
        xr, yr = host_reference_batch(ds.images, ds.labels, order, 11, counter, train)
        torch.cuda.synchronize()
        assert x.shape == xr.shape
        assert torch.allclose(x.cpu(), xr, atol=2e-6, rtol=1e-6)
        assert torch.equal(y.cpu(), yr)
        counter += 64
        batches += 1
    assert batches == (4 if train else 5)             # 280 samples: ragged tail dropped in train, short batch in eval
    if not train:
        assert x.shape[0] == 280 - 4 * 64
    else:
        assert min(order) >= 20                        # stays inside its partition
    # a second epoch draws different crops (the counter keeps advancing) but is reproducible from the same seed
    dl2 = DeviceLoader(ds, 20, 300, 64, train=train, seed=11, device="cuda:0", resident=dl.resident)
    a = next(iter(dl2))[0].clone()
    dl3 = DeviceLoader(ds, 20, 300, 64, train=train, seed=11, device="cuda:0", resident=dl.resident)
    b = next(iter(dl3))[0].clone()
    assert torch.equal(a, b)


@pytest.mark.gpu
def test_worker_trains_and_evaluates_from_device_loader():
    from dps_b200 import Config
    from dps_b200.api import Job
    cfg = Config(workers=1, batch_size=128, train_samples=1024, test_samples=300, lr=0.05)
    job = Job(cfg, rank=0, world=1, device_index=0)
    try:
        tl, vl, part = D.make_loaders(cfg, 0, 1, pin=True, device=job.device)
        assert type(tl).__name__ == "DeviceLoader" and part == (0, 1024)
        rec = job.worker.run_training(tl, vl, epochs=2)
        assert rec["local_steps_completed"] == 16
        assert len(rec["all_accuracies_percent"]) == 2 and 0.0 <= rec["final_test_accuracy_percent"] <= 100.0
        # host loader and device loader evaluate the same test set to the same accuracy (no augmentation in eval)
        _, vh, _ = D.make_loaders(cfg.replace(data_on_device=False), 0, 1, pin=True, device=job.device)
        assert type(vh).__name__ == "BatchLoader"
        acc_h = job.worker.evaluate(vh)
        acc_d = job.worker.evaluate(vl)
        assert acc_h == pytest.approx(acc_d, abs=0.5)
    finally:
        job.close()
