"""W13/W14/W16: model parity with the reference module, partitioning, checkpoint compatibility."""
import pytest
import torch

from dps_b200.config import Config
from dps_b200.models.resnet import ResNet18, build_model
from dps_b200.utils import checkpoint as C
from dps_b200.utils import data as D


def test_state_dict_keys_and_param_count_match_reference(ref_modules):
    ref = ref_modules["server"].ResNet18(num_classes=100)
    ours = ResNet18(100)
    rs, os_ = ref.state_dict(), ours.state_dict()
    assert list(rs.keys()) == list(os_.keys()) and len(os_) == 122
    assert all(rs[k].shape == os_[k].shape and rs[k].dtype == os_[k].dtype for k in rs)
    assert sum(p.numel() for p in ours.parameters()) == 11_220_132
    assert len(list(ours.named_parameters())) == 62 and len(list(ours.named_buffers())) == 60


def test_forward_backward_equal_reference_given_same_weights(ref_modules):
    torch.manual_seed(0)
    ref = ref_modules["server"].ResNet18(num_classes=100)
    ours = ResNet18(100)
    ours.load_state_dict(ref.state_dict(), strict=True)
    x, y = torch.randn(4, 3, 32, 32), torch.randint(0, 100, (4,))
    lr = torch.nn.functional.cross_entropy(ref(x), y)
    lo = torch.nn.functional.cross_entropy(ours(x), y)
    lr.backward(); lo.backward()
    assert torch.allclose(lr, lo, atol=1e-6)
    for (n1, p1), (n2, p2) in zip(ref.named_parameters(), ours.named_parameters()):
        assert n1 == n2 and torch.allclose(p1.grad, p2.grad, atol=1e-5), n1


def test_seedable_init_q13():
    a, b, c = build_model(100, seed=5), build_model(100, seed=5), build_model(100, seed=6)
    assert torch.equal(a.fc.weight, b.fc.weight) and not torch.equal(a.fc.weight, c.fc.weight)


@pytest.mark.parametrize("n,w", [(50000, 4), (50000, 8), (50003, 4), (10, 3)])
def test_partition_disjoint_contiguous_complete(n, w):
    ranges = [D.partition_range(i, w, n) for i in range(w)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (s0, e0), (s1, e1) in zip(ranges, ranges[1:]):
        assert e0 == s1 and e0 - s0 == n // w
    assert D.partition_range(w, w, n) == (n, n)       # Q10: out-of-range id -> empty, not IndexError


def test_loader_shapes_normalisation_determinism():
    cfg = Config(train_samples=40, test_samples=12, batch_size=16, workers=2)
    tl, vl, (s, e) = D.make_loaders(cfg, 1, 2)
    assert (s, e) == (20, 40) and len(tl) == 2 and len(vl) == 1
    xb, yb = next(iter(tl))
    assert xb.shape == (16, 3, 32, 32) and xb.dtype == torch.float32 and yb.dtype == torch.int64
    assert abs(float(xb.mean())) < 1.5
    tl2, _, _ = D.make_loaders(cfg, 1, 2)
    xb2, yb2 = next(iter(tl2))
    assert torch.equal(xb, xb2) and torch.equal(yb, yb2)
    xv, _ = next(iter(vl))
    assert xv.shape == (12, 3, 32, 32)


def test_checkpoint_loads_into_reference_model_strict(ref_modules, tmp_path):
    ours = build_model(100, seed=2)
    path = str(tmp_path / "ckpt.pt")
    mom = {n: torch.zeros_like(p) for n, p in ours.named_parameters()}
    C.save(path, ours.state_dict(), global_step=17, momentum=mom, shard_versions=[17] * 8,
           config=Config().to_dict())
    ref = ref_modules["server"].ResNet18(num_classes=100)
    ref.load_state_dict(torch.load(path, weights_only=True), strict=True)
    assert torch.equal(ref.fc.weight, ours.fc.weight)
    sd, side = C.load(path)
    assert side["global_step"] == 17 and side["shard_versions"] == [17] * 8
    assert set(side["momentum"]) == {n for n, _ in ours.named_parameters()}
    assert len(sd) == 122


def test_loader_prefetch_thread_is_transparent():
    """Background producer (role of the reference's DataLoader worker processes, worker.py:181-187): same batches in
    the same order, and abandoning the iterator early does not hang."""
    import torch
    from dps_b200.utils import data as D
    ds = D.synthetic_dataset(640, 10, 32, 0)
    plain = D.BatchLoader(ds, 0, 640, 64, train=True, seed=5)
    pre = D.BatchLoader(ds, 0, 640, 64, train=True, seed=5, prefetch=3)
    n = 0
    for (x1, y1), (x2, y2) in zip(plain, pre):
        assert torch.equal(x1, x2) and torch.equal(y1, y2)
        n += 1
    assert n == 10
    for i, _ in enumerate(pre):
        if i == 1:
            break
    assert len(list(pre)) == 10
