"""W2: numerics oracle -- our CPU servicer vs the reference servicer, in-process."""
import numpy as np
import pytest

from dps_b200.config import Config
from dps_b200.rpc import codec, proto, servicer as S


def _small_state():
    rng = np.random.default_rng(0)
    return {"a.weight": rng.normal(size=(8, 4)).astype(np.float32),
            "b.bias": rng.normal(size=(8,)).astype(np.float32),
            "bn.num_batches_tracked": np.array(0, dtype=np.int64)}


def _grads(seed, state):
    rng = np.random.default_rng(seed)
    return {k: rng.normal(size=v.shape).astype(np.float32) for k, v in state.items() if v.dtype == np.float32}


@pytest.fixture()
def ref_servicer(ref_modules, capsys):
    def make(mode, workers, lr=0.1):
        s = ref_modules["server"].ParameterServerServicer(mode=mode, total_workers=workers, learning_rate=lr)
        s.parameters = {k: v.copy() for k, v in _small_state().items()}
        return s
    return make


def test_staleness_weight_formula():
    for s, w in [(0, 1.0), (1, 1 / 1.1), (5, 1 / 1.5), (90, 0.1), (1000, 0.1)]:
        assert S.staleness_weight(s) == pytest.approx(w)


def test_sync_round_matches_reference(ref_servicer):
    W = 3
    ref = ref_servicer("sync", W)
    ours = S.ParameterServerServicer("sync", W, 0.1, initial_state=_small_state(),
                                     config=Config(nonblocking_sync=True))
    for rnd in range(3):
        for w in range(W):
            g16 = codec.compress_gradients(_grads(10 * rnd + w, _small_state()))
            ref._handle_sync_gradients(w, ref.decompress_gradients(g16), 0)
            ours._handle_sync_gradients(w, ours.decompress_gradients(g16), 0)
    assert ours.global_step == ref.global_step == 3
    for k, v in ref.parameters.items():
        np.testing.assert_array_equal(ours.parameters[k], v)   # bit-exact: same fp32 op order


def test_async_matches_reference_including_rejection(ref_servicer):
    ref = ref_servicer("async", 4)
    ours = S.ParameterServerServicer("async", 4, 0.1, initial_state=_small_state())
    local_steps = [0, 0, 1, 0, 3, 0, 0, 0, 1]      # the 8th push (local 0, global 7) is stale > 5
    results = []
    for i, ls in enumerate(local_steps):
        g = _grads(i, _small_state())
        r = ref._handle_async_gradients(i % 4, {k: v.copy() for k, v in g.items()}, ls).received
        o = ours._handle_async_gradients(i % 4, g, ls)
        results.append((r, o))
    assert all(r == o for r, o in results) and not all(r for r, _ in results)
    assert ours.global_step == ref.global_step
    for k, v in ref.parameters.items():
        np.testing.assert_allclose(ours.parameters[k], v, rtol=0, atol=1e-6)
    m = ours.server_metrics()
    assert m["rejected_updates"] >= 1 and m["max_staleness_observed"] == max(ours.gradient_staleness.values())


def test_momentum_and_weight_decay_match_torch_sgd():
    import torch
    st = _small_state()
    p = torch.nn.Parameter(torch.from_numpy(st["a.weight"].copy()))
    opt = torch.optim.SGD([p], lr=0.05, momentum=0.9, weight_decay=5e-4)
    ours = S.HostOptimizer({"a.weight": st["a.weight"].copy()}, 0.05, 0.9, 5e-4)
    for i in range(4):
        g = _grads(i, st)["a.weight"]
        p.grad = torch.from_numpy(g.copy())
        opt.step()
        ours.step({"a.weight": g})
    np.testing.assert_allclose(ours.params["a.weight"], p.detach().numpy(), rtol=1e-5, atol=1e-6)


def test_sync_barrier_counts_distinct_ids_q2():
    ours = S.ParameterServerServicer("sync", 2, 0.1, initial_state=_small_state(),
                                     config=Config(nonblocking_sync=True))
    g = _grads(1, _small_state())
    ours._handle_sync_gradients(0, g, 0)
    ours._handle_sync_gradients(0, g, 0)          # same worker twice must NOT trip the barrier
    assert ours.global_step == 0
    ours._handle_sync_gradients(1, g, 0)
    assert ours.global_step == 1


def test_true_barrier_blocks_and_times_out_q1_q3():
    import threading
    import time
    ours = S.ParameterServerServicer("sync", 2, 0.1, initial_state=_small_state(),
                                     config=Config(barrier_timeout_s=0.4))
    g = _grads(1, _small_state())
    t0 = time.time()
    done = []
    th = threading.Thread(target=lambda: done.append(ours._handle_sync_gradients(0, g, 0)))
    th.start()
    time.sleep(0.1)
    assert th.is_alive() and ours.global_step == 0          # blocked: a true barrier
    th.join(2.0)
    assert done == [True] and ours.global_step == 1 and ours.barrier_timeouts == 1
    assert time.time() - t0 >= 0.35


def test_registration_reslots_ids_q10():
    ours = S.ParameterServerServicer("sync", 2, 0.1, initial_state=_small_state())
    a = ours.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None).worker_id
    b = ours.RegisterWorker(proto.RegisterRequest(worker_ip="b"), None).worker_id
    assert (a, b) == (0, 1)
    with pytest.raises(RuntimeError):
        ours.RegisterWorker(proto.RegisterRequest(worker_ip="c"), None)
    ours.evict(0)
    assert ours.RegisterWorker(proto.RegisterRequest(worker_ip="c"), None).worker_id == 0


def test_final_metrics_emitted_once_q11(capsys):
    ours = S.ParameterServerServicer("sync", 1, 0.1, initial_state=_small_state())
    wid = ours.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None).worker_id
    ours.JobFinished(proto.JobFinishedRequest(worker_id=wid), None)
    ours.JobFinished(proto.JobFinishedRequest(worker_id=wid), None)
    out = capsys.readouterr().out
    assert out.count("METRICS_JSON:") == 1 and "SERVER_FINAL_METRICS" in out


def test_finished_worker_does_not_stall_the_barrier():
    """Uneven partitions: a worker that has sent JobFinished is no longer waited for (the true barrier would
    otherwise time out on every remaining round of the others)."""
    import threading
    import time
    ours = S.ParameterServerServicer("sync", 2, 0.1, initial_state=_small_state(),
                                     config=Config(barrier_timeout_s=30.0))
    a = ours.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None).worker_id
    b = ours.RegisterWorker(proto.RegisterRequest(worker_ip="b"), None).worker_id
    g = _grads(1, _small_state())
    # b leaves while a is already waiting in the barrier: a's round completes at once
    done = []
    th = threading.Thread(target=lambda: done.append(ours._handle_sync_gradients(a, g, 0)))
    th.start()
    time.sleep(0.1)
    assert th.is_alive()
    t0 = time.time()
    ours.JobFinished(proto.JobFinishedRequest(worker_id=b), None)
    th.join(5.0)
    assert done == [True] and ours.global_step == 1 and ours.barrier_timeouts == 0 and time.time() - t0 < 2.0
    # and later rounds of a complete immediately
    t0 = time.time()
    assert ours._handle_sync_gradients(a, g, 1) and ours.global_step == 2 and time.time() - t0 < 1.0


def test_liveness_monitor_evicts_silent_worker_and_readmits_it():
    import time
    ours = S.ParameterServerServicer("sync", 2, 0.1, initial_state=_small_state(),
                                     config=Config(barrier_timeout_s=30.0, worker_timeout_s=0.3))
    a = ours.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None).worker_id
    b = ours.RegisterWorker(proto.RegisterRequest(worker_ip="b"), None).worker_id
    ours.start_monitor(period_s=0.05)
    try:
        g = _grads(1, _small_state())
        deadline = time.time() + 3.0
        while b in ours.active_workers and time.time() < deadline:
            ours._touch(a)                                   # a keeps talking, b is silent
            time.sleep(0.05)
        ours._touch(a)                                       # (revives a if a stalled test process let it lapse too)
        assert b not in ours.active_workers and a in ours.active_workers and ours._expected() == 1
        t0 = time.time()
        assert ours._handle_sync_gradients(a, g, 0) and ours.global_step == 1 and time.time() - t0 < 1.0
        # the evicted worker comes back: same id (same data partition), and the barrier waits for it again
        again = ours.RegisterWorker(proto.RegisterRequest(worker_ip="b2"), None).worker_id
        assert again == b and ours._expected() == 2
    finally:
        ours.stop_monitor()


def test_evicted_but_alive_worker_is_revived_by_its_next_rpc():
    ours = S.ParameterServerServicer("sync", 2, 0.1, initial_state=_small_state(), config=Config(barrier_timeout_s=5.0))
    a = ours.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None).worker_id
    b = ours.RegisterWorker(proto.RegisterRequest(worker_ip="b"), None).worker_id
    ours.evict(b)
    assert ours._expected() == 1 and b not in ours.active_workers
    ours.FetchParameters(proto.FetchRequest(worker_id=b), None)           # b was only slow
    assert b in ours.active_workers and ours._expected() == 2 and b not in ours._free_ids
    # ... unless its slot was handed to a replacement in the meantime
    ours.evict(b)
    c = ours.RegisterWorker(proto.RegisterRequest(worker_ip="c"), None).worker_id
    assert c == b
    assert ours._expected() == 2
