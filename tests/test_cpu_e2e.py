"""BASELINE config 1: ResNet-18 CIFAR-100-shape synthetic, 1 server + 2 workers over gRPC
localhost on CPU (plumbing, no GPU).  Also async mode and the sync == large-batch-SGD property."""
import threading

import numpy as np
import pytest
import torch

from dps_b200.config import Config
from dps_b200.models.resnet import build_model
from dps_b200.rpc import servicer as S
from dps_b200.rpc.client import DistributedWorker
from dps_b200.utils import metrics as M


def _run_job(cfg: Config, steps: int, tmp_path):
    server, sv, port = S.build_server(cfg.replace(port=0), bind="127.0.0.1")
    server.start()
    workers, errs = [], []

    def run(i):
        try:
            w = DistributedWorker(f"127.0.0.1:{port}", f"w{i}", config=cfg)
            w.connect_to_server()
            w.setup_model()
            w.setup_data()
            w.max_steps = steps
            workers.append(w)
            w.run_training(evaluate=True, eval_batches=1)
        except Exception as e:  # pragma: no cover
            errs.append(e)

    ths = [threading.Thread(target=run, args=(i,)) for i in range(cfg.workers)]
    [t.start() for t in ths]
    [t.join(300) for t in ths]
    server.stop(0)
    assert not errs, errs
    return sv, workers


@pytest.mark.parametrize("mode", ["sync", "async"])
def test_one_server_two_workers_grpc_cpu(mode, tmp_path):
    torch.set_num_threads(2)
    cfg = Config(mode=mode, workers=2, lr=0.05, batch_size=8, epochs=1, train_samples=64, test_samples=16,
                 metrics_dir=str(tmp_path), barrier_timeout_s=120)
    sv, workers = _run_job(cfg, steps=2, tmp_path=tmp_path)
    assert len(workers) == 2 and sorted(w.worker_id for w in workers) == [0, 1]
    assert {w.partition for w in workers} == {(0, 32), (32, 64)}          # disjoint contiguous (C15)
    assert all(w.local_step_counter == 2 for w in workers)
    assert sv.gradients_processed == 4
    assert sv.global_step == (2 if mode == "sync" else 4 - sv.rejected_updates)
    # metrics: one server record, two worker records, aggregate schema (C23/C24/C29)
    res = M.collect_dir(str(tmp_path), f"{mode}_2workers", str(tmp_path / "out" / "res.json"))
    assert res["server_metrics"]["type"] == "SERVER_FINAL_METRICS"
    assert res["server_metrics"]["mode"] == mode
    agg = res["worker_metrics_aggregated"]
    assert agg["num_workers"] == 2 and agg["total_local_steps"] == 4
    assert {"epoch", "max_time", "avg_time", "min_time"} <= set(agg["epoch_times_by_epoch"][0])
    # parameters actually moved and stayed finite
    init = {k: v.detach().numpy() for k, v in build_model(100, seed=cfg.seed).state_dict().items()}
    cur = sv.state_dict_numpy()
    assert set(cur) == set(init) and len(cur) == 122
    assert not np.allclose(cur["fc.weight"], init["fc.weight"])
    assert all(np.isfinite(v).all() for v in cur.values() if v.dtype == np.float32)
    # BN running stats were maintained on the server (Q5 fix, default policy 'mean')
    assert not np.allclose(cur["bn1.running_mean"], 0.0)


def test_sync_step_equals_large_batch_sgd():
    """SURVEY §4.5 (CPU form): one sync round over W workers == one SGD step on the mean gradient."""
    torch.manual_seed(0)
    W = 2
    model = build_model(100, seed=1)
    init = {k: v.detach().clone() for k, v in model.state_dict().items()}
    xs = [torch.randn(4, 3, 32, 32) for _ in range(W)]
    ys = [torch.randint(0, 100, (4,)) for _ in range(W)]
    sv = S.ParameterServerServicer("sync", W, 0.1, config=Config(nonblocking_sync=True, compress="none"),
                                   initial_state={k: v.numpy() for k, v in init.items()})
    grads_sum = None
    for w in range(W):
        model.load_state_dict(init)
        model.train()
        model.zero_grad()
        torch.nn.functional.cross_entropy(model(xs[w]), ys[w]).backward()
        g = {n: p.grad.numpy().copy() for n, p in model.named_parameters()}
        sv._handle_sync_gradients(w, g, 0)
        grads_sum = g if grads_sum is None else {k: grads_sum[k] + g[k] for k in g}
    for n, p in init.items():
        if n in grads_sum:
            expect = p.numpy() - 0.1 * (grads_sum[n] / W)
            np.testing.assert_allclose(sv.parameters[n], expect, rtol=1e-6, atol=1e-7)


def test_bf16_autocast_worker_keeps_fp32_master(tmp_path):
    """--amp bf16 (gRPC/CPU worker): the local step runs under bf16 autocast + channels_last, gradients and the
    server's master weights stay fp32, and the update is close to the fp32 worker's."""
    torch.set_num_threads(2)
    base = Config(mode="sync", workers=1, lr=0.05, batch_size=8, epochs=1, train_samples=16, test_samples=8,
                  barrier_timeout_s=120, seed=1)
    sv32, w32 = _run_job(base, steps=1, tmp_path=tmp_path)
    sv16, w16 = _run_job(base.replace(amp="bf16"), steps=1, tmp_path=tmp_path)
    assert w16[0].amp and not w32[0].amp
    a, b = sv32.state_dict_numpy(), sv16.state_dict_numpy()
    assert all(v.dtype in (np.float32, np.int64) for v in b.values())
    init = build_model(100, seed=1).state_dict()["fc.weight"].numpy()
    d32, d16 = a["fc.weight"] - init, b["fc.weight"] - init
    assert np.abs(d32).max() > 0 and np.abs(d16).max() > 0
    cos = float((d32 * d16).sum() / (np.linalg.norm(d32) * np.linalg.norm(d16)))
    assert cos > 0.98, cos


def test_pull_compress_fp16_halves_payload_and_keeps_fp32_master(tmp_path):
    from dps_b200.rpc import codec, proto
    cfg = Config(mode="sync", workers=1, pull_compress="fp16")
    sv = S.ParameterServerServicer("sync", 1, 0.1, config=cfg)
    rep = sv.FetchParameters(proto.FetchRequest(worker_id=0), None)
    full = S.ParameterServerServicer("sync", 1, 0.1, config=Config(workers=1)).FetchParameters(
        proto.FetchRequest(worker_id=0), None)
    assert len(rep.parameters) < 0.51 * len(full.parameters)
    arrs, _ = codec.decode_any(rep.parameters, allow_pickle=False)
    assert arrs["conv1.weight"].dtype == np.float16 and arrs["bn1.num_batches_tracked"].dtype == np.int64
    assert sv.parameters["conv1.weight"].dtype == np.float32
    assert np.allclose(arrs["conv1.weight"].astype(np.float32), sv.parameters["conv1.weight"], atol=1e-3)
    # a worker loads the fp16 frame into its fp32 model
    w = DistributedWorker("unused:0", "w", config=cfg)
    w.setup_model()
    w.worker_id = 0

    class _Stub:
        def FetchParameters(self, req):
            return rep
    w.stub = _Stub()
    w.fetch_parameters()
    got = w.model.state_dict()["conv1.weight"]
    assert got.dtype == torch.float32
    assert torch.allclose(got, torch.from_numpy(sv.parameters["conv1.weight"]), atol=1e-3)


def test_server_checkpoint_and_resume(tmp_path, ref_modules):
    """PS recovery on the gRPC path: the reference loses the model when the server restarts.
    Here the server writes a state_dict-compatible checkpoint (periodically and at job end) and a restarted server
    resumes parameters, momentum and the global step from it."""
    from dps_b200.rpc import codec, proto
    from dps_b200.utils import checkpoint as CK
    ck = str(tmp_path / "ps.pt")
    cfg = Config(mode="sync", workers=1, lr=0.05, momentum=0.9, checkpoint=ck, checkpoint_every=2)
    sv = S.ParameterServerServicer("sync", 1, 0.05, config=cfg)
    wid = sv.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None).worker_id
    rng = np.random.default_rng(0)
    names = [k for k, v in sv.parameters.items() if v.dtype == np.float32 and "running" not in k]

    def push(server, step):
        g = {k: (rng.normal(size=server.parameters[k].shape) * 1e-2).astype(np.float32) for k in names}
        payload = codec.encode_any(codec.compress_gradients(g, "fp16"), False)
        assert server.PushGradrients(proto.PushRequest(worker_id=wid, gradients=payload, local_step=step), None).received
        return g

    for s in range(3):
        push(sv, s)
    # periodic checkpoint after step 2 (written by a background thread); the file loads into the REFERENCE model, strict
    sv.wait_for_checkpoints()
    sd, side = CK.load(ck)
    assert side["global_step"] == 2 and set(side["momentum"]) == set(names)
    ref_model = ref_modules["server"].ResNet18(num_classes=100)
    ref_model.load_state_dict(sd, strict=True)
    sv.JobFinished(proto.JobFinishedRequest(worker_id=wid), None)          # final checkpoint at job end
    sd, side = CK.load(ck)
    assert side["global_step"] == 3
    # restart: same parameters, momentum and step; the next update equals the un-interrupted server's
    sv2 = S.ParameterServerServicer("sync", 1, 0.05, config=cfg.replace(resume=ck, checkpoint=None))
    assert sv2.global_step == 3
    for k in names[:5]:
        assert np.array_equal(sv2.parameters[k], sv.parameters[k])
        assert np.array_equal(sv2.opt.velocity[k], sv.opt.velocity[k])
    sv2.RegisterWorker(proto.RegisterRequest(worker_ip="a"), None)
    state = rng.bit_generator.state
    push(sv, 3)
    rng.bit_generator.state = state                                        # the same gradient for the resumed server
    push(sv2, 3)
    assert sv2.global_step == sv.global_step == 4
    for k in names[:5]:
        assert np.array_equal(sv2.parameters[k], sv.parameters[k]), k


def test_injected_worker_crash_is_survived_on_the_grpc_path(tmp_path):
    """kill a rank mid-job.  Worker 1 dies after its first step without JobFinished; the server's
    liveness monitor evicts it (or the barrier times out once) and worker 0 finishes all its steps."""
    import time
    from dps_b200.rpc.client import SimulatedCrash
    torch.set_num_threads(2)
    cfg = Config(mode="sync", workers=2, lr=0.05, batch_size=8, epochs=1, train_samples=64, test_samples=8,
                 barrier_timeout_s=30.0, worker_timeout_s=4.0, inject_fault="1:1")
    server, sv, port = S.build_server(cfg.replace(port=0), bind="127.0.0.1")
    server.start()
    out = {}

    def run(i):
        w = DistributedWorker(f"127.0.0.1:{port}", f"w{i}", config=cfg)
        try:
            w.connect_to_server()
            w.setup_model()
            w.setup_data()
            out[w.worker_id] = w
            w.run_training(evaluate=False)
        except SimulatedCrash:
            out["crashed"] = w.worker_id
        except Exception as e:  # pragma: no cover
            out["error"] = e

    ths = [threading.Thread(target=run, args=(i,)) for i in range(2)]
    t0 = time.time()
    [t.start() for t in ths]
    [t.join(120) for t in ths]
    sv.stop_monitor()
    server.stop(0)
    assert "error" not in out, out.get("error")
    assert out.get("crashed") == 1
    survivor = out[0]
    assert survivor.local_step_counter == 4                       # its whole partition: 32 samples / 8
    # 4 rounds; 5 only if a loaded machine made a start-up pause look like a death and worker 1's single push was
    # applied as a round of its own before worker 0 was revived by its next RPC
    assert 4 <= sv.global_step <= 5 and 1 not in sv.active_workers
    assert sv.barrier_timeouts <= 1                               # eviction, not a timeout per round
    assert time.time() - t0 < 100


def test_concurrent_async_pushes_claim_a_periodic_checkpoint_once(tmp_path):
    """Async mode, checkpoint_every=1, eight pushes racing on eight threads: every due checkpoint is claimed by exactly
    one push under the parameter lock, writers are serialised, tmp names are unique, no push fails, and the file on disk
    ends up at the last claimed step with a matching sidecar."""
    import threading
    from dps_b200.rpc import codec, proto
    from dps_b200.utils import checkpoint as CK
    ck = str(tmp_path / "async.pt")
    cfg = Config(mode="async", workers=8, lr=0.01, checkpoint=ck, checkpoint_every=1, staleness_bound=100)
    sv = S.ParameterServerServicer("async", 8, 0.01, staleness_bound=100, config=cfg)
    names = [k for k, v in sv.parameters.items() if v.dtype == np.float32 and "running" not in k][:4]
    g = {k: np.full(sv.parameters[k].shape, 1e-3, dtype=np.float32) for k in names}
    payload = codec.encode_any(codec.compress_gradients(g, "fp16"), False)
    ids = [sv.RegisterWorker(proto.RegisterRequest(worker_ip=str(i)), None).worker_id for i in range(8)]
    results = []

    def work(wid):
        r = sv.PushGradrients(proto.PushRequest(worker_id=wid, gradients=payload, local_step=0), None)
        results.append(r.received)

    ts = [threading.Thread(target=work, args=(w,)) for w in ids]
    [t.start() for t in ts]
    [t.join() for t in ts]
    sv.wait_for_checkpoints()
    assert results == [True] * 8 and sv.global_step == 8 and sv.checkpoint_errors == 0
    sd, side = CK.load(ck)
    assert "step_mismatch" not in side and 1 <= side["global_step"] <= 8
    assert side["global_step"] == sv._ckpt_written_step
    assert not [f for f in __import__("os").listdir(tmp_path) if ".tmp." in f]          # nothing left behind


def test_heartbeat_ping_is_a_cheap_rpc_and_refreshes_liveness():
    """The client's liveness ping is a FetchParameters with request metadata that makes the server answer with an EMPTY
    payload (the first version fetched all 45 MB every 30 s) and still counts as a sign of life."""
    import time
    from dps_b200.rpc import proto
    cfg = Config(mode="sync", workers=1, lr=0.1)
    server, sv, port = S.build_server(cfg.replace(port=0), bind="127.0.0.1")
    server.start()
    try:
        w = DistributedWorker(f"127.0.0.1:{port}", "w0", config=cfg)
        w.connect_to_server()
        seen0 = sv.registered_workers[w.worker_id]["last_seen"]
        time.sleep(0.05)
        assert w.ping() == 0
        assert sv.registered_workers[w.worker_id]["last_seen"] > seen0
        rep = w.stub.FetchParameters(proto.FetchRequest(worker_id=w.worker_id), metadata=(("dps-ping", "1"),))
        assert len(rep.parameters) == 0
        full = w.stub.FetchParameters(proto.FetchRequest(worker_id=w.worker_id))
        assert len(full.parameters) > 40_000_000
        w.cleanup()
    finally:
        server.stop(0)
