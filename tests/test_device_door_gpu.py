"""The gRPC front door onto the on-device store (dps_b200/rpc/device_door.py): the reference's OWN client stub and
this framework's CPU worker talk to the GPU shards of a running job (reference: the four RPCs of
src/parameter_server/server.py:190-318)."""
import pickle
import threading

import numpy as np
import pytest
import torch

from dps_b200.config import Config

pytestmark = pytest.mark.gpu


def _job(**kw):
    from dps_b200.api import Job
    cfg = Config(mode="sync", workers=1, batch_size=16, lr=0.01, serve_port=0, barrier_timeout_s=20.0, **kw)
    return Job(cfg, rank=0, world=1, device_index=0)


def _drive(job, stop, x, y):
    """The on-box worker keeps stepping (zero learning signal is fine) until the remote side is done."""
    n = 0
    while not stop.is_set() and n < 4000:
        job.worker.train_step(x, y).item()
        n += 1
    return n


def test_reference_stub_registers_fetches_and_pushes_against_a_running_gpu_job(ref_modules):
    import grpc
    job = _job(compat_pickle=True)
    assert job.cfg.remote_slots == 1 and job.door_port
    pb2, pb2_grpc = ref_modules["ps_pb2"], ref_modules["ps_pb2_grpc"]
    g = torch.Generator().manual_seed(0)
    x = torch.randn(16, 3, 32, 32, generator=g).pin_memory()
    y = torch.randint(0, 100, (16,), generator=g).pin_memory()
    job.worker.train_step(x, y).item()      # capture the step graph before the remote attaches
    result, stop = {}, threading.Event()

    def remote():
        try:
            ch = grpc.insecure_channel(f"127.0.0.1:{job.door_port}",
                                       options=[("grpc.max_send_message_length", 500 * 1024 * 1024),
                                                ("grpc.max_receive_message_length", 500 * 1024 * 1024)])
            stub = pb2_grpc.ParameterServerStub(ch)
            reg = stub.RegisterWorker(pb2.RegisterRequest(worker_ip="ref-client"))
            result["reg"] = (reg.worker_id, reg.total_workers)
            rep = stub.FetchParameters(pb2.FetchRequest(worker_id=reg.worker_id))
            sd0 = pickle.loads(rep.parameters)                  # the reference unpickles (worker.py:241)
            result["sd0"], result["step0"] = sd0, rep.global_step
            model = ref_modules["server"].ResNet18(num_classes=100)
            model.load_state_dict({k: torch.tensor(v) for k, v in sd0.items()}, strict=True)
            names = [n for n, _ in model.named_parameters()]
            rng = np.random.default_rng(1)
            grads = {n: rng.normal(size=sd0[n].shape).astype(np.float16) for n in names}
            result["grads"] = grads
            ok = stub.PushGradrients(pb2.PushRequest(worker_id=reg.worker_id, gradients=pickle.dumps(grads),
                                                     local_step=rep.global_step))
            result["received"] = ok.received
            rep1 = stub.FetchParameters(pb2.FetchRequest(worker_id=reg.worker_id))
            result["sd1"], result["step1"] = pickle.loads(rep1.parameters), rep1.global_step
            stub.JobFinished(pb2.JobFinishedRequest(worker_id=reg.worker_id))
        except Exception as e:          # pragma: no cover
            result["error"] = repr(e)
        finally:
            stop.set()

    t = threading.Thread(target=remote)
    t.start()
    _drive(job, stop, x, y)
    t.join(60)
    assert "error" not in result, result.get("error")
    assert result["reg"] == (1, 2) and result["received"] is True, job.door.last_reject_reason
    assert len(result["sd0"]) == 122 and result["step1"] > result["step0"]
    # The round that consumed the remote gradient g averaged over W' = 2 sources (the on-box worker and the remote
    # one): that round moved the master by -lr * (g_local + g) / 2; every other round by -lr * g_local.  g is random
    # and large, the local gradients are uncorrelated with it, so the regression coefficient of the total update on g
    # is -lr / W' = -0.005.
    name = "layer4.1.conv2.weight"
    d = result["sd1"][name].astype(np.float64) - result["sd0"][name].astype(np.float64)
    gr = result["grads"][name].astype(np.float64)
    coef = float((d * gr).sum() / (gr * gr).sum())
    assert abs(coef + 0.005) < 0.00025, coef
    assert job.ps.error_bits() == 0
    job.close()


def test_cpu_rpc_worker_trains_alongside_one_gpu_worker_for_five_sync_rounds():
    from dps_b200.rpc.client import DistributedWorker
    job = _job()
    g = torch.Generator().manual_seed(0)
    x = torch.randn(16, 3, 32, 32, generator=g).pin_memory()
    y = torch.randint(0, 100, (16,), generator=g).pin_memory()
    job.worker.train_step(x, y).item()
    result, stop = {}, threading.Event()

    def remote():
        try:
            cfg = Config(mode="sync", workers=2, batch_size=8, lr=0.1, train_samples=64, test_samples=16)
            w = DistributedWorker(f"127.0.0.1:{job.door_port}", "cpu-worker", config=cfg)
            w.connect_to_server()
            w.setup_model()
            w.setup_data()
            w.max_steps = 5
            w.run_training(evaluate=False)
            result["steps"] = w.local_step_counter
            result["id"] = w.worker_id
        except Exception as e:          # pragma: no cover
            result["error"] = repr(e)
        finally:
            stop.set()

    t = threading.Thread(target=remote)
    t.start()
    v0 = job.ps.version()
    _drive(job, stop, x, y)
    t.join(120)
    assert "error" not in result, result.get("error")
    assert result["steps"] == 5 and result["id"] == 1
    assert job.door.gradients_processed == 5 and job.door.rejected == 0
    assert job.ps.version() >= v0 + 5 and job.ps.error_bits() == 0
    assert job.door.metrics()["remote_attached"] == 0          # JobFinished detached it
    job.close()
