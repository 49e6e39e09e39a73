"""two-way wire compatibility over real gRPC on localhost.

  * reference ParameterServerStub (generated code)  <->  our servicer
  * our client stub                                  <->  reference ParameterServerServicer
"""
import pickle
from concurrent import futures

import grpc
import numpy as np

from dps_b200.config import Config
from dps_b200.rpc import codec, proto, servicer as S


def _state():
    rng = np.random.default_rng(3)
    return {"fc.weight": rng.normal(size=(10, 6)).astype(np.float32),
            "fc.bias": rng.normal(size=(10,)).astype(np.float32)}


def test_reference_stub_against_our_server(ref_modules):
    pb, pbg = ref_modules["ps_pb2"], ref_modules["ps_pb2_grpc"]
    cfg = Config(mode="async", workers=2, lr=0.5, port=0, compat_pickle=True, staleness_bound=1)
    sv = S.ParameterServerServicer("async", 2, 0.5, 1, config=cfg, initial_state=_state())
    server, sv, port = S.build_server(cfg, sv, bind="127.0.0.1")
    server.start()
    try:
        ch = grpc.insecure_channel(f"127.0.0.1:{port}")
        stub = pbg.ParameterServerStub(ch)
        r = stub.RegisterWorker(pb.RegisterRequest(worker_ip="ref-worker"))
        assert (r.worker_id, r.total_workers) == (0, 2)
        f = stub.FetchParameters(pb.FetchRequest(worker_id=0))
        params = pickle.loads(f.parameters)            # the reference worker uses plain pickle
        assert f.global_step == 0 and set(params) == set(_state())
        g = {k: np.ones_like(v, dtype=np.float16) for k, v in _state().items()}
        p = stub.PushGradrients(pb.PushRequest(worker_id=0, gradients=pickle.dumps(g), local_step=0))
        assert p.received is True
        f2 = stub.FetchParameters(pb.FetchRequest(worker_id=0))
        np.testing.assert_allclose(pickle.loads(f2.parameters)["fc.bias"], _state()["fc.bias"] - 0.5)
        assert f2.global_step == 1
        # stale push (bound=1): global 3 vs local 0 -> received=False, like the reference's reject path
        for _ in range(2):
            stub.PushGradrients(pb.PushRequest(worker_id=0, gradients=pickle.dumps(g), local_step=f2.global_step))
        rej = stub.PushGradrients(pb.PushRequest(worker_id=0, gradients=pickle.dumps(g), local_step=0))
        assert rej.received is False
        assert stub.JobFinished(pb.JobFinishedRequest(worker_id=0)).message == "Acknowledged"
        ch.close()
    finally:
        server.stop(0)


def test_our_stub_against_reference_server(ref_modules, capsys):
    pbg, refsrv = ref_modules["ps_pb2_grpc"], ref_modules["server"]
    ref = refsrv.ParameterServerServicer(mode="sync", total_workers=1, learning_rate=0.25)
    ref.parameters = {k: v.copy() for k, v in _state().items()}
    server = grpc.server(futures.ThreadPoolExecutor(4), options=proto.GRPC_OPTIONS)
    pbg.add_ParameterServerServicer_to_server(ref, server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    try:
        ch = proto.make_channel(f"127.0.0.1:{port}")
        stub = proto.ParameterServerStub(ch)
        r = stub.RegisterWorker(proto.RegisterRequest(worker_ip="ours"))
        assert (r.worker_id, r.total_workers) == (0, 1)
        f = stub.FetchParameters(proto.FetchRequest(worker_id=0))
        params, _ = codec.decode_any(f.parameters, allow_pickle=True)
        np.testing.assert_array_equal(params["fc.weight"], _state()["fc.weight"])
        g = codec.compress_gradients({k: np.full_like(v, 2.0) for k, v in _state().items()})
        p = stub.PushGradrients(proto.PushRequest(worker_id=0, gradients=codec.reference_pickle_dumps(g),
                                                  local_step=f.global_step))
        assert p.received is True
        f2 = stub.FetchParameters(proto.FetchRequest(worker_id=0))
        params2, _ = codec.decode_any(f2.parameters, allow_pickle=True)
        np.testing.assert_allclose(params2["fc.weight"], _state()["fc.weight"] - 0.25 * 2.0, rtol=1e-6)
        assert f2.global_step == 1
        assert stub.JobFinished(proto.JobFinishedRequest(worker_id=0)).message == "Acknowledged"
        ch.close()
    finally:
        server.stop(0)


def test_large_message_limits():
    """500 MB limits are configured on both ends; a 40 MB payload (>4 MB gRPC default) passes."""
    cfg = Config(mode="async", workers=1, port=0)
    big = {"w": np.zeros(10_000_000, np.float32)}
    sv = S.ParameterServerServicer("async", 1, 0.1, config=cfg, initial_state=big)
    server, sv, port = S.build_server(cfg, sv, bind="127.0.0.1")
    server.start()
    try:
        stub = proto.ParameterServerStub(proto.make_channel(f"127.0.0.1:{port}"))
        f = stub.FetchParameters(proto.FetchRequest(worker_id=0))
        assert len(f.parameters) > 39_000_000
        ok = stub.PushGradrients(proto.PushRequest(
            worker_id=0, gradients=codec.encode_frame({"w": np.ones(10_000_000, np.float16)}), local_step=0))
        assert ok.received
    finally:
        server.stop(0)


def test_pickle_refused_without_compat_flag():
    cfg = Config(mode="async", workers=1, port=0, compat_pickle=False)
    sv = S.ParameterServerServicer("async", 1, 0.1, config=cfg, initial_state=_state())
    evil = pickle.dumps({"fc.bias": np.ones(10, np.float16)})
    rep = sv.PushGradrients(proto.PushRequest(worker_id=0, gradients=evil, local_step=0), None)
    assert rep.received is False and sv.global_step == 0
