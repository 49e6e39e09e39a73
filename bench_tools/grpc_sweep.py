"""BASELINE.json config 5, gRPC column: push / pull latency and effective bandwidth over gRPC localhost for payloads of
1 KB - 64 MB, the reference's servicer (pickle payloads, its own generated stubs) vs this framework's servicer (DPS1
frames).  CPU only; the device columns of the same sweep are `bench_tools/pushpull_sweep.py`.

Payload = bytes of fp16 gradient per push (so the parameter tensor has payload/2 fp32 elements and a pull moves
2 x payload bytes), as in `profiles/pushpull_sweep_8gpu.json`.  One client, sync mode with one expected worker, so a
push includes decode + fp16->fp32 + SGD apply on the server, a pull includes encode on the server and decode on the
client -- the same work the reference's RPCs do (reference: server.py:213-262, worker.py:233-311).

    python bench_tools/grpc_sweep.py [--reps 5] [--out profiles/grpc_sweep_cpu.json]
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import pickle
import sys
import time
from concurrent import futures

import grpc
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dps_b200.config import Config                                   # noqa: E402
from dps_b200.rpc import codec, proto, servicer as S                 # noqa: E402

SIZES = [1 << 10, 8 << 10, 64 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20]
OPTS = [("grpc.max_send_message_length", 500 * 1024 * 1024), ("grpc.max_receive_message_length", 500 * 1024 * 1024)]


def _ref_modules():
    for base in ("/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        src = os.path.join(base, "src")
        if os.path.isdir(src):
            if src not in sys.path:
                sys.path.insert(0, src)
            return (importlib.import_module("communication.ps_pb2"),
                    importlib.import_module("communication.ps_pb2_grpc"),
                    importlib.import_module("parameter_server.server"))
    return None


def _time(fn, reps):
    fn()                                                             # warm-up
    ts = []
    for _ in range(reps):
        t = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t)
    return min(ts)


def sweep_ours(reps: int):
    rows = []
    for nbytes in SIZES:
        n = nbytes // 2
        state = {"w": np.zeros(n, np.float32)}
        sv = S.ParameterServerServicer("sync", 1, 0.1, initial_state=state, config=Config(workers=1))
        server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=OPTS)
        proto.add_ParameterServerServicer_to_server(sv, server)
        port = server.add_insecure_port("127.0.0.1:0")
        server.start()
        ch = grpc.insecure_channel(f"127.0.0.1:{port}", options=OPTS)
        stub = proto.ParameterServerStub(ch)
        wid = stub.RegisterWorker(proto.RegisterRequest(worker_ip="sweep")).worker_id
        g = {"w": np.ones(n, np.float32)}

        def push():
            payload = codec.encode_any(codec.compress_gradients(g, "fp16"), False)
            assert stub.PushGradrientsRaw(proto.encode_push_request_wire(wid, payload, 0)).received   # as the client

        def pull():
            raw = stub.FetchParametersRaw(proto.FetchRequest(worker_id=wid))     # what DistributedWorker does
            payload, _ = proto.split_fetch_reply_wire(raw)
            arrs, _ = codec.decode_any(payload, allow_pickle=False)
            return arrs["w"][:1].copy()

        rows.append({"payload_bytes": nbytes, "push_s": _time(push, reps), "pull_s": _time(pull, reps)})
        ch.close()
        server.stop(0)
    return rows


def sweep_reference(reps: int):
    mods = _ref_modules()
    if mods is None:
        return None
    pb2, pb2_grpc, refsrv = mods
    rows = []
    for nbytes in SIZES:
        n = nbytes // 2
        sv = refsrv.ParameterServerServicer(mode="sync", total_workers=1, learning_rate=0.1)
        sv.parameters = {"w": np.zeros(n, np.float32)}
        server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=OPTS)
        pb2_grpc.add_ParameterServerServicer_to_server(sv, server)
        port = server.add_insecure_port("127.0.0.1:0")
        server.start()
        ch = grpc.insecure_channel(f"127.0.0.1:{port}", options=OPTS)
        stub = pb2_grpc.ParameterServerStub(ch)
        wid = stub.RegisterWorker(pb2.RegisterRequest(worker_ip="sweep")).worker_id
        g = {"w": np.ones(n, np.float32)}

        def push():                                                   # what worker.py:264-301 does
            payload = pickle.dumps({k: v.astype(np.float16) for k, v in g.items()})
            assert stub.PushGradrients(pb2.PushRequest(worker_id=wid, gradients=payload, local_step=0)).received

        def pull():                                                   # what worker.py:233-252 does
            rep = stub.FetchParameters(pb2.FetchRequest(worker_id=wid))
            return pickle.loads(rep.parameters)["w"][:1].copy()

        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):               # the reference prints on every RPC
            rows.append({"payload_bytes": nbytes, "push_s": _time(push, reps), "pull_s": _time(pull, reps)})
        ch.close()
        server.stop(0)
    return rows


def main() -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--reps", type=int, default=5)
    p.add_argument("--out", default=None)
    a = p.parse_args()
    ours, ref = sweep_ours(a.reps), sweep_reference(a.reps)
    out = {"host": {"cpus": os.cpu_count()}, "note": "gRPC localhost, one client, min over reps; payload = bytes of "
           "fp16 gradient per push, pull moves 2 x payload (fp32 parameters)", "rows": []}
    for i, r in enumerate(ours):
        row = {"payload_bytes": r["payload_bytes"], "ours_push_ms": round(r["push_s"] * 1e3, 3),
               "ours_pull_ms": round(r["pull_s"] * 1e3, 3),
               "ours_push_GBps": round(r["payload_bytes"] / r["push_s"] / 1e9, 4),
               "ours_pull_GBps": round(2 * r["payload_bytes"] / r["pull_s"] / 1e9, 4)}
        if ref:
            q = ref[i]
            row.update({"ref_push_ms": round(q["push_s"] * 1e3, 3), "ref_pull_ms": round(q["pull_s"] * 1e3, 3),
                        "ref_push_GBps": round(q["payload_bytes"] / q["push_s"] / 1e9, 4),
                        "ref_pull_GBps": round(2 * q["payload_bytes"] / q["pull_s"] / 1e9, 4)})
        out["rows"].append(row)
        print(row, flush=True)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
