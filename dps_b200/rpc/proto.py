"""Programmatic protobuf descriptors + gRPC plumbing for the ParameterServer service.

Replaces the reference's generated modules (reference: src/communication/ps_pb2.py,
src/communication/ps_pb2_grpc.py, regenerate_proto.sh) without needing protoc:
the FileDescriptorProto is assembled with ``descriptor_pb2`` and registered in a
*private* DescriptorPool, so this module can live in the same process as the
reference's own ``ps_pb2`` (needed by the wire-compat tests) without a duplicate
symbol clash in the default pool.

Public surface mirrors the generated code:
  RegisterRequest/RegisterReply/PushRequest/PushReply/FetchRequest/FetchReply/
  JobFinishedRequest/JobFinishedReply  (message classes)
  ParameterServerStub(channel)                      (ps_pb2_grpc.py:28)
  ParameterServerServicer                           (ps_pb2_grpc.py:60)
  add_ParameterServerServicer_to_server(svc, srv)   (ps_pb2_grpc.py:95)
"""
from __future__ import annotations

from typing import Tuple

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

SERVICE_NAME = "ParameterServer"          # no proto package -> '/ParameterServer/<Method>'
_F = descriptor_pb2.FieldDescriptorProto

# (message, [(field, number, type)])
_MESSAGES = [
    ("RegisterRequest", [("worker_ip", 1, _F.TYPE_STRING)]),
    ("RegisterReply", [("worker_id", 1, _F.TYPE_INT32), ("total_workers", 2, _F.TYPE_INT32)]),
    ("PushRequest", [("worker_id", 1, _F.TYPE_INT32), ("gradients", 2, _F.TYPE_BYTES),
                     ("local_step", 3, _F.TYPE_INT32)]),
    ("PushReply", [("received", 1, _F.TYPE_BOOL)]),
    ("FetchRequest", [("worker_id", 1, _F.TYPE_INT32)]),
    ("FetchReply", [("parameters", 1, _F.TYPE_BYTES), ("global_step", 2, _F.TYPE_INT32)]),
    ("JobFinishedRequest", [("worker_id", 1, _F.TYPE_INT32)]),
    ("JobFinishedReply", [("message", 1, _F.TYPE_STRING)]),
]

# (method, request, reply) -- note the misspelt PushGradrients: it is the wire name.
METHODS = [
    ("RegisterWorker", "RegisterRequest", "RegisterReply"),
    ("PushGradrients", "PushRequest", "PushReply"),
    ("FetchParameters", "FetchRequest", "FetchReply"),
    ("JobFinished", "JobFinishedRequest", "JobFinishedReply"),
]


def build_file_descriptor_proto() -> descriptor_pb2.FileDescriptorProto:
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name = "dps_b200/ps.proto"
    fdp.syntax = "proto3"
    for mname, fields in _MESSAGES:
        m = fdp.message_type.add()
        m.name = mname
        for fname, num, ftype in fields:
            f = m.field.add()
            f.name, f.number, f.type = fname, num, ftype
            f.label = _F.LABEL_OPTIONAL
    svc = fdp.service.add()
    svc.name = SERVICE_NAME
    for meth, req, rep in METHODS:
        md = svc.method.add()
        md.name, md.input_type, md.output_type = meth, "." + req, "." + rep
    return fdp


_POOL = descriptor_pool.DescriptorPool()
_FILE = _POOL.AddSerializedFile(build_file_descriptor_proto().SerializeToString())
_CLASSES = {name: message_factory.GetMessageClass(_POOL.FindMessageTypeByName(name))
            for name, _ in _MESSAGES}

RegisterRequest = _CLASSES["RegisterRequest"]
RegisterReply = _CLASSES["RegisterReply"]
PushRequest = _CLASSES["PushRequest"]
PushReply = _CLASSES["PushReply"]
FetchRequest = _CLASSES["FetchRequest"]
FetchReply = _CLASSES["FetchReply"]
JobFinishedRequest = _CLASSES["JobFinishedRequest"]
JobFinishedReply = _CLASSES["JobFinishedReply"]

# 500 MB messages + keepalive, as the reference (server.py:372-378, worker.py:203-209).
GRPC_OPTIONS = [
    ("grpc.max_receive_message_length", 500 * 1024 * 1024),
    ("grpc.max_send_message_length", 500 * 1024 * 1024),
    ("grpc.keepalive_time_ms", 30000),
    ("grpc.keepalive_timeout_ms", 5000),
    ("grpc.keepalive_permit_without_calls", True),
]


def method_path(method: str) -> str:
    return f"/{SERVICE_NAME}/{method}"


def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def encode_fetch_reply_wire(parameters: bytes, global_step: int) -> bytes:
    """Serialized FetchReply{parameters = 1, global_step = 2} built with ONE copy of the payload (the generated
    serializer copies it into the message and again into the output: 2 x 45 MB per pull)."""
    return b"".join([b"\x0a", _varint(len(parameters)), parameters, b"\x10",
                     _varint(global_step & 0xFFFFFFFFFFFFFFFF)])       # int32 is sign-extended to 64 bits on the wire


def split_fetch_reply_wire(raw) -> Tuple[memoryview, int]:
    """(zero-copy view of `parameters`, global_step) from a serialized FetchReply; falls back to the protobuf
    parser for any layout other than the two known fields."""
    mv = memoryview(raw)
    pos, n = 0, len(mv)
    params, step = memoryview(b""), 0
    try:
        while pos < n:
            tag = mv[pos]
            pos += 1
            val, shift = 0, 0
            while True:
                b = mv[pos]
                pos += 1
                val |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            if tag == 0x0A:
                params = mv[pos:pos + val]
                if len(params) != val:
                    raise ValueError("truncated")
                pos += val
            elif tag == 0x10:
                step = val - (1 << 64) if val >= (1 << 63) else val
            else:
                raise ValueError("unknown field")
        return params, int(step)
    except (ValueError, IndexError):
        msg = _CLASSES["FetchReply"].FromString(bytes(raw))
        return memoryview(msg.parameters), int(msg.global_step)


def encode_push_request_wire(worker_id: int, gradients: bytes, local_step: int) -> bytes:
    """Serialized PushRequest{worker_id = 1, gradients = 2, local_step = 3} with one copy of the payload."""
    return b"".join([b"\x08", _varint(worker_id & 0xFFFFFFFFFFFFFFFF), b"\x12", _varint(len(gradients)), gradients,
                     b"\x18", _varint(local_step & 0xFFFFFFFFFFFFFFFF)])


def split_push_request_wire(raw) -> Tuple[int, memoryview, int]:
    """(worker_id, zero-copy view of `gradients`, local_step) from a serialized PushRequest (any field order;
    unknown fields fall back to the protobuf parser)."""
    mv = memoryview(raw)
    pos, n = 0, len(mv)
    wid, grads, step = 0, memoryview(b""), 0
    try:
        while pos < n:
            tag = mv[pos]
            pos += 1
            val, shift = 0, 0
            while True:
                b = mv[pos]
                pos += 1
                val |= (b & 0x7F) << shift
                shift += 7
                if not b & 0x80:
                    break
            if tag == 0x12:
                grads = mv[pos:pos + val]
                if len(grads) != val:
                    raise ValueError("truncated")
                pos += val
            elif tag in (0x08, 0x18):
                val = val - (1 << 64) if val >= (1 << 63) else val
                if tag == 0x08:
                    wid = val
                else:
                    step = val
            else:
                raise ValueError("unknown field")
        return int(wid), grads, int(step)
    except (ValueError, IndexError):
        msg = _CLASSES["PushRequest"].FromString(bytes(raw))
        return int(msg.worker_id), memoryview(msg.gradients), int(msg.local_step)


class ParameterServerStub:
    """Client stub; one unary-unary callable per RPC (same attribute names as generated code), plus
    `FetchParametersRaw`: the same RPC with the reply left as raw bytes for `split_fetch_reply_wire`."""

    def __init__(self, channel: grpc.Channel):
        for meth, req, rep in METHODS:
            setattr(self, meth, channel.unary_unary(
                method_path(meth),
                request_serializer=_CLASSES[req].SerializeToString,
                response_deserializer=_CLASSES[rep].FromString,
            ))
        self.FetchParametersRaw = channel.unary_unary(
            method_path("FetchParameters"),
            request_serializer=_CLASSES["FetchRequest"].SerializeToString,
            response_deserializer=lambda b: b)
        # request = bytes from encode_push_request_wire (no copy of the payload into a message object)
        self.PushGradrientsRaw = channel.unary_unary(
            method_path("PushGradrients"),
            request_serializer=lambda b: b,
            response_deserializer=_CLASSES["PushReply"].FromString)


class ParameterServerServicer:
    """Abstract service; subclasses override the four RPC methods."""

    def _unimplemented(self, context):
        context.set_code(grpc.StatusCode.UNIMPLEMENTED)
        context.set_details("Method not implemented!")
        raise NotImplementedError("Method not implemented!")

    def RegisterWorker(self, request, context):
        self._unimplemented(context)

    def PushGradrients(self, request, context):
        self._unimplemented(context)

    def FetchParameters(self, request, context):
        self._unimplemented(context)

    def JobFinished(self, request, context):
        self._unimplemented(context)


def add_ParameterServerServicer_to_server(servicer, server: grpc.Server) -> None:
    handlers = {}
    for meth, req, rep in METHODS:
        fn, des, ser = getattr(servicer, meth), _CLASSES[req].FromString, _CLASSES[rep].SerializeToString
        wire = getattr(servicer, meth + "Wire", None)
        if wire is not None and meth == "FetchParameters":      # handler returns already-serialized reply bytes
            fn, ser = wire, (lambda b: b)
        elif wire is not None and meth == "PushGradrients":     # handler takes the raw request bytes
            fn, des = wire, (lambda b: b)
        handlers[meth] = grpc.unary_unary_rpc_method_handler(fn, request_deserializer=des, response_serializer=ser)
    server.add_generic_rpc_handlers(
        (grpc.method_handlers_generic_handler(SERVICE_NAME, handlers),))


def make_channel(address: str) -> grpc.Channel:
    return grpc.insecure_channel(address, options=GRPC_OPTIONS)
