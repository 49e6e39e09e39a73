"""W1: the programmatic descriptors are wire-identical to the reference's generated code."""
from google.protobuf import descriptor_pb2

from dps_b200.rpc import proto


def _fields(msg_desc):
    return sorted((f.name, f.number, f.type) for f in msg_desc.fields)


def test_method_paths_keep_historical_spelling():
    assert proto.method_path("PushGradrients") == "/ParameterServer/PushGradrients"
    assert [m for m, _, _ in proto.METHODS] == ["RegisterWorker", "PushGradrients", "FetchParameters", "JobFinished"]


def test_descriptor_matches_reference(ref_modules):
    ref = ref_modules["ps_pb2"].DESCRIPTOR
    for name, _ in proto._MESSAGES:
        ours = proto._POOL.FindMessageTypeByName(name)
        theirs = ref.message_types_by_name[name]
        assert _fields(ours) == _fields(theirs), name
    svc = ref.services_by_name["ParameterServer"]
    ref_methods = sorted((m.name, m.input_type.name, m.output_type.name) for m in svc.methods)
    assert ref_methods == sorted(proto.METHODS)
    assert ref.package == ""


def test_bytes_roundtrip_against_reference_classes(ref_modules):
    pb = ref_modules["ps_pb2"]
    ours = proto.PushRequest(worker_id=3, gradients=b"\x00\x01abc", local_step=17).SerializeToString()
    theirs = pb.PushRequest.FromString(ours)
    assert (theirs.worker_id, theirs.gradients, theirs.local_step) == (3, b"\x00\x01abc", 17)
    back = proto.FetchReply.FromString(pb.FetchReply(parameters=b"xyz", global_step=9).SerializeToString())
    assert (back.parameters, back.global_step) == (b"xyz", 9)
    assert proto.RegisterReply.FromString(pb.RegisterReply(worker_id=1, total_workers=8).SerializeToString()).total_workers == 8


def test_proto_file_is_consistent_with_descriptors():
    import os
    import re
    path = os.path.join(os.path.dirname(proto.__file__), "ps.proto")
    text = open(path).read()
    for meth, req, rep in proto.METHODS:
        assert re.search(rf"rpc {meth}\({req}\) returns \({rep}\)", text)
    fdp = proto.build_file_descriptor_proto()
    assert isinstance(fdp, descriptor_pb2.FileDescriptorProto) and fdp.package == ""


def test_fetch_reply_wire_helpers_match_protobuf():
    """encode_fetch_reply_wire / split_fetch_reply_wire are byte-compatible with the protobuf (and therefore with the
    reference's generated FetchReply class)."""
    from dps_b200.rpc import proto
    payload = bytes(range(256)) * 300
    for step in (0, 1, 127, 128, 300, 2 ** 31 - 1, -1):
        wire = proto.encode_fetch_reply_wire(payload, step)
        msg = proto.FetchReply.FromString(wire)
        assert msg.parameters == payload and msg.global_step == step
        view, got = proto.split_fetch_reply_wire(wire)
        assert bytes(view) == payload and got == step
        # the generated serializer's output (field 2 omitted when zero) splits the same way
        view2, got2 = proto.split_fetch_reply_wire(proto.FetchReply(parameters=payload, global_step=step).SerializeToString())
        assert bytes(view2) == payload and got2 == step
    assert proto.split_fetch_reply_wire(b"") == (memoryview(b""), 0)


def test_push_request_wire_helpers_match_protobuf():
    from dps_b200.rpc import proto
    payload = bytes(range(251)) * 100
    for wid, step in ((0, 0), (3, 7), (31, 2 ** 31 - 1), (200, 128), (1, -1)):
        wire = proto.encode_push_request_wire(wid, payload, step)
        msg = proto.PushRequest.FromString(wire)
        assert (msg.worker_id, msg.gradients, msg.local_step) == (wid, payload, step)
        for w in (wire, proto.PushRequest(worker_id=wid, gradients=payload, local_step=step).SerializeToString()):
            got_wid, view, got_step = proto.split_push_request_wire(w)
            assert (got_wid, bytes(view), got_step) == (wid, payload, step)
