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
