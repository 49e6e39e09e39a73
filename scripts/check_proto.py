#!/usr/bin/env python
"""Contract check for the wire schema (role of the reference's regenerate_proto.sh:12-16).

The reference regenerates ps_pb2.py / ps_pb2_grpc.py from src/communication/ps.proto with grpc_tools.protoc and
pins protobuf/grpcio versions at import.  This framework has no generated code: dps_b200/rpc/proto.py builds the
descriptors programmatically, so there is nothing to regenerate -- but ps.proto stays the human-readable source of
truth.  This script parses dps_b200/rpc/ps.proto (a small recursive scan, no protoc needed) and verifies that every
message, field name, field number, field type and RPC (including the misspelt PushGradrients) agrees with the
programmatic descriptors.  Exit code 0 = in sync.

    python scripts/check_proto.py
"""
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dps_b200.rpc import proto as P                                             # noqa: E402

TYPE_NAMES = {1: "double", 2: "float", 3: "int64", 5: "int32", 8: "bool", 9: "string", 12: "bytes"}


def parse_proto(text: str):
    text = re.sub(r"//[^\n]*", "", text)
    messages, rpcs = {}, {}
    for m in re.finditer(r"message\s+(\w+)\s*\{([^}]*)\}", text):
        fields = {}
        for f in re.finditer(r"(\w+)\s+(\w+)\s*=\s*(\d+)\s*;", m.group(2)):
            fields[f.group(2)] = (f.group(1), int(f.group(3)))
        messages[m.group(1)] = fields
    svc = re.search(r"service\s+(\w+)\s*\{(.*)\}", text, re.S)
    service = svc.group(1) if svc else None
    if svc:
        for r in re.finditer(r"rpc\s+(\w+)\s*\(\s*(\w+)\s*\)\s*returns\s*\(\s*(\w+)\s*\)", svc.group(2)):
            rpcs[r.group(1)] = (r.group(2), r.group(3))
    return service, messages, rpcs


def main() -> int:
    path = os.path.join(os.path.dirname(os.path.abspath(P.__file__)), "ps.proto")
    service, messages, rpcs = parse_proto(open(path).read())
    errors = []
    fd = P._FILE
    for name, fields in messages.items():
        md = fd.message_types_by_name.get(name)
        if md is None:
            errors.append(f"message {name} missing from descriptors")
            continue
        have = {f.name: (TYPE_NAMES.get(f.type, str(f.type)), f.number) for f in md.fields}
        if have != fields:
            errors.append(f"message {name}: proto {fields} != descriptors {have}")
    for name in fd.message_types_by_name:
        if name not in messages:
            errors.append(f"message {name} missing from ps.proto")
    sd = fd.services_by_name.get(service)
    if sd is None:
        errors.append(f"service {service} missing from descriptors")
    else:
        have = {m.name: (m.input_type.name, m.output_type.name) for m in sd.methods}
        if have != rpcs:
            errors.append(f"rpcs: proto {rpcs} != descriptors {have}")
    if "PushGradrients" not in rpcs:
        errors.append("the wire contract spells the push RPC 'PushGradrients'")
    for e in errors:
        print("MISMATCH:", e)
    print(f"{path}: {len(messages)} messages, {len(rpcs)} rpcs, {'OK' if not errors else 'OUT OF SYNC'}")
    return 1 if errors else 0


if __name__ == "__main__":
    raise SystemExit(main())
