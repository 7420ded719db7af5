"""Protobuf message classes for the two PDLP schemas, built at import time from the
schema description exported by the native library (`pdlp_b200_describe_schema`), so the
C++ structs in include/pdlp_b200/pdlp_schema.h stay the single source of truth.

The classes are wire- and text-format compatible with
/root/reference/ortools/pdlp/solvers.proto:20-481 and solve_log.proto:24-450 (package
`operations_research.pdlp`); tests/test_schema.py checks that against descriptors
compiled from the reference .proto files (fixture under tests/golden/).
"""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
from google.protobuf.internal import enum_type_wrapper

_PACKAGE = "operations_research.pdlp"
_FILES = {"solvers": "ortools/pdlp/solvers.proto", "solve_log": "ortools/pdlp/solve_log.proto"}
_GLOP_FILE = "ortools/glop/parameters.proto"

_SCALAR = {
    "double": descriptor_pb2.FieldDescriptorProto.TYPE_DOUBLE,
    "int32": descriptor_pb2.FieldDescriptorProto.TYPE_INT32,
    "int64": descriptor_pb2.FieldDescriptorProto.TYPE_INT64,
    "bool": descriptor_pb2.FieldDescriptorProto.TYPE_BOOL,
    "string": descriptor_pb2.FieldDescriptorProto.TYPE_STRING,
}


def _container(file_proto, nested_index, dotted, kind):
    """Returns the proto container that should hold `dotted` (a possibly nested name)."""
    parts = dotted.split(".")
    parent = nested_index.get(".".join(parts[:-1])) if len(parts) > 1 else None
    if kind == "enum":
        return (parent.enum_type if parent is not None else file_proto.enum_type), parts[-1]
    return (parent.nested_type if parent is not None else file_proto.message_type), parts[-1]


def build(describe_text):
    """Parses the description and returns {message_name: class}, {enum_name: {value_name: number}}, the file
    protos and {enum_name: EnumTypeWrapper} (what a generated _pb2 module exposes: `.Name()`, `.Value()`, `.keys()`)."""
    files = {}
    for key, name in _FILES.items():
        fp = descriptor_pb2.FileDescriptorProto()
        fp.name = name
        fp.package = _PACKAGE
        fp.syntax = "proto2"
        files[key] = fp
    files["solvers"].dependency.append(_GLOP_FILE)
    files["solve_log"].dependency.append(_FILES["solvers"])
    glop = descriptor_pb2.FileDescriptorProto()
    glop.name = _GLOP_FILE
    glop.package = "operations_research.glop"
    glop.syntax = "proto2"
    glop.message_type.add().name = "GlopParameters"  # carried opaquely (unknown fields are preserved)

    blocks = []
    cur = None
    for line in describe_text.splitlines():
        tok = line.split(" ")
        if tok[0] in ("enum", "message"):
            cur = {"kind": tok[0], "name": tok[1], "file": tok[2], "items": []}
        elif tok[0] == "end":
            blocks.append(cur)
            cur = None
        elif tok[0] in ("value", "field"):
            cur["items"].append(tok[1:])

    # Messages that own nested types must exist before their children are placed, and the
    # description lists children first; create all message shells up front, parents first.
    msg_names = sorted((b["name"] for b in blocks if b["kind"] == "message"), key=lambda n: n.count("."))
    file_of = {b["name"]: b["file"] for b in blocks}
    for b in blocks:
        if b["kind"] == "enum" and "." in b["name"]:
            file_of.setdefault(b["name"].rsplit(".", 1)[0], b["file"])
    index = {}
    for name in msg_names:
        cont, short = _container(files[file_of[name]], index, name, "message")
        m = cont.add()
        m.name = short
        index[name] = m
    enums = {}
    for b in blocks:
        if b["kind"] != "enum":
            continue
        cont, short = _container(files[b["file"]], index, b["name"], "enum")
        e = cont.add()
        e.name = short
        enums[b["name"]] = {}
        for vname, vnum in b["items"]:
            v = e.value.add()
            v.name = vname
            v.number = int(vnum)
            enums[b["name"]][vname] = int(vnum)

    F = descriptor_pb2.FieldDescriptorProto
    for b in blocks:
        if b["kind"] != "message":
            continue
        m = index[b["name"]]
        oneofs = {}
        for number, fname, label, ftype, type_name, default, packed, oneof in b["items"]:
            f = m.field.add()
            f.name = fname
            f.number = int(number)
            f.label = F.LABEL_REPEATED if label == "repeated" else F.LABEL_OPTIONAL
            if ftype in _SCALAR:
                f.type = _SCALAR[ftype]
            elif ftype == "enum":
                f.type = F.TYPE_ENUM
                f.type_name = "." + _PACKAGE + "." + type_name
            else:
                f.type = F.TYPE_MESSAGE
                f.type_name = type_name if type_name.startswith(".") else "." + _PACKAGE + "." + type_name
            if default != "-":
                f.default_value = default
            if packed == "1":
                f.options.packed = True
            if oneof != "-":
                if oneof not in oneofs:
                    oneofs[oneof] = len(m.oneof_decl)
                    m.oneof_decl.add().name = oneof
                f.oneof_index = oneofs[oneof]

    pool = descriptor_pool.DescriptorPool()
    pool.Add(glop)
    pool.Add(files["solvers"])
    pool.Add(files["solve_log"])
    classes = {}
    for name in msg_names:
        classes[name] = message_factory.GetMessageClass(pool.FindMessageTypeByName(_PACKAGE + "." + name))
    wrappers = {name: enum_type_wrapper.EnumTypeWrapper(pool.FindEnumTypeByName(_PACKAGE + "." + name)) for name in enums}
    return classes, enums, files, wrappers
