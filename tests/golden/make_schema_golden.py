"""Generates tests/golden/pdlp_schema.json from the reference's own .proto files
(ortools/pdlp/solvers.proto, solve_log.proto): every message, field number, type, label,
default, packed flag and oneof membership, and every enum value.  Run in the build container
(where /root/reference exists); the JSON is committed and is what the tests read.

    python tests/golden/make_schema_golden.py
"""
import json
import os
import re
import sys

REF = "/root/reference/ortools/pdlp"
SCALARS = {"double", "float", "int32", "int64", "uint32", "uint64", "sint32", "sint64", "bool", "string", "bytes", "fixed32", "fixed64"}


def tokenize(text):
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.findall(r'"[^"]*"|[A-Za-z_][\w.]*|-?[\d.]+(?:[eE][-+]?\d+)?|[{}\[\]=;,]', text)


class Parser:
    def __init__(self, tokens):
        self.t, self.i = tokens, 0
        self.enums, self.messages = {}, {}

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else None

    def next(self):
        tok = self.t[self.i]
        self.i += 1
        return tok

    def expect(self, tok):
        got = self.next()
        assert got == tok, (got, tok, self.t[max(0, self.i - 8): self.i + 3])

    def parse_file(self):
        while self.peek() is not None:
            tok = self.next()
            if tok in ("syntax", "package", "import", "option"):
                while self.next() != ";":
                    pass
            elif tok == "enum":
                self.parse_enum("")
            elif tok == "message":
                self.parse_message("")
            else:
                raise AssertionError(tok)

    def parse_enum(self, scope):
        name = scope + self.next()
        self.expect("{")
        values = {}
        while self.peek() != "}":
            key = self.next()
            if key in ("option", "reserved"):
                while self.next() != ";":
                    pass
                continue
            self.expect("=")
            values[key] = int(self.next())
            if self.peek() == "[":
                while self.next() != "]":
                    pass
            self.expect(";")
        self.expect("}")
        self.enums[name] = values

    def parse_options(self):
        opts = {}
        if self.peek() == "[":
            self.next()
            while True:
                key = self.next()
                self.expect("=")
                opts[key] = self.next()
                if self.next() == "]":
                    break
        return opts

    def parse_field(self, label, scope_name, oneof):
        typ = self.next()
        name = self.next()
        self.expect("=")
        number = int(self.next())
        opts = self.parse_options()
        self.expect(";")
        return dict(number=number, name=name, label=label, type=typ, default=opts.get("default"), packed=opts.get("packed") == "true", oneof=oneof)

    def parse_message(self, scope):
        name = scope + self.next()
        self.expect("{")
        fields = []
        while self.peek() != "}":
            tok = self.next()
            if tok == "message":
                self.parse_message(name + ".")
            elif tok == "enum":
                self.parse_enum(name + ".")
            elif tok in ("reserved", "option", "extensions"):
                while self.next() != ";":
                    pass
            elif tok == "oneof":
                oneof = self.next()
                self.expect("{")
                while self.peek() != "}":
                    self.i += 0
                    fields.append(self.parse_field("optional", name, oneof))
                self.expect("}")
            elif tok in ("optional", "repeated", "required"):
                fields.append(self.parse_field(tok, name, None))
            else:
                raise AssertionError((tok, name))
        self.expect("}")
        self.messages[name] = fields


def resolve(parser):
    """Type names -> kind + fully qualified (package-relative) name, defaults -> canonical."""
    out = {}
    for mname, fields in parser.messages.items():
        rows = []
        for f in fields:
            typ = f["type"]
            kind, type_name = typ, None
            if typ not in SCALARS:
                candidates = []
                scope = mname
                while True:
                    candidates.append((scope + "." if scope else "") + typ)
                    if not scope:
                        break
                    scope = scope.rpartition(".")[0]
                found = next((c for c in candidates if c in parser.messages or c in parser.enums), None)
                if found is None:  # another package (glop parameters)
                    kind, type_name = "message", "." + typ.lstrip(".")
                else:
                    kind, type_name = ("message" if found in parser.messages else "enum"), found
            default = f["default"]
            if default is not None and kind in ("double", "float"):
                default = repr(float(default))
            elif default is not None and kind.startswith(("int", "uint", "sint")):
                default = str(int(default))
            rows.append(dict(number=f["number"], name=f["name"], label=f["label"], kind=kind, type_name=type_name, default=default, packed=f["packed"],
                             oneof=f["oneof"]))
        out[mname] = sorted(rows, key=lambda r: r["number"])
    return out


def main():
    parser = Parser([])
    for fname in ("solvers.proto", "solve_log.proto"):
        p = Parser(tokenize(open(os.path.join(REF, fname)).read()))
        p.parse_file()
        parser.enums.update(p.enums)
        parser.messages.update(p.messages)
    golden = dict(source="ortools/pdlp/solvers.proto + solve_log.proto (OR-Tools 9.7)", enums=parser.enums, messages=resolve(parser))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pdlp_schema.json")
    with open(path, "w") as f:
        json.dump(golden, f, indent=1, sort_keys=True)
    print(path, len(golden["messages"]), "messages", len(golden["enums"]), "enums")


if __name__ == "__main__":
    sys.exit(main())
