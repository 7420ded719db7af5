"""Generates tests/golden/linear_solver_schema.json from the reference's
ortools/linear_solver/linear_solver.proto: every message, field number, type, label, default,
packed flag and oneof membership, and every enum value (hex literals as integers).  Run in the
build container (where /root/reference exists); the JSON is committed and is what
tests/test_mp_model.py reads.

    python tests/golden/make_linear_solver_golden.py
"""
import json
import os
import re
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_schema_golden import Parser, resolve, tokenize  # noqa: E402

SRC = "/root/reference/ortools/linear_solver/linear_solver.proto"


def main():
    text = open(SRC).read()
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"0x[0-9a-fA-F]+", lambda m: str(int(m.group(0), 16)), text)
    text = text.replace("-inf", "-1e999")  # the shared tokenizer has no signed identifiers; float("-1e999") is -inf
    # map<K, V> fields (MPModelDeltaProto only) are recorded as repeated V.
    text = re.sub(r"map<\s*\w+\s*,\s*(\w+)\s*>", r"repeated \1", text)
    p = Parser(tokenize(text))
    p.parse_file()
    golden = dict(source="ortools/linear_solver/linear_solver.proto (OR-Tools 9.7)", enums=p.enums, messages=resolve(p))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "linear_solver_schema.json")
    with open(path, "w") as f:
        json.dump(golden, f, indent=1, sort_keys=True)
    print(path, len(golden["messages"]), "messages", len(golden["enums"]), "enums")


if __name__ == "__main__":
    sys.exit(main())
