#!/usr/bin/env python
"""Static SASS comparison of the row-stream kernels with and without PDLP_ROWSTREAM_LEAN.

Compiles or-tools_b200/csrc/pdlp_step.cu twice for sm_100a (no GPU needed), disassembles the K x
(DualStepOp) and K^T y (KtyStepOp) instances and counts instructions: whole kernel, and the
staging stretch of the main path (from the wait on the block's indices to the barrier before the row
reduction; for the lean build the dense branch, which full work blocks take).  Writes a markdown table.
    python scripts/sass_rowstream_compare.py > profiles/r01_sass_rowstream_lean.md
"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "or-tools_b200", "csrc")


def disassemble(flag, tmp):
    obj = os.path.join(tmp, f"step{flag[-1]}.o")
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false", "-Xcompiler", "-fPIC",
           "-I" + os.path.join(ROOT, "include"), flag, "-Xptxas", "-v", "-c", os.path.join(CSRC, "pdlp_step.cu"), "-o", obj]
    log = subprocess.run(cmd, capture_output=True, text=True, check=True).stderr
    elf = subprocess.run(["cuobjdump", "-elf", obj], capture_output=True, text=True).stdout
    out = {}
    for op in ("DualStepOp", "KtyStepOp"):
        sym = sorted(set(re.findall(r"_ZN9pdlp_b20016rowstream_kernelINS_\d+_GLOBAL__N__[0-9a-z_]+\d+%sEEEvT_NS_9RowBlocksEPd" % op, elf)))[0]
        sass = subprocess.run(["cuobjdump", "-sass", "-fun", sym, obj], capture_output=True, text=True).stdout
        ins = [re.sub(r"\s*/\*.*$", "", m.group(2)).strip() for m in re.finditer(r"^\s*/\*([0-9a-f]{4})\*/\s+(.*)$", sass, flags=re.M)]
        regs = re.search(re.escape(sym) + r"'.*?Used (\d+) registers", log, flags=re.S)
        spills = re.search(re.escape(sym) + r"\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores", log)
        out[op] = dict(ins=ins, regs=int(regs.group(1)) if regs else None, spill=int(spills.group(2)) if spills else None)
    return out


def mnemonic(line):
    return re.sub(r"^@!?U?P\w+\s+", "", line).split(" ")[0].split(".")[0].rstrip(";")


def staging_stretch(ins, lean):
    """Instructions between the DEPBAR that opens the main path and the BAR.SYNC before the reduction."""
    depbars = [i for i, s in enumerate(ins) if s.startswith("DEPBAR")]
    bars = [i for i, s in enumerate(ins) if s.startswith("BAR.SYNC")]
    # main path: the DEPBAR followed by the longest barrier-free stretch
    best = max(((next(b for b in bars if b > d) - d, d) for d in depbars if any(b > d for b in bars)), key=lambda t: t[0])
    start = best[1]
    end = start + best[0]
    stretch = ins[start:end]
    if not lean:
        return stretch
    # the lean build has both branches back to back; the dense one runs up to the first unconditional BRA after the products
    for i, s in enumerate(stretch):
        if re.match(r"BRA 0x", s) and any(t.startswith("STS.64") for t in stretch[max(0, i - 12):i]):
            return stretch[:i + 1]
    return stretch


def main():
    with tempfile.TemporaryDirectory() as tmp:
        builds = {"default": disassemble("-DPDLP_ROWSTREAM_LEAN=0", tmp), "lean": disassemble("-DPDLP_ROWSTREAM_LEAN=1", tmp)}
    print("# Row-stream kernels: SASS of the default and the lean staging path (static, sm_100a)\n")
    print("`python scripts/sass_rowstream_compare.py` (nvcc 12.9 cross-compile, no GPU).  \"staging stretch\" = straight-line code from the wait on")
    print("the work block's indices to the barrier in front of the row reduction, executed once per work block (8 nonzeros per thread);")
    print("for the lean build the branch full blocks take, including the shared prologue (bank pad, threads per row, epilogue prefetch).")
    print("Not yet timed on the device (DESIGN.md 8).\n")
    print("| kernel | build | registers | spill bytes | instructions | staging stretch | per nonzero | address / predicate ALU in stretch | memory ops in stretch |")
    print("|---|---|---|---|---|---|---|---|---|")
    for op, label in (("DualStepOp", "K2 `rowstream<DualStepOp>`"), ("KtyStepOp", "K3 `rowstream<KtyStepOp>`")):
        for build in ("default", "lean"):
            b = builds[build][op]
            st = staging_stretch(b["ins"], build == "lean")
            hist = collections.Counter(mnemonic(s) for s in st)
            mem = sum(v for k, v in hist.items() if k in ("LDG", "LDS", "STS", "STG", "LDGSTS", "LDC", "LDCU"))
            alu = sum(v for k, v in hist.items() if k in ("IMAD", "IADD3", "VIADD", "LEA", "SHF", "LOP3", "ISETP", "MOV", "S2R", "SEL", "IABS", "P2R", "PLOP3", "CS2R"))
            print(f"| {label} | {build} | {b['regs']} | {b['spill']} | {len(b['ins'])} | {len(st)} | {len(st) / 8:.0f} | {alu} | {mem} |")
    return 0


if __name__ == "__main__":
    sys.exit(main())
