"""Command-line interface to the B200 PDLP solver: `python -m pdlp_b200.solve --input FILE ...`.

Mirrors the reference's driver, /root/reference/examples/cpp/pdlp_solve.cc:40-135: the same four flags
with the same meaning, verbosity_level 2 unless --params overrides it (:86-91), integrality dropped by
the reader (:93-95), ^C interrupts the solve through `interrupt_solve` (:97-102), the SolveLog written as
.textproto / .pb / .json (:58-75), the primal solution in Miplib .sol format (`=obj= value` then one
`name value` line per variable, `var<i>` when the problem has no names; :112-130).  One flag beyond the
reference: --device (CUDA ordinal, default: the current device).
"""
import argparse
import signal
import sys

import numpy as np
from google.protobuf import json_format, text_format

import pdlp_b200


def write_solve_log(filename, log):
    """WriteSolveLog (pdlp_solve.cc:58-75)."""
    if filename.endswith(".textproto"):
        data = text_format.MessageToString(log).encode()
    elif filename.endswith(".pb"):
        data = log.SerializeToString()
    elif filename.endswith(".json"):
        data = json_format.MessageToJson(log, preserving_proto_field_name=True).encode()
    else:
        raise SystemExit(f"Unrecognized file extension for --solve_log_file: {filename}. Expected .textproto, .pb, or .json")
    with open(filename, "wb") as f:
        f.write(data)


def sol_string(qp, result):
    """The .sol text of pdlp_solve.cc:112-127, or None when the result carries no convergence information."""
    info = [c for c in result.solve_log.solution_stats.convergence_information if c.candidate_type == result.solve_log.solution_type]
    if not info:
        return None
    names = qp.variable_names
    lines = [f"=obj= {_number(info[0].primal_objective)}"]
    for i, v in enumerate(np.asarray(result.primal_solution)):
        lines.append(f"{names[i] if names is not None else 'var' + str(i)} {_number(v)}")
    return "\n".join(lines) + "\n"


def _number(v):
    """absl::StrCat(double): six significant digits (%g)."""
    return "%g" % float(v)


def parse_params(text):
    """The params of pdlp_solve.cc:86-91: verbosity_level 2, then --params merged over it (text format)."""
    params = pdlp_b200.PrimalDualHybridGradientParams()
    params.verbosity_level = 2
    try:
        text_format.Merge(text or "", params)
    except text_format.ParseError as e:
        raise SystemExit(f"Error parsing --params: {e}")
    return params


def solve(input_file, params_text="", solve_log_file="", sol_file="", device=-1):
    """Solve (pdlp_solve.cc:77-131).  Returns the SolverResult."""
    if not input_file:
        raise SystemExit("--input is required")
    params = parse_params(params_text)
    qp = pdlp_b200.read_quadratic_program_or_die(input_file, include_names=True)
    interrupted = np.zeros(1, dtype=np.int32)
    previous = signal.getsignal(signal.SIGINT)
    try:
        signal.signal(signal.SIGINT, lambda *_: interrupted.__setitem__(0, 1))
    except ValueError:  # not the main thread: no handler, as when the reference's Register fails
        previous = None
    try:
        result = pdlp_b200.primal_dual_hybrid_gradient(qp, params, interrupt_solve=interrupted, device=device)
    finally:
        if previous is not None:
            signal.signal(signal.SIGINT, previous)
    if solve_log_file:
        print(f"Writing SolveLog to '{solve_log_file}'.", file=sys.stderr)
        write_solve_log(solve_log_file, result.solve_log)
    if sol_file:
        text = sol_string(qp, result)
        if text is not None:
            print(f"Writing .sol solution to '{sol_file}'.", file=sys.stderr)
            with open(sol_file, "w") as f:
                f.write(text)
    return result


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m pdlp_b200.solve",
                                 description="Run PDLP on the given input file. The following formats are supported: a .mps, .mps.gz file, "
                                             "an MPModelProto [.pb (binary), .textproto (text), *.json, *.json.gz]")
    ap.add_argument("--input", default="", help="REQUIRED: Input file name.")
    ap.add_argument("--params", default="", help="PrimalDualHybridGradientParams in text format")
    ap.add_argument("--solve_log_file", default="", help="If non-empty, writes PDLP's SolveLog here. The extension must be .textproto (text), .pb (binary), or .json.")
    ap.add_argument("--sol_file", default="", help="If non-empty, output the final primal solution in Miplib .sol format.")
    ap.add_argument("--device", type=int, default=-1, help="CUDA device ordinal (default: the current device)")
    args = ap.parse_args(argv)
    solve(args.input, args.params, args.solve_log_file, args.sol_file, args.device)
    return 0


if __name__ == "__main__":
    sys.exit(main())
