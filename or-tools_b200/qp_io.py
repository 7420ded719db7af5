"""Problem file I/O in front of the device solver: the host-side readers
and writers of /root/reference/ortools/pdlp/quadratic_program_io.{h,cc}.

  read_quadratic_program_or_die(filename, include_names=False)   quadratic_program_io.cc:53-70
  read_mps_linear_program(filename, include_names=False)         :356-395 (two-pass MPS reader)
  read_mpmodel_proto_file_or_die(filename, include_names=False)  :72-80  (.pb / .textproto / .json[.gz])
  write_linear_program_to_mps(qp, filename)                      :82-96
  write_quadratic_program_to_mpmodel_proto(qp, filename)         :98-104

The MPS grammar is the one of ortools/lp_data/mps_reader_template.h:715-1160 (sections NAME,
OBJSENSE, ROWS, LAZYCONS, COLUMNS, RHS, RANGES, BOUNDS, ENDATA; fixed format tried first, then
free format, :663-671; the first N row is the objective and minus its RHS the objective offset,
:1031-1058; RANGES as in :1061-1087; bound mnemonics LO UP FX FR MI PL BV LI UI, :1090-1160;
integrality is dropped as quadratic_program_io.cc:312-315 does).  Rows and columns are numbered in
order of first appearance, repeated (row, column) entries are summed
(`SetEigenMatrixFromTriplets`, quadratic_program.cc:315-357), maximisation problems come back
negated with `objective_scaling_factor = -1` (quadratic_program_io.cc:246-255).

"OrDie" functions abort the process in the reference; here they raise `RuntimeError`.  Pure host
code: parsing happens once, before the upload, and is outside the timed path (BASELINE.json keeps
problem I/O on the host).
"""
import bz2
import gzip

import numpy as np
from google.protobuf import json_format, text_format, unknown_fields

try:
    import scipy.sparse as _sp
except ImportError:  # pragma: no cover
    _sp = None

from . import mp_model

_INF = float("inf")
_SECTIONS = {"NAME", "OBJSENSE", "ROWS", "LAZYCONS", "COLUMNS", "RHS", "RANGES", "BOUNDS", "INDICATORS", "ENDATA"}
_FIELD_START = (1, 4, 14, 24, 39, 49)   # mps_reader_template.cc:26-29
_FIELD_LENGTH = (2, 8, 8, 12, 8, 12)
_SPACE_POS = (12, 13, 22, 23, 36, 37, 38, 47, 48, 61, 62, 63)
_MAX_FIXED_LINE = _FIELD_START[-1] + _FIELD_LENGTH[-1]
_BOUND_TYPES = {"LO", "UP", "FX", "FR", "MI", "PL", "BV", "LI", "UI", "SC"}


def combine_repeated_triplets_in_place(triplets):
    """internal::CombineRepeatedTripletsInPlace (quadratic_program.cc:346-364): sums CONSECUTIVE (row, column, value)
    entries with equal row and column, in place; entries that repeat further apart stay separate."""
    if not triplets:
        return
    out = 0
    for k in range(1, len(triplets)):
        row, col, value = triplets[k]
        if triplets[out][0] == row and triplets[out][1] == col:
            triplets[out] = (row, col, triplets[out][2] + value)
        else:
            out += 1
            triplets[out] = (row, col, value)
    del triplets[out + 1:]


def set_matrix_from_triplets(rows, cols, values, shape):
    """SetEigenMatrixFromTriplets (quadratic_program.cc:315-343): the compressed column-major matrix of the triplets
    sorted by (column, row) with repeated entries summed; an entry whose values sum to zero stays stored, as Eigen's
    insert() keeps it."""
    rows = np.asarray(rows, dtype=np.int64).reshape(-1)
    cols = np.asarray(cols, dtype=np.int64).reshape(-1)
    values = np.asarray(values, dtype=np.float64).reshape(-1)
    order = np.lexsort((rows, cols))  # stable: equal entries keep their order, so the sums associate as the reference's
    rows, cols, values = rows[order], cols[order], values[order]
    if len(rows):
        first = np.ones(len(rows), dtype=bool)
        first[1:] = (rows[1:] != rows[:-1]) | (cols[1:] != cols[:-1])
        starts = np.flatnonzero(first)
        group = np.cumsum(first) - 1
        summed = np.zeros(len(starts))
        np.add.at(summed, group, values)  # left to right within a group
        rows, cols, values = rows[starts], cols[starts], summed
    indptr = np.zeros(shape[1] + 1, dtype=np.int64)
    np.add.at(indptr, cols + 1, 1)
    return _sp.csc_matrix((values, rows, np.cumsum(indptr)), shape=shape)


class MpsParseError(ValueError):
    pass


def _open_text(filename):
    if filename.endswith(".gz"):
        return gzip.open(filename, "rt")
    if filename.endswith(".bz2"):
        return bz2.open(filename, "rt")
    return open(filename, "r")


def _read_bytes(filename):
    if filename.endswith(".gz"):
        with gzip.open(filename, "rb") as f:
            return f.read()
    with open(filename, "rb") as f:
        return f.read()


class _Parser:
    """One pass of MPSReaderTemplate::ProcessLine over the file in one format."""

    def __init__(self, free_form, include_names):
        self.free = free_form
        self.include_names = include_names
        self.section = None
        self.line_num = 0
        self.line = ""
        self.name = None
        self.maximize = False
        self.objective_name = ""
        self.objective_offset = 0.0
        self.rows = {}   # name -> index
        self.cols = {}
        self.clb, self.cub = [], []
        self.vlb, self.vub = [], []
        self.obj = {}
        self.binary_default = []
        self.trip_r, self.trip_c, self.trip_v = [], [], []
        self.in_integer = False

    def fail(self, message):
        raise MpsParseError(f'{message} Line {self.line_num}: "{self.line}".')

    def number(self, text):
        try:
            return float(text)
        except ValueError:
            self.fail(f"Failed to convert \"{text}\" to double.")

    def row(self, name):
        i = self.rows.get(name)
        if i is None:
            i = self.rows[name] = len(self.rows)
            self.clb.append(0.0)
            self.cub.append(0.0)
        return i

    def col(self, name):
        j = self.cols.get(name)
        if j is None:
            j = self.cols[name] = len(self.cols)
            self.vlb.append(0.0)
            self.vub.append(_INF)
            self.binary_default.append(False)
        return j

    def is_fixed_format(self, line):
        if len(line) > _MAX_FIXED_LINE:
            return False
        for i in _SPACE_POS:
            if i >= len(line):
                break
            if line[i] != " ":
                return False
        return True

    def split(self, line):
        if self.free:
            fields = line.split()
            if len(fields) > 6:
                self.fail("Found too many fields.")
            return fields
        fields = []
        for start, length in zip(_FIELD_START, _FIELD_LENGTH):
            if start >= len(line):
                break
            fields.append(line[start:start + length].rstrip())
        return fields

    def process(self, raw):
        self.line_num += 1
        line = raw.replace("\r", "").rstrip()
        self.line = line
        if not line or line[0] == "*":
            return
        fields = self.split(line)
        if line[0] != " ":  # a new section (mps_reader_template.h:725-767)
            word = line.split(" ")[0]
            if word not in _SECTIONS:
                self.fail("Unknown section.")
            self.section = word
            if not self.free and not self.is_fixed_format(line) and word != "NAME":
                self.fail("Line is not in fixed format.")
            if word == "NAME":
                free_fields = line.split()
                free_name = free_fields[1] if len(free_fields) >= 2 else ""
                if self.free:
                    if len(free_fields) >= 2:
                        self.name = free_name
                else:
                    fixed_name = fields[2] if len(fields) >= 3 else ""
                    if free_name != fixed_name:
                        self.fail("Fixed form invalid: name differs between free and fixed forms.")
                    self.name = fixed_name
            return
        if not self.free and not self.is_fixed_format(line):
            self.fail("Line is not in fixed format.")
        s = self.section
        if s == "NAME":
            self.fail("Second NAME field.")
        elif s == "OBJSENSE":
            word = line.strip()
            if word not in ("MIN", "MAX"):
                self.fail("Expected objective sense (MAX or MIN).")
            self.maximize = word == "MAX"
        elif s in ("ROWS", "LAZYCONS"):
            self.rows_line(fields)
        elif s == "COLUMNS":
            self.columns_line(fields, line)
        elif s == "RHS":
            self.pairs_line(fields, self.store_rhs)
        elif s == "RANGES":
            self.pairs_line(fields, self.store_range)
        elif s == "BOUNDS":
            self.bounds_line(fields)
        elif s == "INDICATORS":
            raise MpsParseError("unsupported features/sections found (INDICATORS)")
        elif s == "ENDATA":
            pass
        else:
            self.fail("Unknown section.")

    def rows_line(self, fields):
        if len(fields) < 2:
            self.fail("Not enough fields in ROWS section.")
        kind, name = fields[0], fields[1]
        if kind not in ("E", "L", "G", "N"):
            self.fail("Unknown row type.")
        if kind == "N" and not self.objective_name:  # the first N row is the objective
            self.objective_name = name
            return
        i = self.row(name)
        if kind == "L":
            self.clb[i] = -_INF
        elif kind == "G":
            self.cub[i] = _INF
        elif kind == "N":
            self.clb[i], self.cub[i] = -_INF, _INF

    def columns_line(self, fields, line):
        if "'MARKER'" in line:
            if "'INTORG'" in line:
                if self.in_integer:
                    self.fail("Found INTORG inside the integer section.")
                self.in_integer = True
            elif "'INTEND'" in line:
                if not self.in_integer:
                    self.fail("Found INTEND without corresponding INTORG.")
                self.in_integer = False
            return
        start = 0 if self.free else 1
        if len(fields) < start + 3:
            self.fail("Not enough fields in COLUMNS section.")
        j = self.col(fields[start])
        if self.in_integer:  # default bounds of integer variables are [0, 1]
            self.vlb[j], self.vub[j] = 0.0, 1.0
            self.binary_default[j] = True
        else:
            self.vlb[j], self.vub[j] = 0.0, _INF
        self.store_coefficient(j, fields[start + 1], fields[start + 2])
        if len(fields) == start + 4:
            self.fail("Unexpected number of fields.")
        if len(fields) - start > 4:
            self.store_coefficient(j, fields[start + 3], fields[start + 4])

    def store_coefficient(self, j, row_name, text):
        if not row_name or row_name == "$":
            return
        value = self.number(text)
        if value in (_INF, -_INF):
            self.fail("Constraint coefficients cannot be infinity.")
        if value == 0.0:
            return
        if row_name == self.objective_name:
            self.obj[j] = value
        else:
            self.trip_r.append(self.row(row_name))
            self.trip_c.append(j)
            self.trip_v.append(value)

    def pairs_line(self, fields, store):
        start = 0 if self.free else 2
        offset = start + ((len(fields) & 1) if self.free else 0)
        if len(fields) < offset + 2:
            self.fail("Not enough fields in RHS section.")
        store(fields[offset], fields[offset + 1])
        if len(fields) - start >= 4:
            store(fields[offset + 2], fields[offset + 3])

    def store_rhs(self, row_name, text):
        if not row_name:
            return
        value = self.number(text)
        if row_name == self.objective_name:
            self.objective_offset = -value
            return
        i = self.row(row_name)
        self.clb[i] = -_INF if self.clb[i] == -_INF else value
        self.cub[i] = _INF if self.cub[i] == _INF else value

    def store_range(self, row_name, text):
        if not row_name:
            return
        i = self.row(row_name)
        rng = self.number(text)
        lo, hi = self.clb[i], self.cub[i]
        if lo == hi:
            if rng < 0.0:
                lo += rng
            else:
                hi += rng
        if lo == -_INF:
            lo = hi - abs(rng)
        if hi == _INF:
            hi = lo + abs(rng)
        self.clb[i], self.cub[i] = lo, hi

    def bounds_line(self, fields):
        if len(fields) < 3:
            self.fail("Not enough fields in BOUNDS section.")
        kind, name = fields[0], fields[2]
        text = fields[3] if len(fields) >= 4 else ""
        if kind not in _BOUND_TYPES:
            self.fail("Unknown bound type.")
        if kind == "SC":
            raise MpsParseError("unsupported features/sections found (semi-continuous variables)")
        j = self.col(name)
        lo, hi = self.vlb[j], self.vub[j]
        if self.binary_default[j]:
            lo, hi = 0.0, _INF
        if kind in ("LO", "LI"):
            lo = self.number(text)
            if kind == "LI" and lo == 0.0:
                hi = _INF
        elif kind in ("UP", "UI"):
            hi = self.number(text)
        elif kind == "FX":
            lo = hi = self.number(text)
        elif kind == "FR":
            lo, hi = -_INF, _INF
        elif kind == "MI":
            lo = -_INF
        elif kind == "PL":
            hi = _INF
        elif kind == "BV":
            lo, hi = 0.0, 1.0
        self.binary_default[j] = False
        self.vlb[j], self.vub[j] = lo, hi

    def finish(self):
        from ._api import QuadraticProgram

        n, m = len(self.cols), len(self.rows)
        qp = QuadraticProgram()
        qp.resize_and_initialize(n, m)
        qp.constraint_lower_bounds = np.array(self.clb, dtype=np.float64).reshape(m)
        qp.constraint_upper_bounds = np.array(self.cub, dtype=np.float64).reshape(m)
        qp.variable_lower_bounds = np.array(self.vlb, dtype=np.float64).reshape(n)
        qp.variable_upper_bounds = np.array(self.vub, dtype=np.float64).reshape(n)
        c = np.zeros(n)
        for j, v in self.obj.items():
            c[j] = v
        qp.constraint_matrix = set_matrix_from_triplets(self.trip_r, self.trip_c, self.trip_v, (m, n))  # quadratic_program_io.cc:257
        qp.objective_offset = self.objective_offset
        if self.maximize:
            qp.objective_scaling_factor = -1.0
            qp.objective_offset *= -1
            c = -c
        qp.objective_vector = c
        if self.include_names:
            qp.problem_name = self.name
            qp.variable_names = [None] * n
            qp.constraint_names = [None] * m
            for name, j in self.cols.items():
                qp.variable_names[j] = name
            for name, i in self.rows.items():
                qp.constraint_names[i] = name
        return qp


def parse_mps(text, include_names=False):
    """`MPSReaderTemplate::ParseString` with kAutoDetect (fixed format first, then free) on the
    PDLP data wrapper.  Raises MpsParseError (a ValueError) with the reference's message."""
    lines = text.split("\n")
    last_error = None
    for free in (False, True):
        parser = _Parser(free, include_names)
        try:
            for line in lines:
                parser.process(line)
        except MpsParseError as e:
            last_error = e
            continue
        return parser.finish()
    raise last_error


def read_mps_linear_program(filename, include_names=False):
    """quadratic_program_io.cc:356-395."""
    try:
        with _open_text(filename) as f:
            text = f.read()
    except OSError as e:
        raise MpsParseError(f"Could not read or parse file `{filename}` as an MPS file: {e}") from e
    try:
        return parse_mps(text, include_names)
    except MpsParseError as e:
        raise MpsParseError(f"{e} Could not read or parse file `{filename}` as an MPS file") from e


def read_mps_linear_program_or_die(filename, include_names=False):
    try:
        return read_mps_linear_program(filename, include_names)
    except MpsParseError as e:
        raise RuntimeError(f"Error reading MPS Linear Program from {filename}: {e}") from e


def read_file_to_proto(filename, proto):
    """`ReadFileToProto` (ortools/util/file_util.cc): binary, text or JSON, optionally gzipped, chosen by
    content the way the reference does (binary first, then text, then JSON); the suffix is only a hint."""
    data = _read_bytes(filename)
    order = ("binary", "text", "json")
    if filename.endswith((".textproto", ".txt")):
        order = ("text", "binary", "json")
    elif filename.endswith((".json", ".json.gz")):
        order = ("json", "text", "binary")
    for kind in order:
        proto.Clear()
        try:
            if kind == "binary":
                proto.ParseFromString(data)
                # Any byte string of unknown fields "parses": require something recognised, like the
                # reference's check that the binary parse leaves no unknown fields behind.
                if data and not proto.ListFields():
                    continue
                if len(unknown_fields.UnknownFieldSet(proto)) > 0:
                    continue
            elif kind == "text":
                text_format.Parse(data.decode("utf-8"), proto)
            else:
                json_format.Parse(data.decode("utf-8"), proto)
            return True
        except Exception:  # noqa: BLE001 - try the next encoding
            continue
    proto.Clear()
    return False


def read_mpmodel_proto_file_or_die(filename, include_names=False):
    """quadratic_program_io.cc:72-80: integer variables are relaxed."""
    proto = mp_model.MPModelProto()
    if not read_file_to_proto(filename, proto):
        raise RuntimeError(f"Check failed: ReadFileToProto(mpmodel_proto_file, &lp_proto) {filename}")
    try:
        return mp_model.qp_from_mpmodel_proto(proto, relax_integer_variables=True, include_names=include_names)
    except ValueError as e:
        raise RuntimeError(str(e)) from e


def read_quadratic_program_or_die(filename, include_names=False):
    """quadratic_program_io.cc:53-70; mirrored in the pybind module (python/pdlp.cc:122-123)."""
    if filename.endswith((".mps", ".mps.gz", ".mps.bz2")):
        return read_mps_linear_program_or_die(filename, include_names)
    if filename.endswith((".pb", ".textproto", ".json", ".json.gz")):
        return read_mpmodel_proto_file_or_die(filename, include_names)
    raise RuntimeError(f"Invalid filename suffix in {filename}. Valid suffixes are .mps, .mps.gz, .pb, .textproto,.json, and .json.gz")


def _mps_number(v):
    return repr(float(v))


def export_model_as_mps(proto):
    """Free-format MPS text of a linear MPModelProto.  The reference goes through
    `ExportModelAsMpsFormat` (ortools/linear_solver/model_exporter.cc), whose column layout and name
    mangling are not restated: this writer emits the same sections with the same meaning (objective row
    `COST`, offset as minus its RHS, `OBJSENSE MAX` for maximisation, RANGES for two-sided rows, one
    BOUNDS line per non-default bound) and full-precision numbers, so reading the file back gives the
    same problem bit for bit."""
    n, m = len(proto.variable), len(proto.constraint)
    vnames = [v.name if v.name else f"V{j}" for j, v in enumerate(proto.variable)]
    cnames = [c.name if c.name else f"C{i}" for i, c in enumerate(proto.constraint)]
    if len(set(vnames)) != n or len(set(cnames)) != m or "COST" in cnames:
        vnames = [f"V{j}" for j in range(n)]
        cnames = [f"C{i}" for i in range(m)]
    out = [f"NAME {proto.name}" if proto.name else "NAME"]
    if proto.maximize:
        out += ["OBJSENSE", " MAX"]
    out.append("ROWS")
    out.append(" N COST")
    kinds = []
    for i, c in enumerate(proto.constraint):
        lo, hi = c.lower_bound, c.upper_bound
        if lo == hi:
            kind = "E"
        elif lo == -_INF and hi == _INF:
            kind = "N"
        elif lo == -_INF:
            kind = "L"
        elif hi == _INF:
            kind = "G"
        else:
            kind = "R"  # ranged: written as L with a RANGES entry
        kinds.append(kind)
        out.append(f" {'L' if kind == 'R' else kind} {cnames[i]}")
    columns = [[] for _ in range(n)]
    for i, c in enumerate(proto.constraint):
        for j, v in zip(c.var_index, c.coefficient):
            columns[j].append((cnames[i], v))
    out.append("COLUMNS")
    for j, v in enumerate(proto.variable):
        entries = ([("COST", v.objective_coefficient)] if v.objective_coefficient != 0.0 else []) + columns[j]
        if not entries:
            entries = [("COST", 0.0)]  # keeps the column (and its position) in the file
        for row, value in entries:
            out.append(f" {vnames[j]} {row} {_mps_number(value)}")
    out.append("RHS")
    if proto.objective_offset != 0.0:
        out.append(f" RHS COST {_mps_number(-proto.objective_offset)}")
    for i, c in enumerate(proto.constraint):
        rhs = {"E": c.lower_bound, "L": c.upper_bound, "R": c.upper_bound, "G": c.lower_bound, "N": 0.0}[kinds[i]]
        if rhs != 0.0:
            out.append(f" RHS {cnames[i]} {_mps_number(rhs)}")
    if "R" in kinds:
        out.append("RANGES")
        for i, c in enumerate(proto.constraint):
            if kinds[i] == "R":
                out.append(f" RNG {cnames[i]} {_mps_number(c.upper_bound - c.lower_bound)}")
    out.append("BOUNDS")
    for j, v in enumerate(proto.variable):
        lo, hi = v.lower_bound, v.upper_bound
        if lo == 0.0 and hi == _INF:
            continue
        if lo == -_INF and hi == _INF:
            out.append(f" FR BND {vnames[j]}")
        elif lo == hi:
            out.append(f" FX BND {vnames[j]} {_mps_number(lo)}")
        else:
            if lo == -_INF:
                out.append(f" MI BND {vnames[j]}")
            elif lo != 0.0:
                out.append(f" LO BND {vnames[j]} {_mps_number(lo)}")
            if hi != _INF:
                out.append(f" UP BND {vnames[j]} {_mps_number(hi)}")
    out.append("ENDATA")
    return "\n".join(out) + "\n"


def write_linear_program_to_mps(qp, filename):
    """quadratic_program_io.cc:82-96."""
    if qp.objective_matrix is not None:
        raise ValueError("'linear_program' has a quadratic objective")
    proto = mp_model.qp_to_mpmodel_proto(qp)
    text = export_model_as_mps(proto)
    opener = gzip.open if filename.endswith(".gz") else open
    with opener(filename, "wt") as f:
        f.write(text)


def write_quadratic_program_to_mpmodel_proto(qp, filename):
    """quadratic_program_io.cc:98-104: binary MPModelProto."""
    proto = mp_model.qp_to_mpmodel_proto(qp)
    with open(filename, "wb") as f:
        f.write(proto.SerializeToString())
