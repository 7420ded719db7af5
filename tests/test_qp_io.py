"""Problem file readers / writers: or-tools_b200/qp_io.py against the
semantics of /root/reference/ortools/lp_data/mps_reader_template.h:715-1160 and
ortools/pdlp/quadratic_program_io.cc:53-395.  The reference ships no test file for this module in
this tree; the cases below restate the documented MPS rules (fixed / free format, objective RHS,
RANGES table, bound mnemonics, integer markers) on the reference's TestLp (test_util.cc:35-49).
"""
import gzip
import os

import numpy as np
import pytest
from google.protobuf import json_format, text_format

import fixtures
import pdlp_b200
from pdlp_b200 import mp_model, qp_io

INF = np.inf

# TestLp (test_util.cc:35-49) in fixed-format MPS: columns 2-3 type, 5-12 name, 15-22 row, 25-36 value,
# 40-47 row, 50-61 value.
FIXED = """\
NAME          TESTLP
ROWS
 N  COST
 E  C0
 L  C1
 G  C2
 L  C3
COLUMNS
    X0        COST               5.5   C0                 2.0
    X0        C1                 1.0   C2                 4.0
    X1        COST              -2.0   C0                 1.0
    X2        COST              -1.0   C0                 1.0
    X2        C1                 1.0   C3                 1.5
    X3        COST               1.0   C0                 2.0
    X3        C3                -1.0
RHS
    RHS       COST              14.0   C0                12.0
    RHS       C1                 7.0   C2                -4.0
    RHS       C3                 1.0
RANGES
    RNG       C3                 2.0
BOUNDS
 FR BND       X0
 LO BND       X1                -2.0
 MI BND       X2
 UP BND       X2                 6.0
 LO BND       X3                 2.5
 UP BND       X3                 3.5
ENDATA
"""

FREE = """\
* the same LP, free format, entries in a different order
NAME TESTLP_WITH_A_LONG_NAME
ROWS
 N COST
 E C0
 L C1
 G C2
 L C3
COLUMNS
 X0 COST 5.5 C0 2
 X0 C1 1 C2 4
 X1 COST -2 C0 1
 X2 COST -1 C0 1
 X2 C1 1 C3 1.5
 X3 COST 1 C0 2
 X3 C3 -1
RHS
 RHS COST 14 C0 12
 RHS C1 7 C2 -4
 RHS C3 1
RANGES
 RNG C3 2
BOUNDS
 FR BND X0
 LO BND X1 -2
 MI BND X2
 UP BND X2 6
 LO BND X3 2.5
 UP BND X3 3.5
ENDATA
"""


def assert_is_test_lp(qp, sign=1.0):
    ref = fixtures.make_test_lp()
    assert qp.objective_scaling_factor == sign
    np.testing.assert_array_equal(sign * qp.objective_vector, ref.objective_vector)
    assert sign * qp.objective_offset == ref.objective_offset
    np.testing.assert_array_equal(qp.constraint_matrix.toarray(), ref.constraint_matrix.toarray())
    np.testing.assert_array_equal(qp.constraint_lower_bounds, ref.constraint_lower_bounds)
    np.testing.assert_array_equal(qp.constraint_upper_bounds, ref.constraint_upper_bounds)
    np.testing.assert_array_equal(qp.variable_lower_bounds, ref.variable_lower_bounds)
    np.testing.assert_array_equal(qp.variable_upper_bounds, ref.variable_upper_bounds)
    assert qp.objective_matrix is None
    K = qp.constraint_matrix
    assert K.format == "csc" and K.has_sorted_indices


@pytest.mark.parametrize("text,name", [(FIXED, "TESTLP"), (FREE, "TESTLP_WITH_A_LONG_NAME")], ids=["fixed", "free"])
def test_parse_mps(text, name):
    qp = qp_io.parse_mps(text)
    assert_is_test_lp(qp)
    assert qp.problem_name is None and qp.variable_names is None
    named = qp_io.parse_mps(text, include_names=True)
    assert named.problem_name == name
    assert named.variable_names == ["X0", "X1", "X2", "X3"] and named.constraint_names == ["C0", "C1", "C2", "C3"]


def test_objsense_max_negates():
    text = FREE.replace("ROWS\n", "OBJSENSE\n MAX\nROWS\n", 1)
    qp = qp_io.parse_mps(text)
    # quadratic_program_io.cc:246-255: offset and objective vector are negated, scaling factor -1.
    ref = fixtures.make_test_lp()
    assert qp.objective_scaling_factor == -1.0
    np.testing.assert_array_equal(qp.objective_vector, -ref.objective_vector)
    assert qp.objective_offset == 14.0


def _tiny(rows, columns, rhs="", ranges="", bounds=""):
    text = "NAME T\nROWS\n N OBJ\n" + rows + "COLUMNS\n" + columns
    if rhs:
        text += "RHS\n" + rhs
    if ranges:
        text += "RANGES\n" + ranges
    if bounds:
        text += "BOUNDS\n" + bounds
    return text + "ENDATA\n"


def test_ranges_table():
    # mps_reader_template.h:1061-1087: E with R < 0 -> [rhs + R, rhs]; E with R >= 0 -> [rhs, rhs + R];
    # L -> [rhs - |R|, rhs]; G -> [rhs, rhs + |R|].
    qp = qp_io.parse_mps(_tiny(" E e_neg\n E e_pos\n L l\n G g\n", " x e_neg 1 e_pos 1\n x l 1 g 1\n",
                               rhs=" r e_neg 10 e_pos 10\n r l 10 g 10\n", ranges=" R e_neg -3 e_pos 3\n R l -4 g 5\n"))
    np.testing.assert_array_equal(qp.constraint_lower_bounds, [7, 10, 6, 10])
    np.testing.assert_array_equal(qp.constraint_upper_bounds, [10, 13, 10, 15])


def test_bound_mnemonics_and_integer_markers():
    columns = (" a c 1\n b c 1\n d c 1\n e c 1\n f c 1\n g c 1\n"
               " M1 'MARKER' 'INTORG'\n i1 c 1\n i2 c 1\n i3 c 1\n i4 c 1\n M2 'MARKER' 'INTEND'\n h c 1\n")
    bounds = (" UP B a 4\n LO B b -1\n FX B d 2.5\n FR B e\n MI B f\n BV B g\n"
              " UP B i2 7\n LI B i3 0\n UI B i4 9\n PL B h\n")
    qp = qp_io.parse_mps(_tiny(" L c\n", columns, bounds=bounds), include_names=True)
    lo = dict(zip(qp.variable_names, qp.variable_lower_bounds))
    hi = dict(zip(qp.variable_names, qp.variable_upper_bounds))
    assert (lo["a"], hi["a"]) == (0, 4)
    assert (lo["b"], hi["b"]) == (-1, INF)
    assert (lo["d"], hi["d"]) == (2.5, 2.5)
    assert (lo["e"], hi["e"]) == (-INF, INF)
    assert (lo["f"], hi["f"]) == (-INF, INF)
    assert (lo["g"], hi["g"]) == (0, 1)
    assert (lo["h"], hi["h"]) == (0, INF)
    assert (lo["i1"], hi["i1"]) == (0, 1)       # integer by marker: binary by default (:889-893)
    assert (lo["i2"], hi["i2"]) == (0, 7)       # any bound resets the binary default (:1109-1115)
    assert (lo["i3"], hi["i3"]) == (0, INF)     # LI 0: general integer without upper bound (:1120-1124)
    assert (lo["i4"], hi["i4"]) == (0, 9)


def test_second_n_row_is_a_free_constraint_and_duplicates_are_summed():
    qp = qp_io.parse_mps(_tiny(" N free\n E c\n", " x OBJ 1 free 2\n x c 3\n x c 4\n y c 1 $ 0\n y free 0\n"), include_names=True)
    assert qp.constraint_names == ["free", "c"]
    np.testing.assert_array_equal(qp.constraint_lower_bounds, [-INF, 0])
    np.testing.assert_array_equal(qp.constraint_upper_bounds, [INF, 0])
    np.testing.assert_array_equal(qp.constraint_matrix.toarray(), [[2, 0], [7, 1]])
    np.testing.assert_array_equal(qp.objective_vector, [1, 0])


@pytest.mark.parametrize("text,fragment", [
    ("NAME T\nROWZ\n", "Unknown section."),
    (_tiny(" Q c\n", " x c 1\n"), "Unknown row type."),
    (_tiny(" E c\n", " x c\n"), "Not enough fields in COLUMNS section."),
    (_tiny(" E c\n", " x c 1 c\n"), "Unexpected number of fields."),
    (_tiny(" E c\n", " x c abc\n"), "to double"),
    (_tiny(" E c\n", " x c inf\n"), "Constraint coefficients cannot be infinity."),
    (_tiny(" E c\n", " x c 1\n", bounds=" ZZ B x 1\n"), "Unknown bound type."),
    (_tiny(" E c\n", " M 'MARKER' 'INTEND'\n x c 1\n"), "Found INTEND without corresponding INTORG."),
    ("NAME T\nOBJSENSE\n UP\n", "Expected objective sense (MAX or MIN)."),
])
def test_parse_errors(text, fragment):
    with pytest.raises(ValueError, match=fragment.replace("(", "\\(").replace(")", "\\)")):
        qp_io.parse_mps(text)


def test_error_names_the_line():
    with pytest.raises(ValueError, match='Line 5: " x c abc"'):
        qp_io.parse_mps("NAME T\nROWS\n E c\nCOLUMNS\n x c abc\nENDATA\n")


def test_files_round_trip(tmp_path):
    ref = fixtures.make_test_lp()
    ref.problem_name = "testlp"
    # MPS (plain and gzipped) written by write_linear_program_to_mps
    for name in ("a.mps", "a.mps.gz"):
        path = str(tmp_path / name)
        qp_io.write_linear_program_to_mps(ref, path)
        assert_is_test_lp(qp_io.read_quadratic_program_or_die(path))
        assert qp_io.read_quadratic_program_or_die(path, include_names=True).problem_name == "testlp"
    # hand-written fixed-format file, gzip and bz2
    import bz2
    with gzip.open(str(tmp_path / "f.mps.gz"), "wt") as f:
        f.write(FIXED)
    with bz2.open(str(tmp_path / "f.mps.bz2"), "wt") as f:
        f.write(FIXED)
    assert_is_test_lp(qp_io.read_quadratic_program_or_die(str(tmp_path / "f.mps.gz")))
    assert_is_test_lp(qp_io.read_quadratic_program_or_die(str(tmp_path / "f.mps.bz2")))
    # MPModelProto files: binary, text, JSON, gzipped JSON
    proto = mp_model.qp_to_mpmodel_proto(ref)
    pb = str(tmp_path / "m.pb")
    qp_io.write_quadratic_program_to_mpmodel_proto(ref, pb)
    with open(str(tmp_path / "m.textproto"), "w") as f:
        f.write(text_format.MessageToString(proto))
    with open(str(tmp_path / "m.json"), "w") as f:
        f.write(json_format.MessageToJson(proto))
    with gzip.open(str(tmp_path / "m.json.gz"), "wt") as f:
        f.write(json_format.MessageToJson(proto))
    for name in ("m.pb", "m.textproto", "m.json", "m.json.gz"):
        assert_is_test_lp(pdlp_b200.read_quadratic_program_or_die(str(tmp_path / name)))


def test_maximisation_round_trip_through_mps(tmp_path):
    ref = fixtures.make_test_lp()
    ref.objective_scaling_factor = -1.0
    path = str(tmp_path / "max.mps")
    qp_io.write_linear_program_to_mps(ref, path)
    assert "OBJSENSE" in open(path).read()
    back = qp_io.read_quadratic_program_or_die(path)
    assert back.objective_scaling_factor == -1.0
    np.testing.assert_array_equal(back.objective_vector, ref.objective_vector)
    assert back.objective_offset == ref.objective_offset


def test_integer_variables_in_proto_files_are_relaxed(tmp_path):
    proto = mp_model.qp_to_mpmodel_proto(fixtures.make_test_lp())
    proto.variable[1].is_integer = True
    path = str(tmp_path / "int.pb")
    with open(path, "wb") as f:
        f.write(proto.SerializeToString())
    assert_is_test_lp(qp_io.read_quadratic_program_or_die(path))


def test_or_die_errors(tmp_path):
    with pytest.raises(RuntimeError, match="Invalid filename suffix"):
        qp_io.read_quadratic_program_or_die(str(tmp_path / "x.lp"))
    with pytest.raises(RuntimeError, match="Error reading MPS Linear Program"):
        qp_io.read_quadratic_program_or_die(str(tmp_path / "missing.mps"))
    bad = str(tmp_path / "bad.mps")
    with open(bad, "w") as f:
        f.write("NAME T\nROWZ\n")
    with pytest.raises(RuntimeError, match="Unknown section"):
        qp_io.read_quadratic_program_or_die(bad)
    garbage = str(tmp_path / "g.textproto")
    with open(garbage, "w") as f:
        f.write("this is not a proto {")
    with pytest.raises(RuntimeError, match="ReadFileToProto"):
        qp_io.read_quadratic_program_or_die(garbage)
    qp = fixtures.diagonal_qp1()
    with pytest.raises(ValueError, match="has a quadratic objective"):
        qp_io.write_linear_program_to_mps(qp, str(tmp_path / "q.mps"))


def test_read_then_solve(backend, tmp_path):
    path = str(tmp_path / "t.mps")
    with open(path, "w") as f:
        f.write(FIXED)
    qp = pdlp_b200.read_quadratic_program_or_die(path)
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1e-8
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1e-8
    r = backend.solve(qp, p)
    assert r.solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    np.testing.assert_allclose(r.primal_solution, [-1, 8, 1, 2.5], atol=1e-6)


# ---- quadratic_program_test.cc:426-487: the triplet helpers behind the MPS reader's matrix ----

@pytest.mark.parametrize("triplets,expected", [
    ([], []),
    ([(1, 2, 3.0)], [(1, 2, 3.0)]),
    ([(1, 2, 3.0), (2, 1, 1.0), (1, 1, 0.0)], [(1, 2, 3.0), (2, 1, 1.0), (1, 1, 0.0)]),
    ([(1, 2, 3.0), (1, 2, -1.0), (1, 1, 0.0)], [(1, 2, 2.0), (1, 1, 0.0)]),
    ([(1, 2, 3.0), (2, 1, 1.0), (2, 1, 1.0)], [(1, 2, 3.0), (2, 1, 2.0)]),
    ([(1, 2, 3.0), (1, 2, 1.0), (1, 2, 2.0)], [(1, 2, 6.0)]),
])
def test_combine_repeated_triplets_in_place(triplets, expected):
    pdlp_b200.combine_repeated_triplets_in_place(triplets)
    assert triplets == expected


def test_set_matrix_from_triplets_handles_an_empty_matrix():
    K = pdlp_b200.set_matrix_from_triplets([], [], [], (2, 2))
    np.testing.assert_array_equal(K.toarray(), [[0, 0], [0, 0]])


def test_set_matrix_from_triplets_is_correct_for_a_tiny_matrix():
    triplets = [(0, 0, 1.0), (1, 0, -1.0), (0, 0, 0.0), (1, 1, 1.0), (0, 0, 1.0)]
    K = pdlp_b200.set_matrix_from_triplets([t[0] for t in triplets], [t[1] for t in triplets], [t[2] for t in triplets], (2, 2))
    np.testing.assert_array_equal(K.toarray(), [[2, 0], [-1, 1]])
    assert K.has_sorted_indices and K.nnz == 3 and K.indptr.tolist() == [0, 2, 3]


# ---- The reference's own MPS files (data re-typed; the expected values follow mps_reader_template.h:200-260 by hand) ----

MAXIMIZATION_MPS = """\
* Generated by MPModelProtoExporter
*   Name             : SupportedMaximizationProblem
*   Format           : Free
*   Constraints      : 0
*   Variables        : 1
*     Binary         : 0
*     Integer        : 0
*     Continuous     : 1
NAME          SupportedMaximizationProblem
OBJSENSE
  MAX
ROWS
 N  COST
COLUMNS
    X_ONE   COST         1
BOUNDS
 UP BOUND   X_ONE        4
ENDATA
"""  # ortools/linear_solver/testdata/maximization.mps (model_builder_test.py:167-173 reads it and checks the name)

EXAMPLE_MPS = """\
NAME          EXAMPLE
ROWS
 N  OBJ
 G  ROW01
 L  ROW02
 E  ROW03
 G  ROW04
 L  ROW05
COLUMNS
    COL01     OBJ                1.0
    COL01     ROW01              3.0   ROW05              5.6
    COL02     ROW01              1.0   ROW02              2.0
    COL03     ROW02              1.1   ROW03              1.0
    COL04     ROW01             -2.0   ROW04              2.8
    COL05     OBJ                2.0
    COL05     ROW01             -1.0   ROW05              1.0
    COL06     ROW03              1.0
    COL07     ROW04             -1.2
    COL08     OBJ               -1.0
    COL08     ROW01             -1.0   ROW05              1.9
RHS
    RHS1      ROW01              2.5
    RHS1      ROW02              2.1
    RHS1      ROW03              4.0
    RHS1      ROW04              1.8
    RHS1      ROW05             15.0
RANGES
    RNG1      ROW04              3.2
    RNG1      ROW05             12.0
BOUNDS
 LO BND1      COL01              2.5
 UP BND1      COL02              4.1
 LO BND1      COL05              0.5
 UP BND1      COL05              4.0
 UP BND1      COL08              4.3
ENDATA
"""  # examples/cpp/test2.mps: fixed format, RANGES on a G and an L row

MIN_SIZE_MPS = """\
NAME          MIN_SIZE_MAX_FEATURES

ROWS
 N  COST
 E  ROW1
 L  ROW2
 G  ROW3
 E  ROW4
 E  ROW5
 G  ROW6
 E  ROW7
 L  ROW8
 N  UNCONST.

COLUMNS
    X1        ROW1              1      UNCONST.            1
    X2        ROW1              1      UNCONST.            2
    X1        COST              1
    X2        COST              2
    X1        ROW2              3
    X2        ROW2              4
    X1        ROW3              1
    X2        ROW3              2
    X3        ROW3              3      ROW7                1
    X4        ROW3              4      ROW7                1
    X5        ROW3              5      ROW4                2
    X6        ROW3              6      ROW4                13
    X7        ROW5              1      ROW6                7
    X8        ROW5              3      ROW6                2
    X7        ROW8              1
    X8        ROW8              1


RHS
              ROW1              2
              ROW2              6
              ROW3              -85
              ROW4              2
              ROW5              8
              ROW6              0
              ROW7              50
              ROW8              50

BOUNDS
 FR           X1
 UP           X2                125.
 FX           X3                12
 MI           X4
 LO           X5                5.
 PL           X6

RANGES
    RANGE1    ROW2              20.
              ROW3              25.
              ROW4              12.
              ROW7              -100.
ENDATA

"""  # examples/cpp/test.mps: fixed-format lines with empty name fields, a second N row, negative range on an E row


def test_reference_maximization_mps():
    qp = qp_io.parse_mps(MAXIMIZATION_MPS, include_names=True)
    assert qp.problem_name == "SupportedMaximizationProblem"
    assert qp.objective_scaling_factor == -1.0 and list(qp.objective_vector) == [-1.0]  # max x  ==  -(min -x)
    assert qp.variable_names == ["X_ONE"] and list(qp.variable_lower_bounds) == [0.0] and list(qp.variable_upper_bounds) == [4.0]
    assert qp.constraint_matrix.shape == (0, 1)


def test_reference_example_mps():
    qp = qp_io.parse_mps(EXAMPLE_MPS, include_names=True)
    inf = np.inf
    assert qp.problem_name == "EXAMPLE" and qp.objective_scaling_factor == 1.0 and qp.objective_offset == 0.0
    assert qp.variable_names == ["COL0%d" % j for j in range(1, 9)] and qp.constraint_names == ["ROW0%d" % i for i in range(1, 6)]
    np.testing.assert_array_equal(qp.objective_vector, [1, 0, 0, 0, 2, 0, 0, -1])
    np.testing.assert_array_equal(qp.variable_lower_bounds, [2.5, 0, 0, 0, 0.5, 0, 0, 0])
    np.testing.assert_array_equal(qp.variable_upper_bounds, [inf, 4.1, inf, inf, 4.0, inf, inf, 4.3])
    # G row with range R: [rhs, rhs + |R|]; L row: [rhs - |R|, rhs]  (mps_reader_template.h:1061-1087)
    np.testing.assert_array_equal(qp.constraint_lower_bounds, [2.5, -inf, 4.0, 1.8, 3.0])
    np.testing.assert_array_equal(qp.constraint_upper_bounds, [inf, 2.1, 4.0, 5.0, 15.0])
    np.testing.assert_array_equal(qp.constraint_matrix.toarray(), [
        [3.0, 1.0, 0, -2.0, -1.0, 0, 0, -1.0],
        [0, 2.0, 1.1, 0, 0, 0, 0, 0],
        [0, 0, 1.0, 0, 0, 1.0, 0, 0],
        [0, 0, 0, 2.8, 0, 0, -1.2, 0],
        [5.6, 0, 0, 0, 1.0, 0, 0, 1.9]])


def test_reference_min_size_mps_is_rejected_in_both_formats():
    # Fixed format: the name is longer than its 8-column field and "-100." reaches column 37; free format: " FR X1" has
    # two fields.  ParseString tries fixed, then free, and reports the second error (mps_reader_template.h:697-701).
    with pytest.raises(qp_io.MpsParseError) as e:
        qp_io.parse_mps(MIN_SIZE_MPS)
    assert str(e.value) == 'Not enough fields in BOUNDS section. Line 45: " FR           X1".'


def test_reference_min_size_mps_made_fixed_format():
    """The same file with an 8-column name and the range written as -100: every line is fixed format, so the empty set
    names (RHS, RANGES, BOUNDS) are empty fields and not missing ones."""
    inf = np.inf
    qp = qp_io.parse_mps(MIN_SIZE_MPS.replace("MIN_SIZE_MAX_FEATURES", "MINSIZE").replace("-100.", "-100"), include_names=True)
    assert qp.problem_name == "MINSIZE"
    assert qp.constraint_names == ["ROW%d" % i for i in range(1, 9)] + ["UNCONST."]  # the second N row is a free constraint
    np.testing.assert_array_equal(qp.objective_vector, [1, 2, 0, 0, 0, 0, 0, 0])
    np.testing.assert_array_equal(qp.variable_lower_bounds, [-inf, 0, 12, -inf, 5, 0, 0, 0])  # FR, UP, FX, MI, LO, PL
    np.testing.assert_array_equal(qp.variable_upper_bounds, [inf, 125, 12, inf, inf, inf, inf, inf])
    # ranges: L 20 -> [6 - 20, 6]; G 25 -> [-85, -60]; E +12 -> [2, 14]; E -100 -> [-50, 50]
    np.testing.assert_array_equal(qp.constraint_lower_bounds, [2, -14, -85, 2, 8, 0, -50, -inf, -inf])
    np.testing.assert_array_equal(qp.constraint_upper_bounds, [2, 6, -60, 14, 8, inf, 50, 50, inf])
    np.testing.assert_array_equal(qp.constraint_matrix.toarray(), [
        [1, 1, 0, 0, 0, 0, 0, 0],
        [3, 4, 0, 0, 0, 0, 0, 0],
        [1, 2, 3, 4, 5, 6, 0, 0],
        [0, 0, 0, 0, 2, 13, 0, 0],
        [0, 0, 0, 0, 0, 0, 1, 3],
        [0, 0, 0, 0, 0, 0, 7, 2],
        [0, 0, 1, 1, 0, 0, 0, 0],
        [0, 0, 0, 0, 0, 0, 1, 1],
        [1, 2, 0, 0, 0, 0, 0, 0]])


def test_reference_example_mps_solves_to_the_highs_optimum(backend):
    """External validity on a file the reference ships: PDLP's objective on examples/cpp/test2.mps against HiGHS."""
    from scipy.optimize import linprog

    qp = qp_io.parse_mps(EXAMPLE_MPS)
    K = qp.constraint_matrix.toarray()
    lo, hi = np.asarray(qp.constraint_lower_bounds), np.asarray(qp.constraint_upper_bounds)
    A = np.vstack([K[np.isfinite(hi)], -K[np.isfinite(lo)]])
    b = np.concatenate([hi[np.isfinite(hi)], -lo[np.isfinite(lo)]])
    bounds = [(l if np.isfinite(l) else None, u if np.isfinite(u) else None) for l, u in zip(qp.variable_lower_bounds, qp.variable_upper_bounds)]
    highs = linprog(np.asarray(qp.objective_vector), A_ub=A, b_ub=b, bounds=bounds, method="highs")
    assert highs.status == 0
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1e-8
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1e-8
    p.termination_criteria.iteration_limit = 50000
    out = backend.solve(qp, p)
    assert out.solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    objective = float(np.dot(qp.objective_vector, out.primal_solution)) + qp.objective_offset
    assert abs(objective - highs.fun) <= 1e-6 * max(1.0, abs(highs.fun)), (objective, highs.fun)


SOLOMON = "/root/reference/ortools/routing/testdata/solomon_bp_c101.mps"


@pytest.mark.skipif(not os.path.exists(SOLOMON), reason="the reference tree is only present in the build container")
def test_a_real_model_of_the_reference_tree_read_and_solved_on_the_oracle():
    """External validity beyond synthetic inputs: the LP relaxation of the 151 x 2550 set-partitioning
    model under ortools/routing/testdata (free-format MPS with integrality markers, which the reader drops), PDLP
    objective against HiGHS.  CPU only: the file is not copied into this repository."""
    import backends
    import scipy.sparse as sp
    from scipy.optimize import linprog

    qp = qp_io.read_quadratic_program_or_die(SOLOMON)
    K = qp.constraint_matrix.tocsr()
    assert K.shape == (151, 2550) and K.nnz == 7600
    lo, hi = np.asarray(qp.constraint_lower_bounds), np.asarray(qp.constraint_upper_bounds)
    eq = lo == hi
    ub, lb = np.isfinite(hi) & ~eq, np.isfinite(lo) & ~eq
    highs = linprog(np.asarray(qp.objective_vector), A_ub=sp.vstack([K[ub], -K[lb]]), b_ub=np.concatenate([hi[ub], -lo[lb]]), A_eq=K[eq], b_eq=lo[eq],
                    bounds=list(zip(qp.variable_lower_bounds, qp.variable_upper_bounds)), method="highs")
    assert highs.status == 0
    p = pdlp_b200.PrimalDualHybridGradientParams()
    p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1e-6
    p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1e-6
    p.termination_criteria.iteration_limit = 200000
    out = backends.make("oracle").solve(qp, p)
    assert out.solve_log.termination_reason == pdlp_b200.TerminationReason.TERMINATION_REASON_OPTIMAL
    objective = float(np.dot(qp.objective_vector, out.primal_solution)) + qp.objective_offset
    assert abs(objective - highs.fun) <= 1e-5 * max(1.0, abs(highs.fun)), (objective, highs.fun)
