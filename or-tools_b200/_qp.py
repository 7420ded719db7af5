"""`QuadraticProgram`, the problem container of the reference's Python module
(/root/reference/ortools/pdlp/python/pdlp.cc:48-103, quadratic_program.h:53-143).  Plain numpy / scipy
data with no dependency on the native library, so host-side tools (problem generators, file readers,
bench.py's CPU reference arm) can build problems without loading it."""
import math

import numpy as np

try:
    import scipy.sparse as _sp
except ImportError:  # pragma: no cover
    _sp = None

_INF = math.inf


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))


class QuadraticProgram:
    """quadratic_program.h:53-143.  `constraint_matrix` is a scipy.sparse matrix (any
    format; converted to CSC) of shape (num_constraints, num_variables)."""

    def __init__(self):
        self.objective_vector = np.zeros(0)
        self.constraint_matrix = _sp.csc_matrix((0, 0)) if _sp is not None else None
        self.constraint_lower_bounds = np.zeros(0)
        self.constraint_upper_bounds = np.zeros(0)
        self.variable_lower_bounds = np.zeros(0)
        self.variable_upper_bounds = np.zeros(0)
        self.problem_name = None
        self.variable_names = None
        self.constraint_names = None
        self.objective_offset = 0.0
        self.objective_scaling_factor = 1.0
        self._objective_matrix_diagonal = None

    @property
    def objective_matrix(self):
        return self._objective_matrix_diagonal

    def set_objective_matrix_diagonal(self, diagonal):
        self._objective_matrix_diagonal = _f64(diagonal)

    def clear_objective_matrix(self):
        self._objective_matrix_diagonal = None

    def resize_and_initialize(self, num_variables, num_constraints):
        self.objective_vector = np.zeros(num_variables)
        self.variable_lower_bounds = np.full(num_variables, -_INF)
        self.variable_upper_bounds = np.full(num_variables, _INF)
        self.constraint_lower_bounds = np.full(num_constraints, -_INF)
        self.constraint_upper_bounds = np.full(num_constraints, _INF)
        self.constraint_matrix = _sp.csc_matrix((num_constraints, num_variables))

    def apply_objective_scaling_and_offset(self, objective):
        return self.objective_scaling_factor * (objective + self.objective_offset)
