"""Two interchangeable drivers for the reference's own test cases.

`OracleBackend` runs the CPU restatement under oracle/ (checker; CPU tests pin it against
the reference's golden values).  `DeviceBackend` runs the product: pdlp_b200 through the C
ABI on cuda:0.  Both expose the same small surface so that one test body, re-typed from
the reference's *_test.cc files, checks both -- the oracle on the CPU box (`-m "not gpu"`)
and the CUDA path on the B200 (`-m gpu`).
"""
import numpy as np

import oracle_binding as ob
import pdlp_b200

P = pdlp_b200.PrimalDualHybridGradientParams
TR = pdlp_b200.TerminationReason


def _dimension_error(qp):
    nv, nc = len(qp.variable_lower_bounds), len(qp.constraint_lower_bounds)
    rows, cols = qp.constraint_matrix.shape
    ok = (nv == len(qp.variable_upper_bounds) == len(qp.objective_vector) == cols and nc == len(qp.constraint_upper_bounds) == rows
          and (qp.objective_matrix is None or len(qp.objective_matrix) == nv))
    return not ok


class _Result:
    pass


class OracleBackend:
    name = "oracle"

    def solve(self, qp, params, initial_solution=None, callback=None, interrupt=None):
        """callback(iteration_type:int, IterationStats proto); interrupt: np.int32[1]."""
        if ob.validate_params(params):
            return self._error(TR.TERMINATION_REASON_INVALID_PARAMETER, ob.validate_params(params))
        if _dimension_error(qp):
            return self._error(TR.TERMINATION_REASON_INVALID_PROBLEM, "Inconsistent dimensions")
        init = None
        if initial_solution is not None:
            x0, y0 = np.asarray(initial_solution[0], dtype=float), np.asarray(initial_solution[1], dtype=float)
            if x0.size != len(qp.objective_vector) or y0.size != len(qp.constraint_lower_bounds):
                return self._error(TR.TERMINATION_REASON_INVALID_INITIAL_SOLUTION, "size mismatch")
            init = (x0, y0)
        cb = None
        if callback is not None:
            cb = lambda t, d: callback(t, ob.to_proto(d, pdlp_b200.IterationStats()))
        r = ob.solve(qp, params, init, cb, interrupt)
        if qp.problem_name:
            r.solve_log.instance_name = qp.problem_name
        return r

    @staticmethod
    def _error(reason, text):
        r = _Result()
        r.primal_solution = r.dual_solution = r.reduced_costs = np.zeros(0)
        r.solve_log = pdlp_b200.SolveLog()
        r.solve_log.termination_reason = reason
        r.solve_log.termination_string = text
        return r

    def problem(self, qp):
        return _OracleProblem(qp)

    def validate_params(self, params):
        return ob.validate_params(params)


class _OracleProblem:
    def __init__(self, qp):
        self.h = ob.Problem(qp, num_threads=2, num_shards=2)
        self.n, self.m = self.h.n, self.h.m

    def matvec(self, which, v):
        return self.h.matvec(which, v)

    def col_norms(self, l2, row_scale, col_scale):
        return self.h.col_norms(l2, row_scale, col_scale)

    def stats(self):
        return ob.to_proto(self.h.stats(), pdlp_b200.QuadraticProgramStats())

    def rescale(self, ruiz, l2):
        rs, cs = self.h.rescale(ruiz, l2)
        return rs, cs, self.h.get()["val"]

    def kkt_stats(self, params, x, y):
        ones_n, ones_m = np.ones(self.n), np.ones(self.m)
        st = pdlp_b200.IterationStats()
        ob.to_proto(self.h.convergence_info(params, ones_n, ones_m, x, y), st.convergence_information.add())
        ob.to_proto(self.h.infeasibility_info(params, ones_n, ones_m, x, y), st.infeasibility_information.add())
        return st

    def reduced_costs(self, params, x, y, use_zero_primal_objective=False):
        return self.h.reduced_costs(params, x, y, use_zero_primal_objective)

    def localized_bounds(self, x, y, primal_weight, radius, use_diagonal=False, tolerance=1e-8, max_norm=False):
        return self.h.localized_bounds(x, y, 0 if max_norm else 1, primal_weight, radius, use_diagonal, tolerance)


class DeviceBackend:
    name = "device"

    def solve(self, qp, params, initial_solution=None, callback=None, interrupt=None):
        init = None
        if initial_solution is not None:
            init = pdlp_b200.PrimalAndDualSolution(np.asarray(initial_solution[0], dtype=float), np.asarray(initial_solution[1], dtype=float))
        cb = None
        if callback is not None:
            cb = lambda info: callback(info.iteration_type, info.iteration_stats)
        return pdlp_b200.primal_dual_hybrid_gradient(qp, params, initial_solution=init, interrupt_solve=interrupt, iteration_stats_callback=cb)

    def problem(self, qp):
        return _DeviceProblem(qp)

    def validate_params(self, params):
        return pdlp_b200.validate_params(params)


class _DeviceProblem:
    def __init__(self, qp):
        self.h = pdlp_b200.DeviceProblem(qp)
        self.n, self.m = self.h.n, self.h.m

    def matvec(self, which, v):
        return self.h.matvec(which, v)[0]

    def col_norms(self, l2, row_scale, col_scale):
        return self.h.scaled_norms(0, l2, row_scale, col_scale)

    def stats(self):
        return self.h.stats()

    def rescale(self, ruiz, l2):
        return self.h.rescale(ruiz, l2)

    def kkt_stats(self, params, x, y):
        return self.h.kkt_stats(params, x, y)

    def reduced_costs(self, params, x, y, use_zero_primal_objective=False):
        return self.h.reduced_costs(params, x, y, use_zero_primal_objective)

    def localized_bounds(self, x, y, primal_weight, radius, use_diagonal=False, tolerance=1e-8, max_norm=False):
        return self.h.localized_bounds(x, y, primal_weight, radius, use_diagonal, tolerance, max_norm=max_norm)


def make(name):
    return OracleBackend() if name == "oracle" else DeviceBackend()
