"""ctypes binding of include/pdlp_b200.h with the reference's Python names
(/root/reference/ortools/pdlp/python/pdlp.cc:48-150)."""
import ctypes
import enum
import math

import numpy as np

try:
    import scipy.sparse as _sp
except ImportError:  # pragma: no cover
    _sp = None

from . import (
    CALLBACK_TYPE, IterationStats, PrimalDualHybridGradientParams, SolveLog, _CCounters, _CProblem, _check, _classes,
    _c_double_p, _c_int64_p, _lib, _take_bytes,
)

_INF = math.inf


def _f64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))


def _ptr(a, typ=_c_double_p):
    return a.ctypes.data_as(typ) if a is not None and a.size > 0 else ctypes.cast(None, typ)


def _vptr(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None and a.size > 0 else ctypes.c_void_p(None)


from ._qp import QuadraticProgram  # noqa: E402,F401  (defined without the native library)


def is_linear_program(qp):
    return qp.objective_matrix is None


def validate_quadratic_program_dimensions(qp):
    """Raises ValueError with the reference's message (quadratic_program.cc:37-80)."""
    msg = _dimension_error(qp)
    if msg:
        raise ValueError(msg)


def _dimension_error(qp):
    nv = len(qp.variable_lower_bounds)
    nc = len(qp.constraint_lower_bounds)
    rows, cols = qp.constraint_matrix.shape
    if nv != len(qp.variable_upper_bounds):
        return f"Inconsistent dimensions: variable lower bound vector has size {nv} while variable upper bound vector has size {len(qp.variable_upper_bounds)}"
    if nv != len(qp.objective_vector):
        return f"Inconsistent dimensions: variable lower bound vector has size {nv} while objective vector has size {len(qp.objective_vector)}"
    if nv != cols:
        return f"Inconsistent dimensions: variable lower bound vector has size {nv} while constraint matrix has {cols} columns"
    if qp.objective_matrix is not None and nv != len(qp.objective_matrix):
        return f"Inconsistent dimensions: variable lower bound vector has size {nv} while objective matrix has {len(qp.objective_matrix)} rows"
    if nc != len(qp.constraint_upper_bounds):
        return f"Inconsistent dimensions: constraint lower bound vector has size {nc} while constraint upper bound vector has size {len(qp.constraint_upper_bounds)}"
    if nc != rows:
        return f"Inconsistent dimensions: constraint lower bound vector has size {nc} while constraint matrix has {rows} rows "
    return ""


class PrimalAndDualSolution:
    def __init__(self, primal_solution=None, dual_solution=None):
        self.primal_solution = np.zeros(0) if primal_solution is None else primal_solution
        self.dual_solution = np.zeros(0) if dual_solution is None else dual_solution


class SolverResult:
    def __init__(self):
        self.primal_solution = np.zeros(0)
        self.dual_solution = np.zeros(0)
        self.reduced_costs = np.zeros(0)
        self.solve_log = SolveLog()


class IterationCallbackInfo:
    """primal_dual_hybrid_gradient.h:62-93 (iteration_type is the IterationType number)."""

    def __init__(self, iteration_type, iteration_stats, bound_norms):
        self.iteration_type = iteration_type
        self.iteration_stats = iteration_stats
        self.bound_norms = bound_norms


class _Marshalled:
    """Keeps the numpy arrays behind a _CProblem alive."""

    def __init__(self, qp):
        K = qp.constraint_matrix
        if _sp is not None and not _sp.isspmatrix_csc(K):
            K = _sp.csc_matrix(K)
        if not K.has_canonical_format:
            K = K.copy()
            K.sum_duplicates()
        m, n = K.shape
        # scipy keeps 32-bit offsets / indices below 2^31 nonzeros: hand them over as they are.
        narrow = K.indptr.dtype == np.int32 and K.indices.dtype == np.int32
        itype = np.int32 if narrow else np.int64
        self.col_ptr = np.ascontiguousarray(K.indptr, dtype=itype)
        self.row_idx = np.ascontiguousarray(K.indices, dtype=itype)
        self.values = np.ascontiguousarray(K.data, dtype=np.float64)
        self.c = _f64(qp.objective_vector)
        self.q = None if qp.objective_matrix is None else _f64(qp.objective_matrix)
        self.vlb = _f64(qp.variable_lower_bounds)
        self.vub = _f64(qp.variable_upper_bounds)
        self.clb = _f64(qp.constraint_lower_bounds)
        self.cub = _f64(qp.constraint_upper_bounds)
        self.name = qp.problem_name.encode() if qp.problem_name else None
        self.n, self.m = n, m
        i32 = ctypes.POINTER(ctypes.c_int32)
        self.c_problem = _CProblem(
            n, m, self.values.size,
            ctypes.cast(None, _c_int64_p) if narrow else self.col_ptr.ctypes.data_as(_c_int64_p),
            ctypes.cast(None, _c_int64_p) if narrow else _ptr(self.row_idx, _c_int64_p), _ptr(self.values),
            _ptr(self.c), _ptr(self.q), _ptr(self.vlb), _ptr(self.vub), _ptr(self.clb), _ptr(self.cub),
            float(qp.objective_offset), float(qp.objective_scaling_factor), self.name,
            self.col_ptr.ctypes.data_as(i32) if narrow else ctypes.cast(None, i32),
            (self.row_idx.ctypes.data_as(i32) if self.row_idx.size else ctypes.cast(None, i32)) if narrow else ctypes.cast(None, i32),
        )


def _params_bytes(params):
    if params is None:
        return b""
    if isinstance(params, (bytes, bytearray)):
        return bytes(params)
    return params.SerializeToString()


def validate_params(params):
    """"" when valid, else the reference's status text (solvers_proto_validation.cc:171-298)."""
    b = _params_bytes(params)
    msg = ctypes.c_void_p()
    _check(_lib.pdlp_b200_validate_params(b, len(b), ctypes.byref(msg)))
    if not msg.value:
        return ""
    text = ctypes.string_at(msg).decode()
    _lib.pdlp_b200_free(msg)
    return text


def check_termination(params, stats, bound_norms, which="iterate", force_numerical_termination=False, interrupt=None):
    """termination.cc:161-219 through the C ABI.  Returns (reason, point_type, relative_residuals) with
    reason = point_type = 0 when the reference would return nullopt.  `which`: "iterate" or "simple"."""
    pb = _params_bytes(params)
    sb = stats.SerializeToString() if stats is not None else b""
    norms = _f64(bound_norms)
    reason, ptype = ctypes.c_int(0), ctypes.c_int(0)
    rel = np.zeros(5)
    flag = ctypes.c_void_p(interrupt.ctypes.data) if interrupt is not None else ctypes.c_void_p(None)
    _check(_lib.pdlp_b200_check_termination(pb, len(pb), sb, len(sb), _vptr(norms), 0 if which == "iterate" else 1, int(force_numerical_termination), flag,
                                            ctypes.byref(reason), ctypes.byref(ptype), _vptr(rel)))
    return reason.value, ptype.value, rel


def work_blocks(offsets, cut=-1):
    """The work-block partition the row-stream kernels run on (host only, pdlp_b200_work_blocks):
    returns (block_start, cut_block, max_row_nnz) for `rows + 1` compressed offsets."""
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    rows = off.size - 1
    if rows < 0:
        raise ValueError("offsets must have rows + 1 entries")
    starts = np.zeros(rows + 1, dtype=np.int32)
    nb, cb, mx = ctypes.c_int64(0), ctypes.c_int32(-1), ctypes.c_int32(0)
    _check(_lib.pdlp_b200_work_blocks(ctypes.c_void_p(off.ctypes.data), ctypes.c_int64(rows), ctypes.c_int64(int(cut)), ctypes.c_void_p(starts.ctypes.data),
                                      ctypes.byref(nb), ctypes.byref(cb), ctypes.byref(mx)))
    return starts[:nb.value + 1].copy(), cb.value, mx.value


_last_counters = {}


def last_solve_counters():
    """Kernel launches, trial steps and loop timings of the last solve (bench.py)."""
    return dict(_last_counters)


def primal_dual_hybrid_gradient(qp, params, initial_solution=None, interrupt_solve=None, iteration_stats_callback=None, device=-1, shard=None, _presolve=None):
    """primal_dual_hybrid_gradient.h:140-156.  `shard` (sharding.ColumnShard): `qp` is this rank's column block.  `interrupt_solve`, if given, is a
    numpy int32 array of one element (non-zero = stop); `iteration_stats_callback`
    receives an IterationCallbackInfo."""
    dim_error = _dimension_error(qp)
    result = SolverResult()
    if dim_error:
        perr = validate_params(params)
        if perr:
            result.solve_log.termination_reason = 10  # INVALID_PARAMETER, checked first (pdhg.cc:2792)
            result.solve_log.termination_string = perr
        else:
            result.solve_log.termination_reason = 9  # INVALID_PROBLEM
            result.solve_log.termination_string = "INVALID_ARGUMENT: " + dim_error
        return result
    mar = _Marshalled(qp)
    n, m = mar.n, mar.m
    pbytes = _params_bytes(params)
    x0 = y0 = None
    if initial_solution is not None:
        x0 = _f64(initial_solution.primal_solution)
        y0 = _f64(initial_solution.dual_solution)
        if x0.size != n or y0.size != m:
            perr = validate_params(params)
            if perr:
                result.solve_log.termination_reason = 10
                result.solve_log.termination_string = perr
                return result
            result.solve_log.termination_reason = 13  # INVALID_INITIAL_SOLUTION (pdhg.cc:899-947)
            if x0.size != n:
                result.solve_log.termination_string = f"Initial primal solution has size {x0.size} which differs from problem primal size {n}"
            else:
                result.solve_log.termination_string = f"Initial dual solution has size {y0.size} which differs from problem dual size {m}"
            return result
        if x0.size == 0:
            x0 = np.zeros(1)
        if y0.size == 0:
            y0 = np.zeros(1)
    # Filled by the library straight from the device (no zero-fill, no copies on the way out).
    primal = np.empty(max(n, 1))
    dual = np.empty(max(m, 1))
    reduced = np.empty(max(n, 1))
    has_solution = ctypes.c_int(0)
    log_ptr = ctypes.c_void_p()
    log_len = ctypes.c_size_t(0)
    counters = _CCounters()
    cb = None
    cb_error = []
    if iteration_stats_callback is not None:
        def _trampoline(itype, data, length, norms, _user):
            try:
                stats = IterationStats()
                stats.ParseFromString(ctypes.string_at(data, length))
                iteration_stats_callback(IterationCallbackInfo(itype, stats, [norms[i] for i in range(4)]))
            except BaseException as e:  # noqa: BLE001 - must not unwind through C
                cb_error.append(e)
        cb = CALLBACK_TYPE(_trampoline)
    interrupt_ptr = ctypes.c_void_p(None)
    if interrupt_solve is not None:
        assert interrupt_solve.dtype == np.int32 and interrupt_solve.size >= 1
        interrupt_ptr = ctypes.c_void_p(interrupt_solve.ctypes.data)
    tail = (pbytes, len(pbytes), _vptr(x0), _vptr(y0), interrupt_ptr, ctypes.cast(cb, ctypes.c_void_p) if cb is not None else None, None, int(device),
            _vptr(primal), _vptr(dual), _vptr(reduced), ctypes.byref(has_solution), ctypes.byref(log_ptr), ctypes.byref(log_len), ctypes.byref(counters))
    if _presolve is not None:
        presolved_mar, presolve_status, post = _presolve
        status = _lib.pdlp_b200_solve_presolved(ctypes.byref(mar.c_problem), ctypes.byref(presolved_mar.c_problem) if presolved_mar is not None else None,
                                                int(presolve_status), ctypes.cast(post, ctypes.c_void_p), None, *tail)
    else:
        status = _lib.pdlp_b200_solve_sharded(ctypes.byref(mar.c_problem), ctypes.byref(shard.c_shard) if shard is not None else None, *tail)
    _check(status)
    if cb_error:
        raise cb_error[0]
    result.solve_log.ParseFromString(_take_bytes(log_ptr, log_len.value))
    if has_solution.value:
        result.primal_solution = primal[:n]
        result.dual_solution = dual[:m]
        result.reduced_costs = reduced[:n]
    _last_counters.clear()
    _last_counters.update(
        kernel_launches=counters.kernel_launches, trial_steps=counters.trial_steps, setup_seconds=counters.setup_seconds,
        loop_seconds=counters.loop_seconds, device_loop_seconds=counters.device_loop_seconds, collectives=counters.collectives, upload_seconds=counters.upload_seconds, symmetric_buffers=counters.symmetric_buffers, peer_exchanges=counters.peer_exchanges, interrupts_seen_on_device=counters.interrupts_seen_on_device,
    )
    return result


class VariableStatus(enum.IntEnum):
    """Glop's VariableStatus with Glop's numbering (ortools/lp_data/lp_types.h:201-219)."""
    BASIC = 0
    FIXED_VALUE = 1
    AT_LOWER_BOUND = 2
    AT_UPPER_BOUND = 3
    FREE = 4


class ConstraintStatus(enum.IntEnum):
    """Glop's ConstraintStatus (lp_types.h:235-241)."""
    BASIC = 0
    FIXED_VALUE = 1
    AT_LOWER_BOUND = 2
    AT_UPPER_BOUND = 3
    FREE = 4


class GlopProblemStatus(enum.IntEnum):
    """Glop's ProblemStatus with Glop's numbering (ortools/lp_data/lp_types.h:105-160)."""
    OPTIMAL = 0
    PRIMAL_INFEASIBLE = 1
    DUAL_INFEASIBLE = 2
    INFEASIBLE_OR_UNBOUNDED = 3
    PRIMAL_UNBOUNDED = 4
    DUAL_UNBOUNDED = 5
    INIT = 6
    PRIMAL_FEASIBLE = 7
    DUAL_FEASIBLE = 8
    ABNORMAL = 9
    INVALID_PROBLEM = 10
    IMPRECISE = 11


def glop_status_to_termination_reason(glop_status):
    """GlopStatusToTerminationReason (primal_dual_hybrid_gradient.cc:1123-1143): a preprocessor's verdict as the
    TerminationReason number `primal_dual_hybrid_gradient_presolved(presolve_status=...)` takes."""
    from . import TerminationReason  # (defined after this module is imported)
    s = GlopProblemStatus(glop_status)
    if s == GlopProblemStatus.OPTIMAL:
        return TerminationReason.TERMINATION_REASON_OPTIMAL
    if s == GlopProblemStatus.INVALID_PROBLEM:
        return TerminationReason.TERMINATION_REASON_INVALID_PROBLEM
    if s in (GlopProblemStatus.ABNORMAL, GlopProblemStatus.IMPRECISE):
        return TerminationReason.TERMINATION_REASON_NUMERICAL_ERROR
    if s in (GlopProblemStatus.PRIMAL_INFEASIBLE, GlopProblemStatus.DUAL_INFEASIBLE, GlopProblemStatus.INFEASIBLE_OR_UNBOUNDED,
             GlopProblemStatus.DUAL_UNBOUNDED, GlopProblemStatus.PRIMAL_UNBOUNDED):
        return TerminationReason.TERMINATION_REASON_PRIMAL_OR_DUAL_INFEASIBLE
    return TerminationReason.TERMINATION_REASON_OTHER


def compute_statuses(qp, primal_solution, dual_solution):
    """internal::ComputeStatuses (primal_dual_hybrid_gradient.h:164, .cc:2835-2883): the Glop basis statuses of a PDLP
    point, (constraint_statuses, variable_statuses) as lists of ConstraintStatus / VariableStatus.  Host code (no GPU):
    what a postsolve callback of primal_dual_hybrid_gradient_presolved needs before it calls a Glop-style
    `RecoverSolution`."""
    mar = _Marshalled(qp)
    x, y = _f64(primal_solution), _f64(dual_solution)
    if x.size != mar.n or y.size != mar.m:
        raise ValueError(f"solution sizes {x.size}, {y.size} do not match the problem ({mar.n} variables, {mar.m} constraints)")
    vs, cs = np.zeros(max(mar.n, 1), dtype=np.int8), np.zeros(max(mar.m, 1), dtype=np.int8)
    _check(_lib.pdlp_b200_compute_statuses(ctypes.byref(mar.c_problem), x.ctypes.data if x.size else None, y.ctypes.data if y.size else None,
                                           vs.ctypes.data, cs.ctypes.data))
    return [ConstraintStatus(int(v)) for v in cs[: mar.m]], [VariableStatus(int(v)) for v in vs[: mar.n]]


POSTSOLVE_TYPE = ctypes.CFUNCTYPE(None, _c_double_p, _c_double_p, _c_double_p, _c_double_p, ctypes.c_void_p)


def primal_dual_hybrid_gradient_presolved(original_qp, presolved_qp, postsolve, params, presolve_status=0, initial_solution=None, interrupt_solve=None,
                                          iteration_stats_callback=None, device=-1):
    """The presolve hand-off (include/pdlp_b200.h: pdlp_b200_solve_presolved; the reference's Glop branch,
    primal_dual_hybrid_gradient.cc:1145-1202, :986-1034, :1288-1341, with the caller's presolver).

    PDHG iterates on `presolved_qp`; statistics, termination tests and the returned vectors are those of
    `original_qp` at the point `postsolve(x_presolved, y_presolved) -> (x_original, y_original)` recovers (numpy
    arrays, unscaled).  presolved_qp None: the presolver solved the problem by itself and `presolve_status` is its
    verdict (a TerminationReason number); postsolve is then called with two empty arrays."""
    for name, q in (("original", original_qp), ("presolved", presolved_qp)):
        if q is not None:
            err = _dimension_error(q)
            if err:
                raise ValueError(f"{name} problem: {err}")
    n_o, m_o = len(original_qp.objective_vector), len(original_qp.constraint_lower_bounds)
    presolved_mar = _Marshalled(presolved_qp) if presolved_qp is not None else None
    n_p = presolved_mar.n if presolved_mar is not None else 0
    m_p = presolved_mar.m if presolved_mar is not None else 0
    errors = []

    def _post(px, py, ox, oy, _user):
        try:
            x = np.ctypeslib.as_array(px, shape=(n_p,)).copy() if n_p else np.zeros(0)
            y = np.ctypeslib.as_array(py, shape=(m_p,)).copy() if m_p else np.zeros(0)
            rx, ry = postsolve(x, y)
            rx, ry = _f64(rx), _f64(ry)
            if rx.size != n_o or ry.size != m_o:
                raise ValueError(f"postsolve returned vectors of sizes {rx.size}, {ry.size}; the original problem has {n_o} variables and {m_o} constraints")
            if n_o:
                np.ctypeslib.as_array(ox, shape=(n_o,))[:] = rx
            if m_o:
                np.ctypeslib.as_array(oy, shape=(m_o,))[:] = ry
        except BaseException as e:  # noqa: BLE001 - must not unwind through C; the outputs stay as they are
            errors.append(e)

    result = primal_dual_hybrid_gradient(original_qp, params, initial_solution, interrupt_solve, iteration_stats_callback, device,
                                         _presolve=(presolved_mar, presolve_status, POSTSOLVE_TYPE(_post)))
    if errors:
        raise errors[0]
    return result


class DeviceProblem:
    """Kernel-level handle (pdlp_b200_problem_create and friends): a problem resident on
    the device in both matrix orientations, for unit parity tests and profiling."""

    def __init__(self, qp, device=-1, shard=None):
        self._mar = _Marshalled(qp)
        self.n, self.m = self._mar.n, self._mar.m
        self.nnz = self._mar.values.size
        self._h = ctypes.c_void_p()
        self._shard = shard
        _check(_lib.pdlp_b200_problem_create_sharded(ctypes.byref(self._mar.c_problem), ctypes.byref(shard.c_shard) if shard is not None else None,
                                                     int(device), ctypes.byref(self._h)))

    def close(self):
        if self._h:
            _lib.pdlp_b200_problem_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def matvec(self, which, v, repeat=1):
        """which = 0: K^T v (v of size m); 1: K v (v of size n).  Returns (out, seconds_per_launch)."""
        v = _f64(v)
        out = np.zeros(max(self.n if which == 0 else self.m, 1))
        sec = ctypes.c_double(0)
        _check(_lib.pdlp_b200_matvec(self._h, int(which), _vptr(v if v.size else np.zeros(1)), _vptr(out), int(repeat), ctypes.byref(sec)))
        return out[: self.n if which == 0 else self.m], sec.value

    def scaled_norms(self, which, l2, row_scale, col_scale):
        rs, cs = _f64(row_scale), _f64(col_scale)
        size = self.n if which == 0 else self.m
        out = np.zeros(max(size, 1))
        _check(_lib.pdlp_b200_scaled_norms(self._h, int(which), int(l2), _vptr(rs if rs.size else np.zeros(1)), _vptr(cs if cs.size else np.zeros(1)), _vptr(out)))
        return out[:size]

    def stats(self):
        p = ctypes.c_void_p()
        ln = ctypes.c_size_t()
        _check(_lib.pdlp_b200_stats(self._h, ctypes.byref(p), ctypes.byref(ln)))
        log = SolveLog()
        log.ParseFromString(_take_bytes(p, ln.value))
        return log.original_problem_stats

    def rescale(self, l_inf_ruiz_iterations, l2_norm_rescaling):
        """Returns (row_scale, col_scale, scaled CSC values)."""
        rs, cs, vals = np.zeros(max(self.m, 1)), np.zeros(max(self.n, 1)), np.zeros(max(self.nnz, 1))
        _check(_lib.pdlp_b200_rescale(self._h, int(l_inf_ruiz_iterations), int(bool(l2_norm_rescaling)), _vptr(rs), _vptr(cs), _vptr(vals)))
        return rs[: self.m], cs[: self.n], vals[: self.nnz]

    def kkt_stats(self, params, x, y):
        """IterationStats with one ConvergenceInformation, InfeasibilityInformation and PointMetadata."""
        b = _params_bytes(params)
        x, y = _f64(x), _f64(y)
        p = ctypes.c_void_p()
        ln = ctypes.c_size_t()
        _check(_lib.pdlp_b200_kkt_stats(self._h, b, len(b), _vptr(x if x.size else np.zeros(1)), _vptr(y if y.size else np.zeros(1)), ctypes.byref(p), ctypes.byref(ln)))
        st = IterationStats()
        st.ParseFromString(_take_bytes(p, ln.value))
        return st

    def reduced_costs(self, params, x, y, use_zero_primal_objective=False):
        b = _params_bytes(params)
        x, y = _f64(x), _f64(y)
        rc = np.zeros(max(self.n, 1))
        _check(_lib.pdlp_b200_reduced_costs(self._h, b, len(b), _vptr(x if x.size else np.zeros(1)), _vptr(y if y.size else np.zeros(1)), int(use_zero_primal_objective), _vptr(rc)))
        return rc[: self.n]

    def localized_bounds(self, x, y, primal_weight, radius, use_diagonal_qp_trust_region_solver=False, tolerance=1e-8, max_norm=False):
        x, y = _f64(x), _f64(y)
        out = np.zeros(4)
        if max_norm:  # PrimalDualNorm::kMaxNorm (trust_region.cc:797-884)
            _check(_lib.pdlp_b200_localized_bounds_max_norm(self._h, _vptr(x if x.size else np.zeros(1)), _vptr(y if y.size else np.zeros(1)),
                                                            ctypes.c_double(primal_weight), ctypes.c_double(radius), _vptr(out)))
            return dict(lagrangian_value=out[0], lower_bound=out[1], upper_bound=out[2], radius=out[3])
        _check(_lib.pdlp_b200_localized_bounds(self._h, _vptr(x if x.size else np.zeros(1)), _vptr(y if y.size else np.zeros(1)), ctypes.c_double(primal_weight),
                                               ctypes.c_double(radius), int(use_diagonal_qp_trust_region_solver), ctypes.c_double(tolerance), _vptr(out)))
        return dict(lagrangian_value=out[0], lower_bound=out[1], upper_bound=out[2], radius=out[3])

    def trial_steps(self, steps, step_size, primal_weight, x=None, y=None):
        """Runs `steps` adaptive-rule trial steps from (x, y); returns (x, y, state dict, seconds, launches)."""
        x = np.zeros(max(self.n, 1)) if x is None else _f64(x).copy()
        y = np.zeros(max(self.m, 1)) if y is None else _f64(y).copy()
        st = np.zeros(8)
        sec = ctypes.c_double(0)
        launches = ctypes.c_longlong(0)
        _check(_lib.pdlp_b200_trial_steps(self._h, int(steps), ctypes.c_double(step_size), ctypes.c_double(primal_weight), _vptr(x), _vptr(y), _vptr(st),
                                          ctypes.byref(sec), ctypes.byref(launches)))
        state = dict(step_size=st[0], iterations_completed=int(st[1]), num_rejected_steps=int(st[2]), last_movement=st[3], last_nonlinearity=st[4],
                     total_trials=int(st[5]), pending_rejected_steps=int(st[6]), forced_numerical_termination=bool(st[7]))
        return x[: self.n], y[: self.m], state, sec.value, launches.value

    def trial_steps_profiled(self, steps, step_size, primal_weight, x=None, y=None):
        """Average seconds per launch of the five kernels of a trial step (CUDA events around each)."""
        x = np.zeros(max(self.n, 1)) if x is None else _f64(x)
        y = np.zeros(max(self.m, 1)) if y is None else _f64(y)
        ks = np.zeros(5)
        _check(_lib.pdlp_b200_trial_steps_profiled(self._h, int(steps), ctypes.c_double(step_size), ctypes.c_double(primal_weight), _vptr(x), _vptr(y), _vptr(ks)))
        return dict(zip(("primal_step", "dual_step_Kx", "kty_step_KTy", "decide", "average"), ks.tolist()))
