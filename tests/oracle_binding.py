"""ctypes driver of the CPU oracle (oracle/oracle_capi.cc).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference leg."""
import ctypes
import json
import os
import subprocess

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "_build", "libpdlp_oracle.so")


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])


def load():
    if not os.path.exists(ORACLE_LIB):
        build()
    lib = ctypes.CDLL(ORACLE_LIB)
    lib.oracle_sigma_max.restype = ctypes.c_double
    lib.oracle_sigma_max_at.restype = ctypes.c_double
    lib.oracle_blas1.restype = ctypes.c_double
    lib.oracle_problem_create.restype = ctypes.c_void_p
    return lib


def flatten_params(msg, prefix=""):
    """protobuf message -> 'a.b=value' lines holding exactly the explicitly set fields."""
    lines = []
    if msg is None:
        return ""
    for fd, value in msg.ListFields():
        name = prefix + fd.name
        if fd.type == fd.TYPE_MESSAGE:
            if fd.name == "glop_parameters":
                continue
            lines.append(name + "=")
            sub = flatten_params(value, name + ".")
            if sub:
                lines.append(sub)
        elif fd.is_repeated if hasattr(fd, "is_repeated") else fd.label == fd.LABEL_REPEATED:
            lines.append(name + "=" + ",".join(str(v) for v in value))
        elif fd.type == fd.TYPE_BOOL:
            lines.append(name + "=" + ("1" if value else "0"))
        elif fd.type == fd.TYPE_DOUBLE:
            lines.append(name + "=" + repr(float(value)))
        else:
            lines.append(name + "=" + str(int(value)))
    return "\n".join(lines)


def _f(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))


def _p(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None and a.size else ctypes.c_void_p(None)


class Marshalled:
    def __init__(self, qp):
        K = sp.csc_matrix(qp.constraint_matrix)
        K.sum_duplicates()
        self.m, self.n = K.shape
        self.cp = np.ascontiguousarray(K.indptr, dtype=np.int64)
        self.ri = np.ascontiguousarray(K.indices, dtype=np.int64)
        self.v = np.ascontiguousarray(K.data, dtype=np.float64)
        f = lambda a: np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1))
        self.c = f(qp.objective_vector)
        self.q = None if qp.objective_matrix is None else f(qp.objective_matrix)
        self.vlb, self.vub = f(qp.variable_lower_bounds), f(qp.variable_upper_bounds)
        self.clb, self.cub = f(qp.constraint_lower_bounds), f(qp.constraint_upper_bounds)
        self.offset, self.scale = float(qp.objective_offset), float(qp.objective_scaling_factor)

    def args(self):
        return [ctypes.c_int64(self.n), ctypes.c_int64(self.m), ctypes.c_int64(self.v.size), _p(self.cp), _p(self.ri), _p(self.v), _p(self.c),
                _p(self.q), _p(self.vlb), _p(self.vub), _p(self.clb), _p(self.cub), ctypes.c_double(self.offset), ctypes.c_double(self.scale)]


CALLBACK = ctypes.CFUNCTYPE(None, ctypes.c_int, ctypes.c_char_p, ctypes.c_void_p)


class OracleResult:
    pass


def to_proto(d, msg):
    """JSON dict printed by the oracle (proto field names as keys) -> protobuf message."""
    fields = msg.DESCRIPTOR.fields_by_name
    for k, v in d.items():
        fd = fields[k]
        rep = fd.is_repeated if hasattr(fd, "is_repeated") else fd.label == fd.LABEL_REPEATED
        if fd.type == fd.TYPE_MESSAGE:
            if rep:
                for item in v:
                    to_proto(item, getattr(msg, k).add())
            else:
                sub = getattr(msg, k)
                sub.SetInParent()
                to_proto(v, sub)
        elif rep:
            getattr(msg, k).extend(v)
        elif fd.type == fd.TYPE_STRING and v == "" and k == "instance_name":
            continue  # unnamed problems leave the field unset (primal_dual_hybrid_gradient_test.cc:746)
        else:
            setattr(msg, k, v)
    return msg


def _solve_log_proto(d, params):
    import pdlp_b200
    log = to_proto(d, pdlp_b200.SolveLog())
    if params is not None and log.termination_reason not in (9, 10, 13):
        log.params.CopyFrom(params)  # the oracle echoes nothing; SolverResult.solve_log.params is the caller's
    return log


def _params_text(params):
    """Params as the oracle's 'a.b=value' lines: a protobuf message, or such text already (callers that
    must not import the product package, i.e. bench.py --impl reference, write the lines themselves)."""
    return (params if isinstance(params, str) else flatten_params(params)).encode()


def time_steps(qp, params, warmup, steps):
    """Seconds of `steps` iterations of the oracle's step loop alone (after `warmup` untimed ones; no
    termination checks, no restarts): dict(seconds, accepted_steps, rejected_steps, preprocessing_seconds)."""
    lib = load()
    mar = Marshalled(qp)
    out = np.zeros(4)
    rc = lib.oracle_time_steps(*mar.args(), _params_text(params), int(warmup), int(steps), _p(out))
    if rc != 0:
        raise RuntimeError("oracle_time_steps: preprocessing rejected the problem or a step forced numerical termination")
    return dict(seconds=float(out[0]), accepted_steps=int(out[1]), rejected_steps=int(out[2]), preprocessing_seconds=float(out[3]))


def solve(qp, params=None, initial_solution=None, callback=None, interrupt=None):
    lib = load()
    mar = Marshalled(qp)
    text = _params_text(params)
    x = np.zeros(max(mar.n, 1)); y = np.zeros(max(mar.m, 1)); rc = np.zeros(max(mar.n, 1))
    sizes = np.zeros(3, dtype=np.int64)
    log = ctypes.c_char_p()
    x0 = y0 = None
    if initial_solution is not None:
        x0 = np.ascontiguousarray(initial_solution[0], dtype=np.float64)
        y0 = np.ascontiguousarray(initial_solution[1], dtype=np.float64)
        if x0.size == 0: x0 = np.zeros(1)
        if y0.size == 0: y0 = np.zeros(1)
    cb = None
    if callback is not None:
        cb = CALLBACK(lambda t, s, u: callback(t, json.loads(s)))
    lib.oracle_solve(*mar.args(), text, _p(x0), _p(y0), _p(interrupt) if interrupt is not None else None, cb, None, _p(x), _p(y), _p(rc), _p(sizes),
                     ctypes.byref(log))
    r = OracleResult()
    r.solve_log_dict = json.loads(log.value)
    r.solve_log = None if isinstance(params, str) else _solve_log_proto(r.solve_log_dict, params)
    r.primal_solution = x[: sizes[0]].copy()
    r.dual_solution = y[: sizes[1]].copy()
    r.reduced_costs = rc[: sizes[2]].copy()
    return r


class Problem:
    """Handle on an oracle ShardedQp for the unit entry points."""

    def __init__(self, qp, num_threads=1, num_shards=1):
        self.lib = load()
        self.mar = Marshalled(qp)
        self.h = ctypes.c_void_p(self.lib.oracle_problem_create(*self.mar.args(), num_threads, num_shards))
        self.n, self.m = self.mar.n, self.mar.m

    def __del__(self):
        try:
            self.lib.oracle_problem_destroy(self.h)
        except Exception:
            pass

    def matvec(self, which, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        out = np.zeros(max(self.n if which == 0 else self.m, 1))
        self.lib.oracle_matvec(self.h, which, _p(v), _p(out))
        return out[: self.n if which == 0 else self.m]

    def col_norms(self, l2, row_scale, col_scale):
        rs = np.ascontiguousarray(row_scale, dtype=np.float64); cs = np.ascontiguousarray(col_scale, dtype=np.float64)
        out = np.zeros(max(self.n, 1))
        self.lib.oracle_col_norms(self.h, int(l2), _p(rs), _p(cs), _p(out))
        return out[: self.n]

    def stats(self):
        s = ctypes.c_char_p()
        self.lib.oracle_stats(self.h, ctypes.byref(s))
        return json.loads(s.value)

    def rescale(self, ruiz, l2):
        rs = np.zeros(max(self.m, 1)); cs = np.zeros(max(self.n, 1))
        self.lib.oracle_rescale(self.h, int(ruiz), int(l2), _p(rs), _p(cs))
        return rs[: self.m], cs[: self.n]

    def rescale_with(self, col_scale, row_scale):
        """RescaleQuadraticProgram (sharded_quadratic_program.cc:134-167) with the given scaling vectors."""
        col_scale, row_scale = _f(col_scale), _f(row_scale)  # named: the arrays must outlive the call
        self.lib.oracle_rescale_with(self.h, _p(col_scale), _p(row_scale))

    def replace_large_constraint_bounds_with_infinity(self, threshold):
        self.lib.oracle_replace_large_constraint_bounds(self.h, ctypes.c_double(threshold))

    def transposed_values(self):
        nnz = self.mar.v.size
        out = np.zeros(max(nnz, 1))
        self.lib.oracle_problem_get_transposed_values(self.h, _p(out))
        return out[:nnz]

    def random_projections(self, x, y, seeds):
        """SetRandomProjections (iteration_stats.cc:598-610): (primal projections, dual projections), one per seed."""
        x, y = _f(x), _f(y)
        s = np.ascontiguousarray(seeds, dtype=np.int32)
        po, do = np.zeros(max(len(s), 1)), np.zeros(max(len(s), 1))
        self.lib.oracle_random_projections(self.h, _p(x), _p(y), ctypes.c_void_p(s.ctypes.data), len(s), _p(po), _p(do))
        return po[: len(s)], do[: len(s)]

    def project(self, x=None, y=None):
        """ProjectToPrimalVariableBounds / ProjectToDualVariableBounds (sharded_optimization_utils.cc:701-750): copies."""
        px = None if x is None else _f(x).copy()
        py = None if y is None else _f(y).copy()
        self.lib.oracle_project(self.h, _p(px) if px is not None else None, _p(py) if py is not None else None)
        return px, py

    def scaling_vectors(self, which, iters, row_scale, col_scale):
        rs = np.ascontiguousarray(row_scale, dtype=np.float64).copy(); cs = np.ascontiguousarray(col_scale, dtype=np.float64).copy()
        self.lib.oracle_scaling_vectors(self.h, int(which), int(iters), _p(rs), _p(cs))
        return rs, cs

    def get(self):
        nnz = self.mar.v.size
        out = dict(val=np.zeros(max(nnz, 1)), c=np.zeros(max(self.n, 1)), q=np.zeros(max(self.n, 1)), vlb=np.zeros(max(self.n, 1)),
                   vub=np.zeros(max(self.n, 1)), clb=np.zeros(max(self.m, 1)), cub=np.zeros(max(self.m, 1)))
        self.lib.oracle_problem_get(self.h, _p(out["val"]), _p(out["c"]), _p(out["q"]), _p(out["vlb"]), _p(out["vub"]), _p(out["clb"]), _p(out["cub"]))
        sizes = dict(val=nnz, c=self.n, q=self.n, vlb=self.n, vub=self.n, clb=self.m, cub=self.m)
        return {k: v[: sizes[k]] for k, v in out.items()}

    def convergence_info(self, params, col_scale, row_scale, x, y, cw_primal_offset=1.0, cw_dual_offset=1.0, candidate_type=1):
        col_scale, row_scale, x, y = map(_f, (col_scale, row_scale, x, y))  # named: the arrays must outlive the call
        s = ctypes.c_char_p()
        self.lib.oracle_convergence_info(self.h, _params_text(params), _p(col_scale), _p(row_scale), _p(x), _p(y),
                                         ctypes.c_double(cw_primal_offset), ctypes.c_double(cw_dual_offset), candidate_type, ctypes.byref(s))
        return json.loads(s.value)

    def infeasibility_info(self, params, col_scale, row_scale, x, y, candidate_type=1):
        col_scale, row_scale, x, y = map(_f, (col_scale, row_scale, x, y))
        s = ctypes.c_char_p()
        self.lib.oracle_infeasibility_info(self.h, _params_text(params), _p(col_scale), _p(row_scale), _p(x), _p(y), candidate_type,
                                           ctypes.byref(s))
        return json.loads(s.value)

    def reduced_costs(self, params, x, y, zero_objective=False):
        x, y = _f(x), _f(y)
        rc = np.zeros(max(self.n, 1))
        self.lib.oracle_reduced_costs(self.h, _params_text(params), _p(x), _p(y), int(zero_objective), _p(rc))
        return rc[: self.n]

    def lagrangian_gradients(self, x, y):
        x, y = _f(x), _f(y)
        pg = np.zeros(max(self.n, 1)); dg = np.zeros(max(self.m, 1)); vals = np.zeros(2)
        self.lib.oracle_lagrangian_gradients(self.h, _p(x), _p(y), _p(pg), _p(dg), _p(vals))
        return pg[: self.n], dg[: self.m], vals

    def localized_bounds(self, x, y, norm, primal_weight, radius, use_diagonal=False, tol=1e-8):
        x, y = _f(x), _f(y)
        out = np.zeros(4)
        self.lib.oracle_localized_bounds(self.h, _p(x), _p(y), int(norm), ctypes.c_double(primal_weight), ctypes.c_double(radius), int(use_diagonal),
                                         ctypes.c_double(tol), _p(out))
        return dict(lagrangian_value=out[0], lower_bound=out[1], upper_bound=out[2], radius=out[3])

    def localized_bounds_cached(self, x, y, kx, kty, primal_weight, radius, use_diagonal=False, tol=1e-8, max_norm=False):
        """ComputeLocalizedLagrangianBounds with the caller's primal_product = K x and dual_product = K^T y."""
        x, y, kx, kty = _f(x), _f(y), _f(kx), _f(kty)
        out = np.zeros(4)
        self.lib.oracle_localized_bounds_cached(self.h, _p(x), _p(y), _p(kx), _p(kty), 0 if max_norm else 1, ctypes.c_double(primal_weight), ctypes.c_double(radius),
                                                int(use_diagonal), ctypes.c_double(tol), _p(out))
        return dict(lagrangian_value=out[0], lower_bound=out[1], upper_bound=out[2], radius=out[3])

    def sigma_max(self, rel_err=0.2, fail_prob=0.0005, seed=1):
        it = ctypes.c_int(0)
        v = self.lib.oracle_sigma_max(self.h, ctypes.c_double(rel_err), ctypes.c_double(fail_prob), int(seed), ctypes.byref(it))
        return v, it.value

    def sigma_max_at(self, x=None, y=None, rel_err=0.2, fail_prob=0.0005, seed=1):
        """EstimateMaximumSingularValueOfConstraintMatrix on the subspaces that x and / or y leave active."""
        it = ctypes.c_int(0)
        x = None if x is None else _f(x)
        y = None if y is None else _f(y)
        v = self.lib.oracle_sigma_max_at(self.h, _p(x) if x is not None else None, _p(y) if y is not None else None, ctypes.c_double(rel_err),
                                         ctypes.c_double(fail_prob), int(seed), ctypes.byref(it))
        return v, it.value


def trust_region(obj, lb, ub, center, w, radius, diag=None, num_shards=1, tol=1e-8):
    lib = load()
    obj, lb, ub, center, w = map(_f, (obj, lb, ub, center, w))
    diag = None if diag is None else _f(diag)
    sol = np.zeros(max(obj.size, 1)); out = np.zeros(2)
    lib.oracle_trust_region(ctypes.c_int64(obj.size), _p(obj), _p(diag) if diag is not None else None, _p(lb), _p(ub), _p(center), _p(w),
                            ctypes.c_double(radius), int(num_shards), ctypes.c_double(tol), _p(sol), _p(out))
    return sol[: obj.size], out[0], out[1]


def blas1(op, n, a=None, b=None, scalar=0.0, num_shards=2):
    """A sharded BLAS-1 function of sharder.h by its reference name: (scalar result, vector result)."""
    va = None if a is None else _f(a)
    vb = None if b is None else _f(b)
    out = np.zeros(max(n, 1))
    r = load().oracle_blas1(op.encode(), ctypes.c_int64(n), int(num_shards), _p(va) if va is not None else None, _p(vb) if vb is not None else None,
                            ctypes.c_double(scalar), _p(out))
    return r, out[:n]


def weighted_average(vectors, weights, num_shards=1, clear_after_add=-1):
    """ShardedWeightedAverage after Add(vectors[k], weights[k]) for every k: (average, has_nonzero_weight, num_terms)."""
    v = np.ascontiguousarray(np.atleast_2d(np.asarray(vectors, dtype=np.float64)))
    w = _f(weights)
    n = v.shape[1]
    count = 0 if len(w) == 0 else v.shape[0]
    out, meta = np.zeros(max(n, 1)), np.zeros(2)
    load().oracle_weighted_average(ctypes.c_int64(n), int(num_shards), int(count), _p(v.reshape(-1)), _p(w if w.size else np.zeros(1)), int(clear_after_add), _p(out), _p(meta))
    return out[:n], bool(meta[0]), int(meta[1])


def shard_starts(n, num_shards, masses=None):
    lib = load()
    starts = np.zeros(min(num_shards, n) + 2 + n, dtype=np.int64)
    ms = None if masses is None else np.ascontiguousarray(masses, dtype=np.int64)
    k = lib.oracle_shard_starts(ctypes.c_int64(n), int(num_shards), _p(ms) if ms is not None else None, _p(starts))
    return starts[: k + 1]


def validate_params(params):
    lib = load()
    s = ctypes.c_char_p()
    lib.oracle_validate_params(_params_text(params), ctypes.byref(s))
    return s.value.decode()


def check_termination(params, stats, bound_norms, which="iterate", force_numerical_termination=False, interrupt=None):
    """Same contract as pdlp_b200.check_termination, on the oracle."""
    lib = load()
    conv = np.array([[c.candidate_type, c.primal_objective, c.dual_objective, c.l_inf_primal_residual, c.l2_primal_residual,
                      c.l_inf_componentwise_primal_residual, c.l_inf_dual_residual, c.l2_dual_residual, c.l_inf_componentwise_dual_residual]
                     for c in stats.convergence_information], dtype=np.float64).reshape(-1)
    inf = np.array([[c.candidate_type, c.max_primal_ray_infeasibility, c.primal_ray_linear_objective, c.primal_ray_quadratic_norm,
                     c.max_dual_ray_infeasibility, c.dual_ray_objective] for c in stats.infeasibility_information], dtype=np.float64).reshape(-1)
    work = np.array([stats.iteration_number, stats.cumulative_kkt_matrix_passes, stats.cumulative_time_sec], dtype=np.float64)
    norms = _f(bound_norms)
    out = np.zeros(2, dtype=np.int32)
    rel = np.zeros(5)
    lib.oracle_check_termination(_params_text(params), 0 if which == "iterate" else 1, int(force_numerical_termination),
                                 _p(interrupt) if interrupt is not None else None, _p(norms), len(stats.convergence_information), _p(conv),
                                 len(stats.infeasibility_information), _p(inf), _p(work), _p(out), _p(rel))
    return int(out[0]), int(out[1]), rel
