// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// QuadraticProgram, ShardedQuadraticProgram, weighted average, problem stats,
// Ruiz / L2 rescaling, Lagrangian gradients, projections, sigma_max power method.
//   follows ortools/pdlp/quadratic_program.h:53-143, quadratic_program.cc:37-90,
//           ortools/pdlp/sharded_quadratic_program.cc:63-175,
//           ortools/pdlp/sharded_optimization_utils.cc:43-750.
#ifndef PDLP_ORACLE_QP_H_
#define PDLP_ORACLE_QP_H_

#include <memory>
#include <optional>
#include <random>
#include <string>

#include "oracle_core.h"

namespace oracle {

// quadratic_program.h:53-143.  `q` is the diagonal of the objective matrix.
struct Qp {
  Vec c;
  std::optional<Vec> q;
  Csc K;  // m x n
  Vec clb, cub, vlb, vub;
  double objective_offset = 0.0;
  double objective_scaling_factor = 1.0;
  std::optional<std::string> name;
  bool IsLp() const { return !q.has_value(); }
  double ApplyObjectiveScalingAndOffset(double v) const {
    return objective_scaling_factor * (v + objective_offset);
  }
};

// quadratic_program.cc:37-80.  Returns "" when consistent.
inline std::string ValidateDimensions(const Qp& qp) {
  const int64_t nv = qp.vlb.size(), nc = qp.clb.size();
  auto S = [](int64_t v) { return std::to_string(v); };
  if (nv != (int64_t)qp.vub.size())
    return "Inconsistent dimensions: variable lower bound vector has size " + S(nv) +
           " while variable upper bound vector has size " + S(qp.vub.size());
  if (nv != (int64_t)qp.c.size())
    return "Inconsistent dimensions: variable lower bound vector has size " + S(nv) +
           " while objective vector has size " + S(qp.c.size());
  if (nv != qp.K.cols)
    return "Inconsistent dimensions: variable lower bound vector has size " + S(nv) +
           " while constraint matrix has " + S(qp.K.cols) + " columns";
  if (qp.q.has_value() && nv != (int64_t)qp.q->size())
    return "Inconsistent dimensions: variable lower bound vector has size " + S(nv) +
           " while objective matrix has " + S(qp.q->size()) + " rows";
  if (nc != (int64_t)qp.cub.size())
    return "Inconsistent dimensions: constraint lower bound vector has size " + S(nc) +
           " while constraint upper bound vector has size " + S(qp.cub.size());
  if (nc != qp.K.rows)
    return "Inconsistent dimensions: constraint lower bound vector has size " + S(nc) +
           " while constraint matrix has " + S(qp.K.rows) + " rows ";
  return "";
}

// quadratic_program.cc:82-90.
inline bool HasValidBounds(const Qp& qp) {
  for (size_t i = 0; i < qp.clb.size(); ++i) if (!(qp.clb[i] <= qp.cub[i])) return false;
  for (size_t i = 0; i < qp.vlb.size(); ++i) if (!(qp.vlb[i] <= qp.vub[i])) return false;
  return true;
}

// sharded_quadratic_program.cc:63-175.
class ShardedQp {
 public:
  ShardedQp(Qp qp, int num_threads, int num_shards)
      : qp_(std::move(qp)),
        kt_(qp_.K.Transposed()),
        pool_(num_threads == 1 ? nullptr : std::make_unique<Pool>(num_threads)),
        k_sharder_(qp_.K, num_shards, pool_.get()),
        kt_sharder_(kt_, num_shards, pool_.get()),
        primal_sharder_(static_cast<int64_t>(qp_.vlb.size()), num_shards, pool_.get()),
        dual_sharder_(static_cast<int64_t>(qp_.clb.size()), num_shards, pool_.get()) {
    ORACLE_CHECK(num_threads >= 1);
    ORACLE_CHECK(num_shards >= num_threads);
  }
  const Qp& qp() const { return qp_; }
  const Csc& Kt() const { return kt_; }
  const Sharder& KSharder() const { return k_sharder_; }
  const Sharder& KtSharder() const { return kt_sharder_; }
  const Sharder& PrimalSharder() const { return primal_sharder_; }
  const Sharder& DualSharder() const { return dual_sharder_; }
  int64_t PrimalSize() const { return qp_.K.cols; }
  int64_t DualSize() const { return qp_.K.rows; }

  void SwapVariableBounds(Vec& lb, Vec& ub) { qp_.vlb.swap(lb); qp_.vub.swap(ub); }
  void SwapConstraintBounds(Vec& lb, Vec& ub) { qp_.clb.swap(lb); qp_.cub.swap(ub); }
  void SwapObjectiveVector(Vec& c) { qp_.c.swap(c); }

  // sharded_quadratic_program.cc:134-167.
  void Rescale(const Vec& col_scale, const Vec& row_scale) {
    primal_sharder_.ForEach([&](int, int64_t b, int64_t e) {
      for (int64_t j = b; j < e; ++j) {
        ORACLE_CHECK(col_scale[j] > 0.0);
        qp_.c[j] = qp_.c[j] * col_scale[j];
        qp_.vlb[j] = qp_.vlb[j] / col_scale[j];
        qp_.vub[j] = qp_.vub[j] / col_scale[j];
        if (!qp_.IsLp()) (*qp_.q)[j] = (*qp_.q)[j] * (col_scale[j] * col_scale[j]);
      }
    });
    dual_sharder_.ForEach([&](int, int64_t b, int64_t e) {
      for (int64_t i = b; i < e; ++i) {
        ORACLE_CHECK(row_scale[i] > 0.0);
        qp_.clb[i] = qp_.clb[i] * row_scale[i];
        qp_.cub[i] = qp_.cub[i] * row_scale[i];
      }
    });
    ScaleMatrix(col_scale, row_scale, k_sharder_, qp_.K);
    ScaleMatrix(row_scale, col_scale, kt_sharder_, kt_);
  }

  // sharded_quadratic_program.cc:119-130,169-175.
  void ReplaceLargeConstraintBoundsWithInfinity(double threshold) {
    for (Vec* v : {&qp_.clb, &qp_.cub}) {
      dual_sharder_.ForEach([&](int, int64_t b, int64_t e) {
        for (int64_t i = b; i < e; ++i) {
          if ((*v)[i] <= -threshold) (*v)[i] = -kInf;
          if ((*v)[i] >= threshold) (*v)[i] = kInf;
        }
      });
    }
  }

 private:
  // matrix[i,j] *= row_scale[i] * col_scale[j] (sharded_quadratic_program.cc:97-117).
  static void ScaleMatrix(const Vec& col_scale, const Vec& row_scale, const Sharder& sh, Csc& m) {
    sh.ForEach([&](int, int64_t b, int64_t e) {
      for (int64_t j = b; j < e; ++j)
        for (int64_t k = m.ptr[j]; k < m.ptr[j + 1]; ++k) m.val[k] *= row_scale[m.idx[k]] * col_scale[j];
    });
  }
  Qp qp_;
  Csc kt_;
  std::unique_ptr<Pool> pool_;
  Sharder k_sharder_, kt_sharder_, primal_sharder_, dual_sharder_;
};

// Running weighted mean, update avg += w/(W+w) (x - avg) (sharded_optimization_utils.cc:43-79).
class WeightedAverage {
 public:
  explicit WeightedAverage(const Sharder* sh) : sh_(sh) { avg_ = ZeroVector(*sh_); }
  void Add(const Vec& x, double w) {
    ORACLE_CHECK(w >= 0.0);
    if (w > 0.0) {
      const double ratio = w / (sum_w_ + w);
      sh_->ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) avg_[i] += ratio * (x[i] - avg_[i]); });
      sum_w_ += w;
    }
    ++num_terms_;
  }
  void Clear() { SetZero(*sh_, avg_); sum_w_ = 0.0; num_terms_ = 0; }
  bool HasNonzeroWeight() const { return sum_w_ > 0.0; }
  double Weight() const { return sum_w_; }
  int NumTerms() const { return num_terms_; }
  Vec ComputeAverage() const { return CloneVector(avg_, *sh_); }

 private:
  const Sharder* sh_;
  Vec avg_;
  double sum_w_ = 0.0;
  int num_terms_ = 0;
};

// ---- problem statistics (sharded_optimization_utils.cc:83-343; solve_log.proto:24-98) ----
struct QpStats {
  int64_t num_variables = 0, num_constraints = 0;
  double constraint_matrix_col_min_l_inf_norm = 0, constraint_matrix_row_min_l_inf_norm = 0;
  int64_t constraint_matrix_num_nonzeros = 0;
  double constraint_matrix_abs_max = 0, constraint_matrix_abs_min = 0, constraint_matrix_abs_avg = 0, constraint_matrix_l2_norm = 0;
  double combined_bounds_max = 0, combined_bounds_min = 0, combined_bounds_avg = 0, combined_bounds_l2_norm = 0;
  double combined_variable_bounds_max = 0, combined_variable_bounds_min = 0, combined_variable_bounds_avg = 0, combined_variable_bounds_l2_norm = 0;
  int64_t variable_bound_gaps_num_finite = 0;
  double variable_bound_gaps_max = 0, variable_bound_gaps_min = 0, variable_bound_gaps_avg = 0, variable_bound_gaps_l2_norm = 0;
  double objective_vector_abs_max = 0, objective_vector_abs_min = 0, objective_vector_abs_avg = 0, objective_vector_l2_norm = 0;
  int64_t objective_matrix_num_nonzeros = 0;
  double objective_matrix_abs_max = 0, objective_matrix_abs_min = 0, objective_matrix_abs_avg = 0, objective_matrix_l2_norm = 0;
};

namespace detail {
inline double CombineBounds(double v1, double v2) {  // :83-92
  double mx = 0.0;
  if (std::abs(v1) < kInf) mx = std::abs(v1);
  if (std::abs(v2) < kInf) mx = std::max(mx, std::abs(v2));
  return mx;
}
struct VInfo { int64_t nfn = 0, ninf = 0, nzero = 0; double largest = 0, smallest = 0, average = 0, l2 = 0; };
struct Acc {  // VectorInfoAccumulator :117-177
  int64_t ninf = 0, nzero = 0, nfn = 0;
  double mx = -kInf, mn = kInf, sum = 0, sumsq = 0;
  void Add(double v) {
    if (std::isinf(v)) ++ninf;
    else if (v == 0) ++nzero;
    else { ++nfn; const double a = std::abs(v); mx = std::max(mx, a); mn = std::min(mn, a); sum += a; sumsq += a * a; }
  }
  void Add(const Acc& o) { ninf += o.ninf; nzero += o.nzero; nfn += o.nfn; mx = std::max(mx, o.mx); mn = std::min(mn, o.mn); sum += o.sum; sumsq += o.sumsq; }
  VInfo Finish() const {
    VInfo r;
    r.nfn = nfn; r.ninf = ninf; r.nzero = nzero;
    r.largest = nfn > 0 ? mx : 0.0;
    r.smallest = nfn > 0 ? mn : 0.0;
    r.average = (nfn + nzero > 0) ? sum / (nfn + nzero) : std::numeric_limits<double>::quiet_NaN();
    r.l2 = std::sqrt(sumsq);
    return r;
  }
};
template <typename ElemFn>
inline VInfo AccumulateOver(const Sharder& sh, ElemFn elem) {
  std::vector<Acc> local(sh.NumShards());
  sh.ForEach([&](int s, int64_t b, int64_t e) { Acc a; for (int64_t i = b; i < e; ++i) elem(i, a); local[s] = a; });
  Acc total;
  for (const Acc& a : local) total.Add(a);
  return total.Finish();
}
}  // namespace detail

inline QpStats ComputeStats(const ShardedQp& sq) {
  using detail::Acc; using detail::VInfo;
  const Qp& qp = sq.qp();
  const Vec ones_p = OnesVector(sq.PrimalSharder()), ones_d = OnesVector(sq.DualSharder());
  const Vec row_norms = ScaledColLInfNorm(sq.Kt(), ones_p, ones_d, sq.KtSharder());
  const Vec col_norms = ScaledColLInfNorm(qp.K, ones_d, ones_p, sq.KSharder());
  const VInfo rn = detail::AccumulateOver(sq.DualSharder(), [&](int64_t i, Acc& a) { a.Add(row_norms[i]); });
  const VInfo cn = detail::AccumulateOver(sq.PrimalSharder(), [&](int64_t i, Acc& a) { a.Add(col_norms[i]); });
  const VInfo mat = detail::AccumulateOver(sq.KSharder(), [&](int64_t j, Acc& a) { for (int64_t k = qp.K.ptr[j]; k < qp.K.ptr[j + 1]; ++k) a.Add(qp.K.val[k]); });
  const VInfo cb = detail::AccumulateOver(sq.DualSharder(), [&](int64_t i, Acc& a) { a.Add(detail::CombineBounds(qp.cub[i], qp.clb[i])); });
  const VInfo vb = detail::AccumulateOver(sq.PrimalSharder(), [&](int64_t i, Acc& a) { a.Add(detail::CombineBounds(qp.vub[i], qp.vlb[i])); });
  const VInfo ov = detail::AccumulateOver(sq.PrimalSharder(), [&](int64_t i, Acc& a) { a.Add(qp.c[i]); });
  const VInfo gaps = detail::AccumulateOver(sq.PrimalSharder(), [&](int64_t i, Acc& a) { a.Add(qp.vub[i] - qp.vlb[i]); });
  QpStats s;
  s.num_variables = sq.PrimalSize();
  s.num_constraints = sq.DualSize();
  s.constraint_matrix_col_min_l_inf_norm = cn.smallest;
  s.constraint_matrix_row_min_l_inf_norm = rn.smallest;
  s.constraint_matrix_num_nonzeros = mat.nfn;
  s.constraint_matrix_abs_max = mat.largest; s.constraint_matrix_abs_min = mat.smallest;
  s.constraint_matrix_abs_avg = mat.average; s.constraint_matrix_l2_norm = mat.l2;
  s.combined_bounds_max = cb.largest; s.combined_bounds_min = cb.smallest; s.combined_bounds_avg = cb.average; s.combined_bounds_l2_norm = cb.l2;
  s.combined_variable_bounds_max = vb.largest; s.combined_variable_bounds_min = vb.smallest;
  s.combined_variable_bounds_avg = vb.average; s.combined_variable_bounds_l2_norm = vb.l2;
  s.variable_bound_gaps_num_finite = gaps.nfn + gaps.nzero;
  s.variable_bound_gaps_max = gaps.largest; s.variable_bound_gaps_min = gaps.smallest; s.variable_bound_gaps_avg = gaps.average; s.variable_bound_gaps_l2_norm = gaps.l2;
  s.objective_vector_abs_max = ov.largest; s.objective_vector_abs_min = ov.smallest; s.objective_vector_abs_avg = ov.average; s.objective_vector_l2_norm = ov.l2;
  if (qp.IsLp()) {
    s.objective_matrix_abs_avg = std::numeric_limits<double>::quiet_NaN();
  } else {
    const VInfo om = detail::AccumulateOver(sq.PrimalSharder(), [&](int64_t i, Acc& a) { a.Add((*qp.q)[i]); });
    s.objective_matrix_num_nonzeros = om.nfn;
    s.objective_matrix_abs_max = om.largest; s.objective_matrix_abs_min = om.smallest; s.objective_matrix_abs_avg = om.average; s.objective_matrix_l2_norm = om.l2;
  }
  return s;
}

// ---- rescaling (sharded_optimization_utils.cc:347-444) ----
struct ScalingVectors { Vec row, col; };

namespace detail {
inline void DivideBySqrtOfDivisor(const Vec& divisor, const Sharder& sh, Vec& v) {  // :354-365
  sh.ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) if (divisor[i] != 0) v[i] /= std::sqrt(divisor[i]); });
}
inline void ScalingIterations(const ShardedQp& sq, int iters, bool l2, Vec& row, Vec& col) {  // :367-405
  const Qp& qp = sq.qp();
  for (int it = 0; it < iters; ++it) {
    Vec col_norm, row_norm;
    if (l2) {
      col_norm = ScaledColL2Norm(qp.K, row, col, sq.KSharder());
      row_norm = ScaledColL2Norm(sq.Kt(), col, row, sq.KtSharder());
    } else {
      col_norm = ScaledColLInfNorm(qp.K, row, col, sq.KSharder());
      row_norm = ScaledColLInfNorm(sq.Kt(), col, row, sq.KtSharder());
    }
    DivideBySqrtOfDivisor(col_norm, sq.PrimalSharder(), col);
    DivideBySqrtOfDivisor(row_norm, sq.DualSharder(), row);
  }
}
}  // namespace detail

inline void LInfRuizRescaling(const ShardedQp& sq, int iters, Vec& row, Vec& col) { detail::ScalingIterations(sq, iters, false, row, col); }
inline void L2NormRescaling(const ShardedQp& sq, Vec& row, Vec& col) { detail::ScalingIterations(sq, 1, true, row, col); }

inline ScalingVectors ApplyRescaling(int l_inf_ruiz_iterations, bool l2_norm_rescaling, ShardedQp& sq) {  // :423-444
  ScalingVectors s{OnesVector(sq.DualSharder()), OnesVector(sq.PrimalSharder())};
  bool any = false;
  if (l_inf_ruiz_iterations > 0) { any = true; LInfRuizRescaling(sq, l_inf_ruiz_iterations, s.row, s.col); }
  if (l2_norm_rescaling) { any = true; L2NormRescaling(sq, s.row, s.col); }
  if (any) sq.Rescale(s.col, s.row);
  return s;
}

// ---- Lagrangian gradients (sharded_optimization_utils.cc:446-527) ----
struct LagrangianPart { double value = 0.0; Vec gradient; };

inline LagrangianPart ComputePrimalGradient(const ShardedQp& sq, const Vec& x, const Vec& dual_product) {
  const Qp& qp = sq.qp();
  LagrangianPart r;
  r.gradient.resize(sq.PrimalSize());
  r.value = sq.PrimalSharder().Sum([&](int, int64_t b, int64_t e) {
    double v = 0.0;
    if (qp.IsLp()) {
      for (int64_t j = b; j < e; ++j) { r.gradient[j] = qp.c[j] - dual_product[j]; v += x[j] * r.gradient[j]; }
    } else {
      for (int64_t j = b; j < e; ++j) {
        const double qx = (*qp.q)[j] * x[j];
        r.gradient[j] = qp.c[j] + qx - dual_product[j];
        v += x[j] * (r.gradient[j] - 0.5 * qx);
      }
    }
    return v;
  });
  return r;
}

inline double DualSubgradientCoefficient(double lb, double ub, double dual, double primal_product) {  // :476-500
  if (dual < 0.0) return ub;
  if (dual > 0.0) return lb;
  if (std::isfinite(lb) && std::isfinite(ub)) {
    if (primal_product < lb) return lb;
    if (primal_product > ub) return ub;
    return primal_product;
  }
  if (std::isfinite(lb)) return lb;
  if (std::isfinite(ub)) return ub;
  return 0.0;
}

inline LagrangianPart ComputeDualGradient(const ShardedQp& sq, const Vec& y, const Vec& primal_product) {
  const Qp& qp = sq.qp();
  LagrangianPart r;
  r.gradient.resize(sq.DualSize());
  r.value = sq.DualSharder().Sum([&](int, int64_t b, int64_t e) {
    double v = 0.0;
    for (int64_t i = b; i < e; ++i) {
      r.gradient[i] = DualSubgradientCoefficient(qp.clb[i], qp.cub[i], y[i], primal_product[i]);
      v += r.gradient[i] * y[i];
    }
    for (int64_t i = b; i < e; ++i) r.gradient[i] -= primal_product[i];
    return v;
  });
  return r;
}

// ---- projections (sharded_optimization_utils.cc:721-750) ----
inline void ProjectToPrimalVariableBounds(const ShardedQp& sq, Vec& x) {
  const Qp& qp = sq.qp();
  sq.PrimalSharder().ForEach([&](int, int64_t b, int64_t e) { for (int64_t j = b; j < e; ++j) x[j] = std::max(std::min(x[j], qp.vub[j]), qp.vlb[j]); });
}
inline void ProjectToDualVariableBounds(const ShardedQp& sq, Vec& y) {
  const Qp& qp = sq.qp();
  sq.DualSharder().ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t i = b; i < e; ++i) {
      if (!std::isfinite(qp.cub[i])) y[i] = std::max(y[i], 0.0);
      if (!std::isfinite(qp.clb[i])) y[i] = std::min(y[i], 0.0);
    }
  });
}

// ---- sigma_max power method (sharded_optimization_utils.cc:536-699) ----
// NOTE parity unpinned for exact values: the reference draws the start vector with
// absl::Gaussian(mt19937(1)), whose algorithm is not in /root/reference; this uses
// std::normal_distribution on the same engine.  The reference's own tests pin only
// tolerance bands (sharded_optimization_utils_test.cc:516-594).
struct SingularValueAndIterations { double singular_value; int num_iterations; double estimated_relative_error; };

inline double PowerMethodFailureProbability(int64_t dimension, double epsilon, int k) {  // :550-557
  if (k < 2 || epsilon <= 0.0) return 1.0;
  return std::min(0.824, 0.354 / std::sqrt(epsilon * (k - 1))) * std::sqrt((double)dimension) * std::pow(1.0 - epsilon, k - 0.5);
}

inline SingularValueAndIterations EstimateMaximumSingularValue(const ShardedQp& sq, const Vec* primal_active, const Vec* dual_active,
                                                               double desired_relative_error, double failure_probability, std::mt19937& gen) {
  const Qp& qp = sq.qp();
  const int64_t dim = qp.K.cols;
  Vec ev(dim);
  std::normal_distribution<double> gauss(0.0, 1.0);
  for (double& v : ev) v = gauss(gen);
  auto normalize = [&](Vec& v) { const double nrm = Norm(v, sq.PrimalSharder()); if (nrm != 0.0) sq.PrimalSharder().ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) v[i] /= nrm; }); return nrm; };
  if (primal_active) CwiseProductInPlace(*primal_active, sq.PrimalSharder(), ev);
  normalize(ev);
  double estimate = 0.0;
  int iters = 0;
  const double eps = 1.0 - (1.0 - desired_relative_error) * (1.0 - desired_relative_error);
  while (PowerMethodFailureProbability(dim, eps, iters) > failure_probability) {
    Vec dual_ev = TransposedMatVec(sq.Kt(), ev, sq.KtSharder());
    if (dual_active) CwiseProductInPlace(*dual_active, sq.DualSharder(), dual_ev);
    Vec next = TransposedMatVec(qp.K, dual_ev, sq.KSharder());
    if (primal_active) CwiseProductInPlace(*primal_active, sq.PrimalSharder(), next);
    estimate = Dot(ev, next, sq.PrimalSharder());
    ev = std::move(next);
    ++iters;
    normalize(ev);
  }
  return {std::sqrt(estimate), iters, desired_relative_error};
}

// ComputePrimalActiveSetIndicator / ComputeDualActiveSetIndicator (sharded_optimization_utils.cc:617-672): 1 where the
// coordinate is not held by a bound (for y: the implicit sign bound of a one-sided or free row), else 0.
inline Vec PrimalActiveSetIndicator(const ShardedQp& sq, const Vec& x) {
  const Qp& qp = sq.qp();
  Vec ind(sq.PrimalSize());
  sq.PrimalSharder().ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t j = b; j < e; ++j) ind[j] = (x[j] == qp.vlb[j] || x[j] == qp.vub[j]) ? 0.0 : 1.0;
  });
  return ind;
}
inline Vec DualActiveSetIndicator(const ShardedQp& sq, const Vec& y) {
  const Qp& qp = sq.qp();
  Vec ind(sq.DualSize());
  sq.DualSharder().ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t i = b; i < e; ++i) ind[i] = (y[i] == 0.0 && (std::isinf(qp.clb[i]) || std::isinf(qp.cub[i]))) ? 0.0 : 1.0;
  });
  return ind;
}
// EstimateMaximumSingularValueOfConstraintMatrix (sharded_optimization_utils.cc:676-699): the public form, on the
// columns / rows that the given points leave free.
inline SingularValueAndIterations EstimateMaximumSingularValueOfConstraintMatrix(const ShardedQp& sq, const Vec* primal_solution, const Vec* dual_solution,
                                                                                double desired_relative_error, double failure_probability, std::mt19937& gen) {
  Vec primal_indicator, dual_indicator;
  if (primal_solution) primal_indicator = PrimalActiveSetIndicator(sq, *primal_solution);
  if (dual_solution) dual_indicator = DualActiveSetIndicator(sq, *dual_solution);
  return EstimateMaximumSingularValue(sq, primal_solution ? &primal_indicator : nullptr, dual_solution ? &dual_indicator : nullptr, desired_relative_error,
                                      failure_probability, gen);
}

}  // namespace oracle
#endif  // PDLP_ORACLE_QP_H_
