// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// Bound-constrained trust-region solves and localized Lagrangian bounds used by the
// restart logic.
//   follows ortools/pdlp/trust_region.h:32-125,193-417 and
//           ortools/pdlp/trust_region.cc:37-1016.
#ifndef PDLP_ORACLE_TR_H_
#define PDLP_ORACLE_TR_H_

#include "oracle_stats.h"

namespace oracle {

struct TrustRegionResult { double solution_step_size = 0, objective_value = 0; Vec solution; };
struct LocalizedLagrangianBounds { double lagrangian_value = 0, lower_bound = 0, upper_bound = 0, radius = 0; };
inline double BoundGap(const LocalizedLagrangianBounds& b) { return b.upper_bound - b.lower_bound; }
enum class PrimalDualNorm { kMaxNorm, kEuclideanNorm };

namespace tr {

// trust_region.h:193-228.
template <typename P> double DistanceAtCriticalStepSize(const P& p, int64_t i) {
  if (p.Objective(i) == 0.0) return 0.0;
  return p.Objective(i) > 0.0 ? p.LowerBound(i) - p.CenterPoint(i) : p.UpperBound(i) - p.CenterPoint(i);
}
template <typename P> double CriticalStepSize(const P& p, int64_t i) {
  if (p.Objective(i) == 0.0) return kInf;
  return -p.NormWeight(i) * DistanceAtCriticalStepSize(p, i) / p.Objective(i);
}
template <typename P> double ProjectedValue(const P& p, int64_t i, double step) {
  const double full = p.CenterPoint(i) - step * p.Objective(i) / p.NormWeight(i);
  return std::min(std::max(full, p.LowerBound(i)), p.UpperBound(i));
}

// Median by nth_element, off by one half element for even sizes (trust_region.h:234-244).
template <typename P> double EasyMedianOfCritical(const P& p, std::vector<int64_t> idx) {
  ORACLE_CHECK(!idx.empty());
  auto mid = idx.begin() + idx.size() / 2;
  std::nth_element(idx.begin(), mid, idx.end(), [&](int64_t a, int64_t b) { return CriticalStepSize(p, a) < CriticalStepSize(p, b); });
  return CriticalStepSize(p, *mid);
}
inline double EasyMedian(std::vector<double> v) {
  ORACLE_CHECK(!v.empty());
  auto mid = v.begin() + v.size() / 2;
  std::nth_element(v.begin(), mid, v.end());
  return *mid;
}

struct StepResult { double solution_step_size, objective_value; };

// Exact linear-time solve by repeated median pivoting on the critical step sizes
// (trust_region.cc:332-449 with helpers :177-309 and trust_region.h:250-337).
template <typename P> StepResult SolveTrustRegionStepSize(const P& p, double target_radius, const Sharder& sh) {
  ORACLE_CHECK(target_radius >= 0.0);
  ORACLE_CHECK(sh.All([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) if (p.NormWeight(i) <= 0.0) return false; return true; }));
  if (target_radius == 0.0) return {0.0, 0.0};
  if (sh.All([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) if (p.Objective(i) != 0.0) return false; return true; })) return {0.0, 0.0};

  const int ns = sh.NumShards();
  std::vector<std::vector<int64_t>> undecided(ns);
  double variable_radius_coefficient = sh.Sum([&](int s, int64_t b, int64_t e) {
    double coef = 0.0;
    undecided[s].clear();
    for (int64_t i = b; i < e; ++i) {
      if (std::isfinite(CriticalStepSize(p, i))) undecided[s].push_back(i);
      else coef += p.Objective(i) * p.Objective(i) / p.NormWeight(i);
    }
    return coef;
  });
  double fixed_radius_squared = 0.0;
  auto num_undecided = [&] { int64_t c = 0; for (auto& u : undecided) c += u.size(); return c; };

  while (num_undecided() > 0) {
    // Median of shard medians (trust_region.cc:177-203).
    std::vector<std::optional<double>> shard_medians(ns);
    sh.ForEach([&](int s, int64_t, int64_t) { if (!undecided[s].empty()) shard_medians[s] = EasyMedianOfCritical(p, undecided[s]); });
    std::vector<double> medians;
    for (auto& m : shard_medians) if (m.has_value()) medians.push_back(*m);
    const double threshold = EasyMedian(medians);

    const double undecided_r2 = sh.Sum([&](int s, int64_t, int64_t) {
      double sum = 0.0;
      for (int64_t i : undecided[s]) { const double d = ProjectedValue(p, i, threshold) - p.CenterPoint(i); sum += p.NormWeight(i) * (d * d); }
      return sum;
    });
    const double r2_at_threshold = undecided_r2 + fixed_radius_squared + variable_radius_coefficient * (threshold * threshold);
    if (r2_at_threshold > target_radius * target_radius) {
      variable_radius_coefficient += sh.Sum([&](int s, int64_t, int64_t) {
        double coef = 0.0;
        auto& u = undecided[s];
        for (int64_t i : u) if (CriticalStepSize(p, i) >= threshold) coef += p.Objective(i) * p.Objective(i) / p.NormWeight(i);
        u.erase(std::remove_if(u.begin(), u.end(), [&](int64_t i) { return CriticalStepSize(p, i) >= threshold; }), u.end());
        return coef;
      });
    } else {
      fixed_radius_squared += sh.Sum([&](int s, int64_t, int64_t) {
        double r2 = 0.0;
        auto& u = undecided[s];
        for (int64_t i : u) if (CriticalStepSize(p, i) <= threshold) { const double d = DistanceAtCriticalStepSize(p, i); r2 += p.NormWeight(i) * (d * d); }
        u.erase(std::remove_if(u.begin(), u.end(), [&](int64_t i) { return CriticalStepSize(p, i) <= threshold; }), u.end());
        return r2;
      });
    }
  }
  double step = 0.0;
  if (variable_radius_coefficient > 0.0) step = std::sqrt((target_radius * target_radius - fixed_radius_squared) / variable_radius_coefficient);
  else step = std::numeric_limits<double>::max();
  const double objective = sh.Sum([&](int, int64_t b, int64_t e) {
    double v = 0.0;
    for (int64_t i = b; i < e; ++i) v += p.Objective(i) * (ProjectedValue(p, i, step) - p.CenterPoint(i));
    return v;
  });
  return {step, objective};
}

template <typename P> Vec ComputeSolution(const P& p, double step, const Sharder& sh) {  // :280-292
  Vec sol(sh.NumElements());
  sh.ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) sol[i] = ProjectedValue(p, i, step); });
  return sol;
}

// Problem adaptors (trust_region.cc:75-162,480-607; trust_region.h:347-417).
struct VectorProblem {
  const Vec &obj, &lb, &ub, &center, &w;
  double Objective(int64_t i) const { return obj[i]; }
  double LowerBound(int64_t i) const { return lb[i]; }
  double UpperBound(int64_t i) const { return ub[i]; }
  double CenterPoint(int64_t i) const { return center[i]; }
  double NormWeight(int64_t i) const { return w[i]; }
};
struct DiagonalVectorProblem : VectorProblem {
  const Vec& diag;
  double ObjectiveMatrixDiagonalAt(int64_t i) const { return diag[i]; }
};
struct PrimalProblem {
  const Qp& qp; const Vec& x; const Vec& grad; double w = 1.0;
  double Objective(int64_t i) const { return grad[i]; }
  double LowerBound(int64_t i) const { return qp.vlb[i]; }
  double UpperBound(int64_t i) const { return qp.vub[i]; }
  double CenterPoint(int64_t i) const { return x[i]; }
  double NormWeight(int64_t) const { return w; }
};
struct DualProblem {
  const Qp& qp; const Vec& y; const Vec& grad; double w = 1.0;
  double Objective(int64_t i) const { return -grad[i]; }
  double LowerBound(int64_t i) const { return std::isfinite(qp.cub[i]) ? -kInf : 0.0; }
  double UpperBound(int64_t i) const { return std::isfinite(qp.clb[i]) ? kInf : 0.0; }
  double CenterPoint(int64_t i) const { return y[i]; }
  double NormWeight(int64_t) const { return w; }
};
struct JointProblem {
  const Qp& qp; const Vec &x, &y, &gx, &gy; double primal_weight; int64_t n;
  double Objective(int64_t i) const { return i < n ? gx[i] : -gy[i - n]; }
  double LowerBound(int64_t i) const { return i < n ? qp.vlb[i] : (std::isfinite(qp.cub[i - n]) ? -kInf : 0.0); }
  double UpperBound(int64_t i) const { return i < n ? qp.vub[i] : (std::isfinite(qp.clb[i - n]) ? kInf : 0.0); }
  double CenterPoint(int64_t i) const { return i < n ? x[i] : y[i - n]; }
  double NormWeight(int64_t i) const { return i < n ? 0.5 * primal_weight : 0.5 / primal_weight; }
  double ObjectiveMatrixDiagonalAt(int64_t i) const { return (qp.q.has_value() && i < n) ? (*qp.q)[i] : 0.0; }
};

// Diagonal-QP variant by bisection on the ball multiplier (trust_region.cc:615-753).
template <typename P> double ProjectedValueOfScaledDifference(const P& p, int64_t i, double scaling) {
  const double w = p.NormWeight(i);
  return std::min(std::max((-p.Objective(i) / std::sqrt(w)) / (p.ObjectiveMatrixDiagonalAt(i) / w + scaling), std::sqrt(w) * (p.LowerBound(i) - p.CenterPoint(i))),
                  std::sqrt(w) * (p.UpperBound(i) - p.CenterPoint(i)));
}
template <typename P> double NormOfDeltaProjection(const P& p, const Sharder& sh, double scaling) {
  return std::sqrt(sh.Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) { const double v = ProjectedValueOfScaledDifference(p, i, scaling); s += v * v; } return s; }));
}
template <typename P> double FindScalingFactor(const P& p, const Sharder& sh, double target_radius, double tol) {
  double lo = 0.0, hi = 1.0;
  while (NormOfDeltaProjection(p, sh, hi) >= target_radius) { lo = hi; hi *= 2; }
  while ((hi - lo) >= tol * std::max(1.0, lo)) {
    const double mid = (lo + hi) / 2.0;
    if (NormOfDeltaProjection(p, sh, mid) <= target_radius) hi = mid; else lo = mid;
  }
  return (hi + lo) / 2.0;
}
template <typename P> TrustRegionResult SolveDiagonalTrustRegionProblem(const P& p, const Sharder& sh, double target_radius, double tol) {
  ORACLE_CHECK(target_radius >= 0.0);
  TrustRegionResult r;
  r.solution.resize(sh.NumElements());
  if (target_radius == 0.0) {
    sh.ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) r.solution[i] = p.CenterPoint(i); });
    return r;
  }
  const double scaling = FindScalingFactor(p, sh, target_radius, tol);
  sh.ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t i = b; i < e; ++i) r.solution[i] = p.CenterPoint(i) + std::sqrt(1 / p.NormWeight(i)) * ProjectedValueOfScaledDifference(p, i, scaling);
  });
  r.objective_value = sh.Sum([&](int, int64_t b, int64_t e) {
    double s = 0;
    for (int64_t i = b; i < e; ++i) { const double d = r.solution[i] - p.CenterPoint(i); s += 0.5 * d * p.ObjectiveMatrixDiagonalAt(i) * d + d * p.Objective(i); }
    return s;
  });
  r.solution_step_size = scaling;
  return r;
}

}  // namespace tr

// trust_region.cc:453-471.
inline TrustRegionResult SolveTrustRegion(const Vec& obj, const Vec& lb, const Vec& ub, const Vec& center, const Vec& w, double radius, const Sharder& sh) {
  tr::VectorProblem p{obj, lb, ub, center, w};
  const tr::StepResult s = tr::SolveTrustRegionStepSize(p, radius, sh);
  return {s.solution_step_size, s.objective_value, tr::ComputeSolution(p, s.solution_step_size, sh)};
}
// trust_region.cc:755-766.
inline TrustRegionResult SolveDiagonalTrustRegion(const Vec& obj, const Vec& diag, const Vec& lb, const Vec& ub, const Vec& center, const Vec& w, double radius,
                                                  const Sharder& sh, double tol) {
  tr::DiagonalVectorProblem p{{obj, lb, ub, center, w}, diag};
  return tr::SolveDiagonalTrustRegionProblem(p, sh, radius, tol);
}

// trust_region.cc:797-884 (max norm) and :886-974 (Euclidean norm), dispatcher :978-1016.
inline LocalizedLagrangianBounds ComputeLocalizedLagrangianBounds(const ShardedQp& sq, const Vec& x, const Vec& y, PrimalDualNorm norm, double primal_weight,
                                                                  double radius, const Vec* primal_product, const Vec* dual_product,
                                                                  bool use_diagonal_qp_solver, double diagonal_qp_tol) {
  const Qp& qp = sq.qp();
  Vec kx_store, kty_store;
  if (primal_product == nullptr) { kx_store = TransposedMatVec(sq.Kt(), x, sq.KtSharder()); primal_product = &kx_store; }
  if (dual_product == nullptr) { kty_store = TransposedMatVec(qp.K, y, sq.KSharder()); dual_product = &kty_store; }
  const LagrangianPart pp = ComputePrimalGradient(sq, x, *dual_product);
  const LagrangianPart dp = ComputeDualGradient(sq, y, *primal_product);
  const double lagrangian = pp.value + dp.value;

  if (norm == PrimalDualNorm::kMaxNorm) {
    const double primal_radius = std::sqrt(2) * radius / std::sqrt(primal_weight);
    const double dual_radius = std::sqrt(2) * radius * std::sqrt(primal_weight);
    tr::PrimalProblem p1{qp, x, pp.gradient};
    tr::DualProblem p2{qp, y, dp.gradient};
    const tr::StepResult r1 = tr::SolveTrustRegionStepSize(p1, primal_radius, sq.PrimalSharder());
    const tr::StepResult r2 = tr::SolveTrustRegionStepSize(p2, dual_radius, sq.DualSharder());
    return {lagrangian, lagrangian + r1.objective_value, lagrangian + (-r2.objective_value), radius};
  }

  const int64_t n = sq.PrimalSize(), m = sq.DualSize();
  Sharder joint(sq.PrimalSharder(), n + m);
  tr::JointProblem jp{qp, x, y, pp.gradient, dp.gradient, primal_weight, n};
  Vec sol;
  if (use_diagonal_qp_solver) sol = tr::SolveDiagonalTrustRegionProblem(jp, joint, radius, diagonal_qp_tol).solution;
  else sol = tr::ComputeSolution(jp, tr::SolveTrustRegionStepSize(jp, radius, joint).solution_step_size, joint);

  double primal_delta = sq.PrimalSharder().Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t j = b; j < e; ++j) s += pp.gradient[j] * (sol[j] - x[j]); return s; });
  if (use_diagonal_qp_solver && qp.q.has_value()) {
    primal_delta += sq.PrimalSharder().Sum([&](int, int64_t b, int64_t e) {
      double s = 0; for (int64_t j = b; j < e; ++j) { const double d = sol[j] - x[j]; s += 0.5 * (*qp.q)[j] * (d * d); } return s; });
  }
  const double dual_delta = sq.DualSharder().Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) s += dp.gradient[i] * (sol[n + i] - y[i]); return s; });
  return {lagrangian, lagrangian + primal_delta, lagrangian + dual_delta, radius};
}

}  // namespace oracle
#endif  // PDLP_ORACLE_TR_H_
