// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under or-tools_b200/ may include, link
// or call this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// `--impl reference` leg use it, and only as the checker / the timed CPU baseline.
//
// CPU restatement (C++17, std only) of the reference PDLP CPU path.  Parity status:
// pinned against the reference's own golden vectors (tests/test_oracle_*.py re-type
// the fixtures and expectations of ortools/pdlp/*_test.cc).  The reference itself
// cannot be compiled here (Eigen 3.4.0, abseil 20230125.3, protobuf 23.3 absent).
//
// This header: sparse CSC container, thread pool, Sharder and the sharded BLAS-1 /
// SpMV-transpose / column-norm ops.
//   follows ortools/base/threadpool.cc:19-88, ortools/pdlp/sharder.h:34-231,
//           ortools/pdlp/sharder.cc:36-337.
#ifndef PDLP_ORACLE_CORE_H_
#define PDLP_ORACLE_CORE_H_

#include <algorithm>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <limits>
#include <list>
#include <mutex>
#include <thread>
#include <vector>

namespace oracle {

using Vec = std::vector<double>;
constexpr double kInf = std::numeric_limits<double>::infinity();

#define ORACLE_CHECK(cond)                                                     \
  do {                                                                         \
    if (!(cond)) {                                                             \
      std::fprintf(stderr, "oracle CHECK failed: %s (%s:%d)\n", #cond,         \
                   __FILE__, __LINE__);                                        \
      std::abort();                                                            \
    }                                                                          \
  } while (0)

// Compressed sparse column matrix, fp64 values and int64 indices like
// Eigen::SparseMatrix<double, ColMajor, int64_t> (quadratic_program.h:130).
struct Csc {
  int64_t rows = 0, cols = 0;
  std::vector<int64_t> ptr{0};  // cols + 1
  std::vector<int64_t> idx;     // row index per nonzero
  Vec val;
  int64_t nnz() const { return static_cast<int64_t>(val.size()); }
  int64_t col_nnz(int64_t j) const { return ptr[j + 1] - ptr[j]; }

  // Explicit transpose, stable in the original column order (what
  // `qp_.constraint_matrix.transpose()` materialises, sharded_quadratic_program.cc:67).
  Csc Transposed() const {
    Csc t;
    t.rows = cols;
    t.cols = rows;
    t.ptr.assign(rows + 1, 0);
    for (int64_t k = 0; k < nnz(); ++k) ++t.ptr[idx[k] + 1];
    for (int64_t i = 0; i < rows; ++i) t.ptr[i + 1] += t.ptr[i];
    t.idx.resize(nnz());
    t.val.resize(nnz());
    std::vector<int64_t> next(t.ptr.begin(), t.ptr.end() - 1);
    for (int64_t j = 0; j < cols; ++j) {
      for (int64_t k = ptr[j]; k < ptr[j + 1]; ++k) {
        const int64_t p = next[idx[k]]++;
        t.idx[p] = j;
        t.val[p] = val[k];
      }
    }
    return t;
  }
};

// FIFO task list under one mutex, workers woken with notify_all per Schedule
// (ortools/base/threadpool.cc:53-87).
class Pool {
 public:
  explicit Pool(int num_workers) {
    for (int i = 0; i < num_workers; ++i) {
      workers_.emplace_back([this] { Run(); });
    }
  }
  ~Pool() {
    {
      std::unique_lock<std::mutex> l(mu_);
      done_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  void Schedule(std::function<void()> fn) {
    {
      std::unique_lock<std::mutex> l(mu_);
      tasks_.push_back(std::move(fn));
    }
    cv_.notify_all();
  }

 private:
  void Run() {
    for (;;) {
      std::function<void()> fn;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [this] { return done_ || !tasks_.empty(); });
        if (tasks_.empty()) return;
        fn = std::move(tasks_.front());
        tasks_.pop_front();
      }
      fn();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_;
  std::list<std::function<void()>> tasks_;
  std::vector<std::thread> workers_;
  bool done_ = false;
};

// Stand-in for absl::BlockingCounter (sharder.cc:112,134).
class Latch {
 public:
  explicit Latch(int n) : n_(n) {}
  void CountDown() {
    std::unique_lock<std::mutex> l(mu_);
    if (--n_ == 0) cv_.notify_all();
  }
  void Wait() {
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [this] { return n_ == 0; });
  }

 private:
  std::mutex mu_;
  std::condition_variable cv_;
  int n_;
};

// Contiguous index ranges with balanced mass (sharder.cc:36-107).
class Sharder {
 public:
  using MassFn = std::function<int64_t(int64_t)>;
  using RangeFn = std::function<void(int shard, int64_t begin, int64_t end)>;

  // Mass-balanced boundaries (sharder.cc:36-70).
  Sharder(int64_t num_elements, int num_shards, Pool* pool, const MassFn& mass)
      : pool_(pool) {
    ORACLE_CHECK(num_elements >= 0);
    starts_.push_back(0);
    if (num_elements == 0) return;
    ORACLE_CHECK(num_shards >= 1);
    int64_t overall = 0;
    for (int64_t e = 0; e < num_elements; ++e) overall += mass(e);
    int64_t this_mass = mass(0);
    for (int64_t e = 1; e < num_elements; ++e) {
      const int64_t em = mass(e);
      if (this_mass + (em / 2) >= overall / num_shards) {
        masses_.push_back(this_mass);
        starts_.push_back(e);
        this_mass = em;
      } else {
        this_mass += em;
      }
    }
    starts_.push_back(num_elements);
    masses_.push_back(this_mass);
  }

  // Uniform boundaries (sharder.cc:72-101).
  Sharder(int64_t num_elements, int num_shards, Pool* pool) : pool_(pool) {
    ORACLE_CHECK(num_elements >= 0);
    if (num_elements == 0) {
      starts_.push_back(0);
      return;
    }
    ORACLE_CHECK(num_shards >= 1);
    if (num_shards >= num_elements) {
      for (int64_t e = 0; e < num_elements; ++e) {
        starts_.push_back(e);
        masses_.push_back(1);
      }
    } else {
      for (int s = 0; s < num_shards; ++s) {
        const int64_t b = (num_elements * s) / num_shards;
        const int64_t e = (num_elements * (s + 1)) / num_shards;
        if (e - b > 0) {
          starts_.push_back(b);
          masses_.push_back(e - b);
        }
      }
    }
    starts_.push_back(num_elements);
  }

  // Column-mass sharder of a matrix: element mass 1 + nnz(col) (sharder.h:161-165).
  Sharder(const Csc& m, int num_shards, Pool* pool)
      : Sharder(m.cols, num_shards, pool,
                [&m](int64_t c) { return 1 + m.col_nnz(c); }) {}

  // Same shard count, different length (sharder.cc:103-107).
  Sharder(const Sharder& other, int64_t num_elements)
      : Sharder(num_elements, std::max(1, other.NumShards()), other.pool_) {}

  int NumShards() const { return static_cast<int>(starts_.size()) - 1; }
  int64_t NumElements() const { return starts_.back(); }
  int64_t ShardStart(int s) const { return starts_[s]; }
  int64_t ShardSize(int s) const { return starts_[s + 1] - starts_[s]; }
  int64_t ShardMass(int s) const { return masses_[s]; }
  const std::vector<int64_t>& Starts() const { return starts_; }

  // Fork-join over shards (sharder.cc:109-141).
  void ForEach(const RangeFn& fn) const {
    const int ns = NumShards();
    if (pool_ != nullptr) {
      Latch latch(ns);
      for (int s = 0; s < ns; ++s) {
        pool_->Schedule([&, s] {
          fn(s, starts_[s], starts_[s + 1]);
          latch.CountDown();
        });
      }
      latch.Wait();
    } else {
      for (int s = 0; s < ns; ++s) fn(s, starts_[s], starts_[s + 1]);
    }
  }
  // Per-shard partials combined serially (sharder.cc:143-150).
  double Sum(const std::function<double(int, int64_t, int64_t)>& fn) const {
    Vec local(NumShards(), 0.0);
    ForEach([&](int s, int64_t b, int64_t e) { local[s] = fn(s, b, e); });
    double total = 0.0;
    for (double v : local) total += v;
    return total;
  }
  double Max(const std::function<double(int, int64_t, int64_t)>& fn) const {
    Vec local(NumShards(), 0.0);
    ForEach([&](int s, int64_t b, int64_t e) { local[s] = fn(s, b, e); });
    double mx = 0.0;
    for (double v : local) mx = std::max(mx, std::abs(v));
    return mx;
  }
  bool All(const std::function<bool(int, int64_t, int64_t)>& fn) const {
    std::vector<int> local(NumShards(), 1);
    ForEach([&](int s, int64_t b, int64_t e) { local[s] = fn(s, b, e) ? 1 : 0; });
    return std::all_of(local.begin(), local.end(), [](int v) { return v != 0; });
  }

 private:
  std::vector<int64_t> starts_;
  std::vector<int64_t> masses_;
  Pool* pool_;
};

// out[j] = sum_k val[k] * v[row[k]] over column j, i.e. M^T v (sharder.cc:163-176).
inline Vec TransposedMatVec(const Csc& m, const Vec& v, const Sharder& sh) {
  ORACLE_CHECK(static_cast<int64_t>(v.size()) == m.rows);
  Vec out(m.cols);
  sh.ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t j = b; j < e; ++j) {
      double acc = 0.0;
      for (int64_t k = m.ptr[j]; k < m.ptr[j + 1]; ++k) acc += m.val[k] * v[m.idx[k]];
      out[j] = acc;
    }
  });
  return out;
}

// ---- sharded BLAS-1 (sharder.cc:178-289) ----
inline void SetZero(const Sharder& sh, Vec& dest) {
  dest.resize(sh.NumElements());
  sh.ForEach([&](int, int64_t b, int64_t e) { std::fill(dest.begin() + b, dest.begin() + e, 0.0); });
}
inline Vec ZeroVector(const Sharder& sh) { Vec r; SetZero(sh, r); return r; }
inline Vec OnesVector(const Sharder& sh) {
  Vec r(sh.NumElements());
  sh.ForEach([&](int, int64_t b, int64_t e) { std::fill(r.begin() + b, r.begin() + e, 1.0); });
  return r;
}
inline void AddScaledVector(double scale, const Vec& inc, const Sharder& sh, Vec& dest) {
  sh.ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) dest[i] += scale * inc[i]; });
}
inline void AssignVector(const Vec& v, const Sharder& sh, Vec& dest) {
  dest.resize(v.size());
  sh.ForEach([&](int, int64_t b, int64_t e) { std::copy(v.begin() + b, v.begin() + e, dest.begin() + b); });
}
inline Vec CloneVector(const Vec& v, const Sharder& sh) { Vec d; AssignVector(v, sh, d); return d; }
inline void CwiseProductInPlace(const Vec& scale, const Sharder& sh, Vec& dest) {
  sh.ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) dest[i] = dest[i] * scale[i]; });
}
inline void CwiseQuotientInPlace(const Vec& scale, const Sharder& sh, Vec& dest) {
  sh.ForEach([&](int, int64_t b, int64_t e) { for (int64_t i = b; i < e; ++i) dest[i] = dest[i] / scale[i]; });
}
inline double Dot(const Vec& a, const Vec& c, const Sharder& sh) {
  return sh.Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) s += a[i] * c[i]; return s; });
}
inline double LInfNorm(const Vec& v, const Sharder& sh) {
  return sh.Max([&](int, int64_t b, int64_t e) { double m = 0; for (int64_t i = b; i < e; ++i) m = std::max(m, std::abs(v[i])); return m; });
}
inline double L1Norm(const Vec& v, const Sharder& sh) {
  return sh.Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) s += std::abs(v[i]); return s; });
}
inline double SquaredNorm(const Vec& v, const Sharder& sh) {
  return sh.Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) s += v[i] * v[i]; return s; });
}
inline double Norm(const Vec& v, const Sharder& sh) { return std::sqrt(SquaredNorm(v, sh)); }
inline double SquaredDistance(const Vec& a, const Vec& c, const Sharder& sh) {
  return sh.Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) { const double d = a[i] - c[i]; s += d * d; } return s; });
}
inline double Distance(const Vec& a, const Vec& c, const Sharder& sh) { return std::sqrt(SquaredDistance(a, c, sh)); }
inline double ScaledLInfNorm(const Vec& v, const Vec& scale, const Sharder& sh) {
  return sh.Max([&](int, int64_t b, int64_t e) { double m = 0; for (int64_t i = b; i < e; ++i) m = std::max(m, std::abs(v[i] * scale[i])); return m; });
}
inline double ScaledSquaredNorm(const Vec& v, const Vec& scale, const Sharder& sh) {
  return sh.Sum([&](int, int64_t b, int64_t e) { double s = 0; for (int64_t i = b; i < e; ++i) { const double p = v[i] * scale[i]; s += p * p; } return s; });
}
inline double ScaledNorm(const Vec& v, const Vec& scale, const Sharder& sh) { return std::sqrt(ScaledSquaredNorm(v, scale, sh)); }

// Column norms of diag(row_scale) * M * diag(col_scale) (sharder.cc:291-335).
inline Vec ScaledColLInfNorm(const Csc& m, const Vec& row_scale, const Vec& col_scale, const Sharder& sh) {
  Vec out(m.cols);
  sh.ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t j = b; j < e; ++j) {
      double mx = 0.0;
      for (int64_t k = m.ptr[j]; k < m.ptr[j + 1]; ++k) mx = std::max(mx, std::abs(m.val[k] * row_scale[m.idx[k]]));
      out[j] = mx * std::abs(col_scale[j]);
    }
  });
  return out;
}
inline Vec ScaledColL2Norm(const Csc& m, const Vec& row_scale, const Vec& col_scale, const Sharder& sh) {
  Vec out(m.cols);
  sh.ForEach([&](int, int64_t b, int64_t e) {
    for (int64_t j = b; j < e; ++j) {
      double ss = 0.0;
      for (int64_t k = m.ptr[j]; k < m.ptr[j + 1]; ++k) { const double t = m.val[k] * row_scale[m.idx[k]]; ss += t * t; }
      out[j] = std::sqrt(ss) * std::abs(col_scale[j]);
    }
  });
  return out;
}

}  // namespace oracle
#endif  // PDLP_ORACLE_CORE_H_
