// Device side of the PDHG trial step (overview and reference lines: pdlp_step.cu): the ops of K1 - K3
// and K5 and the one-CTA decision kernels.  Pure device code over pdlp_skeletons.cuh and StepState,
// included by pdlp_step.cu only - and, with PDLP_EMULATE, by tests/kernel_emu/ to run the step on CPU.
#ifndef PDLP_B200_STEP_KERNELS_CUH_
#define PDLP_B200_STEP_KERNELS_CUH_

#include "pdlp_skeletons.cuh"
#include "pdlp_step_state.h"

namespace pdlp_b200 {

namespace {

constexpr double kDivergentMovement = 1.0e100;  // primal_dual_hybrid_gradient.cc:553

// The caller's interrupt flag as the device sees it: one 4-byte read of host-mapped memory by one thread per
// accepted step (the host forwards the caller's flag into it while it waits for the device, Solver::PullState).
__host__ __device__ inline bool InterruptRaised(const StepState& s) {
  return s.interrupt_flag != nullptr && *reinterpret_cast<const volatile int*>(s.interrupt_flag) != 0;
}
constexpr int kMaxInnerTrials = 60;             // :2231, :2299

struct OneSum {
  static constexpr int kNumSlots = 1;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
};

// K1.  ComputeNextPrimalSolution (:1659-1705) + the extrapolation of :1716-1722.
struct PrimalStepOp : OneSum {
  const StepState* st;
  double* x0; double* x1;            // the two x buffers
  const double* kty0; const double* kty1;
  const double* c; const double* q;  // q null for an LP
  const double* lb; const double* ub;
  double* xbar;
  int write_xbar;                    // Malitsky-Pock extrapolates per dual trial instead
  double* avg_x;                     // deferred averaging (null: K5 does it)
  // resolved by Begin()
  const double* x; double* xn; const double* kty; double tau;
  int avg_pending; double avg_ratio;
  __device__ bool Begin() {
    if (st->active == 0) return false;
    const int cur = st->cur;
    avg_pending = avg_x != nullptr ? st->avg_pending : 0;
    avg_ratio = st->avg_ratio;
    tau = (st->mp_tau > 0.0) ? st->mp_tau : st->step_size / st->primal_weight;
    x = cur ? x1 : x0;
    xn = cur ? x0 : x1;
    kty = cur ? kty1 : kty0;
    return true;
  }
  __device__ void Apply(int64_t j, Slots<1>& acc) const {
    // Every load first, every store last: the stores may alias the loads as far as the compiler
    // knows, and a store between them would split the element into two dependent memory round trips.
    const double xj = x[j];
    const double cj = c[j], kj = kty[j], uj = ub[j], lj = lb[j];
    const double qj = q != nullptr ? q[j] : 0.0;
    const double aj = avg_pending ? avg_x[j] : 0.0;
    double t = xj - tau * (cj - kj);
    if (q != nullptr) t = t / (tau * qj + 1.0);
    const double v = fmax(fmin(t, uj), lj);
    const double d = v - xj;
    xn[j] = v;
    if (write_xbar) xbar[j] = v + 1.0 * d;
    // ShardedWeightedAverage::Add of the iterate accepted by the previous trial (sou.cc:54-66)
    if (avg_pending) avg_x[j] = aj + avg_ratio * (xj - aj);
    acc.v[0] += d * d;
  }
};

// Malitsky-Pock: xbar = x' + theta (x' - x) for the current dual trial (:1716-1722).
struct ExtrapolateOp : NoSlots {
  const StepState* st;
  const double* x0; const double* x1; double* xbar;
  const double* x; const double* xn; double theta;
  __device__ bool Begin() {
    if (st->active == 0) return false;
    const int cur = st->cur;
    x = cur ? x1 : x0;
    xn = cur ? x0 : x1;
    theta = st->mp_new_tau / st->mp_tau;
    return true;
  }
  __device__ void Apply(int64_t j, Slots<0>&) const {
    const double v = xn[j];
    xbar[j] = v + theta * (v - x[j]);
  }
};

// K2.  ComputeNextDualSolution (:1726-1746) on the CSR copy.
struct DualStepOp : OneSum {
  const StepState* st;
  const double* val; const int* idx; const double* xbar;
  double* y0; double* y1;
  const double* clb; const double* cub;
  int skip_below;  // rows below it are shared with other ranks and counted on rank 0 only
  double* avg_y;   // deferred averaging (null: K5 does it)
  const double* y; double* yn; double sigma;
  int avg_pending; double avg_ratio;
  __device__ bool Begin() {
    if (st->active == 0) return false;
    const int cur = st->cur;
    avg_pending = avg_y != nullptr ? st->avg_pending : 0;
    avg_ratio = st->avg_ratio_dual;
    sigma = (st->mp_tau > 0.0) ? (st->primal_weight * st->primal_weight) * st->mp_new_tau : st->step_size * st->primal_weight;
    y = cur ? y1 : y0;
    yn = cur ? y0 : y1;
    return true;
  }
  __device__ const int* IndexAddress(int64_t k) const { return idx + k; }
  __device__ const double* ValueAddress(int64_t k) const { return val + k; }
  __device__ double Gather(int j) const { return xbar[j]; }
  __device__ double Product(double v, double g) const { return v * g; }
  __device__ double Load(int64_t k) const { return val[k] * xbar[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  struct Pre { double y, lb, ub, avg; };
  __device__ Pre RowPrefetch(int i) const { return Pre{y[i], clb[i], cub[i], avg_pending ? avg_y[i] : 0.0}; }
  __device__ void RowFinish(int i, double kx, const Pre& pre, Slots<1>& red) const {
    const double yi = pre.y;
    if (avg_pending) avg_y[i] = pre.avg + avg_ratio * (yi - pre.avg);
    const double temp = yi - sigma * kx;
    const double v = fmax(fmin(0.0, temp + sigma * pre.ub), temp + sigma * pre.lb);
    yn[i] = v;
    const double d = v - yi;
    if (i >= skip_below) red.v[0] += d * d;
  }
};

// K3.  K^T y' on the CSC copy (:2323-2325) fused with ComputeNonlinearity (:1761-1771)
// and, for Malitsky-Pock, |K^T y' - K^T y|^2 (:2247-2250).
struct KtyStepOp {
  static constexpr int kNumSlots = 2;
  __host__ __device__ static constexpr int SlotKindOf(int) { return kSum; }
  const StepState* st;
  const double* val; const int* idx;
  const double* y0; const double* y1;
  const double* x0; const double* x1;
  double* kty0; double* kty1;
  const double* yn; const double* x; const double* xn; const double* kty; double* ktyn;
  __device__ bool Begin() {
    if (st->active == 0) return false;
    const int cur = st->cur;
    yn = cur ? y0 : y1;
    x = cur ? x1 : x0;
    xn = cur ? x0 : x1;
    kty = cur ? kty1 : kty0;
    ktyn = cur ? kty0 : kty1;
    return true;
  }
  __device__ const int* IndexAddress(int64_t k) const { return idx + k; }
  __device__ const double* ValueAddress(int64_t k) const { return val + k; }
  __device__ double Gather(int i) const { return yn[i]; }
  __device__ double Product(double v, double g) const { return v * g; }
  __device__ double Load(int64_t k) const { return val[k] * yn[idx[k]]; }
  __device__ double RowInit() const { return 0.0; }
  __device__ double RowAdd(double acc, double staged, int) const { return acc + staged; }
  __device__ double RowMerge(double a, double b) const { return a + b; }
  struct Pre { double kty, dx; };
  __device__ Pre RowPrefetch(int j) const { return Pre{kty[j], xn[j] - x[j]}; }
  __device__ void RowFinish(int j, double v, const Pre& pre, Slots<2>& red) const {
    ktyn[j] = v;
    const double dk = v - pre.kty;
    red.v[0] += pre.dx * dk;
    red.v[1] += dk * dk;
  }
};

// K4 for the adaptive and constant rules.  One CTA; thread 0 carries the scalar logic
// of :2305-2374 (adaptive) or :2382-2405 (constant).
// `peer` (world > 1): a column-sharded solve in peer mode - the four step-test sums of every rank sit in this
// rank's scalar slots (fold_and_push_step_partials_kernel); the kernel waits for them and adds them in rank
// order instead of reading the partial tables.  The wait comes first and is unconditional (pdlp_peer.cuh).
__global__ void __launch_bounds__(kBlock) decide_adaptive_kernel(StepState* st, const double* p1, int g1, const double* p2, int g2, const double* p3, int g3,
                                                                int constant_step, PeerView peer, unsigned long long peer_epoch) {
  __shared__ double buf[kBlock];
  __shared__ double peer_sums[4];
  if (peer.world > 1) {
    PeerCombineSumsCta(peer, peer_epoch, peer_sums, 4, false);
  }
  if (st->active == 0) {
    // avg_pending and reset_average stay as they are: the batch-end flush consumes them.
    if (threadIdx.x == 0) st->accepted_last = 0;
    return;
  }
  const double dx2 = peer.world > 1 ? peer_sums[0] : ReducePartialsSum(p1, g1, 1, 0, buf);
  const double dy2 = peer.world > 1 ? peer_sums[1] : ReducePartialsSum(p2, g2, 1, 0, buf);
  const double dot = peer.world > 1 ? peer_sums[2] : ReducePartialsSum(p3, g3, 2, 0, buf);
  if (threadIdx.x != 0) return;
  StepState s = *st;
  s.total_trials += 1;
  s.accepted_last = 0;
  s.avg_pending = 0;  // this trial's K1 / K2 applied the previous accepted iterate
  const double movement = (0.5 * s.primal_weight) * dx2 + (0.5 / s.primal_weight) * dy2;
  const double nonlinearity = -dot;
  s.last_movement = movement;
  s.last_nonlinearity = nonlinearity;
  if (movement == 0.0 || movement > kDivergentMovement) {
    // :2313-2322 / :2387-2395.  The inner loop breaks without counting this trial.
    if (movement == 0.0) { s.reset_average = 1; s.avg_weight_sum = 1.0; s.avg_weight_sum_dual = 1.0; s.avg_num_terms = 1; }
    s.force_numerical = 1;
    s.active = 0;
    if (!constant_step) s.num_rejected_steps += s.inner - 1;
    s.inner = 0;
    // The reference's `for (iterations_completed_ = 0;; ++iterations_completed_)` (:2716) counts the
    // iteration whose step forced numerical termination before the final check sees it.
    s.iterations_completed += 1;
    *st = s;
    return;
  }
  if (constant_step) {
    s.cur ^= 1;
    s.avg_ratio = s.step_size / (s.avg_weight_sum + s.step_size);
    s.avg_weight_sum += s.step_size;
    s.avg_ratio_dual = s.step_size / (s.avg_weight_sum_dual + s.step_size);
    s.avg_weight_sum_dual += s.step_size;
    s.avg_num_terms += 1;
    s.accepted_last = 1;
    s.avg_pending = 1;
    s.iterations_completed += 1;
  } else {
    const double limit = nonlinearity > 0.0 ? movement / nonlinearity : INFINITY;
    const bool accept = s.step_size <= limit;
    const double total_steps_attempted = static_cast<double>(s.num_rejected_steps + s.inner + s.iterations_completed + 1);
    if (accept) {
      s.cur ^= 1;
      s.avg_ratio = s.step_size / (s.avg_weight_sum + s.step_size);
      s.avg_weight_sum += s.step_size;
      s.avg_ratio_dual = s.step_size / (s.avg_weight_sum_dual + s.step_size);
      s.avg_weight_sum_dual += s.step_size;
      s.avg_num_terms += 1;
      s.accepted_last = 1;
      s.avg_pending = 1;
    }
    const double first = isinf(limit) ? limit : (1.0 - pow(total_steps_attempted + 1.0, -s.reduction_exponent)) * limit;
    const double second = (1.0 + pow(total_steps_attempted + 1.0, -s.growth_exponent)) * s.step_size;
    s.step_size = fmin(first, second);
    if (accept) {
      s.num_rejected_steps += s.inner;
      s.inner = 0;
      s.iterations_completed += 1;
    } else {
      s.inner += 1;
      if (s.inner >= kMaxInnerTrials) {  // :2299-2304
        s.reset_average = 1; s.avg_weight_sum = 1.0; s.avg_weight_sum_dual = 1.0; s.avg_num_terms = 1;
        s.force_numerical = 1;
        s.active = 0;
        s.num_rejected_steps += s.inner - 1;
        s.inner = 0;
        s.iterations_completed += 1;  // counted by the reference's for loop (:2716)
      }
    }
  }
  if (s.accepted_last) {
    if (s.iterations_completed >= s.target_iteration) s.active = 0;
    const double passes = s.iterations_completed + s.passes_per_rejected_step * s.num_rejected_steps;
    if (passes >= s.kkt_pass_limit) s.active = 0;
    if (s.active != 0 && InterruptRaised(s)) { s.active = 0; s.stopped_by_interrupt = 1; }
  }
  *st = s;
}

// K5.  ShardedWeightedAverage::Add for x and y in one launch over n + m elements
// (sou.cc:54-66), plus ResetAverageToCurrent (:2192-2197) when requested.
struct AverageOp : NoSlots {
  const StepState* st;
  const double* x0; const double* x1; const double* y0; const double* y1;
  double* avg_x; double* avg_y; int64_t n;
  int deferred;  // 1: the batch-end flush of the adaptive / constant rules (keyed on avg_pending)
  const double* xs; const double* xold; const double* ys;
  double ratio_x, ratio_y, ratio_old_x; int reset, first_mp;
  __device__ bool Begin() {
    reset = st->reset_average;
    first_mp = st->first_mp_add_pending;
    if ((deferred ? st->avg_pending : st->accepted_last) == 0 && reset == 0) return false;
    const int cur = st->cur;
    xs = cur ? x1 : x0;
    xold = cur ? x0 : x1;
    ys = cur ? y1 : y0;
    ratio_x = st->avg_ratio;
    ratio_y = st->avg_ratio_dual;
    ratio_old_x = st->avg_ratio_old_x;
    return true;
  }
  __device__ void Apply(int64_t i, Slots<0>&) const {
    if (i < n) {
      if (reset) { avg_x[i] = xs[i]; return; }
      double a = avg_x[i];
      // Malitsky-Pock first accepted step (:2260-2264): the old x enters the empty
      // average first; after the flip it sits in the other buffer.
      if (first_mp) a += ratio_old_x * (xold[i] - a);
      a += ratio_x * (xs[i] - a);
      avg_x[i] = a;
    } else {
      const int64_t k = i - n;
      if (reset) { avg_y[k] = ys[k]; return; }
      const double a = avg_y[k];
      avg_y[k] = a + ratio_y * (ys[k] - a);
    }
  }
};

// After the batch-end flush: nothing is pending any more.
__global__ void finish_average_kernel(StepState* st) {
  st->avg_pending = 0;
  st->reset_average = 0;
}

// K4 for Malitsky-Pock (:2237-2291).
__global__ void __launch_bounds__(kBlock) decide_mp_kernel(StepState* st, const double* p1, int g1, const double* p2, int g2, const double* p3, int g3, PeerView peer,
                                                          unsigned long long peer_epoch) {
  __shared__ double buf[kBlock];
  __shared__ double peer_sums[4];
  if (peer.world > 1) {
    PeerCombineSumsCta(peer, peer_epoch, peer_sums, 4, false);
  }
  if (st->active == 0) {
    if (threadIdx.x == 0) { st->accepted_last = 0; st->reset_average = 0; st->first_mp_add_pending = 0; }
    return;
  }
  const double dx2 = peer.world > 1 ? peer_sums[0] : ReducePartialsSum(p1, g1, 1, 0, buf);
  const double dy2 = peer.world > 1 ? peer_sums[1] : ReducePartialsSum(p2, g2, 1, 0, buf);
  const double dkty2 = peer.world > 1 ? peer_sums[3] : ReducePartialsSum(p3, g3, 2, 1, buf);
  if (threadIdx.x != 0) return;
  StepState s = *st;
  s.total_trials += 1;
  s.accepted_last = 0;
  s.first_mp_add_pending = 0;
  const double new_ratio = s.mp_new_tau / s.mp_tau;
  const double dy_norm = sqrt(dy2), dkty_norm = sqrt(dkty2);
  if (s.primal_weight * s.mp_new_tau * dkty_norm <= s.mp_contraction * dy_norm) {
    s.step_size = s.mp_new_tau * s.primal_weight;
    s.ratio_last_two_step_sizes = new_ratio;
    if (!(s.avg_weight_sum > 0.0)) {
      // x_avg_.Add(old x, new_tau * ratio) into the empty average: ratio w/(0+w) = 1.
      const double w0 = s.mp_new_tau * new_ratio;
      if (w0 > 0.0) { s.first_mp_add_pending = 1; s.avg_ratio_old_x = w0 / (0.0 + w0); s.avg_weight_sum = w0; }
      // num_terms of the primal average counts this add too; the restart logic reads
      // the primal counter (:1911), so it is the one mirrored here.
      s.avg_num_terms += 1;
    }
    s.cur ^= 1;
    s.avg_ratio = s.mp_new_tau / (s.avg_weight_sum + s.mp_new_tau);
    s.avg_weight_sum += s.mp_new_tau;
    s.avg_ratio_dual = s.mp_new_tau / (s.avg_weight_sum_dual + s.mp_new_tau);
    s.avg_weight_sum_dual += s.mp_new_tau;
    s.avg_num_terms += 1;
    s.accepted_last = 1;
    const double movement = (0.5 * s.primal_weight) * dx2 + (0.5 / s.primal_weight) * dy2;
    s.last_movement = movement;
    if (movement == 0.0) { s.reset_average = 1; s.avg_weight_sum = 1.0; s.avg_weight_sum_dual = 1.0; s.avg_num_terms = 1; s.force_numerical = 1; s.active = 0; }
    else if (movement > kDivergentMovement) { s.force_numerical = 1; s.active = 0; }
    s.num_rejected_steps += s.inner;
    s.inner = 0;
    s.iterations_completed += 1;
    s.mp_tau = -1.0;  // next launch of K1 starts a new outer iteration
    if (s.iterations_completed >= s.target_iteration) s.active = 0;
    const double passes = s.iterations_completed + s.passes_per_rejected_step * s.num_rejected_steps;
    if (passes >= s.kkt_pass_limit) s.active = 0;
    if (s.active != 0 && InterruptRaised(s)) { s.active = 0; s.stopped_by_interrupt = 1; }
  } else {
    s.mp_new_tau = s.mp_downscaling * s.mp_new_tau;
    s.inner += 1;
    if (s.inner >= kMaxInnerTrials) {  // :2231-2236
      s.reset_average = 1; s.avg_weight_sum = 1.0; s.avg_weight_sum_dual = 1.0; s.avg_num_terms = 1;
      s.force_numerical = 1;
      s.active = 0;
      s.num_rejected_steps += s.inner;
      s.inner = 0;
      s.mp_tau = -1.0;
      s.iterations_completed += 1;  // counted by the reference's for loop (:2716)
    }
  }
  *st = s;
}

// Malitsky-Pock: starts an outer iteration when none is in flight (:2213-2222).
__global__ void mp_begin_kernel(StepState* st) {
  if (st->active == 0 || st->mp_tau > 0.0) return;
  const double tau = st->step_size / st->primal_weight;
  const double dilating = 1.0 + (st->mp_interpolation * (sqrt(1.0 + st->ratio_last_two_step_sizes) - 1.0));
  st->mp_tau = tau;
  st->mp_new_tau = tau * dilating;
}

// Column-sharded solve: |dx|^2 (K1), |dy|^2 over the rows this rank owns (K2) and the two K3 sums are
// per-rank partials.  They are folded to four doubles and all-reduced (the second, tiny collective of a trial step; the first is the
// row-sum all-reduce inside K2); the decision kernel then reads them as one-block "tables".  Every
// rank takes the same decision from the same numbers.
// The four step-test sums of this rank, reduced in the same fixed order the decision kernel uses.
__global__ void __launch_bounds__(kBlock) fold_step_partials_kernel(const double* p1, int g1, const double* p2, int g2, const double* p3, int g3,
                                                                   double* out) {
  __shared__ double buf[kBlock];
  const double a = ReducePartialsSum(p1, g1, 1, 0, buf);
  const double b = ReducePartialsSum(p2, g2, 1, 0, buf);
  const double c = ReducePartialsSum(p3, g3, 2, 0, buf);
  const double d = ReducePartialsSum(p3, g3, 2, 1, buf);
  if (threadIdx.x == 0) { out[0] = a; out[1] = b; out[2] = c; out[3] = d; }
}

// The same fold in peer mode: the four sums go straight into every rank's scalar slot (pdlp_peer.cuh).
__global__ void __launch_bounds__(kBlock) fold_and_push_step_partials_kernel(const double* p1, int g1, const double* p2, int g2, const double* p3, int g3,
                                                                            PeerView view, unsigned long long epoch) {
  __shared__ double buf[kBlock];
  double v[4];
  v[0] = ReducePartialsSum(p1, g1, 1, 0, buf);
  v[1] = ReducePartialsSum(p2, g2, 1, 0, buf);
  v[2] = ReducePartialsSum(p3, g3, 2, 0, buf);
  v[3] = ReducePartialsSum(p3, g3, 2, 1, buf);
  PeerPushSumsCta(view, epoch, v, 4);
}

}  // namespace

}  // namespace pdlp_b200
#endif  // PDLP_B200_STEP_KERNELS_CUH_
