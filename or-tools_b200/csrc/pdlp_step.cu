// The PDHG trial step on the device: the replacement of Solver::TakeAdaptiveStep,
// TakeConstantSizeStep and TakeMalitskyPockStep and the helpers they call
//   ortools/pdlp/primal_dual_hybrid_gradient.cc:1659-1771 (next primal / dual, movement,
//   nonlinearity), :2211-2406 (the three step rules), sharded_optimization_utils.cc:54-66
//   (weighted average).
//
// One trial is the launches below, none of which returns to the host (ops and decision kernels:
// pdlp_step_kernels.cuh; this file holds the launch sequences):
//   K1 primal_step   x' = proj(x - tau (c - K^T y)) [/ (1 + tau Q)], xbar = 2x' - x, |dx|^2
//   K2 dual_step     CSR rows: (K xbar)_i fused with the dual prox, writes y', |dy|^2
//   K3 kty_step      CSC cols: (K^T y')_j, fused with dx_j (K^T y' - K^T y)_j
//   K4 decide        movement / nonlinearity / accept / next step size, on one CTA
//   K5 average       on accept: avg += w/(W+w) (z - avg) for x and y
// Under the adaptive and constant rules (four launches per trial) K5 is folded into the NEXT trial's K1 and K2, which read the
// accepted x and y anyway (8 (n + m) bytes and one launch less per step): the decision kernel marks
// the update pending, K1 / K2 apply it with the same arithmetic, and a flush at the end of every
// enqueued batch applies what is still pending, so the host always sees complete averages.
// All kernels read their scalars (step size, primal weight, which buffer is current)
// from the device-resident StepState and exit early once the state says stop, so the
// host can enqueue a whole major iteration's worth of trials and sync once.
#include "pdlp_launch.cuh"
#include "pdlp_step_kernels.cuh"

namespace pdlp_b200 {

namespace {

struct StepSums { const double* p1; int g1; const double* p2; int g2; const double* p3; int g3; PeerView peer; unsigned long long peer_epoch; };
StepSums CombineStepPartials(Workspace& ws, int g1, int g2, int g3) {
  if (!IsSharded(ws)) return {ws.partials[0].get(), g1, ws.partials[1].get(), g2, ws.partials[2].get(), g3, PeerView{}, 0};
  double* t = ws.exchange_sums;
  if (ws.peer) {
    // Every rank pushes its four sums to all ranks; each then adds the N contributions in rank order.
    const PeerView view = ws.comm->peer_view();
    const unsigned long long epoch = ws.comm->NextSumEpoch();
    // The decision kernel waits for the N contributions and adds them in rank order itself.
    fold_and_push_step_partials_kernel<<<1, kBlock, 0, ws.stream>>>(ws.partials[0].get(), g1, ws.partials[1].get(), g2, ws.partials[2].get(), g3, view, epoch);
    ++ws.launches;
    return {t, 1, t + 1, 1, t + 2, 1, view, epoch};
  }
  fold_step_partials_kernel<<<1, kBlock, 0, ws.stream>>>(ws.partials[0].get(), g1, ws.partials[1].get(), g2, ws.partials[2].get(), g3, t);
  ++ws.launches;
  ws.comm->AllReduce(t, 4, false, ws.stream);
  return {t, 1, t + 1, 1, t + 2, 1, PeerView{}, 0};
}

}  // namespace

void EnqueueAdaptiveTrials(Workspace& ws, const DeviceProblem& p, Iterate& it, int trials, bool constant_step) {
  StepState* st = it.state.get();
  const SparseRows& csr = p.K.by_row;
  const SparseRows& csc = p.K.by_col;
  PrimalStepOp k1{{}, st, it.x[0].get(), it.x[1].get(), it.kty[0].get(), it.kty[1].get(), p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(), it.xbar.get(), 1,
                  it.avg_x.get()};
  DualStepOp k2{{}, st, csr.val.get(), csr.idx.get(), it.xbar.get(), it.y[0].get(), it.y[1].get(), p.clb.get(), p.cub.get(), static_cast<int>(ws.dual_skip_below),
                it.avg_y.get()};
  KtyStepOp k3{st, csc.val.get(), csc.idx.get(), it.y[0].get(), it.y[1].get(), it.x[0].get(), it.x[1].get(), it.kty[0].get(), it.kty[1].get()};
  AverageOp flush{{}, st, it.x[0].get(), it.x[1].get(), it.y[0].get(), it.y[1].get(), it.avg_x.get(), it.avg_y.get(), p.n, 1};
  auto mark = [&]() {
    if (ws.profile == nullptr) return;
    cudaEvent_t e;
    PDLP_CUDA_CHECK(cudaEventCreate(&e));
    PDLP_CUDA_CHECK(cudaEventRecord(e, ws.stream));
    ws.profile->push_back(e);
  };
  for (int t = 0; t < trials; ++t) {
    if (t > 0 && ws.stop_enqueue != nullptr && *ws.stop_enqueue != 0) break;  // interrupt raised: the device stops at its next accepted step, do not queue a batch of no-ops behind it
    mark();
    const int g1 = EnqueueMap(ws, k1, p.n, 0);
    mark();
    const int g2 = EnqueueRows(ws, k2, csr, 1);
    mark();
    const int g3 = EnqueueRows(ws, k3, csc, 2);
    mark();
    const StepSums sums = CombineStepPartials(ws, g1, g2, g3);
    decide_adaptive_kernel<<<1, kBlock, 0, ws.stream>>>(st, sums.p1, sums.g1, sums.p2, sums.g2, sums.p3, sums.g3, constant_step ? 1 : 0, sums.peer, sums.peer_epoch);
    ++ws.launches;
    mark();
    mark();  // (the averaging slot of the per-kernel profile: folded into the next K1 / K2)
  }
  // The last accepted iterate (or a requested reset) enters the average here.
  if (trials > 0) {
    EnqueueMap(ws, flush, p.n + p.m, 0);
    finish_average_kernel<<<1, 1, 0, ws.stream>>>(st);
    ++ws.launches;
  }
  PDLP_CUDA_CHECK(cudaGetLastError());
}

void EnqueueMalitskyPockTrials(Workspace& ws, const DeviceProblem& p, Iterate& it, int trials) {
  StepState* st = it.state.get();
  const SparseRows& csr = p.K.by_row;
  const SparseRows& csc = p.K.by_col;
  PrimalStepOp k1{{}, st, it.x[0].get(), it.x[1].get(), it.kty[0].get(), it.kty[1].get(), p.c.get(), p.q_or_null(), p.vlb.get(), p.vub.get(), it.xbar.get(), 0, nullptr};
  ExtrapolateOp kx{{}, st, it.x[0].get(), it.x[1].get(), it.xbar.get()};
  DualStepOp k2{{}, st, csr.val.get(), csr.idx.get(), it.xbar.get(), it.y[0].get(), it.y[1].get(), p.clb.get(), p.cub.get(), static_cast<int>(ws.dual_skip_below),
                nullptr};
  KtyStepOp k3{st, csc.val.get(), csc.idx.get(), it.y[0].get(), it.y[1].get(), it.x[0].get(), it.x[1].get(), it.kty[0].get(), it.kty[1].get()};
  AverageOp k5{{}, st, it.x[0].get(), it.x[1].get(), it.y[0].get(), it.y[1].get(), it.avg_x.get(), it.avg_y.get(), p.n, 0};
  for (int t = 0; t < trials; ++t) {
    if (t > 0 && ws.stop_enqueue != nullptr && *ws.stop_enqueue != 0) break;  // see EnqueueAdaptiveTrials
    mp_begin_kernel<<<1, 1, 0, ws.stream>>>(st);
    ++ws.launches;
    // The primal step is recomputed per dual trial (same tau, same result): keeping it
    // inside the trial keeps every trial the same launch sequence; its |dx|^2 partials
    // land in region 0 and are read by the decision kernel.
    const int g1 = EnqueueMap(ws, k1, p.n, 0);
    EnqueueMap(ws, kx, p.n, 1);
    const int g2 = EnqueueRows(ws, k2, csr, 1);
    const int g3 = EnqueueRows(ws, k3, csc, 2);
    const StepSums sums = CombineStepPartials(ws, g1, g2, g3);
    decide_mp_kernel<<<1, kBlock, 0, ws.stream>>>(st, sums.p1, sums.g1, sums.p2, sums.g2, sums.p3, sums.g3, sums.peer, sums.peer_epoch);
    ++ws.launches;
    EnqueueMap(ws, k5, p.n + p.m, 0);
  }
  PDLP_CUDA_CHECK(cudaGetLastError());
}

}  // namespace pdlp_b200
