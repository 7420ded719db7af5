// The device-resident state of the PDHG step loop (plain data; shared by the host driver, the step
// kernels and the CPU emulation tests).
#ifndef PDLP_B200_STEP_STATE_H_
#define PDLP_B200_STEP_STATE_H_

namespace pdlp_b200 {

// Device-side loop control and step-size state of Solver (primal_dual_hybrid_gradient.cc
// :504-709 members step_size_, primal_weight_, num_rejected_steps_, iterations_completed_,
// the averages' weight/term counters).  Lives in HBM; the host reads it back only at
// major-iteration / termination-check / restart boundaries.
struct StepState {
  int active;                // device keeps taking trial steps while 1
  int cur;                   // which of the two iterate buffers holds the current iterate
  int iterations_completed;
  int target_iteration;      // stop when iterations_completed reaches this
  int num_rejected_steps;
  int inner;                 // 0-based trial index inside the current inner loop
  int force_numerical;       // InnerStepOutcome::kForceNumericalTermination
  int reset_average;         // ResetAverageToCurrent() requested
  int accepted_last;         // last trial accepted (consumed by the averaging kernel)
  int avg_num_terms;
  int first_mp_add_pending;  // Malitsky-Pock: old x must be added to the empty average
  int avg_pending;           // adaptive / constant rule: the accepted iterate has not entered the average yet (applied by the next K1 / K2 or the batch-end flush)
  int stopped_by_interrupt;   // the decision kernel saw *interrupt_flag != 0 after an accepted step and switched the loop off
  long long total_trials;
  // Host-mapped pinned flag (termination.cc:178 polls interrupt_solve at every iteration): read by the decision
  // kernel after every accepted step.  Null when the solve has no interrupt flag or is sharded.
  const int* interrupt_flag;
  double step_size, primal_weight;
  double avg_weight_sum, avg_weight_sum_dual;  // differ only under Malitsky-Pock (:2260-2272)
  double avg_ratio, avg_ratio_dual;            // w / (W + w) of the pending average update
  double avg_ratio_old_x;    // Malitsky-Pock first-step extra term
  double ratio_last_two_step_sizes;
  double kkt_pass_limit, passes_per_rejected_step;
  double reduction_exponent, growth_exponent;
  double mp_downscaling, mp_contraction, mp_interpolation;
  double mp_tau, mp_new_tau;  // Malitsky-Pock inner-loop scalars
  double last_movement, last_nonlinearity;
};

}  // namespace pdlp_b200
#endif  // PDLP_B200_STEP_STATE_H_
