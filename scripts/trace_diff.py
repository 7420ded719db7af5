"""Debug aid: per-iteration trace of the device solver against the oracle."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import pdlp_b200, fixtures, oracle_binding as ob

P = pdlp_b200.PrimalDualHybridGradientParams
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 2
limit = int(sys.argv[2]) if len(sys.argv) > 2 else 3500
which = sys.argv[3] if len(sys.argv) > 3 else "random"
p = P()
p.termination_criteria.simple_optimality_criteria.eps_optimal_absolute = 1e-6
p.termination_criteria.simple_optimality_criteria.eps_optimal_relative = 1e-6
p.termination_criteria.iteration_limit = limit
p.record_iteration_stats = True
qp = fixtures.random_lp(400, 800, 5, seed) if which == "random" else getattr(fixtures, which)()
dev = pdlp_b200.primal_dual_hybrid_gradient(qp, p)
ora = ob.solve(qp, p)
ds, os_ = dev.solve_log.iteration_stats, ora.solve_log_dict["iteration_stats"]
print("dev", dev.solve_log.termination_reason, dev.solve_log.iteration_count, "ora", ora.solve_log_dict["termination_reason"], ora.solve_log_dict["iteration_count"])
shown = 0
for d, o in zip(ds, os_):
    rel = abs(d.step_size - o["step_size"]) / o["step_size"]
    relw = abs(d.primal_weight - o["primal_weight"]) / o["primal_weight"]
    flag = d.restart_used != o["restart_used"] or d.cumulative_rejected_steps != o["cumulative_rejected_steps"]
    if flag or rel > 1e-7 or relw > 1e-7 or d.restart_used != 1:
        print(d.iteration_number, "restart", d.restart_used, o["restart_used"], "rej", d.cumulative_rejected_steps, o["cumulative_rejected_steps"],
              "step rel", f"{rel:.2e}", "w rel", f"{relw:.2e}", "w", d.primal_weight)
        shown += 1
        if shown > 40: break
