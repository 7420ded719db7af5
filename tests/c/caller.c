/* A plain C (C99) caller of include/pdlp_b200.h: what a cgo / JNI / FFI binding sees.  Built with gcc (not g++) by
 * tests/test_cabi.py: the header must compile as C, and the boundary must be usable with nothing but pointers and sizes.
 *   caller cpu   host-only entry points (no device): defaults, validation, statuses, work blocks, error text
 *   caller gpu   solves TestLp (ortools/pdlp/test_util.cc:35-60: optimum -34 at x = [-1, 8, 1, 2.5]) with default params */
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "pdlp_b200.h"

static int callbacks = 0;
static void count_callback(int iteration_type, const uint8_t* stats, size_t stats_len, const double* bound_norms, void* user) {
  (void)iteration_type; (void)stats; (void)stats_len; (void)bound_norms;
  ++*(int*)user;
}

int main(int argc, char** argv) {
  const double inf = INFINITY;
  const int64_t col_ptr[5] = {0, 3, 4, 7, 9};
  const int64_t row_idx[9] = {0, 1, 2, 0, 0, 1, 3, 0, 3};
  const double values[9] = {2, 1, 4, 1, 1, 1, 1.5, 2, -1};
  const double c[4] = {5.5, -2, -1, 1};
  const double vlb[4] = {-inf, -2, -inf, 2.5}, vub[4] = {inf, inf, 6, 3.5};
  const double clb[4] = {12, -inf, -4, -1}, cub[4] = {12, 7, inf, 1};
  pdlp_b200_problem lp;
  memset(&lp, 0, sizeof lp);
  lp.num_variables = 4; lp.num_constraints = 4; lp.num_nonzeros = 9;
  lp.col_ptr = col_ptr; lp.row_idx = row_idx; lp.values = values;
  lp.objective_vector = c;
  lp.variable_lower_bounds = vlb; lp.variable_upper_bounds = vub;
  lp.constraint_lower_bounds = clb; lp.constraint_upper_bounds = cub;
  lp.objective_offset = -14; lp.objective_scaling_factor = 1;

  if (argc > 1 && strcmp(argv[1], "gpu") == 0) {
    double x[4], y[4], rc[4];
    int has_solution = 0;
    uint8_t* log = NULL;
    size_t log_len = 0;
    pdlp_b200_counters counters;
    const pdlp_b200_status st = pdlp_b200_solve(&lp, NULL, 0, NULL, NULL, NULL, count_callback, &callbacks, -1, x, y, rc, &has_solution, &log, &log_len, &counters);
    if (st != PDLP_B200_OK) { fprintf(stderr, "solve: %s\n", pdlp_b200_last_error()); return 10; }
    if (!has_solution || log == NULL || log_len == 0) return 11;
    const double want[4] = {-1, 8, 1, 2.5};
    double objective = -14;
    for (int j = 0; j < 4; ++j) {
      if (fabs(x[j] - want[j]) > 1e-4) return 12;
      objective += c[j] * x[j];
    }
    if (fabs(objective + 34.0) > 1e-4 || callbacks < 2 || counters.kernel_launches <= 0) return 13;
    pdlp_b200_free(log);
    printf("gpu ok: objective %.6f, %d callbacks, %lld kernel launches, %zu log bytes\n", objective, callbacks, counters.kernel_launches, log_len);
    return 0;
  }

  /* defaults round-trip through the validator */
  uint8_t* params = NULL;
  size_t params_len = 0;
  char* message = NULL;
  if (pdlp_b200_default_params(&params, &params_len) != PDLP_B200_OK || params == NULL) return 1; /* no field set: length 0 */
  message = (char*)"unset";
  if (pdlp_b200_validate_params(params, params_len, &message) != PDLP_B200_OK || message != NULL) return 2; /* valid: no text */
  pdlp_b200_free(params);
  /* bytes that are not a PrimalDualHybridGradientParams message: an ABI-level error with a text */
  const uint8_t junk[3] = {0xff, 0xff, 0xff};
  message = NULL;
  if (pdlp_b200_validate_params(junk, sizeof junk, &message) != PDLP_B200_BAD_ARGUMENT || strlen(pdlp_b200_last_error()) == 0) return 3;
  /* internal::ComputeStatuses at the optimum (primal_dual_hybrid_gradient_test.cc:2076-2092) */
  const double x_opt[4] = {-1, 8, 1, 2.5}, y_opt[4] = {-2, 0, 2.375, 2.0 / 3};
  int8_t vs[4], cs[4];
  if (pdlp_b200_compute_statuses(&lp, x_opt, y_opt, vs, cs) != PDLP_B200_OK) return 4;
  const int8_t want_vs[4] = {0, 0, 0, 2}, want_cs[4] = {1, 0, 2, 2}; /* BASIC x3, AT_LOWER_BOUND; FIXED_VALUE, BASIC, AT_LOWER_BOUND x2 */
  if (memcmp(vs, want_vs, 4) != 0 || memcmp(cs, want_cs, 4) != 0) return 5;
  /* the work-block partition of the column offsets: one block holds all four columns */
  int32_t block_start[5], cut_block = 0, max_row_nnz = 0;
  int64_t num_blocks = 0;
  if (pdlp_b200_work_blocks(col_ptr, 4, -1, block_start, &num_blocks, &cut_block, &max_row_nnz) != PDLP_B200_OK) return 6;
  if (num_blocks != 1 || block_start[0] != 0 || block_start[1] != 4 || max_row_nnz != 3 || cut_block != -1) return 7;
  /* a NULL problem is a bad argument, not a crash */
  if (pdlp_b200_compute_statuses(NULL, x_opt, y_opt, vs, cs) != PDLP_B200_BAD_ARGUMENT) return 8;
  printf("cpu ok\n");
  return 0;
}
