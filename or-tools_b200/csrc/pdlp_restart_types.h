// Result of the localized Lagrangian bounds (trust_region.h:32-56 LocalizedLagrangianBounds).
#ifndef PDLP_B200_RESTART_TYPES_H_
#define PDLP_B200_RESTART_TYPES_H_

namespace pdlp_b200 {

struct LocalizedBounds { double lagrangian_value, lower_bound, upper_bound, radius; };

}  // namespace pdlp_b200
#endif  // PDLP_B200_RESTART_TYPES_H_
