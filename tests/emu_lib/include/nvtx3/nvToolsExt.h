// TEST INFRASTRUCTURE ONLY (see ../cuda_runtime.h): NVTX ranges do nothing in the emulated build.
#ifndef PDLP_TESTS_EMU_NVTX_H_
#define PDLP_TESTS_EMU_NVTX_H_
inline int nvtxRangePushA(const char*) { return 0; }
inline int nvtxRangePop() { return 0; }
#endif
