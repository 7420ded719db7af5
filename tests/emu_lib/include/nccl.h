// TEST INFRASTRUCTURE ONLY (see cuda_runtime.h in this directory): the NCCL types and enumerators pdlp_comm.h names.
// The library resolves NCCL's functions with dlopen at the first sharded call; the emulated build is used with one
// rank only, so none is ever called.
#ifndef PDLP_TESTS_EMU_NCCL_H_
#define PDLP_TESTS_EMU_NCCL_H_
#include <cstddef>
#include "cuda_runtime.h"
typedef struct ncclComm* ncclComm_t;
typedef struct ncclWindow* ncclWindow_t;
typedef enum { ncclSuccess = 0, ncclUnhandledCudaError = 1, ncclSystemError = 2, ncclInternalError = 3, ncclInvalidArgument = 4, ncclInvalidUsage = 5 } ncclResult_t;
#define NCCL_UNIQUE_ID_BYTES 128
typedef struct { char internal[NCCL_UNIQUE_ID_BYTES]; } ncclUniqueId;
typedef enum { ncclInt8 = 0, ncclChar = 0, ncclUint8 = 1, ncclInt32 = 2, ncclInt = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5, ncclFloat16 = 6, ncclHalf = 6,
               ncclFloat32 = 7, ncclFloat = 7, ncclFloat64 = 8, ncclDouble = 8 } ncclDataType_t;
typedef enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3, ncclAvg = 4 } ncclRedOp_t;
#define NCCL_WIN_COLL_SYMMETRIC 0x01
#endif  // PDLP_TESTS_EMU_NCCL_H_
