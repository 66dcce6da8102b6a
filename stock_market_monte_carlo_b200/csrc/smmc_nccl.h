// smmc_nccl.h -- NCCL, bound at run time.
//
// libsmmc.so does not link libnccl: the library is dlopen'ed on first use ("libnccl.so.2", then "libnccl.so"; inside a
// process that already loaded NCCL -- e.g. PyTorch -- that resolves to the same loaded library).  When it cannot be
// found the multi-GPU collective entry points fail with SMMC_ERR_UNSUPPORTED; there is no host-side stand-in.
// Only the handful of entry points the path needs are declared (signatures of nccl.h, NCCL >= 2.4).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace smmc {

struct NcclUniqueId {
  char internal[128];  // NCCL_UNIQUE_ID_BYTES
};
typedef struct ncclComm *NcclComm;

constexpr int kNcclUint64 = 5;   // ncclUint64
constexpr int kNcclFloat64 = 8;  // ncclFloat64 / ncclDouble
constexpr int kNcclSum = 0;      // ncclSum

struct NcclApi {
  int (*GetVersion)(int *);
  int (*GetUniqueId)(NcclUniqueId *);
  int (*CommInitRank)(NcclComm *, int, NcclUniqueId, int);
  int (*CommInitAll)(NcclComm *, int, const int *);
  int (*CommDestroy)(NcclComm);
  int (*CommCount)(const NcclComm, int *);
  int (*CommUserRank)(const NcclComm, int *);
  int (*CommCuDevice)(const NcclComm, int *);
  int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t);
  int (*GroupStart)();
  int (*GroupEnd)();
  const char *(*GetErrorString)(int);
};

// nullptr when libnccl cannot be loaded (why: nccl_load_error()).  Thread-safe; loads once.
const NcclApi *nccl_api();
const char *nccl_load_error();

}  // namespace smmc
