// abi.cu - version / error plumbing of the C ABI (include/pbhc_b200.h).
#include <cstdarg>
#include <cstdio>
#include "common.cuh"

namespace pbhc {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return PBHC_E_CUDA;
  }
  return PBHC_OK;
}
}  // namespace pbhc

extern "C" int pbhc_abi_version(void) { return PBHC_ABI_VERSION; }
extern "C" const char* pbhc_last_error(void) { return pbhc::g_err; }
extern "C" int64_t pbhc_struct_size(int32_t which) {
  switch (which) {
    case 0: return sizeof(PbhcTableDesc);
    case 1: return sizeof(PbhcLookupOut);
    case 2: return sizeof(PbhcTree);
    case 3: return sizeof(PbhcTrackCfg);
    case 4: return sizeof(PbhcTrackIO);
    case 5: return sizeof(PbhcHistDst);
    case 6: return sizeof(PbhcCurriculum);
    case 7: return sizeof(PbhcGatherKey);
    case 8: return sizeof(PbhcActuationCfg);
    default: return -1;
  }
}
