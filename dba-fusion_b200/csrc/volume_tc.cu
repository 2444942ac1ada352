// placeholder until the tcgen05 kernel lands
#include "common.cuh"
namespace dbaf {
size_t corr_volume_tc_workspace(int, int, int, int) { return 256; }
int corr_volume_tc(const __half*, const __half*, int, int, int, int, int, __half* const*, void*, size_t,
                   cudaStream_t, bool* handled) {
  *handled = false;
  return DBAF_OK;
}
}  // namespace dbaf
