// tcgen05 / TMEM extend kernel -- placeholder until the kernel lands: reports "unsupported" so AUTO
// routes everything to the register-MMA kernel.
#include "extend.cuh"

namespace b200attn {
bool extend_tc_supported(const ExtendParams&, int, int) { return false; }
int extend_tc_launch(const ExtendParams&, bool, cudaStream_t) {
  set_error("extend: the tcgen05 kernel is not built in this version");
  return B200ATTN_ERR_UNSUPPORTED;
}
}  // namespace b200attn
