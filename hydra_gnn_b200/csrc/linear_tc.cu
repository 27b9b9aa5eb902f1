// K6 tensor-core path (tcgen05 kind::tf32, 3-term compensated) -- placeholder until the kernel lands.
#include "hg_common.cuh"

namespace hg {
int linear_tc_supported(const hg_linear_args&) { return 0; }
size_t linear_tc_workspace_bytes(const hg_linear_args&) { return 0; }
int linear_tc_launch(const hg_linear_args&, void*, size_t, cudaStream_t) { return HG_EUNSUPPORTED; }
}  // namespace hg
