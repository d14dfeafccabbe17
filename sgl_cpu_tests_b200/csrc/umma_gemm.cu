// K3/K4 (prefill regime): tcgen05 / TMEM / TMA grouped GEMM.  (placeholder until the tensor-core
// path lands; AUTO dispatch never selects it while umma_supported() returns 0)
#include "common.cuh"
namespace sglb {
struct MoeWorkspace;
int umma_supported(int, int64_t, int, int, int, int) { return 0; }
int launch_umma_moe(const sglb_moe_problem*, const MoeWorkspace&, cudaStream_t) {
  set_error("tcgen05 path not built");
  return SGLB_ERR_INVALID;
}
}  // namespace sglb
