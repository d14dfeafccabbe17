// Instantiations of the weight-streaming grouped GEMM for one weight format (see stream_gemm.cuh).
#include "stream_gemm.cuh"
namespace sglb {
SGLB_INSTANTIATE_STREAM(SGLB_FMT_INT8_W8A8)
}
