// tcgen05 family — placeholder until the tensor-core kernel lands.
#include "rq_common.cuh"
#include "rq_kernels.h"

namespace rq {
bool tc_supported(int, int, int, int, int) { return false; }
int quantize_forward_tc(const FwdParams&, int, cudaStream_t) {
    set_error("tcgen05 family not built");
    return RQ_ERR_UNSUPPORTED_SHAPE;
}
}  // namespace rq
