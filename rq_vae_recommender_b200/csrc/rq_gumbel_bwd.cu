// Gumbel-softmax backward — placeholder.
#include "rq_common.cuh"
#include "rq_kernels.h"

namespace rq {
int quantize_backward_gumbel(const BwdParams&, int, float, const float*, uint64_t, uint64_t, cudaStream_t) {
    set_error("GUMBEL_SOFTMAX backward not built yet");
    return RQ_ERR_UNSUPPORTED_SHAPE;
}
}  // namespace rq
