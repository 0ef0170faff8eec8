// tcgen05 EdgeConv (placeholder until the tensor-core kernel lands): reports unsupported.
#include "net_layout.cuh"
namespace gm {
int edgeconv_tc_forward(const float *, const int *, const int *, int, const float *, const float *, int, int, float *,
                        int *, float *, int *, cudaStream_t) {
    return GM_ERR_UNSUPPORTED;
}
}  // namespace gm
