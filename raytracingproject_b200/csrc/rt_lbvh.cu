// rt_lbvh.cu -- GPU LBVH builder (placeholder body until the Morton/Karras kernels land).
#include "rt_lbvh.h"

namespace rt {

cudaError_t build_lbvh_gpu(const HostBox*, int, float4*, int*, int*, int*, cudaStream_t) {
    return cudaErrorNotSupported;
}

}  // namespace rt
