// rt_lbvh.h -- GPU LBVH builder (Morton codes -> radix sort -> Karras hierarchy -> refit).
#pragma once

#include <cuda_runtime.h>

#include "rt_bvh_host.h"

namespace rt {

// Builds the node records (rt_trace.cuh layout) on the device from device-resident boxes.
// d_nodes: n-1 records of 16 floats; d_order: slot -> box index (leaf order); leaves hold one
// primitive.  Returns the root ref and tree depth through the pointers (host).
cudaError_t build_lbvh_gpu(const HostBox* d_boxes, int n, float4* d_nodes, int* d_order,
                           int* root_ref, int* depth, cudaStream_t st);

}  // namespace rt
