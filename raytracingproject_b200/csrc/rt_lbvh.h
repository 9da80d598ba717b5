// rt_lbvh.h -- GPU LBVH builder (Morton codes -> radix sort -> Karras hierarchy -> refit).
#pragma once

#include <cuda_runtime.h>

#include "rt_bvh_host.h"

namespace rt {

// spheres per LBVH leaf when the caller does not say (rt_bvh_opts.max_leaf = 0).  Measured on C5 (4K, 8 spp, 1 B200):
// 10 k / 100 k / 1 M spheres render in 20.8 / 24.1 / 27.7 ms with 1 sphere per leaf, 21.7 / 25.7 / 29.6 ms with <= 4 and
// 22.8 / 26.3 / 32.6 ms with <= 8 (profiles/r1_lbvh_leaf_size_sweep.txt): the leaf loop diverges more than the two levels save
#define RT_LBVH_DEFAULT_LEAF 1

// Builds the node records (rt_trace.cuh layout) on the device from device-resident boxes.
// d_nodes: n-1 records of 16 floats; d_order: slot -> box index (leaf order); leaves hold 1..max_leaf
// primitives (subtrees of that size are referred to as one leaf).  Returns the root ref and tree depth
// through the pointers (host).  mapped_word: one int of pinned host memory (device-addressable) or null -- with it the depth
// reaches the host by a store from the device instead of a device-to-host copy, which would queue on the copy engine behind
// whatever bulk copy another stream has in flight (the previous frame's read-back, rt_readback_begin).
cudaError_t build_lbvh_gpu(const HostBox* d_boxes, int n, int max_leaf, float4* d_nodes, int* d_order,
                           int* root_ref, int* depth, cudaStream_t st, int* mapped_word = nullptr);

}  // namespace rt
