// rt_scene.h -- device-side scene preparation for large worlds (GPU LBVH path of rt_upload_scene):
// sphere boxes and the tagged SoA slots in leaf order, made from the caller's AoS arrays on the device.
#pragma once

#include <cuda_runtime.h>

#include "../../include/rt_b200.h"
#include "rt_bvh_host.h"

namespace rt {

// boxes[k] = sphere_box(spheres[ids[k]]) -- bit-identical to the host function (rt_bvh_host.cpp)
cudaError_t launch_scene_boxes(const rt_sphere* spheres, const int* ids, int n, HostBox* boxes, cudaStream_t st);

// slot s holds sphere ids[order[s]]: a = (center1, r), b = (center_vec, r^2), c = (albedo, param), in double and
// float, meta = list_index << 3 | moving << 2 | material tag -- the same values the host fill of rt_upload_scene writes
cudaError_t launch_scene_fill(const rt_sphere* spheres, const rt_material* materials, const int* ids, const int* order, int n,
                              float4* a32, float4* b32, float4* c32, double4* a64, double4* b64, double4* c64, int* meta,
                              cudaStream_t st, int* slot_sphere = nullptr);
// (order == nullptr: slot s holds sphere ids[s]; slot_sphere, optional: receives the sphere index of every slot)

// Refit (dynamic scenes): every box of the tree recomputed bottom-up from new per-slot boxes, topology unchanged.
// parent and flags: n_nodes ints of scratch each.
cudaError_t launch_tree_refit(float4* nodes, int n_nodes, const HostBox* slot_boxes, int* parent, int* flags, cudaStream_t st);

// Renumbers the inner nodes so that the first min(k, n_nodes, RT_TOP_MAX) of them in breadth-first order from the
// root become indices 0, 1, 2, ... (the root becomes node 0): nodes_out = the permuted records with renamed child refs.
// newpos (n_nodes ints) and top_ids (RT_TOP_MAX ints) are scratch.  Nothing is done when the root is a leaf.
#define RT_TOP_MAX 1024
cudaError_t launch_top_first(const float4* nodes, int n_nodes, int root_ref, int k, int* newpos, int* top_ids, float4* nodes_out, cudaStream_t st);

// 4-wide records (layout: rt_trace.cuh) of a binary tree: record i = the grandchildren of binary node i
cudaError_t launch_widen(const float4* nodes2, int n_nodes, float4* nodes4, cudaStream_t st);

}  // namespace rt
