// rt_bvh_host.h -- host-side BVH builder (SAH).  The reference's bvh.h has no builder at all
// (src/bvh.h:12-14 is "To be implemented later"); only its traversal contract (:16-24) exists.
#pragma once

#include <cstdint>
#include <vector>

namespace rt {

struct HostBox { float lo[3], hi[3]; };

struct HostBVH {
    std::vector<float> nodes;        // 16 floats per inner node, layout in rt_trace.cuh, BFS order
    std::vector<int>   order;        // slot -> index into the input box array (leaf order)
    int root_ref = (int)0x80000000;  // RT_SENTINEL when empty
    int depth = 0;
    int n_inner = 0;
};

// Conservative float box of a (possibly moving) sphere given in double: union of the boxes at
// time 0 and 1 (src/sphere.h:12-13, :22-25), padded by 2^-20 of the coordinate magnitude and
// rounded outwards.
HostBox sphere_box(const double c1[3], const double cv[3], double radius, bool moving);

// Binary SAH build over `boxes`; leaves hold 1..max_leaf primitives.
void build_bvh_sah(const std::vector<HostBox>& boxes, int max_leaf, HostBVH& out);

// Flatten an explicit binary tree (children arrays) into the node records.  Used by the GPU LBVH
// path too: left/right >= 0 inner index, < 0 leaf ~(first<<3 | count-1) over `order`.
int bvh_depth(const HostBVH& b);

}  // namespace rt
