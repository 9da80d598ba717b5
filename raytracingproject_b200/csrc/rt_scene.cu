// rt_scene.cu -- the world on the device without a per-sphere host loop (C5: 10 k .. 1 M spheres).
// The hittable_list of shared_ptr<sphere> (src/hittable_list.h:13, src/sphere.h:60-66, src/material.h:27-28,43-45,73-74)
// arrives as two AoS arrays; one kernel makes the conservative float boxes the LBVH is built over
// (src/sphere.h:12-13, :22-25: box at time 0 united with the box at time 1), one gathers the spheres into the
// SoA slots in leaf order once the tree exists.  Built with --fmad=false: same values as the host path.
#include "rt_scene.h"

namespace rt {

namespace {

__global__ void __launch_bounds__(256) scene_boxes_kernel(const rt_sphere* __restrict__ spheres, const int* __restrict__ ids, int n,
                                                          HostBox* __restrict__ boxes) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const rt_sphere s = spheres[ids[k]];
    const double r = fabs(s.radius);
    HostBox b;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        double lo = s.center1[a] - r, hi = s.center1[a] + r;
        if (s.is_moving) { lo = fmin(lo, s.center1[a] + s.center_vec[a] - r); hi = fmax(hi, s.center1[a] + s.center_vec[a] + r); }
        const double pad = fmax(fmax(fabs(lo), fabs(hi)), r) * 9.5367431640625e-07 + 1e-30;      // ldexp(.., -20): exact
        b.lo[a] = __double2float_rd(lo - pad);                                                   // rounded outwards
        b.hi[a] = __double2float_ru(hi + pad);
    }
    boxes[k] = b;
}

__global__ void __launch_bounds__(256) scene_fill_kernel(const rt_sphere* __restrict__ spheres, const rt_material* __restrict__ materials,
                                                         const int* __restrict__ ids, const int* __restrict__ order, int n,
                                                         float4* __restrict__ a32, float4* __restrict__ b32, float4* __restrict__ c32,
                                                         double4* __restrict__ a64, double4* __restrict__ b64, double4* __restrict__ c64,
                                                         int* __restrict__ meta, int* __restrict__ slot_sphere) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    const int id = ids[order ? order[slot] : slot];
    if (slot_sphere) slot_sphere[slot] = id;
    const rt_sphere s = spheres[id];
    const rt_material mt = materials[s.material];
    const bool mov = s.is_moving != 0;
    const double4 a = make_double4(s.center1[0], s.center1[1], s.center1[2], s.radius);
    const double4 b = make_double4(mov ? s.center_vec[0] : 0.0, mov ? s.center_vec[1] : 0.0, mov ? s.center_vec[2] : 0.0, s.radius * s.radius);
    double param = mt.param;
    if (mt.type == RT_METAL) param = param < 1 ? param : 1;                                     // src/material.h:33
    const double4 c = make_double4(mt.albedo[0], mt.albedo[1], mt.albedo[2], param);
    a64[slot] = a; b64[slot] = b; c64[slot] = c;
    const float rf = float(a.w);
    a32[slot] = make_float4(float(a.x), float(a.y), float(a.z), rf);
    b32[slot] = make_float4(float(b.x), float(b.y), float(b.z), rf * rf);
    c32[slot] = make_float4(float(c.x), float(c.y), float(c.z), float(c.w));
    meta[slot] = (id << 3) | (mov ? 4 : 0) | mt.type;
}

// 4-wide record of binary node i (layout: rt_trace.cuh): an inner child is replaced by ITS two children
// (boxes and refs read from the child's own record), a leaf child stands for itself; free places hold an
// inverted box and RT_SENTINEL.  One thread per binary node, no dependence between threads.
struct WideChild { float lo[3], hi[3]; int ref; };

__device__ __forceinline__ void child_of(const float4& q0, const float4& q1, const float4& q2, const float4& q3, int side, WideChild& c) {
    const float4& xy = side ? q1 : q0;
    c.lo[0] = xy.x; c.lo[1] = xy.y; c.hi[0] = xy.z; c.hi[1] = xy.w;
    c.lo[2] = side ? q2.y : q2.x; c.hi[2] = side ? q2.w : q2.z;
    c.ref = __float_as_int(side ? q3.y : q3.x);
}

__global__ void widen_kernel(const float4* __restrict__ n2, int n_nodes, float4* __restrict__ n4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const float4 q0 = n2[4 * i + 0], q1 = n2[4 * i + 1], q2 = n2[4 * i + 2], q3 = n2[4 * i + 3];
    WideChild ch[4];
    int n = 0;
#pragma unroll
    for (int side = 0; side < 2; ++side) {
        WideChild c;
        child_of(q0, q1, q2, q3, side, c);
        if (c.ref >= 0) {
            const float4 r0 = n2[4 * c.ref + 0], r1 = n2[4 * c.ref + 1], r2 = n2[4 * c.ref + 2], r3 = n2[4 * c.ref + 3];
            child_of(r0, r1, r2, r3, 0, ch[n++]);
            child_of(r0, r1, r2, r3, 1, ch[n++]);
        } else {
            ch[n++] = c;
        }
    }
    for (; n < 4; ++n) {
        for (int a = 0; a < 3; ++a) { ch[n].lo[a] = 1e30f; ch[n].hi[a] = -1e30f; }
        ch[n].ref = int(0x80000000);
    }
    float4* o = n4 + 8 * size_t(i);
#pragma unroll
    for (int c = 0; c < 4; ++c) o[c] = make_float4(ch[c].lo[0], ch[c].lo[1], ch[c].hi[0], ch[c].hi[1]);
    o[4] = make_float4(ch[0].lo[2], ch[1].lo[2], ch[0].hi[2], ch[1].hi[2]);
    o[5] = make_float4(ch[2].lo[2], ch[3].lo[2], ch[2].hi[2], ch[3].hi[2]);
    o[6] = make_float4(__int_as_float(ch[0].ref), __int_as_float(ch[1].ref), __int_as_float(ch[2].ref), __int_as_float(ch[3].ref));
    o[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

// ---- top of the tree first (north_star item 2: "staged in shared memory for the top levels") ----
// A tree too large for shared memory is walked in global memory, but its first levels are fetched by EVERY ray.
// This pass renumbers the inner nodes so that the first k of them in breadth-first order from the root are node
// indices 0 .. k-1 (any prefix of a breadth-first order is closed under "parent of"), which lets a render kernel
// copy nodes [0, n_top) next to its path pools and walk `cur < n_top` from shared memory, the rest from L2 / HBM.
//   top_select_kernel   one block: level-synchronous BFS over at most k inner nodes -> newpos[] (old index -> new index):
//                       top node j goes to j, the nodes that sat in 0 .. k-1 before move to the places the top nodes left
//   top_permute_kernel  every record moves to newpos[i] with its two child refs renamed
__global__ void iota_kernel(int* __restrict__ a, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = i;
}

#define RT_TOP_BLOCK 1024
__global__ void __launch_bounds__(RT_TOP_BLOCK) top_select_kernel(const float4* __restrict__ nodes, int n_nodes, int root, int k,
                                                                  int* __restrict__ newpos, int* __restrict__ top_ids) {
    __shared__ int frontier[2][2 * RT_TOP_BLOCK];
    __shared__ int cnt[RT_TOP_BLOCK + 1];
    __shared__ int n_front, n_top;
    const int t = threadIdx.x;
    if (t == 0) { frontier[0][0] = root; n_front = 1; n_top = 0; }
    __syncthreads();
    int cur = 0;
    while (true) {
        const int nf = n_front, base = n_top;
        if (nf == 0 || base >= k) break;
        const int take = min(nf, k - base);                       // this level's nodes that still fit, in frontier order
        int cl = -1, cr = -1;
        if (t < take) {
            const int node = frontier[cur][t];
            top_ids[base + t] = node;
            const float4 q3 = nodes[4 * size_t(node) + 3];
            cl = __float_as_int(q3.x); cr = __float_as_int(q3.y);
        }
        cnt[t] = (cl >= 0) + (cr >= 0);
        __syncthreads();
        if (t == 0) {                                              // exclusive prefix over <= 1024 counts
            int acc = 0;
            for (int i = 0; i < take; ++i) { const int c = cnt[i]; cnt[i] = acc; acc += c; }
            n_front = min(acc, 2 * RT_TOP_BLOCK); n_top = base + take;
        }
        __syncthreads();
        if (t < take) {
            int o = cnt[t];
            if (cl >= 0 && o < 2 * RT_TOP_BLOCK) frontier[cur ^ 1][o++] = cl;
            if (cr >= 0 && o < 2 * RT_TOP_BLOCK) frontier[cur ^ 1][o++] = cr;
        }
        cur ^= 1;
        __syncthreads();
        if (n_front > RT_TOP_BLOCK) { if (t == 0) n_front = RT_TOP_BLOCK; __syncthreads(); }   // k <= RT_TOP_BLOCK: the rest never fits anyway
    }
    __syncthreads();
    const int kk = n_top;                                          // top nodes found (<= k)
    // old indices 0 .. kk-1 that are not top nodes trade places with the top nodes that sat at >= kk
    int* mark = frontier[0];                                       // mark[i] = 1: old index i (< kk) is a top node
    for (int i = t; i < kk; i += RT_TOP_BLOCK) mark[i] = 0;
    __syncthreads();
    for (int j = t; j < kk; j += RT_TOP_BLOCK) { const int o = top_ids[j]; if (o < kk) mark[o] = 1; }
    __syncthreads();
    if (t == 0) {
        int a = 0;                                                 // next displaced old index
        for (int j = 0; j < kk; ++j) {
            const int o = top_ids[j];
            newpos[o] = j;
            if (o >= kk) {
                while (mark[a]) ++a;
                newpos[a++] = o;
            }
        }
    }
}

__global__ void __launch_bounds__(256) top_permute_kernel(const float4* __restrict__ src, const int* __restrict__ newpos, int n_nodes,
                                                          float4* __restrict__ dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const float4 q0 = src[4 * size_t(i) + 0], q1 = src[4 * size_t(i) + 1], q2 = src[4 * size_t(i) + 2];
    float4 q3 = src[4 * size_t(i) + 3];
    const int cl = __float_as_int(q3.x), cr = __float_as_int(q3.y);
    if (cl >= 0) q3.x = __int_as_float(newpos[cl]);
    if (cr >= 0) q3.y = __int_as_float(newpos[cr]);
    float4* o = dst + 4 * size_t(newpos[i]);
    o[0] = q0; o[1] = q1; o[2] = q2; o[3] = q3;
}

// ---- refit: new boxes for an unchanged tree (dynamic scenes) ----
// The spheres of a scene move (other centres, other motion vectors, other radii) but keep their slots: the tree's TOPOLOGY
// stays, every box is recomputed bottom-up.  Works on the final node array of either builder, whatever its numbering:
//   tree_parents_kernel  parent[c] = 2 * node + side for every inner child c (the root keeps -1)
//   tree_refit_kernel    one thread per node: boxes of its LEAF children from the new slot boxes; the thread that completes
//                        a node (second arrival, atomic counter) unites its two child boxes into the parent's record and
//                        climbs.  Boxes other threads wrote are read with ld.cg (L2): an L1 line may hold a stale copy.
__global__ void __launch_bounds__(256) tree_parents_kernel(const float4* __restrict__ nodes, int n_nodes, int* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const float4 q3 = nodes[4 * size_t(i) + 3];
    const int cl = __float_as_int(q3.x), cr = __float_as_int(q3.y);
    if (cl >= 0) parent[cl] = 2 * i;
    if (cr >= 0) parent[cr] = 2 * i + 1;
}

__device__ __forceinline__ void store_child_box(float4* nodes, int node, int side, const float lo[3], const float hi[3]) {
    float* rec = reinterpret_cast<float*>(nodes + 4 * size_t(node));
    reinterpret_cast<float4*>(rec)[side] = make_float4(lo[0], lo[1], hi[0], hi[1]);      // q0 (left) / q1 (right)
    rec[8 + side] = lo[2];                                                                   // q2 = (Lmin.z, Rmin.z, Lmax.z, Rmax.z)
    rec[10 + side] = hi[2];
}

__global__ void __launch_bounds__(256) tree_refit_kernel(float4* nodes, int n_nodes, const int* __restrict__ parent, int* flags,
                                                         const HostBox* __restrict__ slot_boxes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const float4 q3 = __ldcg(nodes + 4 * size_t(i) + 3);
    const int c[2] = { __float_as_int(q3.x), __float_as_int(q3.y) };
    int nleaf = 0;
#pragma unroll
    for (int side = 0; side < 2; ++side) {
        if (c[side] >= 0) continue;
        const int leaf = ~c[side], first = leaf >> 3, count = (leaf & 7) + 1;
        float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY };
        for (int k = 0; k < count; ++k) {
            const HostBox b = slot_boxes[first + k];
#pragma unroll
            for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], b.lo[a]); hi[a] = fmaxf(hi[a], b.hi[a]); }
        }
        store_child_box(nodes, i, side, lo, hi);
        ++nleaf;
    }
    if (nleaf == 0) return;
    __threadfence();
    int node = i;
    int arrived = atomicAdd(&flags[node], nleaf) + nleaf;
    while (arrived == 2) {                                      // both subtrees of `node` are done: its box goes up
        __threadfence();
        const int p = parent[node];
        if (p < 0) break;                                       // the root
        const float4 q0 = __ldcg(nodes + 4 * size_t(node) + 0), q1 = __ldcg(nodes + 4 * size_t(node) + 1), q2 = __ldcg(nodes + 4 * size_t(node) + 2);
        const float lo[3] = { fminf(q0.x, q1.x), fminf(q0.y, q1.y), fminf(q2.x, q2.y) };
        const float hi[3] = { fmaxf(q0.z, q1.z), fmaxf(q0.w, q1.w), fmaxf(q2.z, q2.w) };
        store_child_box(nodes, p >> 1, p & 1, lo, hi);
        __threadfence();
        node = p >> 1;
        arrived = atomicAdd(&flags[node], 1) + 1;
    }
}

}  // namespace

cudaError_t launch_top_first(const float4* nodes, int n_nodes, int root_ref, int k, int* newpos, int* top_ids, float4* nodes_out, cudaStream_t st) {
    if (n_nodes <= 0 || root_ref < 0) return cudaSuccess;
    k = k < n_nodes ? k : n_nodes;
    if (k > RT_TOP_BLOCK) k = RT_TOP_BLOCK;
    iota_kernel<<<(n_nodes + 255) / 256, 256, 0, st>>>(newpos, n_nodes);
    top_select_kernel<<<1, RT_TOP_BLOCK, 0, st>>>(nodes, n_nodes, root_ref, k, newpos, top_ids);
    top_permute_kernel<<<(n_nodes + 255) / 256, 256, 0, st>>>(nodes, newpos, n_nodes, nodes_out);
    return cudaGetLastError();
}

cudaError_t launch_widen(const float4* nodes2, int n_nodes, float4* nodes4, cudaStream_t st) {
    if (n_nodes <= 0) return cudaSuccess;
    widen_kernel<<<(n_nodes + 255) / 256, 256, 0, st>>>(nodes2, n_nodes, nodes4);
    return cudaGetLastError();
}

cudaError_t launch_scene_boxes(const rt_sphere* spheres, const int* ids, int n, HostBox* boxes, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    scene_boxes_kernel<<<(n + 255) / 256, 256, 0, st>>>(spheres, ids, n, boxes);
    return cudaGetLastError();
}

cudaError_t launch_scene_fill(const rt_sphere* spheres, const rt_material* materials, const int* ids, const int* order, int n,
                              float4* a32, float4* b32, float4* c32, double4* a64, double4* b64, double4* c64, int* meta, cudaStream_t st,
                              int* slot_sphere) {
    if (n <= 0) return cudaSuccess;
    scene_fill_kernel<<<(n + 255) / 256, 256, 0, st>>>(spheres, materials, ids, order, n, a32, b32, c32, a64, b64, c64, meta, slot_sphere);
    return cudaGetLastError();
}

cudaError_t launch_tree_refit(float4* nodes, int n_nodes, const HostBox* slot_boxes, int* parent, int* flags, cudaStream_t st) {
    if (n_nodes <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(parent, 0xff, sizeof(int) * size_t(n_nodes), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(flags, 0, sizeof(int) * size_t(n_nodes), st);
    if (e != cudaSuccess) return e;
    tree_parents_kernel<<<(n_nodes + 255) / 256, 256, 0, st>>>(nodes, n_nodes, parent);
    tree_refit_kernel<<<(n_nodes + 255) / 256, 256, 0, st>>>(nodes, n_nodes, parent, flags, slot_boxes);
    return cudaGetLastError();
}

}  // namespace rt
