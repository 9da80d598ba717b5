// rt_trace.cuh -- scene view, BVH traversal and surface reconstruction (device side).
//
// Replaces hittable_list::hit (src/hittable_list.h:25-39) and the bvh_node::hit contract
// (src/bvh.h:16-24: box test, near child, then far child with t_max narrowed) with an iterative,
// short-stack traversal over flat 64-byte node records, four 128-bit loads per visit.
#pragma once

#include "rt_device.cuh"

namespace rt {

#define RT_SENTINEL ((int)0x80000000)
#define RT_STACK 64
#ifndef RT_STACK_T
#define RT_STACK_T 0                   // 1: stack entries carry the entry distance of their box, so that a popped node which the closest hit found meanwhile
#endif                                 // has put out of reach is dropped at the pop instead of costing a node visit.  Parity-clean, but measured 1.8 % (C3) to
                                       // 4-6 % (C5) SLOWER (profiles/r2_stack_entry_distance_ab.txt): such pops are rare here -- a far child is only pushed if
                                       // its box is within the t_best of that moment, which the BIG stage has already set to the ground hit -- and the 64-bit
                                       // entries + the pop loop cost more than the few visits saved.  make ST=1 builds it
#ifndef RT_PREFETCH_FAR
#define RT_PREFETCH_FAR 0               // tree in global memory: prefetch the pushed far child into L1 (1: its first 32 bytes, 2: all 64)
#endif

// Node record (one per inner node, 64 B = 4 x float4), boxes of BOTH children, padded outwards:
//   q0 = (Lmin.x, Lmin.y, Lmax.x, Lmax.y)   q1 = (Rmin.x, Rmin.y, Rmax.x, Rmax.y)
//   q2 = (Lmin.z, Rmin.z, Lmax.z, Rmax.z)   q3 = (bits left_ref, bits right_ref, -, -)
// -- every aligned pair of floats is one operand of a packed FFMA2 of the slab test (node_boxes below).
// child ref >= 0: inner node index; < 0: leaf, ~ref = first_slot << 3 | (count - 1).
template <typename T> struct SceneView {
    typedef typename Quad<T>::type Q;
    const float4* nodes;       // global
    const float4* nodes4;      // global, or null: the same tree as 4-wide records (scenes walked in global memory)
    const Q* sa;               // (center1, radius)
    const Q* sb;               // (center_vec, radius^2)
    const Q* sc;               // (albedo, param)
    const int* meta;           // list_index << 3 | moving << 2 | material tag
    // big spheres: outside the tree, always double, tested first
    const double4* big_a;
    const double4* big_b;
    const double4* big_c;
    const int* big_meta;
    int n_slots, n_nodes, n_big, root_ref;
    int n_top;                 // wavefront, scene in global memory: nodes [0, n_top) are also kept in shared memory (the tree's first levels)
    int off_sa, off_sb, off_sc, off_meta;   // byte offsets of the staged copies in shared memory (nodes at 0)
};

// Slab test of BOTH child boxes of a node, src/aabb.h:35-53 with 1/d hoisted.
//
// float rays: t = plane * inv + noi with noi = -(o * inv) hoisted per ray -- an FMA per plane instead of
// FADD + FMUL.  The near / far plane of an axis is known from the sign of the direction, so the reciprocal
// is kept split into its positive part p and negative part q (one of them is 0; the product with the zero
// part is exact): near = l*p + (h*q + noi), far = h*p + (l*q + noi), and no min / max per axis is needed.
// The planes of a record are laid out so that every such FMA is one half of a PACKED fp32 FMA of sm_100a
// (fma.rn.f32x2 = FFMA2: two lanes of one 64-bit register pair, one issue slot): 12 FFMA2 + 8 min / max
// per node where the min / max form took 12 FFMA + 20 FMNMX.  This kernel is bound by issue slots.
// The price of the hoisted noi is an ABSOLUTE error of ~ulp(o * inv) on every t (plus 1 ulp relative from
// the approximate reciprocal), so the test is made conservative with a relative factor AND a per-ray
// additive slack `delta` = 2^-22 max|o * inv| (trav_setup, which also keeps inv finite: a zero direction
// component switches its axis off through delta).  A box can only be accepted wrongly, never rejected
// wrongly, and acceptance of a HIT is decided by the sphere tests alone.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// per-ray constants of the slab test
template <typename T> struct SlabRay;
template <> struct SlabRay<float> {
    f32x2 pxy, qxy, oxy, pzz, qzz, ozz;       // (p.x, p.y) (q.x, q.y) (noi.x, noi.y) (p.z, p.z) (q.z, q.z) (noi.z, noi.z)
    float delta;
};
template <> struct SlabRay<double> {
    Vec<double> o, inv;                        // parity mode: the reference's form, (plane - o) * inv
};

__device__ __forceinline__ void node_boxes(const float4& q0, const float4& q1, const float4& q2, const SlabRay<float>& s,
                                           float tmin, float tmax, bool& hl, bool& hr, float& tl, float& tr) {
    const f32x2 Ll = pack2(q0.x, q0.y), Lh = pack2(q0.z, q0.w), Rl = pack2(q1.x, q1.y), Rh = pack2(q1.z, q1.w);
    const f32x2 Zl = pack2(q2.x, q2.y), Zh = pack2(q2.z, q2.w);
    float lnx, lny, lfx, lfy, rnx, rny, rfx, rfy, lnz, rnz, lfz, rfz;
    unpack2(fma2(Ll, s.pxy, fma2(Lh, s.qxy, s.oxy)), lnx, lny);
    unpack2(fma2(Lh, s.pxy, fma2(Ll, s.qxy, s.oxy)), lfx, lfy);
    unpack2(fma2(Rl, s.pxy, fma2(Rh, s.qxy, s.oxy)), rnx, rny);
    unpack2(fma2(Rh, s.pxy, fma2(Rl, s.qxy, s.oxy)), rfx, rfy);
    unpack2(fma2(Zl, s.pzz, fma2(Zh, s.qzz, s.ozz)), lnz, rnz);
    unpack2(fma2(Zh, s.pzz, fma2(Zl, s.qzz, s.ozz)), lfz, rfz);
    tl = fmaxf(fmaxf(lnx, lny), fmaxf(lnz, tmin));
    tr = fmaxf(fmaxf(rnx, rny), fmaxf(rnz, tmin));
    const float fl = fminf(fminf(lfx, lfy), fminf(lfz, tmax)), fr = fminf(fminf(rfx, rfy), fminf(rfz, tmax));
    hl = tl <= fmaf(fl, 1.000001f, s.delta);
    hr = tr <= fmaf(fr, 1.000001f, s.delta);
}

// double rays against the float boxes (exactly representable in double): the boxes' outward
// padding (>= 2^-20 relative) dwarfs double rounding.
__device__ __forceinline__ bool slab(float lx, float ly, float lz, float hx, float hy, float hz,
                                     const Vec<double>& o, const Vec<double>& inv, double tmin, double tmax, double& tnear) {
    const double t0x = (double(lx) - o.x) * inv.x, t1x = (double(hx) - o.x) * inv.x;
    const double t0y = (double(ly) - o.y) * inv.y, t1y = (double(hy) - o.y) * inv.y;
    const double t0z = (double(lz) - o.z) * inv.z, t1z = (double(hz) - o.z) * inv.z;
    const double tn = fmax(fmax(fmin(t0x, t1x), fmin(t0y, t1y)), fmax(fmin(t0z, t1z), tmin));
    const double tf = fmin(fmin(fmax(t0x, t1x), fmax(t0y, t1y)), fmin(fmax(t0z, t1z), tmax));
    tnear = tn;
    return tn <= tf;
}
__device__ __forceinline__ void node_boxes(const float4& q0, const float4& q1, const float4& q2, const SlabRay<double>& s,
                                           double tmin, double tmax, bool& hl, bool& hr, double& tl, double& tr) {
    hl = slab(q0.x, q0.y, q2.x, q0.z, q0.w, q2.z, s.o, s.inv, tmin, tmax, tl);
    hr = slab(q1.x, q1.y, q2.y, q1.z, q1.w, q2.w, s.o, s.inv, tmin, tmax, tr);
}

template <typename Tie>
__device__ __forceinline__ bool leaf_sphere(const float4& a, const float4& b, const Ray<float>& r, float a_len2, float inv_a,
                                            float tmin, float tbest, const Tie& tie, float& t) {
    (void)a_len2;
    return sphere_hit_f32(r.o, r.d, inv_a, sphere_center<float>(a, b, r.tm), b.w, tmin, tbest, tie, t);
}
template <typename Tie>
__device__ __forceinline__ bool leaf_sphere(const double4& a, const double4& b, const Ray<double>& r, double a_len2, double inv_a,
                                            double tmin, double tbest, const Tie& tie, double& t) {
    (void)inv_a;
    return sphere_hit_ref(r.o, r.d, a_len2, sphere_center<double>(a, b, r.tm), b.w, tmin, tbest, tie, t);
}

// Traversal state of one ray.  The traversal is cut into three steps -- init (big spheres, root),
// node visit, leaf visit -- so that the render kernel can schedule them per warp (rt_kernels.cuh);
// trace_closest() below simply runs them to completion for one ray.
//   cur >= 0: inner node to visit; cur < 0 and != RT_SENTINEL: leaf to test; RT_SENTINEL: finished.
#if RT_STACK_T
typedef int2 StackEntry;     // (child ref, float bits of a lower bound of the distance at which the ray enters the child's box)
#else
typedef int StackEntry;
#endif

template <typename T> struct Trav {
    SlabRay<T> slab;         // per-ray constants of the box test
    T qa;                    // fast mode: 1 / |d|^2; parity mode: |d|^2 (what each sphere test wants)
    T tbest;
    int best_slot;           // tree slot, or n_slots + k for big sphere k, of the closest hit so far (-1: none)
    int cur, sp;
    int tos;                 // copy of stack[sp] (the top entry; the sentinel of stack[0] when the stack is empty): a pop takes it from the register and reloads it off the critical path
    float tos_t;             // RT_STACK_T: the entry distance stored with it (-inf below the bottom of the stack)
};

// The far child of a visit waits on the stack while the near one is walked; by the time it is popped the closest hit has
// often moved in front of its box (a ray that hits a sphere in its first leaf leaves 3-4 such entries behind).  Without the
// stored distance each of them cost a whole node visit -- both of ITS children fail against the new t_best -- before the next
// pop.  The test at the pop is the box test of the visit that pushed the entry, repeated with today's t_best (same slack in
// fast mode; in parity mode the stored float is rounded down and t_best up), so nothing that can hold a closer hit, or an equal
// one of a lower list index, is dropped: hits, images and ray counts are unchanged.
__device__ __forceinline__ float cull_limit(const Trav<float>& tv) { return fmaf(tv.tbest, 1.000001f, tv.slab.delta); }
__device__ __forceinline__ float cull_limit(const Trav<double>& tv) { return __double2float_ru(tv.tbest); }
__device__ __forceinline__ float entry_bound(float t) { return t; }
__device__ __forceinline__ float entry_bound(double t) { return __double2float_rd(t); }

// The stack of a walk.  stack[0] holds a sentinel for good (stack_init, once per thread), entries live at 1 .. sp, and the top
// entry is mirrored in a register (tos == stack[sp]): a push is `stack[++sp] = ref`, a pop `cur = tos; sp = max(sp - 1, 0);
// tos = stack[sp]` -- no test for an empty stack, no branch (the first form took 11 instructions and a branch per pop, and a pop
// happens on almost every second visit).  Nothing checks for overflow here: rt_upload_scene refuses trees deeper than
// RT_STACK - 2 (a binary walk pushes at most one entry per level).
__device__ __forceinline__ void stack_init(StackEntry* stack) {
#if RT_STACK_T
    stack[0] = make_int2(RT_SENTINEL, __float_as_int(-INFINITY));
#else
    stack[0] = RT_SENTINEL;
#endif
}

template <typename T>
__device__ __forceinline__ void trav_push(Trav<T>& tv, StackEntry* stack, int ref, T t_entry) {
#if RT_STACK_T
    const float t = entry_bound(t_entry);
    stack[++tv.sp] = make_int2(ref, __float_as_int(t)); tv.tos = ref; tv.tos_t = t;
#else
    (void)t_entry;
    stack[++tv.sp] = ref; tv.tos = ref;
#endif
}

// next node / leaf to walk, or RT_SENTINEL when the stack is empty
template <typename T>
__device__ __forceinline__ int trav_pop(Trav<T>& tv, const StackEntry* stack) {
#if RT_STACK_T
    const float lim = cull_limit(tv);
    int next;
    for (;;) {
        next = tv.tos;
        const float t = tv.tos_t;
        tv.sp = max(tv.sp - 1, 0);
        const StackEntry e = stack[tv.sp]; tv.tos = e.x; tv.tos_t = __int_as_float(e.y);
        if (!(t > lim)) break;               // (the sentinel carries -inf)
    }
    return next;
#else
    const int next = tv.tos;
    tv.sp = max(tv.sp - 1, 0);
    tv.tos = stack[tv.sp];
    return next;
#endif
}

// big spheres first, in double in both modes (the r = 1000 ground sphere): a tight t_max for the
// tree and no float cancellation in |oc|^2 - r^2
template <typename T>
__device__ __forceinline__ void trav_big(const SceneView<T>& sv, const Ray<T>& r, T tmin, T tmax, T& tbest, int& best_slot) {
    tbest = tmax; best_slot = -1;
    if (sv.n_big > 0) {
        const Vec<double> od = mk<double>(r.o.x, r.o.y, r.o.z), dd = mk<double>(r.d.x, r.d.y, r.d.z);
        const double ad = len2(dd);
        const double inv_ad = sizeof(T) == 4 ? rcp_newton(ad) : 0.0;
        double tb = double(tbest);
        for (int k = 0; k < sv.n_big; ++k) {
            const double4 a = sv.big_a[k], b = sv.big_b[k];
            const TieBySlot tie = { sv.meta, sv.big_meta, sv.n_slots, sv.n_slots + k, best_slot };
            double t;
            const bool hit = sizeof(T) == 4
                ? sphere_hit_big(od, dd, ad, inv_ad, sphere_center<double>(a, b, double(r.tm)), b.w, double(tmin), tb, tie, t)
                : sphere_hit_ref(od, dd, ad, sphere_center<double>(a, b, double(r.tm)), b.w, double(tmin), tb, tie, t);
            if (hit) { tb = t; best_slot = sv.n_slots + k; }
        }
        tbest = T(tb);
    }
}

// per-ray constants of the box test
__device__ __forceinline__ void slab_setup(const Ray<float>& r, SlabRay<float>& s) {
    Vec<float> inv = mk<float>(rt_rcp(r.d.x), rt_rcp(r.d.y), rt_rcp(r.d.z));
    // keep the reciprocals finite: with +-inf the FMA form gives plane * inf - o * inf = NaN or an infinity
    // whose sign depends on the plane alone.  At 1e18 the axis is effectively switched off through delta
    // (boxes are accepted conservatively, as the NaN-dropping min/max of the reference's form would)
    const float big = 1e18f;
    if (!(fabsf(inv.x) <= big)) inv.x = copysignf(big, r.d.x);
    if (!(fabsf(inv.y) <= big)) inv.y = copysignf(big, r.d.y);
    if (!(fabsf(inv.z) <= big)) inv.z = copysignf(big, r.d.z);
    const Vec<float> noi = mk<float>(-(r.o.x * inv.x), -(r.o.y * inv.y), -(r.o.z * inv.z));
    s.delta = 2.384185791015625e-07f * fmaxf(fmaxf(fabsf(noi.x), fabsf(noi.y)), fabsf(noi.z));
    const bool nx = inv.x < 0.0f, ny = inv.y < 0.0f, nz = inv.z < 0.0f;
    s.pxy = pack2(nx ? 0.0f : inv.x, ny ? 0.0f : inv.y);
    s.qxy = pack2(nx ? inv.x : 0.0f, ny ? inv.y : 0.0f);
    s.oxy = pack2(noi.x, noi.y);
    s.pzz = pack2(nz ? 0.0f : inv.z, nz ? 0.0f : inv.z);
    s.qzz = pack2(nz ? inv.z : 0.0f, nz ? inv.z : 0.0f);
    s.ozz = pack2(noi.z, noi.z);
}
__device__ __forceinline__ void slab_setup(const Ray<double>& r, SlabRay<double>& s) {
    s.o = r.o;
    s.inv = mk<double>(1.0 / r.d.x, 1.0 / r.d.y, 1.0 / r.d.z);
}
template <typename T> __device__ __forceinline__ void slab_clear(SlabRay<T>& s);
template <> __device__ __forceinline__ void slab_clear<float>(SlabRay<float>& s) { s.pxy = s.qxy = s.oxy = s.pzz = s.qzz = s.ozz = 0ull; s.delta = 0.0f; }
template <> __device__ __forceinline__ void slab_clear<double>(SlabRay<double>& s) { s.o = mk<double>(0, 0, 0); s.inv = mk<double>(0, 0, 0); }

// per-ray constants of the tree walk; tbest / best_slot are already set (trav_big)
template <typename T>
__device__ __forceinline__ void trav_setup(const SceneView<T>& sv, const Ray<T>& r, Trav<T>& tv) {
    const T a_len2 = len2(r.d);
    tv.qa = sizeof(T) == 4 ? rt_rcp(a_len2) : a_len2;
    slab_setup(r, tv.slab);
    tv.cur = sv.root_ref; tv.sp = 0; tv.tos = RT_SENTINEL; tv.tos_t = -INFINITY;
}

template <typename T>
__device__ __forceinline__ void trav_init(const SceneView<T>& sv, const Ray<T>& r, T tmin, T tmax, Trav<T>& tv) {
    trav_big<T>(sv, r, tmin, tmax, tv.tbest, tv.best_slot);
    trav_setup<T>(sv, r, tv);
}

// one inner node: both child boxes, near child next, far child pushed (src/bvh.h:16-24 contract)
template <typename T>
__device__ __forceinline__ void trav_node(const float4* __restrict__ nodes, const Ray<T>& r, T tmin, Trav<T>& tv, StackEntry* stack) {
    const int cur = tv.cur;
    const float4 q0 = nodes[4 * cur + 0], q1 = nodes[4 * cur + 1], q2 = nodes[4 * cur + 2], q3 = nodes[4 * cur + 3];
    T tl, tr;
    bool hl, hr;
    node_boxes(q0, q1, q2, tv.slab, tmin, tv.tbest, hl, hr, tl, tr);
    const int cl = __float_as_int(q3.x), cr = __float_as_int(q3.y);
    // near child next, far child pushed, nothing hit -> pop: written with selects and two short predicated
    // blocks instead of a four-way branch, so the lanes of a warp stay together through the step
    const bool left_first = !hr || (hl && tl <= tr);
    const int near = left_first ? cl : cr, far = left_first ? cr : cl;
    if (hl && hr) trav_push<T>(tv, stack, far, left_first ? tr : tl);
    int next = near;
    if (!(hl || hr)) next = trav_pop<T>(tv, stack);
    tv.cur = next;
}

// The same visit for a tree in global memory whose first n_top nodes (breadth-first from the root, rt_scene.cu
// launch_top_first) are ALSO resident in shared memory: the upper levels, which every ray fetches, come from `top`
// (LDS latency), the rest from L2 / HBM.  One generic-address load path for both, no divergence.
template <typename T>
__device__ __forceinline__ void trav_node_top(const float4* __restrict__ nodes, const float4* top, int n_top, const Ray<T>& r, T tmin,
                                              Trav<T>& tv, StackEntry* stack) {
    const int cur = tv.cur;
    const float4* rec = (cur < n_top ? top : nodes) + 4 * size_t(cur);
    const float4 q0 = rec[0], q1 = rec[1], q2 = rec[2], q3 = rec[3];
    T tl, tr;
    bool hl, hr;
    node_boxes(q0, q1, q2, tv.slab, tmin, tv.tbest, hl, hr, tl, tr);
    const int cl = __float_as_int(q3.x), cr = __float_as_int(q3.y);
    const bool left_first = !hr || (hl && tl <= tr);
    const int near = left_first ? cl : cr, far = left_first ? cr : cl;
#if RT_PREFETCH_FAR
    // a far child that is pushed will be fetched many visits later, after a pop: ask for its record now, so that the pop finds
    // it in L1 instead of paying an L2 round trip on the walk's critical path (A/B knob, make PF=1 / 2)
    if (hl && hr && far >= n_top) {
        const float4* f = nodes + 4 * size_t(far);
        asm volatile("prefetch.global.L1 [%0];" :: "l"(f));
        if (RT_PREFETCH_FAR >= 2) asm volatile("prefetch.global.L1 [%0];" :: "l"(f + 2));
    }
#endif
    if (hl && hr) trav_push<T>(tv, stack, far, left_first ? tr : tl);
    int next = near;
    if (!(hl || hr)) next = trav_pop<T>(tv, stack);
    tv.cur = next;
}

// ---- 4-wide records (wider nodes for trees that live in global memory) ----
// A tree that does not fit shared memory is walked at L2 latency: the walk is a chain of dependent 64-byte
// fetches, one per level.  The 4-wide record of binary node i holds the boxes of i's GRANDCHILDREN (a child
// that is a leaf stands for itself), so one 128-byte fetch -- one L2 line -- advances the walk two levels:
//   q0..q3 = child c: (min.x, min.y, max.x, max.y)       every aligned pair = one operand of the packed FMAs
//   q4 = (min.z c0, min.z c1, max.z c0, max.z c1)   q5 = the same for c2, c3
//   q6 = bits of the four child refs (same encoding as the binary record; RT_SENTINEL: no such child)
// Records are indexed by the binary node they were made from (rt_scene.cu widen_kernel), so a ref means the
// same thing in both forms; the records of the odd levels are simply never visited from the root.
__device__ __forceinline__ void node4_boxes(const float4* q, const SlabRay<float>& s, float tmin, float tmax, float* tn, bool* hit) {
    float nz[4], fz[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const f32x2 Zl = pack2(q[4 + h].x, q[4 + h].y), Zh = pack2(q[4 + h].z, q[4 + h].w);
        unpack2(fma2(Zl, s.pzz, fma2(Zh, s.qzz, s.ozz)), nz[2 * h], nz[2 * h + 1]);
        unpack2(fma2(Zh, s.pzz, fma2(Zl, s.qzz, s.ozz)), fz[2 * h], fz[2 * h + 1]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const f32x2 Cl = pack2(q[c].x, q[c].y), Ch = pack2(q[c].z, q[c].w);
        float nx, ny, fx, fy;
        unpack2(fma2(Cl, s.pxy, fma2(Ch, s.qxy, s.oxy)), nx, ny);
        unpack2(fma2(Ch, s.pxy, fma2(Cl, s.qxy, s.oxy)), fx, fy);
        tn[c] = fmaxf(fmaxf(nx, ny), fmaxf(nz[c], tmin));
        const float tf = fminf(fminf(fx, fy), fminf(fz[c], tmax));
        hit[c] = tn[c] <= fmaf(tf, 1.000001f, s.delta);
    }
}
__device__ __forceinline__ void node4_boxes(const float4* q, const SlabRay<double>& s, double tmin, double tmax, double* tn, bool* hit) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float4 z = q[4 + (c >> 1)];
        const float lz = (c & 1) ? z.y : z.x, hz = (c & 1) ? z.w : z.z;
        hit[c] = slab(q[c].x, q[c].y, lz, q[c].z, q[c].w, hz, s.o, s.inv, tmin, tmax, tn[c]);
    }
}

// one 4-wide node: the children the ray enters, nearest next, the others pushed farthest first
template <typename T>
__device__ __forceinline__ void trav_node4(const float4* __restrict__ nodes4, const Ray<T>& r, T tmin, Trav<T>& tv, StackEntry* stack) {
    const float4* rec = nodes4 + 8 * size_t(tv.cur);
    float4 q[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) q[k] = rec[k];
    T tn[4];
    bool hit[4];
    node4_boxes(q, tv.slab, tmin, tv.tbest, tn, hit);
    const int ref[4] = { __float_as_int(q[6].x), __float_as_int(q[6].y), __float_as_int(q[6].z), __float_as_int(q[6].w) };
    // sort keys: entry distance (a positive float orders like its bits) with the child number in the two low bits
    int key[4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
        key[c] = (hit[c] && ref[c] != RT_SENTINEL) ? ((__float_as_int(float(tn[c])) & ~3) | c) : 0x7fffffff;
#define RT_CSWAP(a, b) { const int lo_ = min(key[a], key[b]), hi_ = max(key[a], key[b]); key[a] = lo_; key[b] = hi_; }
    RT_CSWAP(0, 1) RT_CSWAP(2, 3) RT_CSWAP(0, 2) RT_CSWAP(1, 3) RT_CSWAP(1, 2)
#undef RT_CSWAP
#define RT_REF_OF(k) ((key[k] & 2) ? ((key[k] & 1) ? ref[3] : ref[2]) : ((key[k] & 1) ? ref[1] : ref[0]))
#define RT_TN_OF(k) ((key[k] & 2) ? ((key[k] & 1) ? tn[3] : tn[2]) : ((key[k] & 1) ? tn[1] : tn[0]))
#pragma unroll
    for (int k = 3; k >= 1; --k)
        if (key[k] != 0x7fffffff) trav_push<T>(tv, stack, RT_REF_OF(k), RT_TN_OF(k));
    if (key[0] != 0x7fffffff) tv.cur = RT_REF_OF(0);
    else tv.cur = trav_pop<T>(tv, stack);       // nothing entered
#undef RT_TN_OF
#undef RT_REF_OF
}

// one leaf: its 1..8 sphere slots
template <typename T>
__device__ __forceinline__ void trav_leaf(const SceneView<T>& sv, const typename Quad<T>::type* __restrict__ sa, const typename Quad<T>::type* __restrict__ sb,
                                          const int* __restrict__ meta, const Ray<T>& r, T tmin, Trav<T>& tv, const StackEntry* stack) {
    const int leaf = ~tv.cur;
    const int first = leaf >> 3, count = (leaf & 7) + 1;
#pragma unroll 1                        // (the compiler unrolled it four times: 4 KB of sphere tests in a kernel that misses its instruction cache)
    for (int k = 0; k < count; ++k) {
        const int slot = first + k;
        const typename Quad<T>::type a = sa[slot], b = sb[slot];
        const TieBySlot tie = { meta, sv.big_meta, sv.n_slots, slot, tv.best_slot };
        T t;
        if (leaf_sphere(a, b, r, tv.qa, tv.qa, tmin, tv.tbest, tie, t)) {
            tv.tbest = t; tv.best_slot = slot;
        }
    }
    tv.cur = trav_pop<T>(tv, stack);
}

// Closest hit in (tmin, tmax).  slot_out: tree slot, n_slots + k for big sphere k, or -1.
// nodes/sa/sb/meta may point to shared memory (staged scene) or global memory.
template <typename T>
__device__ __forceinline__ void trace_closest(const SceneView<T>& sv, const float4* __restrict__ nodes,
                                              const typename Quad<T>::type* __restrict__ sa,
                                              const typename Quad<T>::type* __restrict__ sb,
                                              const int* __restrict__ meta,
                                              const Ray<T>& r, T tmin, T tmax, T& t_out, int& slot_out) {
    Trav<T> tv;
    StackEntry stack[RT_STACK];
    stack_init(stack);
    trav_init<T>(sv, r, tmin, tmax, tv);
    if (sv.nodes4 && nodes == sv.nodes) {          // tree in global memory with 4-wide records
        while (tv.cur != RT_SENTINEL) {
            if (tv.cur >= 0) trav_node4<T>(sv.nodes4, r, tmin, tv, stack);
            else trav_leaf<T>(sv, sa, sb, meta, r, tmin, tv, stack);
        }
    } else {
        while (tv.cur != RT_SENTINEL) {
            if (tv.cur >= 0) trav_node<T>(nodes, r, tmin, tv, stack);
            else trav_leaf<T>(sv, sa, sb, meta, r, tmin, tv, stack);
        }
    }
    t_out = tv.tbest;
    slot_out = tv.best_slot;
}

// Linear closest hit over every sphere in list order (device-side ground truth,
// src/hittable_list.h:25-39).  Tree slots are visited in slot order, so the index tie rule is
// applied explicitly.
template <typename T>
__device__ __forceinline__ void brute_closest(const SceneView<T>& sv, const Ray<T>& r, T tmin, T tmax, T& t_out, int& slot_out) {
    T tbest;
    int best_slot;
    trav_big<T>(sv, r, tmin, tmax, tbest, best_slot);
    const T a_len2 = len2(r.d);
    const T inv_a = T(1) / a_len2;
    for (int slot = 0; slot < sv.n_slots; ++slot) {
        const typename Quad<T>::type a = sv.sa[slot], b = sv.sb[slot];
        const TieBySlot tie = { sv.meta, sv.big_meta, sv.n_slots, slot, best_slot };
        T t;
        if (leaf_sphere(a, b, r, a_len2, inv_a, tmin, tbest, tie, t)) {
            tbest = t; best_slot = slot;
        }
    }
    t_out = tbest;
    slot_out = best_slot;
}

// What shading needs about the surface that was hit (replaces hit_record, src/hittable.h:7-22:
// only (slot, t) leave the traversal; p / normal / front_face are rebuilt here,
// src/sphere.h:50-54 + src/hittable.h:15-21).
template <typename T> struct Surface {
    Vec<T> p, normal, albedo;
    T param;
    int type, id;
    bool front_face;
};

// Fast mode only: the float traversal picked (slot, t); the winning sphere's quadratic is solved once
// more in double on the same float inputs, so that t, p and the normal are good to ~1e-7 instead of
// the ~1e-6 * |o|/r that float leaves after the cancellation in oc.d (north_star asks 1e-5 on normals
// of r = 0.2 spheres seen from 13 units away).  One double solve per HIT, not per test.
__device__ __forceinline__ void refine_hit(const Vec<double>& o, const Vec<double>& d, const Vec<double>& center, double rr,
                                           double& t) {
    const Vec<double> oc = o - center;
    const double a = len2(d), half_b = dot(oc, d), c = len2(oc) - rr;
    const double disc = half_b * half_b - a * c;
    if (disc < 0) return;                               // grazing: keep the float root
    const double sq = sqrt_newton(disc), inv_a = rcp_newton(a);
    const double r1 = (-half_b - sq) * inv_a, r2 = (-half_b + sq) * inv_a;
    t = fabs(r1 - t) <= fabs(r2 - t) ? r1 : r2;
}

template <typename T>
__device__ __forceinline__ Surface<T> surface_at(const SceneView<T>& sv, const typename Quad<T>::type* __restrict__ sa,
                                                 const typename Quad<T>::type* __restrict__ sb,
                                                 const typename Quad<T>::type* __restrict__ sc,
                                                 const int* __restrict__ meta, const Ray<T>& r, T& t, int slot) {
    Surface<T> s;
    // the sphere that was hit, as doubles (exact for both the float slots and the double big spheres)
    double4 A, B;
    int m;
    if (slot < sv.n_slots) {
        const typename Quad<T>::type a = sa[slot], b = sb[slot], c = sc[slot];
        m = meta[slot];
        A = make_double4(a.x, a.y, a.z, a.w); B = make_double4(b.x, b.y, b.z, 0.0);
        s.albedo = mk<T>(c.x, c.y, c.z); s.param = c.w;
    } else {
        const int k = slot - sv.n_slots;
        A = sv.big_a[k]; B = sv.big_b[k];
        const double4 c = sv.big_c[k];
        m = sv.big_meta[k];
        s.albedo = mk<T>(T(c.x), T(c.y), T(c.z)); s.param = T(c.w);
    }
    const Vec<double> center = mk<double>(A.x, A.y, A.z) + double(r.tm) * mk<double>(B.x, B.y, B.z);   // src/sphere.h:68-72
    Vec<T> outward;
    if (sizeof(T) == 4) {
        // fast mode: one double solve of the winning sphere on the same float inputs (refine_hit)
        const Vec<double> od = mk<double>(r.o.x, r.o.y, r.o.z), dd = mk<double>(r.d.x, r.d.y, r.d.z);
        double td = double(t);
        refine_hit(od, dd, center, A.w * A.w, td);
        const Vec<double> pd = od + td * dd;
        const Vec<double> nd = rcp_newton(A.w) * (pd - center);
        t = T(td);
        s.p = mk<T>(T(pd.x), T(pd.y), T(pd.z));
        outward = mk<T>(T(nd.x), T(nd.y), T(nd.z));
    } else {
        s.p = r.o + t * r.d;                                            // src/ray.h:19-21
        const Vec<double> pd = mk<double>(s.p.x, s.p.y, s.p.z);
        const Vec<double> nd = (1.0 / A.w) * (pd - center);             // src/sphere.h:52
        outward = mk<T>(T(nd.x), T(nd.y), T(nd.z));
    }
    s.type = RT_META_TYPE(m);
    s.id = RT_META_ID(m);
    s.front_face = dot(r.d, outward) < T(0);                           // src/hittable.h:19-20
    s.normal = s.front_face ? outward : -outward;
    return s;
}

}  // namespace rt
