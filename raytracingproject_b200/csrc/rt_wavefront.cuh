// rt_wavefront.cuh -- the wavefront bounce loop, SM-resident.
//
// ray_color's recursion (src/camera_cpu.h:8-26) as a queue-driven wavefront: paths live in a POOL,
// stages (trace / shade by material / end-of-path + next sample) work on QUEUES of path ids, every
// stage runs over a compacted queue (warp ballot + prefix popcount), hits are regrouped by material
// before shading.  B200-first twist: the pool and its queues live in SHARED memory (227 KB per SM)
// next to the staged scene, one pool of P paths per warp, so regrouping costs a few LDS/STS instead
// of a round trip of the ray state through HBM and a kernel launch per stage; a warp needs only
// __syncwarp() between its stages, and the other warps of the SM fill the issue slots meanwhile.
//
//   pool slot  = one pixel of the warp's 16 x (P/16) tile (x one sample slice): samples of a pixel
//                are walked in order by its slot, so the per-pixel sum has the reference's order
//                (src/camera.h:41-44) and needs no atomics
//   jobs       a job = one tile x one sample slice.  A warp keeps RT_RING jobs open (a ring of
//                descriptors in shared memory): a slot whose pixel has no samples left in its job
//                moves on to the same pixel position of the next one at once, and when every slot
//                has left the oldest job the next job of the frame is fetched into its place.  The
//                pool does not drain between jobs (a drain is several rounds at a fraction of the
//                lanes).  Two open jobs measured faster than four (profiles/r1_job_ring_ab.txt):
//                the rays of a pool that spans two tiles are more coherent than those of one that
//                spans four
//   TRACE      lanes PULL rays from the trace queue; node / leaf / pull sections are scheduled per
//                warp like in render_sched_kernel, but a lane whose ray ends pulls the next ray at
//                once (cheap: 9 LDS + 4 divisions) instead of waiting for a shade section
//   HIT        finished rays, grouped opaque (lambertian + metal) | glass: surface_at + scatter
//   END        misses (sky) and ended paths: accumulate, start the pixel's next sample (get_ray)
//
// Per-lane arithmetic is the same device functions as the megakernels, so parity mode (T = double)
// stays bit-identical to the oracle.
#pragma once

#include "rt_kernels.cuh"

namespace rt {

#ifndef RT_NODE_VISITS
#define RT_NODE_VISITS 2               // node visits per head-count ballot of the TRACE stage: 2 since the code diet of round 2 (profiles/r2_code_size_ab.txt; round 1 measured 3 best of 2 / 3 / 4 / 6 on a larger kernel: r1_node_visits_ab.txt)
#endif

#ifndef RT_RING
#define RT_RING 2                      // jobs a warp keeps open (2 or 4)
#endif

#ifndef RT_DYNAMIC_ITEMS
#define RT_DYNAMIC_ITEMS 1             // fast mode: pool slots take (pixel, sample) work items of the open jobs dynamically (A/B: make DYN=0)
#endif

#ifndef RT_FETCH_MULT
#define RT_FETCH_MULT 1                // jobs taken from the frame's counter at a time = RT_FETCH_MULT x job_group (consecutive indices: in the bulk
#endif                                 // phase 2 x 2 = the same tile pair, two consecutive layers -- for the RING=4 experiment)

#ifndef RT_PHILOX_TABLE
#define RT_PHILOX_TABLE 0              // 1: Philox round keys precomputed in the parameter block (LDCU into uniform registers) instead of two uniform
#endif                                 // additions per round -- measured 0.7 % SLOWER on C3 / C4 (profiles/r2_ab_merge_table_top.txt), so off; make TABLE=1

#ifndef RT_MERGE_HITS
#define RT_MERGE_HITS 1                // HIT stage: opaque and glass hits of a round shaded as one sequence of batches (A/B: make MERGE=0)
#endif

template <typename T, int P> struct Pool {
    // three arrays [P] of 4 x T (one 128-bit -- 2 x 128-bit in parity mode -- load or store each): a path's origin + time, its
    // direction + closest hit distance, its throughput + closest hit slot (an int carried in the bits of the fourth component).
    // The first layout kept eleven scalar arrays: a shading batch then cost 12 loads and 9 stores per path instead of 3 and 3
    enum { QO, QD, QT, NQ };
    enum { SAMPLE, DEPTH, NH };                                         // 16-bit arrays [P]: ring entry (top 1 or 2 bits) | next sample relative to the job's first (all ones: done with that job); bounces left
    enum { QUEUE, LOPAQUE, LPIXEL, LEND, NL };                          // byte lists [P]; LPIXEL (dynamic mode): the pixel of the tile a slot is working on
    // dynamic mode (fast mode only, see the END stage): slots take (pixel, sample) items of the open jobs instead of owning a pixel
    static constexpr bool dyn = RT_DYNAMIC_ITEMS != 0 && sizeof(T) == 4;
    static constexpr int ring = dyn ? 4 : RT_RING, ring_bytes = dyn ? 16 * ring + 16 : 16 * ring;   // job descriptors (int4) of the warp (+ counters), after the lists
    static constexpr int bytes = NQ * P * 4 * int(sizeof(T)) + NH * P * 2 + NL * P + ring_bytes;
    static constexpr int tile_w = 16, tile_h = P / 16;
    static_assert(P % 32 == 0 && P <= 128, "pool: path id is 7 bits, lists are filled 32 at a time");
};

// pixel sum += contribution of one sample.  Only this pixel's slot touches the element, one sample after the
// other, so the order of the sum is the reference's (src/camera.h:41-44).  Fast mode sends the addition to L2 as
// one 128-bit reduction (RED.ADD.F32x4: nothing comes back, the warp does not wait for the load); parity mode
// keeps the plain read-modify-write.
__device__ __forceinline__ void accum_add(float4* dst, const Vec<float>& c) { atomicAdd(dst, make_float4(c.x, c.y, c.z, 0.0f)); }
__device__ __forceinline__ void accum_add(double4* dst, const Vec<double>& c) {
    double4 v = *dst;
    v.x += c.x; v.y += c.y; v.z += c.z;
    *dst = v;
}

// the closest-hit slot rides in the fourth component of the throughput quad, bit for bit (no arithmetic ever touches it)
__device__ __forceinline__ float  slot_bits(int s, float)  { return __int_as_float(s); }
__device__ __forceinline__ double slot_bits(int s, double) { return __longlong_as_double((long long)s); }
__device__ __forceinline__ int bits_slot(float w)  { return __float_as_int(w); }
__device__ __forceinline__ int bits_slot(double w) { return int(__double_as_longlong(w)); }
template <typename Q, typename T> __device__ __forceinline__ Q quad(T x, T y, T z, T w) { Q q; q.x = x; q.y = y; q.z = z; q.w = w; return q; }

template <typename T, bool STAGED, int P, int BLOCK, bool WIDE = false>
__global__ void __launch_bounds__(BLOCK, 1)
render_pool_kernel(const RenderParams<T> p) {
    static_assert(!(WIDE && STAGED), "4-wide records are for trees walked in global memory");
    typedef Pool<T, P> L;
    typedef typename Quad<T>::type Q;
    extern __shared__ __align__(16) unsigned char smem[];
    const Staged<T> st = stage_scene<T>(p.scene, smem, STAGED);
    const SceneView<T>& sv = p.scene;
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    // a scene that stays in global memory: the first levels of its tree (nodes [0, n_top), breadth-first from the root) are
    // copied in front of the pools and walked from shared memory; the rest of the tree is fetched from L2 / HBM
    const int n_top = (STAGED || WIDE) ? 0 : sv.n_top;
    const float4* const top = reinterpret_cast<const float4*>(smem);
    if (!STAGED && !WIDE && n_top > 0) {
        float4* t = reinterpret_cast<float4*>(smem);
        for (int i = threadIdx.x; i < 4 * n_top; i += BLOCK) t[i] = sv.nodes[i];
        __syncthreads();
    }

    unsigned char* wb = smem + (STAGED ? staged_bytes_aligned<T>(sv.n_nodes, sv.n_slots) : size_t(n_top) * 64) + size_t(warp) * L::bytes;
    Q* const G = reinterpret_cast<Q*>(wb);                          // G[L::QO * P + pid] ...
    unsigned short* const S = reinterpret_cast<unsigned short*>(G + L::NQ * P);
    unsigned char* const B = reinterpret_cast<unsigned char*>(S + L::NH * P);

    int4* const ring = reinterpret_cast<int4*>(B + L::NL * P);   // job descriptors: (bx | by << 16, s_begin, s_end, slice)

    const int W = p.cam.image_width, H = p.cam.image_height;
    // Jobs are fetched `group` at a time: the same layer of `group` horizontally adjacent tiles, so that the jobs a warp
    // holds open lie next to each other in the image (rays of neighbouring tiles walk the same part of the tree)
    // Tile-interleaved sharding over GPUs: of the frame's tile groups this launch owns those whose index is
    // tile_rank modulo tile_world (every one with tile_world = 1); local group gl stands for group gl * tile_world + tile_rank
    const int group = p.job_group, ntiles = p.tiles_x * p.tiles_y;
    const int tw = p.tile_world > 1 ? p.tile_world : 1, tr = p.tile_world > 1 ? p.tile_rank : 0;
    const int ngroups = ((ntiles + group - 1) / group - tr + tw - 1) / tw, ntiles_l = ngroups * group;
    const int nbulk = ntiles_l * p.bulk_layers, njobs = nbulk + (p.slices - p.bulk_layers) * ntiles_l;
    int pend_next = 0, pend_left = 0;
    const size_t layer = size_t(H) * size_t(W);
    const int th_p = p.th_s, th_l = p.th_l, th_n = p.th_n;
    unsigned long long my_rays = 0;
#ifdef RT_POOL_STATS
    unsigned long long stat[24] = {};            // TRACE-stage scheduler counters (profiling build only), see tools/pool_stats.py
#define RT_STAT(k, v) (stat[k] += (unsigned long long)(v))
#else
#define RT_STAT(k, v) ((void)0)
#endif
#ifdef RT_WARP_TIMES
    // profiling build (make TIMES=1): when a warp starts, runs out of jobs and exits, how many rounds it ran and how many rays
    // they traced, before and after the frame ran out of jobs -- the frame's tail, measured (tools/warp_times.py)
    unsigned long long wt_first = 0, wt_exhausted = 0, wt_rounds = 0, wt_nq = 0, wt_rounds_tail = 0, wt_nq_tail = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(wt_first));
#endif
    StackEntry stack[RT_STACK];
    stack_init(stack);

    // Ring state (warp-uniform).  Every slot goes through every job of its warp, in order: `newest` is the
    // entry fetched last, byte e of `joined` the number of slots that have moved into entry e's job.  A slot
    // whose pixel is finished joins the job in the next entry, or waits in the END list when it is in the
    // newest.  Entry f can take the next job of the frame once all P slots have joined the job after it.
    // Start: every slot "done" with a dummy job in the last entry, every entry free.
    enum { R = L::ring, SH = R == 2 ? 15 : 14, DONE = (1 << SH) - 1 };
    int newest = R - 1;
    unsigned joined = unsigned(P) * 0x01010101u;
    bool exhausted = false;
    if (lane < R) ring[lane] = make_int4(0, 0, 0, 0);
    for (int k = lane; k < P; k += 32) {
        S[L::SAMPLE * P + k] = (unsigned short)0xffff; S[L::DEPTH * P + k] = 0;
        B[L::LEND * P + k] = (unsigned char)(k | 0x80);               // 0x80: nothing to add
    }
    int n_q = 0, n_end = P, n_opq = 0, n_gls = 0;
    int dyn_n = 0, dyn_e0 = 0, dyn_e1 = 0, dyn_left0 = 0, dyn_left1 = 0;   // dynamic mode: the (at most two) jobs that still have items to hand out
    if (L::dyn && lane < 4) reinterpret_cast<int*>(ring + 4)[lane] = 0;
    __syncwarp();

    {
        for (;;) {
            // ------------------------------------------------------------------ END stage
            // misses add thr * sky (src/camera_cpu.h:23-25); every entry then starts its pixel's next sample
            // (camera::render's sample loop + get_ray, src/camera.h:41-44,87-100) in its job or in the next one
            bool progress = false;
            int n_idle = 0;
            if (L::dyn) {
            // ---- dynamic mode (fast mode).  A slot does not own a pixel: every free slot takes the next (pixel, sample) ITEM of
            // the open jobs -- item k of a job = sample k / P of the tile's pixel k % P -- so a pool stays full until its jobs
            // run out of items, however unequal the paths of its pixels are (a pixel inside a glass sphere costs 5 x the mean;
            // with slots bound to pixels the other slots of its job had to wait for it: 121 of 128 slots busy at 500 spp, 107 at
            // 62 spp, 89 at 16 spp, and a frame ended on the slowest pixel of its last jobs: tools/warp_times.py).  The sums
            // of a pixel are then formed by whichever slots took its samples, in any order: float RED.ADD, fast mode only --
            // parity mode keeps the slot-per-pixel loop below, whose sum has the reference's order.
            // Up to 4 jobs are open: the two newest hand out items, older ones only wait for their last paths to end.
            int* const pend = reinterpret_cast<int*>(ring + 4);        // per entry: items not completed yet (0: the entry is free)
            #pragma unroll 1                        // (stage loops stay rolled: code size, see trav_leaf)
            for (int b = 0; b < n_end; b += 32) {
                const int idx = b + lane;
                const bool valid = idx < n_end;
                int pid = 0, jb = -1;
                if (valid) {
                    const int e = B[L::LEND * P + idx];
                    pid = e & 0x7f;
                    const int sw = S[L::SAMPLE * P + pid];
                    if (sw != 0xffff) {                                // a path ended here: its item is complete
                        jb = sw >> 14;
                        if (!(e & 0x80)) {                             // a miss: thr * sky goes to its pixel
                            const int4 J = ring[jb];
                            const int pix = B[L::LPIXEL * P + pid];
                            const int i = (J.x & 0xffff) + (pix & 15), j = (J.x >> 16) + (pix >> 4);
                            const Q qd = G[L::QD * P + pid], qt = G[L::QT * P + pid];
                            const Vec<T> d = mk<T>(qd.x, qd.y, qd.z), thr = mk<T>(qt.x, qt.y, qt.z);
                            accum_add(p.accum + size_t(J.w) * layer + size_t(j) * size_t(W) + size_t(i), thr * sky<T>(d));
                        }
                    }
                }
                if (__any_sync(FULL, jb >= 0)) {
                    const int c0 = __popc(__ballot_sync(FULL, jb == 0)), c1 = __popc(__ballot_sync(FULL, jb == 1));
                    const int c2 = __popc(__ballot_sync(FULL, jb == 2)), c3 = __popc(__ballot_sync(FULL, jb == 3));
                    if (lane < 4) pend[lane] -= lane == 0 ? c0 : (lane == 1 ? c1 : (lane == 2 ? c2 : c3));
                    __syncwarp();
                    progress = true;
                }
                // keep two jobs with items to hand out
                while (!exhausted && dyn_n < 2) {
                    const unsigned fm = __ballot_sync(FULL, lane < 4 && pend[lane] == 0) & 0xfu;
                    if (fm == 0u) break;                               // four jobs still have paths in flight
                    const int ef = __ffs(fm) - 1;
                    int job = 0;
                    if (pend_left > 0) { job = pend_next++; --pend_left; }
                    else {
                        if (lane == 0) job = int(p.jobs_shared ? atomicAdd_system(p.job_counter, unsigned(group * p.job_stride)) : atomicAdd(p.job_counter, unsigned(group * p.job_stride)));
                        job = __shfl_sync(FULL, job, 0);
                        pend_next = job + 1; pend_left = group - 1;
                    }
                    if (job >= njobs) {
                        exhausted = true;
#ifdef RT_WARP_TIMES
                        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(wt_exhausted));
#endif
                        break;
                    }
                    int slice, tile;
                    if (job < nbulk) { const int gl = group * p.bulk_layers, g = job / gl, r = job - g * gl; slice = r / group; tile = group * (g * tw + tr) + (r - slice * group); }
                    else { const int r = job - nbulk; slice = r / ntiles_l; const int tl = r - slice * ntiles_l, g = tl / group; tile = group * (g * tw + tr) + (tl - g * group); slice += p.bulk_layers; }
                    progress = true;
                    if (tile >= ntiles) continue;                      // the last group may be short of tiles
                    const int bx = (tile % p.tiles_x) * L::tile_w, by = (tile / p.tiles_x) * L::tile_h;
                    const int s_begin = p.slice_begin[slice], s_end = p.slice_begin[slice + 1];
                    // the job's accumulator: every pixel of the tile gets its sample count (src/camera.h:41 starts from zero)
                    for (int q = lane; q < P; q += 32) {
                        const int i = bx + (q & 15), j = by + (q >> 4);
                        if (i < W && j < H) {
                            Q* dst = p.accum + size_t(slice) * layer + size_t(j) * size_t(W) + size_t(i);
                            Q v;
                            if (p.accumulate) { v = *dst; } else { v.x = 0; v.y = 0; v.z = 0; v.w = 0; }
                            v.w += decltype(v.x)(s_end - s_begin);
                            *dst = v;
                        }
                    }
                    const int items = p.max_depth > 0 ? (s_end - s_begin) * P : 0;     // max_depth <= 0: black (src/camera_cpu.h:12-13)
                    if (items > 0) {
                        if (lane == 0) { ring[ef] = make_int4(bx | (by << 16), s_begin, s_end, slice); pend[ef] = items; }
                        if (dyn_n == 0) { dyn_e0 = ef; dyn_left0 = items; } else { dyn_e1 = ef; dyn_left1 = items; }
                        ++dyn_n;
                    }
                    __syncwarp();
                }
                // hand the next items to the slots of this batch, the older job's first
                const unsigned mv = __ballot_sync(FULL, valid);
                const int rank = __popc(mv & lt), nreq = __popc(mv);
                const int a0 = dyn_n > 0 ? dyn_left0 : 0, a1 = dyn_n > 1 ? dyn_left1 : 0;
                bool push = false; int skipped = -1;
                if (valid) {
                    int ent = -1, left = 0, rr = 0;
                    if (rank < a0) { ent = dyn_e0; left = dyn_left0; rr = rank; }
                    else if (rank < a0 + a1) { ent = dyn_e1; left = dyn_left1; rr = rank - a0; }
                    if (ent >= 0) {
                        const int4 J = ring[ent];
                        const int item = (J.z - J.y) * P - left + rr;
                        const int srel = item / P, pix = item - srel * P;
                        const int i = (J.x & 0xffff) + (pix & 15), j = (J.x >> 16) + (pix >> 4);
                        if (i < W && j < H) {
                            PhiloxT<RT_PHILOX_TABLE != 0> rng; if (RT_PHILOX_TABLE) rng.init(p.keys); else rng.init(p.seed);
                            rng.set(uint32_t(j) * uint32_t(W) + uint32_t(i), uint32_t(J.y + srel), 0u);
                            const Ray<T> ray = get_ray<T>(p.cam, i, j, rng);
                            G[L::QO * P + pid] = quad<Q, T>(ray.o.x, ray.o.y, ray.o.z, ray.tm);
                            G[L::QD * P + pid] = quad<Q, T>(ray.d.x, ray.d.y, ray.d.z, T(0));
                            G[L::QT * P + pid] = quad<Q, T>(T(1), T(1), T(1), slot_bits(-1, T(0)));
                            S[L::SAMPLE * P + pid] = (unsigned short)((ent << 14) | srel); S[L::DEPTH * P + pid] = (unsigned short)p.max_depth;
                            B[L::LPIXEL * P + pid] = (unsigned char)pix;
                            push = true;
                        } else {
                            skipped = ent;                             // a pixel position outside the image: nothing to trace
                        }
                    }
                    if (!push) S[L::SAMPLE * P + pid] = (unsigned short)0xffff;
                }
                {
                    const int t0 = min(nreq, a0), t1 = min(nreq - t0, a1);
                    dyn_left0 -= t0; dyn_left1 -= t1;
                    if (dyn_n == 2 && dyn_left1 == 0) dyn_n = 1;
                    if (dyn_n >= 1 && dyn_left0 == 0) { dyn_e0 = dyn_e1; dyn_left0 = dyn_left1; dyn_left1 = 0; --dyn_n; }
                }
                if (__any_sync(FULL, skipped >= 0)) {
                    const int c0 = __popc(__ballot_sync(FULL, skipped == 0)), c1 = __popc(__ballot_sync(FULL, skipped == 1));
                    const int c2 = __popc(__ballot_sync(FULL, skipped == 2)), c3 = __popc(__ballot_sync(FULL, skipped == 3));
                    __syncwarp();
                    if (lane < 4) pend[lane] -= lane == 0 ? c0 : (lane == 1 ? c1 : (lane == 2 ? c2 : c3));
                    progress = true;
                }
                __syncwarp();                                          // the batch has read its entries: the list is refilled in place below
                const unsigned m = __ballot_sync(FULL, push), mi = __ballot_sync(FULL, valid && !push);
                if (push) B[L::QUEUE * P + n_q + __popc(m & lt)] = (unsigned char)pid;
                if (valid && !push) B[L::LEND * P + n_idle + __popc(mi & lt)] = (unsigned char)(pid | 0x80);
                n_q += __popc(m); n_idle += __popc(mi);
                progress = progress || m != 0u;
            }
            } else {
            for (int b = 0; b < n_end; b += 32) {
                // fetch one job ahead of the first slot, no more: jobs a warp holds unstarted are the frame's tail
                if (!exhausted && ((joined >> (8 * newest)) & 0xffu) != 0u && ((joined >> (8 * ((newest + 2) & (R - 1)))) & 0xffu) == unsigned(P)) {
                    int job = 0;
                    if (pend_left > 0) { job = pend_next++; --pend_left; }
                    else {
                        if (lane == 0) job = int(atomicAdd(p.job_counter, unsigned(group * RT_FETCH_MULT)));
                        job = __shfl_sync(FULL, job, 0);
                        pend_next = job + 1; pend_left = group * RT_FETCH_MULT - 1;
                    }
                    if (job >= njobs) {
                        exhausted = true;
#ifdef RT_WARP_TIMES
                        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(wt_exhausted));
#endif
                    }
                    else {
                        // job order: the bulk layers group of tiles by group of tiles (layer by layer over the whole image -- the
                        // open jobs of a warp 30 tile rows apart -- measured 4 % slower), then the short layers one after the
                        // other, so that the frame ends on small jobs
                        int slice, tile;
                        if (job < nbulk) { const int gl = group * p.bulk_layers, g = job / gl, r = job - g * gl; slice = r / group; tile = group * (g * tw + tr) + (r - slice * group); }
                        else { const int r = job - nbulk; slice = r / ntiles_l; const int tl = r - slice * ntiles_l, g = tl / group; tile = group * (g * tw + tr) + (tl - g * group); slice += p.bulk_layers; }
                        const bool none = tile >= ntiles;                       // the last group may be short of tiles: an empty job outside the image
                        const int bx = none ? 0x7fff : (tile % p.tiles_x) * L::tile_w, by = none ? 0 : (tile / p.tiles_x) * L::tile_h;
                        const int s_begin = none ? 0 : p.slice_begin[slice], s_end = none ? 0 : p.slice_begin[slice + 1];
                        newest = (newest + 1) & (R - 1); joined &= ~(0xffu << (8 * newest)); progress = true;
                        if (lane == 0) ring[newest] = make_int4(bx | (by << 16), s_begin, s_end, slice);
                    }
                    __syncwarp();
                }
                const int idx = b + lane;
                bool push = false, idle = false; int pid = 0, join = -1;
                if (idx < n_end) {
                    const int e = B[L::LEND * P + idx];
                    pid = e & 0x7f;
                    const int sw = S[L::SAMPLE * P + pid];
                    int jb = sw >> SH, sidx = sw & DONE;
                    int4 J = ring[jb];
                    int i = (J.x & 0xffff) + (pid & 15), j = (J.x >> 16) + (pid >> 4);
                    if (!(e & 0x80)) {
                        const Q qd = G[L::QD * P + pid], qt = G[L::QT * P + pid];
                        const Vec<T> d = mk<T>(qd.x, qd.y, qd.z), thr = mk<T>(qt.x, qt.y, qt.z);
                        const Vec<T> c = thr * sky<T>(d);
                        Q* dst = p.accum + size_t(J.w) * layer + size_t(j) * size_t(W) + size_t(i);
                        accum_add(dst, c);
                    }
                    bool start = sidx != DONE && J.y + sidx < J.z;
                    if (!start && jb != newest) {
                        // the same pixel position of the next job: set up its accumulator, then its first sample
                        jb = (jb + 1) & (R - 1); J = ring[jb]; sidx = 0;
                        join = jb;
                        i = (J.x & 0xffff) + (pid & 15); j = (J.x >> 16) + (pid >> 4);
                        const bool in_image = i < W && j < H;
                        if (in_image) {
                            Q* dst = p.accum + size_t(J.w) * layer + size_t(j) * size_t(W) + size_t(i);
                            Q v;
                            if (p.accumulate) { v = *dst; } else { v.x = 0; v.y = 0; v.z = 0; v.w = 0; }
                            v.w += decltype(v.x)(J.z - J.y);
                            *dst = v;
                        }
                        start = in_image && J.y < J.z && p.max_depth > 0;      // max_depth <= 0: black (src/camera_cpu.h:12-13)
                    }
                    if (start) {
                        const int s = J.y + sidx;
                        PhiloxT<RT_PHILOX_TABLE != 0> rng; if (RT_PHILOX_TABLE) rng.init(p.keys); else rng.init(p.seed);
                        rng.set(uint32_t(j) * uint32_t(W) + uint32_t(i), uint32_t(s), 0u);
                        const Ray<T> ray = get_ray<T>(p.cam, i, j, rng);
                        G[L::QO * P + pid] = quad<Q, T>(ray.o.x, ray.o.y, ray.o.z, ray.tm);
                        G[L::QD * P + pid] = quad<Q, T>(ray.d.x, ray.d.y, ray.d.z, T(0));
                        G[L::QT * P + pid] = quad<Q, T>(T(1), T(1), T(1), slot_bits(-1, T(0)));
                        S[L::SAMPLE * P + pid] = (unsigned short)((jb << SH) | (sidx + 1)); S[L::DEPTH * P + pid] = (unsigned short)p.max_depth;
                        push = true;
                    } else {
                        S[L::SAMPLE * P + pid] = (unsigned short)((jb << SH) | DONE);
                        idle = true;                                   // waits in the END list for the next job, or for the end of the frame
                    }
                }
                __syncwarp();                                          // the batch has read its entries: the list is refilled in place below
                const unsigned m = __ballot_sync(FULL, push), mi = __ballot_sync(FULL, idle);
                if (push) B[L::QUEUE * P + n_q + __popc(m & lt)] = (unsigned char)pid;
                if (idle) B[L::LEND * P + n_idle + __popc(mi & lt)] = (unsigned char)(pid | 0x80);
                n_q += __popc(m); n_idle += __popc(mi);
                if (__any_sync(FULL, join >= 0)) {
#pragma unroll
                    for (int e = 0; e < R; ++e) joined += unsigned(__popc(__ballot_sync(FULL, join == e))) << (8 * e);
                    progress = true;
                }
                progress = progress || m != 0u;
            }
            }
            n_end = n_idle;
            __syncwarp();
            if (n_q == 0) { if (progress) continue; break; }   // nothing in flight, nothing left to start: the frame is done
#ifdef RT_WARP_TIMES
            if (exhausted) { ++wt_rounds_tail; wt_nq_tail += n_q; } else { ++wt_rounds; wt_nq += n_q; }
#endif

            // ------------------------------------------------------------------ BIG stage
            // every queued ray against the spheres kept out of the tree (the r = 1000 ground), in
            // double, with a full warp: gives the tree walk a tight t_max from its first node on
            if (sv.n_big > 0) {
                #pragma unroll 1                        // (stage loops stay rolled: code size, see trav_leaf)
                for (int b = 0; b < n_q; b += 32) {
                    const int idx = b + lane;
                    if (idx < n_q) {
                        const int pid = B[L::QUEUE * P + idx];
                        Ray<T> ray;
                        const Q qo = G[L::QO * P + pid], qd = G[L::QD * P + pid];
                        ray.o = mk<T>(qo.x, qo.y, qo.z); ray.d = mk<T>(qd.x, qd.y, qd.z); ray.tm = qo.w;
                        T tb; int bslot;
                        trav_big<T>(sv, ray, T(0.001), T(INFINITY), tb, bslot);
                        G[L::QD * P + pid].w = tb; G[L::QT * P + pid].w = slot_bits(bslot, T(0));
                    }
                }
                __syncwarp();
            }

            // ---------------------------------------------------------------- TRACE stage
            {
                int head = 0, pid = 0;
                Ray<T> ray; ray.o = mk<T>(0, 0, 0); ray.d = mk<T>(0, 0, 1); ray.tm = 0;
                Trav<T> tv; tv.cur = RT_SENTINEL; tv.sp = 0; tv.tos = RT_SENTINEL; tv.tos_t = -INFINITY; tv.tbest = 0; tv.best_slot = -1; tv.qa = 1; slab_clear<T>(tv.slab);
                for (;;) {
                    const bool isN = tv.cur >= 0;
                    const bool isP = tv.cur == RT_SENTINEL;
                    const int nN = __popc(__ballot_sync(FULL, isN));
                    const unsigned mP = __ballot_sync(FULL, isP);
                    const int nP = __popc(mP), nL = 32 - nN - nP;
                    const int avail = n_q - head;
                    if (nP == 32 && avail == 0) break;
                    RT_STAT(0, 1); RT_STAT(1, nN); RT_STAT(2, nL); RT_STAT(3, nP); RT_STAT(4, avail > 0);

                    if (avail > 0 && (nP >= th_p || (nN | nL) == 0)) {
                        // ---- pull: idle lanes take the next rays of the queue
                        const int rank = __popc(mP & lt);
                        RT_STAT(5, 1); RT_STAT(6, min(nP, avail)); RT_STAT(7, nN + nL);
                        if (isP && rank < avail) {
                            pid = B[L::QUEUE * P + head + rank];
                            const Q qo = G[L::QO * P + pid], qd = G[L::QD * P + pid];
                            ray.o = mk<T>(qo.x, qo.y, qo.z); ray.d = mk<T>(qd.x, qd.y, qd.z); ray.tm = qo.w;
                            tv.tbest = T(INFINITY); tv.best_slot = -1;
                            if (sv.n_big > 0) {
                                tv.tbest = qd.w; tv.best_slot = bits_slot(G[L::QT * P + pid].w);
                            }
                            trav_setup<T>(sv, ray, tv);
                            if (tv.cur == RT_SENTINEL) { G[L::QD * P + pid].w = tv.tbest; G[L::QT * P + pid].w = slot_bits(tv.best_slot, T(0)); }   // no tree
                            ++my_rays;                                   // one ray_color entry with depth > 0
                        }
                        head += min(nP, avail);
                    } else if (nL >= th_l || nL > nN) {
                        RT_STAT(8, 1); RT_STAT(9, nL); RT_STAT(10, nN); RT_STAT(11, nP);
                        if (!isN && !isP) {
                            trav_leaf<T>(sv, st.sa, st.sb, st.meta, ray, T(0.001), tv, stack);
                            if (tv.cur == RT_SENTINEL) { G[L::QD * P + pid].w = tv.tbest; G[L::QT * P + pid].w = slot_bits(tv.best_slot, T(0)); }
                        }
                    } else {
                        // node visits, RT_NODE_VISITS per head count, repeated while most of the lanes that started them
                        // still have a node; a lane whose walk ends hands in its result after the loop
                        int now = nN;
                        RT_STAT(12, 1); RT_STAT(13, nN); RT_STAT(14, nL); RT_STAT(15, nP);
                        do {
                            RT_STAT(16, 1); RT_STAT(17, now);
                            if (WIDE) {                    // one 128-byte record = two levels of the tree
                                if (tv.cur >= 0) trav_node4<T>(sv.nodes4, ray, T(0.001), tv, stack);
                            } else {
#pragma unroll
                                for (int v = 0; v < RT_NODE_VISITS; ++v) {
                                    if (STAGED) { if (tv.cur >= 0) trav_node<T>(st.nodes, ray, T(0.001), tv, stack); }
                                    else { if (tv.cur >= 0) trav_node_top<T>(st.nodes, top, n_top, ray, T(0.001), tv, stack); }
                                }
                            }
                            now = __popc(__ballot_sync(FULL, tv.cur >= 0));
                        } while (now > 0 && now * 8 >= nN * th_n);
                        if (isN && tv.cur == RT_SENTINEL) { G[L::QD * P + pid].w = tv.tbest; G[L::QT * P + pid].w = slot_bits(tv.best_slot, T(0)); }
                    }
                }
            }
            __syncwarp();

            // ------------------------------------------------------------- CLASSIFY stage
            // regroup the traced rays by what they hit (warp ballot + prefix popcount): misses go to the END list; hits go to
            // ONE shading list, opaque ones (lambertian + metal: they share the ball sampler) filled from its front, glass from
            // its back.  With RT_MERGE_HITS the HIT stage walks that list as one sequence, so the last few opaque hits and the
            // few glass hits of a round share a batch (and its surface / ray load-store code) instead of taking one batch each.
            #pragma unroll 1                        // (stage loops stay rolled: code size, see trav_leaf)
            for (int b = 0; b < n_q; b += 32) {
                const int idx = b + lane;
                int key = -1, pid = 0;
                if (idx < n_q) {
                    pid = B[L::QUEUE * P + idx];
                    const int slot = bits_slot(G[L::QT * P + pid].w);
                    if (slot < 0) key = 2;
                    else {
                        const int m = slot < sv.n_slots ? st.meta[slot] : sv.big_meta[slot - sv.n_slots];
                        key = RT_META_TYPE(m) == 2 ? 1 : 0;
                    }
                }
                const unsigned m0 = __ballot_sync(FULL, key == 0), m1 = __ballot_sync(FULL, key == 1), m2 = __ballot_sync(FULL, key == 2);
                if (key == 0) B[L::LOPAQUE * P + n_opq + __popc(m0 & lt)] = (unsigned char)pid;
                if (key == 1) B[L::LOPAQUE * P + (P - 1) - n_gls - __popc(m1 & lt)] = (unsigned char)pid;
                if (key == 2) B[L::LEND * P + n_end + __popc(m2 & lt)] = (unsigned char)pid;     // miss: sky is added in END
                n_opq += __popc(m0); n_gls += __popc(m1); n_end += __popc(m2);
            }
            n_q = 0;
            __syncwarp();

            // ------------------------------------------------------------------ HIT stage
            // rec.mat->scatter (src/camera_cpu.h:18-20), opaque hits first, then glass
#if RT_MERGE_HITS
            for (int l = 0; l < 1; ++l) {
                const unsigned char* list = B + L::LOPAQUE * P;
                const int n = n_opq + n_gls;
#else
            for (int l = 0; l < 2; ++l) {
                const unsigned char* list = B + L::LOPAQUE * P;
                const int n = l ? n_gls : n_opq;
#endif
                #pragma unroll 1                        // (stage loops stay rolled: code size, see trav_leaf)
                for (int b = 0; b < n; b += 32) {
                    const int idx = b + lane;
                    bool cont = false, ended = false; int pid = 0;
                    if (idx < n) {
#if RT_MERGE_HITS
                        pid = list[idx < n_opq ? idx : (P - 1) - (idx - n_opq)];
#else
                        pid = list[l ? (P - 1) - idx : idx];
#endif
                        const int sw = S[L::SAMPLE * P + pid];
                        const int4 J = ring[sw >> SH];
                        const int pix = L::dyn ? int(B[L::LPIXEL * P + pid]) : pid;            // dynamic mode: the slot's current pixel
                        const int i = (J.x & 0xffff) + (pix & 15), j = (J.x >> 16) + (pix >> 4);
                        Ray<T> ray;
                        const Q qo = G[L::QO * P + pid], qd = G[L::QD * P + pid], qt = G[L::QT * P + pid];
                        ray.o = mk<T>(qo.x, qo.y, qo.z); ray.d = mk<T>(qd.x, qd.y, qd.z); ray.tm = qo.w;
                        T t = qd.w;
                        const int slot = bits_slot(qt.w), s = J.y + (sw & DONE) + (L::dyn ? 1 : 0);   // (static mode keeps "next sample" here)
                        int depth = S[L::DEPTH * P + pid];
                        const Surface<T> sf = surface_at<T>(sv, st.sa, st.sb, st.sc, st.meta, ray, t, slot);
                        PhiloxT<RT_PHILOX_TABLE != 0> rng; if (RT_PHILOX_TABLE) rng.init(p.keys); else rng.init(p.seed);
                        rng.set(uint32_t(j) * uint32_t(W) + uint32_t(i), uint32_t(s - 1), uint32_t(p.max_depth - depth + 1));   // b-th hit: stage b
                        Vec<T> att, dnew;
                        if (!scatter<T>(sf.type, sf.albedo, sf.param, ray.d, sf.normal, sf.front_face, rng, att, dnew)) {
                            ended = true;                                       // absorbed: contributes 0 (:20)
                        } else if (--depth == 0) {
                            ended = true;                                       // depth exhausted: contributes 0 (:12-13)
                        } else {
                            cont = true;
                            ray.o = sf.p; ray.d = dnew;                         // time is inherited (src/material.h:22)
                            G[L::QO * P + pid] = quad<Q, T>(ray.o.x, ray.o.y, ray.o.z, ray.tm);
                            G[L::QD * P + pid] = quad<Q, T>(ray.d.x, ray.d.y, ray.d.z, T(0));
                            G[L::QT * P + pid] = quad<Q, T>(qt.x * att.x, qt.y * att.y, qt.z * att.z, qt.w);
                            S[L::DEPTH * P + pid] = (unsigned short)depth;
                        }
                    }
                    const unsigned mc = __ballot_sync(FULL, cont), me = __ballot_sync(FULL, ended);
                    if (cont) B[L::QUEUE * P + n_q + __popc(mc & lt)] = (unsigned char)pid;
                    if (ended) B[L::LEND * P + n_end + __popc(me & lt)] = (unsigned char)(pid | 0x80);
                    n_q += __popc(mc); n_end += __popc(me);
                }
            }
            n_opq = 0; n_gls = 0;
            __syncwarp();
        }
    }
    for (int o = 16; o > 0; o >>= 1) my_rays += __shfl_xor_sync(FULL, my_rays, o);
    if (lane == 0 && my_rays) atomicAdd(p.ray_counter, my_rays);
#ifdef RT_POOL_STATS
    if (lane == 0) for (int k = 0; k < 24; ++k) if (stat[k]) atomicAdd(p.ray_counter + 1 + k, stat[k]);
#endif
#ifdef RT_WARP_TIMES
    if (lane == 0) {
        unsigned long long wt_exit;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(wt_exit));
        unsigned long long* o = p.ray_counter + 32 + 8 * (size_t(blockIdx.x) * 32 + warp);
        o[0] = wt_first; o[1] = wt_exhausted; o[2] = wt_exit; o[3] = wt_rounds; o[4] = wt_nq; o[5] = wt_rounds_tail; o[6] = wt_nq_tail; o[7] = 1;
    }
#endif
}

}  // namespace rt
