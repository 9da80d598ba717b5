// rt_bvh_host.cpp -- SAH BVH builder on the host (see rt_bvh_host.h).
#include "rt_bvh_host.h"

#include <algorithm>
#include <atomic>
#ifdef RT_BVH_TIMING
#include <chrono>
#include <cstdio>
#endif
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <numeric>
#include <thread>

namespace rt {

namespace {

float down(double v) { float f = float(v); if (double(f) > v) f = std::nextafterf(f, -INFINITY); return f; }
float up(double v)   { float f = float(v); if (double(f) < v) f = std::nextafterf(f, INFINITY); return f; }

struct Box {
    float lo[3] = { INFINITY, INFINITY, INFINITY }, hi[3] = { -INFINITY, -INFINITY, -INFINITY };
    void grow(const HostBox& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void grow(const Box& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    double area() const {
        const double dx = double(hi[0]) - lo[0], dy = double(hi[1]) - lo[1], dz = double(hi[2]) - lo[2];
        if (dx < 0) return 0;
        return 2.0 * (dx * dy + dy * dz + dz * dx);
    }
};

struct TmpNode { Box box; int left = -1, right = -1; int first = 0, count = 0; };   // leaf iff left < 0

// Top-down SAH builder over index arrays that are sorted ONCE per axis -- by (centroid, index), a total order -- and kept sorted
// by stable partitions: a range [first, first + count) holds the same primitives in all three arrays, each in its axis' order,
// so the exact sweep of a node is three linear passes and the whole build O(n log n).  (The first version re-sorted every
// range four times per node: 0.6 ms for the 485 spheres of the final scene, on every rt_upload_scene.)  Same splits, same leaf
// order -- a leaf's slots are in z order, as the last sweep of the old builder left them -- hence the same tree, bit for bit.
//
// Large scenes are built by several host threads: a node works on ITS range of every array only (the scratch arrays are addressed
// through the range as well, a primitive's right_side byte belongs to the one range that holds it), so the two halves of a split
// are independent and the right one is handed to another thread while there are threads to spare and the range is large.  The
// tree does not depend on who built which part of it: node ids differ, the records are numbered breadth-first afterwards.
struct Builder {
    const std::vector<HostBox>& boxes;
    int max_leaf;
    std::vector<int> idx;                 // out: slot -> primitive
    std::vector<int> ord[3], tmp;
    std::vector<unsigned char> right_side;
    std::vector<TmpNode> nodes;           // sized for the worst case up front (2 n - 1): ids come from a counter, nothing moves
    std::atomic<int> n_nodes{0};
    std::atomic<int> spare_threads{0};
    std::vector<float> cen[3];
    std::vector<double> suffix;
    static const int PAR_MIN = 16384;     // ranges below this are not worth a thread

    static int host_threads(size_t n) {
        if (n < size_t(2 * PAR_MIN)) return 1;
        if (const char* e = std::getenv("RT_BVH_BUILD_THREADS")) return std::max(1, std::min(64, std::atoi(e)));
        return int(std::min(16u, std::max(1u, std::thread::hardware_concurrency())));
    }

    Builder(const std::vector<HostBox>& b, int ml) : boxes(b), max_leaf(ml) {
        const size_t n = b.size();
        const int threads = host_threads(n);
        spare_threads = threads - 1;
        idx.resize(n); tmp.resize(n); right_side.assign(n, 0);
        auto sort_axis = [this, &b, n](int a) {
            cen[a].resize(n);
            for (size_t i = 0; i < n; ++i) cen[a][i] = 0.5f * (b[i].lo[a] + b[i].hi[a]);
            ord[a].resize(n);
            const std::vector<float>& c = cen[a];
            if (n < 8192) {
                std::iota(ord[a].begin(), ord[a].end(), 0);
                std::sort(ord[a].begin(), ord[a].end(), [&c](int x, int y) { return c[x] < c[y] || (c[x] == c[y] && x < y); });
                return;
            }
            // the same total order by a stable radix sort of the centroids' bits (monotone key: sign flipped for positives, all
            // bits for negatives; -0 counts as +0, as it does for operator<): four passes of 8 bits over (key, index) pairs that
            // start in index order, so equal centroids stay in index order
            std::vector<uint64_t> p0(n), p1(n);
            for (size_t i = 0; i < n; ++i) {
                const float f = c[i] == 0.0f ? 0.0f : c[i];
                uint32_t u; std::memcpy(&u, &f, 4);
                u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
                p0[i] = (uint64_t(u) << 32) | uint32_t(i);
            }
            for (int pass = 0; pass < 4; ++pass) {
                const int shift = 32 + 8 * pass;
                size_t hist[257] = { 0 };
                for (size_t i = 0; i < n; ++i) ++hist[((p0[i] >> shift) & 255) + 1];
                for (int k = 0; k < 256; ++k) hist[k + 1] += hist[k];
                for (size_t i = 0; i < n; ++i) p1[hist[(p0[i] >> shift) & 255]++] = p0[i];
                p0.swap(p1);
            }
            for (size_t i = 0; i < n; ++i) ord[a][i] = int(uint32_t(p0[i]));
        };
        if (threads > 1) {
            std::thread ty(sort_axis, 1), tz(sort_axis, 2);
            sort_axis(0);
            ty.join(); tz.join();
        } else {
            for (int a = 0; a < 3; ++a) sort_axis(a);
        }
        suffix.resize(n + 1);
        nodes.resize(2 * n);
    }

    // the primitives with right_side == 0 to the front of every array's range, order kept
    void split_ranges(int first, int count) {
        int* t = tmp.data() + first;
        for (int a = 0; a < 3; ++a) {
            int* o = ord[a].data() + first;
            int nl = 0, nr = 0;
            for (int i = 0; i < count; ++i) { if (right_side[o[i]]) t[nr++] = o[i]; else o[nl++] = o[i]; }
            std::memcpy(o + nl, t, size_t(nr) * sizeof(int));
        }
    }

    int build(int first, int count) {
        const int me = n_nodes.fetch_add(1);
        double* const suffix = this->suffix.data() + first;         // this range's part of the scratch
        Box bb;
        for (int i = first; i < first + count; ++i) bb.grow(boxes[ord[0][i]]);
        nodes[me].box = bb;
        nodes[me].first = first; nodes[me].count = count;
        if (count == 1) { idx[first] = ord[0][first]; return me; }

        // exact SAH sweep on every axis for small ranges, 32-bin SAH above
        const double inv_area = 1.0 / std::max(bb.area(), 1e-300);
        double best_cost = std::numeric_limits<double>::infinity();
        int best_axis = -1, best_split = -1;      // split = number of prims on the left in the axis' order
        float best_plane = 0; bool binned = count > 4096;
        if (!binned) {
            for (int a = 0; a < 3; ++a) {
                const int* o = ord[a].data() + first;
                Box r;
                for (int i = count - 1; i >= 1; --i) { r.grow(boxes[o[i]]); suffix[i] = r.area(); }
                Box l;
                for (int i = 1; i < count; ++i) {
                    l.grow(boxes[o[i - 1]]);
                    const double cost = 1.0 + (l.area() * i + suffix[i] * (count - i)) * inv_area;
                    if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = i; }
                }
            }
        } else {
            const int NB = 32;
            for (int a = 0; a < 3; ++a) {
                const int* o = ord[a].data() + first;
                const float cmin = cen[a][o[0]], cmax = cen[a][o[count - 1]];
                if (!(cmax > cmin)) continue;
                Box bin[NB]; int cnt[NB] = { 0 };
                const float scale = NB / (cmax - cmin);
                for (int i = 0; i < count; ++i) {
                    int k = std::min(NB - 1, std::max(0, int((cen[a][o[i]] - cmin) * scale)));
                    bin[k].grow(boxes[o[i]]); cnt[k]++;
                }
                double ra[NB]; int rc[NB];
                Box r; int c = 0;
                for (int k = NB - 1; k >= 1; --k) { r.grow(bin[k]); c += cnt[k]; ra[k] = r.area(); rc[k] = c; }
                Box l; c = 0;
                for (int k = 1; k < NB; ++k) {
                    l.grow(bin[k - 1]); c += cnt[k - 1];
                    if (c == 0 || rc[k] == 0) continue;
                    const double cost = 1.0 + (l.area() * c + ra[k] * rc[k]) * inv_area;
                    if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = c; best_plane = cmin + k / scale; }
                }
            }
        }
        const double leaf_cost = double(count);
        if (count <= max_leaf && (best_axis < 0 || leaf_cost <= best_cost)) {               // leaf: slots in z order
            std::memcpy(idx.data() + first, ord[2].data() + first, size_t(count) * sizeof(int));
            return me;
        }
        if (best_axis < 0) {                                                                // all centroids equal
            best_axis = 0; best_split = count / 2; binned = false;
        }
        const int* o = ord[best_axis].data() + first;
        if (binned) {                       // left = centroids below the plane (they are a prefix of the axis' order)
            int k = 0;
            while (k < count && cen[best_axis][o[k]] < best_plane) ++k;
            best_split = k;
            if (best_split == 0 || best_split == count) best_split = count / 2;
        }
        for (int i = 0; i < count; ++i) right_side[o[i]] = i >= best_split;
        split_ranges(first, count);
        int l, r;
        if (std::min(best_split, count - best_split) >= PAR_MIN && spare_threads.fetch_sub(1) > 0) {
            std::thread other([this, &r, first, count, best_split]() { r = build(first + best_split, count - best_split); });
            l = build(first, best_split);
            other.join();
            spare_threads.fetch_add(1);
        } else {
            if (std::min(best_split, count - best_split) >= PAR_MIN) spare_threads.fetch_add(1);      // (none was free)
            l = build(first, best_split);
            r = build(first + best_split, count - best_split);
        }
        nodes[me].left = l; nodes[me].right = r;
        return me;
    }
};

int leaf_ref(int first, int count) { return ~((first << 3) | (count - 1)); }

}  // namespace

HostBox sphere_box(const double c1[3], const double cv[3], double radius, bool moving) {
    HostBox b;
    const double r = std::fabs(radius);
    for (int a = 0; a < 3; ++a) {
        double lo = c1[a] - r, hi = c1[a] + r;
        if (moving) { lo = std::min(lo, c1[a] + cv[a] - r); hi = std::max(hi, c1[a] + cv[a] + r); }
        const double pad = std::ldexp(std::max(std::max(std::fabs(lo), std::fabs(hi)), r), -20) + 1e-30;
        b.lo[a] = down(lo - pad);
        b.hi[a] = up(hi + pad);
    }
    return b;
}

void build_bvh_sah(const std::vector<HostBox>& boxes, int max_leaf, HostBVH& out) {
    out = HostBVH();
    const int n = int(boxes.size());
    if (n == 0) return;
    max_leaf = std::min(8, std::max(1, max_leaf));
#ifdef RT_BVH_TIMING
    auto t0 = std::chrono::steady_clock::now();
#endif
    Builder b(boxes, max_leaf);
#ifdef RT_BVH_TIMING
    auto t1 = std::chrono::steady_clock::now();
#endif
    b.build(0, n);
#ifdef RT_BVH_TIMING
    auto t2 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "sorts %.1f ms, build %.1f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count(), std::chrono::duration<double, std::milli>(t2 - t1).count());
#endif
    b.nodes.resize(size_t(b.n_nodes.load()));
    out.order = b.idx;
    // a leaf range longer than 8 cannot be encoded; the builder never makes one (count <= max_leaf)
    if (b.nodes[0].left < 0) {                       // whole scene is one leaf
        out.root_ref = leaf_ref(0, b.nodes[0].count);
        out.depth = 1;
        return;
    }
    // BFS renumbering of inner nodes so the top of the tree is a prefix of the node array
    std::vector<int> bfs, inner_id(b.nodes.size(), -1), level(b.nodes.size(), 0);
    bfs.push_back(0);
    for (size_t q = 0; q < bfs.size(); ++q) {
        const TmpNode& nd = b.nodes[bfs[q]];
        inner_id[bfs[q]] = int(q);
        for (int ch : { nd.left, nd.right })
            if (b.nodes[ch].left >= 0) { level[ch] = level[bfs[q]] + 1; bfs.push_back(ch); }
    }
    out.n_inner = int(bfs.size());
    out.nodes.assign(size_t(out.n_inner) * 16, 0.0f);
    int depth = 0;
    for (size_t q = 0; q < bfs.size(); ++q) depth = std::max(depth, level[bfs[q]] + 2);
    auto write_records = [&](size_t q0, size_t q1) {
      for (size_t q = q0; q < q1; ++q) {
        const TmpNode& nd = b.nodes[bfs[q]];
        const TmpNode& L = b.nodes[nd.left];
        const TmpNode& R = b.nodes[nd.right];
        float* f = &out.nodes[q * 16];
        // record layout (rt_trace.cuh): xy of the left box | xy of the right box | z of both | refs
        f[0] = L.box.lo[0]; f[1] = L.box.lo[1]; f[2] = L.box.hi[0]; f[3] = L.box.hi[1];
        f[4] = R.box.lo[0]; f[5] = R.box.lo[1]; f[6] = R.box.hi[0]; f[7] = R.box.hi[1];
        f[8] = L.box.lo[2]; f[9] = R.box.lo[2]; f[10] = L.box.hi[2]; f[11] = R.box.hi[2];
        const int lref = L.left >= 0 ? inner_id[nd.left] : leaf_ref(L.first, L.count);
        const int rref = R.left >= 0 ? inner_id[nd.right] : leaf_ref(R.first, R.count);
        std::memcpy(&f[12], &lref, 4);
        std::memcpy(&f[13], &rref, 4);
      }
    };
    const int writers = Builder::host_threads(size_t(n));
    if (writers > 1) {
        std::vector<std::thread> th;
        const size_t per = (bfs.size() + writers - 1) / writers;
        for (int t = 0; t < writers; ++t) th.emplace_back(write_records, std::min(bfs.size(), t * per), std::min(bfs.size(), (t + 1) * per));
        for (auto& t : th) t.join();
    } else {
        write_records(0, bfs.size());
    }
    out.root_ref = 0;
    out.depth = depth;
}

int bvh_depth(const HostBVH& b) { return b.depth; }

}  // namespace rt
