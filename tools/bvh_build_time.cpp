// bvh_build_time.cpp -- times the host SAH builder (rt_bvh_host.cpp) on synthetic boxes and prints a hash of the tree, to compare builds
// (other thread counts: RT_BVH_BUILD_THREADS; an older builder: link its source instead).
//   g++ -O3 -std=c++17 -pthread tools/bvh_build_time.cpp raytracingproject_b200/csrc/rt_bvh_host.cpp -o build_bvhtime/t && build_bvhtime/t 1000000 0 2
#include "../raytracingproject_b200/csrc/rt_bvh_host.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
static unsigned long long fnv(const void* p, size_t n, unsigned long long h) { const unsigned char* c = (const unsigned char*)p; for (size_t i = 0; i < n; ++i) { h ^= c[i]; h *= 1099511628211ull; } return h; }
int main(int argc, char** argv) {
    int n = argc > 1 ? atoi(argv[1]) : 1000000;
    int mode = argc > 2 ? atoi(argv[2]) : 0;     // 0 planar random, 1 many duplicates / ties / negative zero, 2 clustered 3D
    int reps = argc > 3 ? atoi(argv[3]) : 2;
    std::mt19937 g(1 + mode); std::uniform_real_distribution<float> u(-500, 500), r(0.1f, 0.3f);
    std::vector<rt::HostBox> b(n);
    for (int i = 0; i < n; ++i) {
        auto& x = b[i];
        float c[3] = {u(g), 0.2f, u(g)}; float rr = r(g);
        if (mode == 1) { c[0] = float(int(c[0] / 50)) * 50; c[2] = (i % 7 == 0) ? 0.0f : float(int(c[2] / 100)) * 100; rr = 0.25f; if (i % 11 == 0) { c[0] = -0.0f; rr = 0.0f; } }
        if (mode == 2) { c[1] = u(g) * 0.1f; if (i % 3) { c[0] = c[0] * 0.01f + 100; c[2] = c[2] * 0.01f - 200; } }
        for (int a = 0; a < 3; ++a) { x.lo[a] = c[a] - rr; x.hi[a] = c[a] + rr; }
    }
    for (int rep = 0; rep < reps; ++rep) {
        rt::HostBVH out;
        auto t0 = std::chrono::steady_clock::now();
        rt::build_bvh_sah(b, 4, out);
        double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        unsigned long long h = fnv(out.nodes.data(), out.nodes.size() * 4, 1469598103934665603ull);
        h = fnv(out.order.data(), out.order.size() * 4, h);
        printf("n %d mode %d build %.1f ms depth %d inner %d root %d hash %016llx\n", n, mode, ms, out.depth, out.n_inner, out.root_ref, h);
    }
}
