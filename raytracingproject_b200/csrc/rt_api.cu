// rt_api.cu -- the C ABI of librt_b200.so (include/rt_b200.h): context, scene upload, render,
// readback and the parity hooks.  Host logic only; every kernel lives in rt_kernels_*.cu,
// rt_frame.cu and rt_lbvh.cu.  No CPU fallback anywhere: a CUDA failure is an error.
#include "../../include/rt_b200.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include <nvtx3/nvToolsExt.h>          // header-only: ranges show up in an Nsight Systems timeline, no-ops otherwise

#include "rt_bvh_host.h"
#include "rt_demo.h"
#include "rt_frame.h"
#include "rt_kernels.cuh"
#include "rt_lbvh.h"
#include "rt_scene.h"
#include "rt_static.h"

#define RT_JOB_COUNTERS 8                // dynamic sharding: job counters used round-robin by frame number (the root clears them 4 frames ahead)

using namespace rt;

static_assert(sizeof(rt_sphere) == 64, "rt_sphere layout");
static_assert(sizeof(rt_material) == 40, "rt_material layout");
static_assert(sizeof(rt_ray) == 56, "rt_ray layout");
static_assert(sizeof(rt_hit) == 64, "rt_hit layout");
static_assert(sizeof(HitRec) == sizeof(rt_hit), "HitRec mirrors rt_hit");
static_assert(sizeof(MatRec) == sizeof(rt_material), "MatRec mirrors rt_material");

namespace {

const size_t RT_DEBUG_WORDS = 32 + 8 * 32 * 256;     // debug counters + 8 words per warp (up to 256 SMs x 32 warps)
const size_t RT_MAX_BIG = 16;          // spheres kept out of the tree (tested by every ray, in double)

std::string g_create_error;

template <typename T> struct DevBuf {
    T* p = nullptr; size_t n = 0;
    cudaError_t resize(size_t count) {
        if (count <= n && p) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; n = 0;
        if (count == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc(&p, count * sizeof(T));
        if (e == cudaSuccess) n = count;
        return e;
    }
    // scratch that lives for one call: stream-ordered allocation from the device's memory pool, which keeps freed blocks
    // (rt_create raises its release threshold), so the 17 temporaries of a 1 M-sphere upload cost no cudaMalloc / cudaFree
    // after the first call (each of those takes a driver lock that eight ranks of a node fight over)
    cudaStream_t pool_stream = nullptr; bool pooled = false;
    cudaError_t scratch(size_t count, cudaStream_t st) {
        release();
        if (count == 0) return cudaSuccess;
        cudaError_t e = cudaMallocAsync(&p, count * sizeof(T), st);
        if (e == cudaSuccess) { n = count; pooled = true; pool_stream = st; }
        return e;
    }
    void release() { if (p) { if (pooled) cudaFreeAsync(p, pool_stream); else cudaFree(p); } p = nullptr; n = 0; pooled = false; }
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }             // scratch buffers on an early error return
};

}  // namespace

struct rt_ctx {
    int device = 0, sm_count = 0;
    size_t smem_optin = 0, smem_per_sm = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = { nullptr, nullptr, nullptr, nullptr };
    std::string err;

    // scene (SoA, both precisions; see rt_device.cuh for the slot layout)
    bool has_scene = false;
    int n_spheres = 0, n_slots = 0, n_big = 0, n_nodes = 0, root_ref = RT_SENTINEL, depth = 0;
    DevBuf<float4> nodes, nodes4, sa32, sb32, sc32;   // nodes4: 4-wide records of the same tree (null unless the tree is walked in global memory)
    bool wide = false, wide_forced = false;
    DevBuf<int> slot_sphere;                      // slot -> index of its sphere in the caller's array (for rt_update_spheres)
    std::vector<int> big_ids;                     // the spheres kept out of the tree, in list order
    double big_radius = 100.0;
    int top_ordered = 0;                          // nodes [0, top_ordered) are the first levels of the tree in breadth-first order (rt_scene.cu)
    DevBuf<double4> sa64, sb64, sc64, big_a, big_b, big_c;
    DevBuf<int> meta, big_meta;
    double build_ms = 0, upload_ms = 0;

    // frame
    int W = 0, H = 0, slices = 0, precision = -1;
    int slice_begin[RT_MAX_SLICES + 1] = {};   // sample ranges of the layers of the render in flight
    int bulk_layers = 0;                       // how many of them are full-size (the rest halve down: the tail of the frame)
    bool folded = false;                    // rt_fold_slices ran since the last render: only layer 0 (here and in the peers) holds data
    long long samples_in_buffer = 0;
    DevBuf<unsigned char> accum;            // Quad<T>[slices][H][W]
    DevBuf<unsigned char> reduced;          // Quad<T>[H][W] (after peer reduce, for rt_accum_buffer consumers)
    DevBuf<unsigned int> job_counter;
    DevBuf<unsigned long long> ray_counter;
    DevBuf<unsigned char> rgb8;
    DevBuf<double> sum;
    DevBuf<float> pfm;                      // 8f-2 views of the resolved frame
    DevBuf<uchar4> rgba8;
    DevBuf<char> text;
    DevBuf<unsigned> enc_sums;
    DevBuf<unsigned long long> enc_offsets;   // [blocks] + total at the end
    bool render_timed = false;
    int* host_words = nullptr;              // 16 ints of pinned host memory the device can store to (small results without a device-to-host copy)
    // split read-back (rt_readback_begin / rt_readback_end): the frame's bytes travel on a stream of their own
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_done = nullptr;
    bool copy_pending = false;              // begun and not ended: rgb8 / text must not be overwritten before copy_done
    size_t copy_text_len = 0;

    // peers (CUDA IPC)
    PeerList peers = {};
    void* peer_mapped[RT_MAX_PEERS] = {};

    // tile-interleaved sharding: the last render's share of the tiles, and the image hand-over
    TileOwner tiles = { 0, 0, 0, 0, 0, 0, 0 };
    bool frame_exported = false, tiles_pending = false;   // root: its image is mapped by peers; a tile-sharded render awaits their bytes
    unsigned frame_target = 0;                            // root: counter value once every peer has delivered the frame in flight
    void* frame_remote = nullptr;                         // peer: the root's image (rgb8 + counter), mapped through CUDA IPC
    int frame_remote_W = 0, frame_remote_H = 0;
    DevBuf<int> frame_timeout;                            // root: set by the wait kernel when the peers' tiles did not arrive

    // dynamic sharding: the job counters of frames f % RT_JOB_COUNTERS, in the root's memory (own) or mapped from it
    DevBuf<unsigned> jobs_own;
    unsigned* jobs_ptr = nullptr;           // what the kernels advance (own.p on the root)
    void* jobs_mapped = nullptr;            // peer: the CUDA IPC mapping to close
    unsigned jobs_frame = 0;                // frames rendered with shared_jobs since the counters were set up
    bool jobs_root = false;

    rt_stats stats = {};
};

namespace {

// NVTX range over a scope (tracing).  Names: rt_upload_scene / rt_update_spheres / rt_render (enqueue) /
// rt_readback (resolve + encode + copies) / rt_fold_slices / rt_resolve_tiles.
struct NvtxRange { explicit NvtxRange(const char* name) { nvtxRangePushA(name); } ~NvtxRange() { nvtxRangePop(); } };

// A/B knobs read from the environment (set and not "0")
bool env_on(const char* name) { const char* e = std::getenv(name); return e && *e && *e != '0'; }

int fail(rt_ctx* c, int code, const std::string& msg) { if (c) c->err = msg; else g_create_error = msg; return code; }
int cuda_fail(rt_ctx* c, cudaError_t e, const char* what) {
    return fail(c, RT_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
// layers the resolve pass has to read: all of them, or layer 0 alone after rt_fold_slices (every rank folds)
int live_slices(const rt_ctx* c) { return c->folded ? 1 : c->slices; }
// the image buffer: rgb8 [npix * 3, padded for the encoder's 32-bit loads], then the hand-over counter of tile-sharded frames
size_t frame_counter_offset(size_t npix) { return (npix * 3 + 16 + 127) & ~size_t(127); }
cudaError_t ensure_rgb8(rt_ctx* c, size_t npix) { return c->rgb8.resize(frame_counter_offset(npix) + 128); }
PeerList live_peers(const rt_ctx* c) {
    PeerList p = c->peers;
    if (c->folded) for (int i = 0; i < p.n; ++i) p.slices[i] = 1;
    return p;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_fail(ctx, e_, #call); } while (0)

// ---- camera::initialize, src/camera.h:52-85, in double exactly as written -------------------
typedef Vec<double> V3;
V3 v3(const double a[3]) { return mk<double>(a[0], a[1], a[2]); }
V3 hunit(V3 v) { return (1 / std::sqrt(len2(v))) * v; }                       // src/vec3.h:117-119
V3 hcross(V3 u, V3 v) { return mk<double>(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }
V3 hdiv(V3 v, double t) { return (1 / t) * v; }                                // src/vec3.h:101-103

int image_height_of(const rt_camera& c) {
    int h = static_cast<int>(c.image_width / c.aspect_ratio);                  // src/camera.h:53-54
    return h < 1 ? 1 : h;
}

CameraFrame<double> camera_frame(const rt_camera& c) {
    CameraFrame<double> f;
    const double pi = 3.1415926535897932385;                                   // src/rtweekend.h:17
    f.image_width = c.image_width;
    f.image_height = image_height_of(c);
    f.center = v3(c.lookfrom);
    const double theta = c.vfov * pi / 180.0;                                  // src/rtweekend.h:21-23
    const double h = std::tan(theta / 2);
    const double viewport_height = 2 * h * c.focus_dist;
    const double viewport_width = viewport_height * (static_cast<double>(f.image_width) / f.image_height);
    const V3 w = hunit(v3(c.lookfrom) - v3(c.lookat));
    const V3 u = hunit(hcross(v3(c.vup), w));
    const V3 v = hcross(w, u);
    const V3 viewport_u = viewport_width * u;
    const V3 viewport_v = viewport_height * (-v);
    f.pixel_delta_u = hdiv(viewport_u, f.image_width);
    f.pixel_delta_v = hdiv(viewport_v, f.image_height);
    const V3 viewport_upper_left = ((f.center - (c.focus_dist * w)) - hdiv(viewport_u, 2)) - hdiv(viewport_v, 2);
    f.pixel00_loc = viewport_upper_left + 0.5 * (f.pixel_delta_u + f.pixel_delta_v);
    const double defocus_radius = c.focus_dist * std::tan((c.defocus_angle / 2) * pi / 180.0);
    f.defocus_disk_u = defocus_radius * u;                                     // operator*(vec3,double) = t*v, src/vec3.h:97-99
    f.defocus_disk_v = defocus_radius * v;
    f.defocus = c.defocus_angle > 0;                                           // src/camera.h:94 tests `<= 0`
    return f;
}

template <typename T> Vec<T> narrow(const V3& a) { return mk<T>(T(a.x), T(a.y), T(a.z)); }
template <typename T> CameraFrame<T> narrow_frame(const CameraFrame<double>& f) {
    CameraFrame<T> g;
    g.center = narrow<T>(f.center); g.pixel00_loc = narrow<T>(f.pixel00_loc);
    g.pixel_delta_u = narrow<T>(f.pixel_delta_u); g.pixel_delta_v = narrow<T>(f.pixel_delta_v);
    g.defocus_disk_u = narrow<T>(f.defocus_disk_u); g.defocus_disk_v = narrow<T>(f.defocus_disk_v);
    g.image_width = f.image_width; g.image_height = f.image_height; g.defocus = f.defocus;
    return g;
}

template <typename T> struct Sel;
template <> struct Sel<float> {
    static const float4* a(const rt_ctx* c) { return c->sa32.p; }
    static const float4* b(const rt_ctx* c) { return c->sb32.p; }
    static const float4* cc(const rt_ctx* c) { return c->sc32.p; }
};
template <> struct Sel<double> {
    static const double4* a(const rt_ctx* c) { return c->sa64.p; }
    static const double4* b(const rt_ctx* c) { return c->sb64.p; }
    static const double4* cc(const rt_ctx* c) { return c->sc64.p; }
};

template <typename T> SceneView<T> scene_view(const rt_ctx* c) {
    SceneView<T> sv;
    sv.nodes = c->nodes.p;
    sv.nodes4 = c->wide ? c->nodes4.p : nullptr;
    sv.sa = Sel<T>::a(c); sv.sb = Sel<T>::b(c); sv.sc = Sel<T>::cc(c);
    sv.meta = c->meta.p;
    sv.big_a = c->big_a.p; sv.big_b = c->big_b.p; sv.big_c = c->big_c.p; sv.big_meta = c->big_meta.p;
    sv.n_slots = c->n_slots; sv.n_nodes = c->n_nodes; sv.n_big = c->n_big; sv.root_ref = c->root_ref;
    sv.n_top = 0;
    const int q = int(sizeof(typename Quad<T>::type));
    sv.off_sa = c->n_nodes * 64; sv.off_sb = sv.off_sa + c->n_slots * q; sv.off_sc = sv.off_sb + c->n_slots * q;
    sv.off_meta = sv.off_sc + c->n_slots * q;
    return sv;
}

// Shared memory one megakernel CTA may use for the staged scene: two CTAs per SM, 1 KB/CTA is
// reserved by the driver, and at least 32 KB per CTA stays L1 for the traversal stacks.
template <typename T> bool can_stage(const rt_ctx* c, size_t* bytes) {
    *bytes = staged_bytes<T>(c->n_nodes, c->n_slots);
    const size_t per_cta = std::min(c->smem_optin, (c->smem_per_sm - 2048) / 2);
    return *bytes + 32 * 1024 <= per_cta && c->n_slots > 0;
}

double ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

extern "C" {

uint32_t rt_abi_version(void) { return RT_ABI_VERSION; }

const char* rt_last_error(const rt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int rt_create(rt_ctx** out, int device) {
    if (!out) return fail(nullptr, RT_E_INVALID, "rt_create: null out pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, RT_E_CUDA, std::string("rt_create: no CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback");
    if (device < 0 || device >= count) return fail(nullptr, RT_E_INVALID, "rt_create: device index out of range");
    rt_ctx* ctx = new rt_ctx;
    ctx->device = device;
    cudaDeviceProp prop;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, RT_E_CUDA, std::string("rt_create: ") + cudaGetErrorString(e));
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    ctx->smem_per_sm = prop.sharedMemPerMultiprocessor;
    for (auto& ev : ctx->ev) cudaEventCreate(&ev);
    {   // scratch of a call is allocated stream-ordered; freed blocks stay in the device's pool for the next call
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess && pool) {
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        cudaGetLastError();
    }
    if (cudaHostAlloc(reinterpret_cast<void**>(&ctx->host_words), 16 * sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
        ctx->host_words = nullptr; cudaGetLastError();      // (the copy path serves then)
    }
    ctx->job_counter.resize(1);
    ctx->frame_timeout.resize(1);
    cudaMemset(ctx->frame_timeout.p, 0, sizeof(int));
    ctx->ray_counter.resize(RT_DEBUG_WORDS);      // [0] rays; [1..31] scheduler counters of a -DRT_POOL_STATS build; [32..] per-warp times of a -DRT_WARP_TIMES build
    *out = ctx;
    return RT_OK;
}

int rt_destroy(rt_ctx* ctx) {
    if (!ctx) return RT_OK;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    rt_ipc_close_peers(ctx);
    rt_frame_close(ctx);
    rt_jobs_close(ctx); ctx->jobs_own.release();
    ctx->nodes.release(); ctx->nodes4.release(); ctx->sa32.release(); ctx->sb32.release(); ctx->sc32.release();
    ctx->sa64.release(); ctx->sb64.release(); ctx->sc64.release();
    ctx->big_a.release(); ctx->big_b.release(); ctx->big_c.release(); ctx->meta.release(); ctx->big_meta.release();
    ctx->accum.release(); ctx->reduced.release(); ctx->job_counter.release(); ctx->ray_counter.release();
    ctx->rgb8.release(); ctx->sum.release(); ctx->text.release(); ctx->enc_sums.release(); ctx->enc_offsets.release();
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->host_words) cudaFreeHost(ctx->host_words);
    if (ctx->copy_done) cudaEventDestroy(ctx->copy_done);
    for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return RT_OK;
}

int rt_image_size(const rt_camera* cam, int* width, int* height) {
    if (!cam || cam->image_width < 1 || !(cam->aspect_ratio > 0)) return RT_E_INVALID;
    if (width) *width = cam->image_width;
    if (height) *height = image_height_of(*cam);
    return RT_OK;
}

int rt_synchronize(rt_ctx* ctx) {
    if (!ctx) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

int rt_stream(rt_ctx* ctx, void** cuda_stream) {
    if (!ctx || !cuda_stream) return RT_E_INVALID;
    *cuda_stream = ctx->stream;
    return RT_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------ scene upload --
extern "C" int rt_upload_scene(rt_ctx* ctx, const rt_sphere* spheres, size_t n_spheres,
                               const rt_material* materials, size_t n_materials, const rt_bvh_opts* opts) {
    NvtxRange nvtx_("rt_upload_scene");
    if (!ctx) return RT_E_INVALID;
    if (n_spheres && (!spheres || !materials)) return fail(ctx, RT_E_INVALID, "rt_upload_scene: null arrays");
    if (opts && opts->abi != RT_ABI_VERSION) return fail(ctx, RT_E_INVALID, "rt_upload_scene: ABI version mismatch");
    if (n_spheres >= (size_t(1) << 28)) return fail(ctx, RT_E_INVALID, "rt_upload_scene: too many spheres");
    const int builder = opts ? opts->builder : RT_BVH_SAH_HOST;
    const int max_leaf = (opts && opts->max_leaf > 0) ? std::min(opts->max_leaf, 8) : 4;
    const int lbvh_leaf = (opts && opts->max_leaf > 0) ? std::min(opts->max_leaf, 8) : RT_LBVH_DEFAULT_LEAF;
    double big_radius = (opts && opts->big_radius != 0.0) ? opts->big_radius : 100.0;
    if (big_radius < 0) big_radius = INFINITY;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    const auto t_start = std::chrono::steady_clock::now();
    // from here on the device buffers are rewritten: until the upload has succeeded as a whole there is NO scene
    // (an error exit must not leave the old counts and root paired with new node / slot data)
    ctx->has_scene = false; ctx->root_ref = RT_SENTINEL; ctx->n_nodes = 0; ctx->n_slots = 0; ctx->n_big = 0;

    // one pass over the caller's arrays: material indices / types are checked and the outsized spheres found (large worlds:
    // in parallel over a few host threads -- three serial passes over 1 M spheres were a third of the upload)
    // split: outsized spheres stay out of the tree (tested first, in double)
    // (every ray tests every one of them: at most RT_MAX_BIG, the LARGEST ones; further candidates go into the tree like
    // any sphere -- same hits, float boxes around a huge sphere just prune less -- and are reported in rt_stats.big_in_tree)
    std::vector<int> tree_ids, big_ids;
    {
        const int nthr = n_spheres >= 65536 ? int(std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency()))) : 1;
        std::vector<std::vector<int>> big_of(nthr);
        std::vector<int> bad(nthr, 0);
        auto scan = [&](int t) {
            const size_t lo = n_spheres * size_t(t) / nthr, hi = n_spheres * size_t(t + 1) / nthr;
            for (size_t i = lo; i < hi; ++i) {
                const int mi = spheres[i].material;
                if (mi < 0 || size_t(mi) >= n_materials) { bad[t] = 1; return; }
                const int ty = materials[mi].type;
                if (ty < RT_LAMBERTIAN || ty > RT_DIELECTRIC) { bad[t] = 2; return; }
                if (std::fabs(spheres[i].radius) >= big_radius) big_of[t].push_back(int(i));
            }
        };
        if (nthr == 1) scan(0);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < nthr; ++t) th.emplace_back(scan, t);
            for (auto& x : th) x.join();
        }
        for (int t = 0; t < nthr; ++t) {
            if (bad[t] == 1) return fail(ctx, RT_E_INVALID, "rt_upload_scene: sphere material index out of range");
            if (bad[t] == 2) return fail(ctx, RT_E_INVALID, "rt_upload_scene: unknown material type");
        }
        for (int t = 0; t < nthr; ++t) big_ids.insert(big_ids.end(), big_of[t].begin(), big_of[t].end());     // list order
    }
    int big_in_tree = 0;
    if (big_ids.size() > RT_MAX_BIG) {
        std::vector<int> by_radius = big_ids;
        std::stable_sort(by_radius.begin(), by_radius.end(), [&](int a, int b) { return std::fabs(spheres[a].radius) > std::fabs(spheres[b].radius); });
        by_radius.resize(RT_MAX_BIG);
        std::sort(by_radius.begin(), by_radius.end());                     // list order again
        big_in_tree = int(big_ids.size() - by_radius.size());
        big_ids.swap(by_radius);
    }
    {
        tree_ids.resize(n_spheres - big_ids.size());
        size_t k = 0, o = 0;
        for (size_t i = 0; i < n_spheres; ++i) {
            if (k < big_ids.size() && big_ids[k] == int(i)) ++k;
            else tree_ids[o++] = int(i);
        }
    }
    const int n_tree = int(tree_ids.size());
    ctx->big_ids = big_ids; ctx->big_radius = big_radius;
    CK(ctx->slot_sphere.resize(std::max(1, n_tree)));
    const size_t nt = std::max(1, n_tree), nb = std::max<size_t>(1, big_ids.size());
    CK(ctx->sa32.resize(nt)); CK(ctx->sb32.resize(nt)); CK(ctx->sc32.resize(nt));
    CK(ctx->sa64.resize(nt)); CK(ctx->sb64.resize(nt)); CK(ctx->sc64.resize(nt)); CK(ctx->meta.resize(nt));
    CK(ctx->big_a.resize(nb)); CK(ctx->big_b.resize(nb)); CK(ctx->big_c.resize(nb)); CK(ctx->big_meta.resize(nb));

    // SoA slots (host form): a = (center1, r), b = (center_vec, r^2), c = (albedo, param), meta
    auto fill = [&](const std::vector<int>& ids, std::vector<double4>& a, std::vector<double4>& b, std::vector<double4>& c, std::vector<int>& m) {
        a.resize(ids.size()); b.resize(ids.size()); c.resize(ids.size()); m.resize(ids.size());
        for (size_t k = 0; k < ids.size(); ++k) {
            const rt_sphere& s = spheres[ids[k]];
            const rt_material& mt = materials[s.material];
            const bool mov = s.is_moving != 0;
            a[k] = make_double4(s.center1[0], s.center1[1], s.center1[2], s.radius);
            b[k] = make_double4(mov ? s.center_vec[0] : 0.0, mov ? s.center_vec[1] : 0.0, mov ? s.center_vec[2] : 0.0, s.radius * s.radius);
            double param = mt.param;
            if (mt.type == RT_METAL) param = param < 1 ? param : 1;             // src/material.h:33
            c[k] = make_double4(mt.albedo[0], mt.albedo[1], mt.albedo[2], param);
            m[k] = (ids[k] << 3) | (mov ? 4 : 0) | mt.type;
        }
    };

    const auto t_build = std::chrono::steady_clock::now();
    if (builder == RT_BVH_LBVH_GPU && n_tree >= 2) {
        // Device path (large scenes): the caller's AoS arrays go to the device as they are; boxes, tree and
        // the SoA slots in leaf order are made there (rt_scene.cu) -- no per-sphere host loop after the split
        DevBuf<rt_sphere> d_sph; DevBuf<rt_material> d_mat; DevBuf<int> d_ids; DevBuf<HostBox> d_boxes; DevBuf<int> d_order;
        CK(d_sph.scratch(n_spheres, ctx->stream)); CK(d_mat.scratch(n_materials, ctx->stream)); CK(d_ids.scratch(n_tree, ctx->stream));
        CK(d_boxes.scratch(n_tree, ctx->stream)); CK(d_order.scratch(n_tree, ctx->stream)); CK(ctx->nodes.resize(size_t(n_tree - 1) * 4));
        CK(cudaMemcpyAsync(d_sph.p, spheres, sizeof(rt_sphere) * n_spheres, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(d_mat.p, materials, sizeof(rt_material) * n_materials, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(d_ids.p, tree_ids.data(), sizeof(int) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
        cudaError_t e = launch_scene_boxes(d_sph.p, d_ids.p, n_tree, d_boxes.p, ctx->stream);
        int root = RT_SENTINEL, depth = 0;
        if (e == cudaSuccess) e = build_lbvh_gpu(d_boxes.p, n_tree, lbvh_leaf, ctx->nodes.p, d_order.p, &root, &depth, ctx->stream, ctx->host_words);
        if (e == cudaSuccess) e = launch_scene_fill(d_sph.p, d_mat.p, d_ids.p, d_order.p, n_tree, ctx->sa32.p, ctx->sb32.p, ctx->sc32.p,
                                                    ctx->sa64.p, ctx->sb64.p, ctx->sc64.p, ctx->meta.p, ctx->stream, ctx->slot_sphere.p);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        d_sph.release(); d_mat.release(); d_ids.release(); d_boxes.release(); d_order.release();
        if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_upload_scene: device scene build");
        if (depth > RT_STACK - 2) return fail(ctx, RT_E_INVALID, "rt_upload_scene: LBVH deeper than the traversal stack");
        ctx->root_ref = root; ctx->depth = depth; ctx->n_nodes = n_tree - 1;
        ctx->build_ms = ms_since(t_build);
    } else {
        std::vector<HostBox> boxes(tree_ids.size());
        for (size_t k = 0; k < tree_ids.size(); ++k) {
            const rt_sphere& s = spheres[tree_ids[k]];
            boxes[k] = sphere_box(s.center1, s.center_vec, s.radius, s.is_moving != 0);
        }
        HostBVH bvh;
        build_bvh_sah(boxes, max_leaf, bvh);
        if (bvh.depth > RT_STACK - 2) return fail(ctx, RT_E_INVALID, "rt_upload_scene: BVH deeper than the traversal stack");
        ctx->root_ref = bvh.root_ref; ctx->depth = bvh.depth; ctx->n_nodes = bvh.n_inner;
        CK(ctx->nodes.resize(std::max<size_t>(1, size_t(bvh.n_inner) * 4)));
        if (bvh.n_inner)
            CK(cudaMemcpyAsync(ctx->nodes.p, bvh.nodes.data(), bvh.nodes.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        ctx->build_ms = ms_since(t_build);

        // SoA slots in leaf order
        std::vector<int> slot_ids(n_tree);
        for (int k = 0; k < n_tree; ++k) slot_ids[k] = tree_ids[bvh.order[k]];
        std::vector<double4> a64, b64, c64;
        std::vector<int> meta;
        fill(slot_ids, a64, b64, c64, meta);
        std::vector<float4> a32(n_tree), b32(n_tree), c32(n_tree);
        for (int k = 0; k < n_tree; ++k) {
            a32[k] = make_float4(float(a64[k].x), float(a64[k].y), float(a64[k].z), float(a64[k].w));
            const float r = a32[k].w;
            b32[k] = make_float4(float(b64[k].x), float(b64[k].y), float(b64[k].z), r * r);
            c32[k] = make_float4(float(c64[k].x), float(c64[k].y), float(c64[k].z), float(c64[k].w));
        }
        if (n_tree) {
            CK(cudaMemcpyAsync(ctx->sa32.p, a32.data(), sizeof(float4) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->sb32.p, b32.data(), sizeof(float4) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->sc32.p, c32.data(), sizeof(float4) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->sa64.p, a64.data(), sizeof(double4) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->sb64.p, b64.data(), sizeof(double4) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->sc64.p, c64.data(), sizeof(double4) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->meta.p, meta.data(), sizeof(int) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->slot_sphere.p, slot_ids.data(), sizeof(int) * n_tree, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));               // the staging vectors go out of scope
        }
    }
    {
        std::vector<double4> ba, bb, bc;
        std::vector<int> bmeta;
        fill(big_ids, ba, bb, bc, bmeta);
        if (!big_ids.empty()) {
            CK(cudaMemcpyAsync(ctx->big_a.p, ba.data(), sizeof(double4) * ba.size(), cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->big_b.p, bb.data(), sizeof(double4) * bb.size(), cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->big_c.p, bc.data(), sizeof(double4) * bc.size(), cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaMemcpyAsync(ctx->big_meta.p, bmeta.data(), sizeof(int) * bmeta.size(), cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
        }
    }
    // top of the tree first: nodes 0 .. k-1 = the first levels in breadth-first order, so that a render whose scene stays in
    // global memory can keep them in shared memory (north_star item 2).  Parity is defined on hits, not on node numbers.
    ctx->top_ordered = 0;
    if (ctx->n_nodes > 1 && ctx->root_ref >= 0 && !env_on("RT_NO_TOP_FIRST")) {       // (the env knob is for A/B measurements)
        DevBuf<int> newpos, top_ids; DevBuf<float4> permuted;
        CK(newpos.scratch(ctx->n_nodes, ctx->stream)); CK(top_ids.scratch(RT_TOP_MAX, ctx->stream)); CK(permuted.scratch(size_t(ctx->n_nodes) * 4, ctx->stream));
        CK(launch_top_first(ctx->nodes.p, ctx->n_nodes, ctx->root_ref, RT_TOP_MAX, newpos.p, top_ids.p, permuted.p, ctx->stream));
        CK(cudaMemcpyAsync(ctx->nodes.p, permuted.p, size_t(ctx->n_nodes) * 64, cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        ctx->root_ref = 0;
        ctx->top_ordered = std::min(ctx->n_nodes, RT_TOP_MAX);
    }
    // 4-wide records (8f-4) for trees walked in global memory: opt-in (opts->wide = 1; the wavefront kernel then leaves
    // the scene in global memory whatever its size).  Measured on C5 (10 k / 100 k / 1 M spheres): 2 % SLOWER than the
    // binary records (profiles/r1_wide_nodes_ab.txt), so nothing turns them on by default.
    {
        const int want = opts ? opts->wide : 0;
        ctx->wide_forced = want > 0;
        ctx->wide = ctx->n_nodes > 0 && want > 0;
        // the 4-wide walk pushes up to 3 refs per record = per two levels of the binary tree: a tree the binary walk
        // can hold (depth <= RT_STACK) may still overflow here
        if (ctx->wide && 3 * ((ctx->depth + 1) / 2) > RT_STACK - 2)
            return fail(ctx, RT_E_INVALID, "rt_upload_scene: tree too deep for the 4-wide walk (3 * ceil(depth / 2) > traversal stack); use wide = 0");
        if (ctx->wide) {
            CK(ctx->nodes4.resize(size_t(ctx->n_nodes) * 8));
            CK(launch_widen(ctx->nodes.p, ctx->n_nodes, ctx->nodes4.p, ctx->stream));
        } else {
            ctx->nodes4.release();
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_spheres = int(n_spheres); ctx->n_slots = n_tree; ctx->n_big = int(big_ids.size());
    ctx->has_scene = true;
    ctx->upload_ms = ms_since(t_start);
    ctx->stats.n_spheres = ctx->n_spheres; ctx->stats.n_big = ctx->n_big; ctx->stats.n_nodes = ctx->n_nodes; ctx->stats.big_in_tree = big_in_tree;
    ctx->stats.bvh_depth = ctx->depth; ctx->stats.build_ms = ctx->build_ms; ctx->stats.upload_ms = ctx->upload_ms;
    return RT_OK;
}

// ------------------------------------------------------------------------ scene update --
// SURVEY.md 8f-4, dynamic scenes: the world of the last rt_upload_scene with its spheres MOVED (centres, motion vectors,
// radii, materials may all change; the count, the order and the split into tree / outsized spheres may not).  The tree keeps
// its topology and is refitted on the device: new SoA slots, new slot boxes, every node box recomputed bottom-up
// (rt_scene.cu tree_refit_kernel) -- no sort, no hierarchy pass, no host builder.  Hits stay exact (the boxes are
// conservative by construction); only the tree's QUALITY ages as spheres drift from where the builder saw them.
extern "C" int rt_update_spheres(rt_ctx* ctx, const rt_sphere* spheres, size_t n_spheres, const rt_material* materials, size_t n_materials) {
    NvtxRange nvtx_("rt_update_spheres");
    if (!ctx) return RT_E_INVALID;
    if (!ctx->has_scene) return fail(ctx, RT_E_STATE, "rt_update_spheres: no scene uploaded");
    if (n_spheres != size_t(ctx->n_spheres)) return fail(ctx, RT_E_STATE, "rt_update_spheres: the sphere count changed; use rt_upload_scene");
    if (n_spheres && (!spheres || !materials)) return fail(ctx, RT_E_INVALID, "rt_update_spheres: null arrays");
    CK(cudaSetDevice(ctx->device));
    const auto t_start = std::chrono::steady_clock::now();
    {
        size_t k = 0;
        for (size_t i = 0; i < n_spheres; ++i) {
            if (spheres[i].material < 0 || size_t(spheres[i].material) >= n_materials)
                return fail(ctx, RT_E_INVALID, "rt_update_spheres: sphere material index out of range");
            const int ty = materials[spheres[i].material].type;
            if (ty < RT_LAMBERTIAN || ty > RT_DIELECTRIC) return fail(ctx, RT_E_INVALID, "rt_update_spheres: unknown material type");
            const bool was_big = k < ctx->big_ids.size() && ctx->big_ids[k] == int(i);
            if (was_big) ++k;
            // (with more than 16 candidates the largest 16 were chosen at upload: a candidate that was in the tree stays there)
            if (was_big && !(std::fabs(spheres[i].radius) >= ctx->big_radius))
                return fail(ctx, RT_E_STATE, "rt_update_spheres: an outsized sphere (kept out of the tree) shrank below big_radius; use rt_upload_scene");
            if (!was_big && std::fabs(spheres[i].radius) >= ctx->big_radius && ctx->big_ids.size() < RT_MAX_BIG)
                return fail(ctx, RT_E_STATE, "rt_update_spheres: a sphere grew to big_radius (outsized spheres are kept out of the tree); use rt_upload_scene");
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->has_scene = false;                                    // until the update has succeeded as a whole
    const int n_tree = ctx->n_slots;
    DevBuf<rt_sphere> d_sph; DevBuf<rt_material> d_mat; DevBuf<HostBox> d_boxes; DevBuf<int> d_parent, d_flags;
    if (n_tree > 0) {
        CK(d_sph.resize(n_spheres)); CK(d_mat.resize(n_materials)); CK(d_boxes.resize(n_tree));
        CK(cudaMemcpyAsync(d_sph.p, spheres, sizeof(rt_sphere) * n_spheres, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(d_mat.p, materials, sizeof(rt_material) * n_materials, cudaMemcpyHostToDevice, ctx->stream));
        CK(launch_scene_fill(d_sph.p, d_mat.p, ctx->slot_sphere.p, nullptr, n_tree, ctx->sa32.p, ctx->sb32.p, ctx->sc32.p,
                             ctx->sa64.p, ctx->sb64.p, ctx->sc64.p, ctx->meta.p, ctx->stream));
        if (ctx->n_nodes > 0) {
            CK(d_parent.resize(ctx->n_nodes)); CK(d_flags.resize(ctx->n_nodes));
            CK(launch_scene_boxes(d_sph.p, ctx->slot_sphere.p, n_tree, d_boxes.p, ctx->stream));
            CK(launch_tree_refit(ctx->nodes.p, ctx->n_nodes, d_boxes.p, d_parent.p, d_flags.p, ctx->stream));
            if (ctx->wide) CK(launch_widen(ctx->nodes.p, ctx->n_nodes, ctx->nodes4.p, ctx->stream));
        }
    }
    if (!ctx->big_ids.empty()) {
        std::vector<double4> ba(ctx->big_ids.size()), bb(ba.size()), bc(ba.size());
        std::vector<int> bmeta(ba.size());
        for (size_t k = 0; k < ba.size(); ++k) {
            const rt_sphere& sp = spheres[ctx->big_ids[k]];
            const rt_material& mt = materials[sp.material];
            const bool mov = sp.is_moving != 0;
            ba[k] = make_double4(sp.center1[0], sp.center1[1], sp.center1[2], sp.radius);
            bb[k] = make_double4(mov ? sp.center_vec[0] : 0.0, mov ? sp.center_vec[1] : 0.0, mov ? sp.center_vec[2] : 0.0, sp.radius * sp.radius);
            double param = mt.param;
            if (mt.type == RT_METAL) param = param < 1 ? param : 1;             // src/material.h:33
            bc[k] = make_double4(mt.albedo[0], mt.albedo[1], mt.albedo[2], param);
            bmeta[k] = (ctx->big_ids[k] << 3) | (mov ? 4 : 0) | mt.type;
        }
        CK(cudaMemcpyAsync(ctx->big_a.p, ba.data(), sizeof(double4) * ba.size(), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->big_b.p, bb.data(), sizeof(double4) * bb.size(), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->big_c.p, bc.data(), sizeof(double4) * bc.size(), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->big_meta.p, bmeta.data(), sizeof(int) * bmeta.size(), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->has_scene = true;
    ctx->upload_ms = ms_since(t_start);
    ctx->build_ms = 0.0;
    ctx->stats.upload_ms = ctx->upload_ms; ctx->stats.build_ms = 0.0;
    return RT_OK;
}

// ------------------------------------------------------------------------------ render --
namespace {

// o.sched bits 16..19: block shape + 1 (0 = default for the precision)
template <typename T> int sched_shape(int sched) {
    const int v = (sched >> 16) & 0xf;
    return v >= 1 && v <= RT_SCHED_SHAPES ? v - 1 : RT_SCHED_DEFAULT_SHAPE(T);
}

// wavefront: o.sched bits 16..19 = pool shape + 1; falls back to smaller pools when the staged scene
// plus the pools do not fit the SM's shared memory, and to an unstaged scene after that
template <typename T> int pool_shape_for(const rt_ctx* c, int sched, bool* staged) {
    const int v = (sched >> 16) & 0xf;
    int shape = v >= 1 && v <= RT_POOL_SHAPES ? v - 1 : RT_POOL_DEFAULT_SHAPE(T);
    const size_t scene = staged_bytes_aligned<T>(c->n_nodes, c->n_slots);
    const size_t budget = c->smem_optin;
    auto need = [&](int sh) { int P, B; pool_shape(sh, &P, &B); return size_t(B / 32) * pool_warp_bytes<T>(P); };
    *staged = c->n_slots > 0 && scene + need(shape) <= budget && !c->wide_forced;
    if (!*staged && c->n_slots > 0 && !c->wide_forced) {   // fewer warps, most first, until scene + pools fit
        const int by_warps[] = { 3, 5, 4, 2, 1, 0 };
        for (int sh : by_warps) {
            int P, B, P0, B0; pool_shape(sh, &P, &B); pool_shape(shape, &P0, &B0);
            if (B < B0 && scene + need(sh) <= budget) { shape = sh; *staged = true; break; }
        }
    }
    // a scene that stays in global memory is walked at L2 latency: the walk is bound by dependent fetches, not
    // by issue slots, and the shared memory the scene does not use buys more warps (32 x 128-path pools)
    if (!*staged && v == 0 && need(RT_POOL_UNSTAGED_SHAPE(T)) <= budget) shape = RT_POOL_UNSTAGED_SHAPE(T);
    if (!*staged && need(shape) > budget) shape = 0;
    return shape;
}

// Sample layers of a wavefront frame (auto mode).  The pool does not drain between jobs (two jobs open per warp), so
// the job size is free to vary: BULK layers of about 1/20 of a warp's share of the frame each (at least 8 samples: a
// slot that finishes the newer job waits for the stragglers of the older one), then layers that halve down to a few
// samples.  The short layers are taken last, so the frame ends on small jobs: the tail is a job or two per warp,
// whatever its size (measured: profiles/r1_job_overhead_ring2.txt, r1_job_sizes_geometric.txt).
// sizes[0..n) sum to sample_count, n <= RT_MAX_SLICES; *n_bulk = how many of them are bulk layers.
int plan_layers(int sample_count, long long tiles, long long resident_warps, int jobs_per_warp, int* sizes, int* n_bulk) {
    if (jobs_per_warp <= 0) jobs_per_warp = 20;                    // 16-24 measured best, flat
    const double tiles_per_warp = double(std::max<long long>(1, tiles)) / double(std::max<long long>(1, resident_warps));
    int bulk = std::max(8, int(std::ceil(sample_count * tiles_per_warp / double(jobs_per_warp))));
    bulk = std::max(bulk, (sample_count + RT_MAX_SLICES - 9) / (RT_MAX_SLICES - 8));       // all of them fit the buffer's 64 layers
    // the frame ends on layers of `small` samples: a job of the last layer on an expensive tile (glass: 5 x the mean rays per
    // sample) is what every other warp waits for (tools/warp_times.py)
    int small_div = 8;
    if (const char* e = std::getenv("RT_PLAN_SMALL_DIV")) small_div = std::max(1, std::atoi(e));      // (A/B knob)
    const int small = std::max(2, bulk / small_div);
    int left = sample_count, n = 0;
    while (left > 2 * bulk && n < RT_MAX_SLICES - 8) { sizes[n++] = bulk; left -= bulk; }
    *n_bulk = n;
    while (left > small && n < RT_MAX_SLICES - 1) { const int h = (left + 1) / 2; sizes[n++] = h; left -= h; }
    if (left > 0) sizes[n++] = left;
    return n;
}

template <typename T>
int render_impl(rt_ctx* ctx, const rt_camera& cam, const rt_render_opts& o, int sample_begin, int sample_count, int slices) {
    typedef typename Quad<T>::type Q;
    const CameraFrame<double> f64 = camera_frame(cam);
    RenderParams<T> p;
    p.cam = narrow_frame<T>(f64);
    p.scene = scene_view<T>(ctx);
    p.accum = reinterpret_cast<Q*>(ctx->accum.p);
    p.job_counter = ctx->job_counter.p;
    p.ray_counter = ctx->ray_counter.p;
    p.seed = o.seed;
    philox_round_keys(o.seed, p.keys);
    p.sample_begin = sample_begin; p.sample_count = sample_count; p.max_depth = cam.max_depth;
    p.slices = slices;
    for (int k = 0; k <= slices; ++k) p.slice_begin[k] = ctx->slice_begin[k];
    p.bulk_layers = std::min(slices, ctx->bulk_layers);
    p.job_group = 2;          // pairs: +0.3-1.5 % over single jobs; 3 and more lose to the tail (profiles/r1_job_group_sweep.txt)
    if (const char* e = std::getenv("RT_JOB_GROUP")) p.job_group = std::max(1, std::atoi(e));      // (A/B knob)
    p.tiles_x = (ctx->W + 7) / 8; p.tiles_y = (ctx->H + 3) / 4;
    p.accumulate = o.accumulate;
    p.tile_rank = 0; p.tile_world = 1;
    size_t smem = 0;
    bool staged = can_stage<T>(ctx, &smem);
    int pshape = 0;
    if (o.variant == RT_WAVEFRONT) {
        pshape = pool_shape_for<T>(ctx, o.sched, &staged);
        int P, B; pool_shape(pshape, &P, &B);
        p.tiles_x = (ctx->W + 15) / 16; p.tiles_y = (ctx->H + P / 16 - 1) / (P / 16);
        if (o.tile_world > 1) { p.tile_rank = o.tile_rank; p.tile_world = o.tile_world; }
        ctx->tiles = TileOwner{ p.tile_world, p.tile_rank, p.tiles_x, 16, P / 16, p.job_group, ctx->W };
        ctx->tiles_pending = p.tile_world > 1;
    } else {
        ctx->tiles = TileOwner{ 0, 0, 0, 0, 0, 0, 0 };
    }
    if (o.variant == RT_WAVEFRONT && !staged && !ctx->wide && ctx->top_ordered > 0) {
        // the shared memory the pools leave free holds the top of the tree (rt_render_opts.top_cap: tuning knob)
        int P, B; pool_shape(pshape, &P, &B);
        const size_t pools = size_t(B / 32) * pool_warp_bytes<T>(P);
        const size_t spare = ctx->smem_optin > pools ? ctx->smem_optin - pools : 0;
        int n_top = int(std::min<size_t>(size_t(ctx->top_ordered), spare / 64));
        if (o.top_cap < 0) n_top = 0; else if (o.top_cap > 0) n_top = std::min(n_top, o.top_cap);
        p.scene.n_top = n_top;
    }
    ctx->stats.scene_in_smem = staged ? 1 : (p.scene.n_top > 0 ? 2 : 0);
    ctx->stats.top_nodes_in_smem = p.scene.n_top;
    p.jobs_shared = 0;
    p.job_stride = 1;
    if (const char* e = std::getenv("RT_JOB_STRIDE")) p.job_stride = std::max(1, std::atoi(e));      // (diagnostic knob)
    if (o.shared_jobs) {
        // the frame's jobs come from the counter shared with the other GPUs; this rank's buffer holds only what it takes
        p.jobs_shared = 1;
        p.job_counter = ctx->jobs_ptr + (ctx->jobs_frame % RT_JOB_COUNTERS);
        if (ctx->jobs_root)
            CK(cudaMemsetAsync(ctx->jobs_ptr + ((ctx->jobs_frame + RT_JOB_COUNTERS / 2) % RT_JOB_COUNTERS), 0, sizeof(unsigned), ctx->stream));
        // this rank's buffer holds only the jobs it takes: everything else must read zero in the combine
        if (!o.accumulate) CK(cudaMemsetAsync(ctx->accum.p, 0, size_t(slices) * ctx->W * ctx->H * sizeof(Q), ctx->stream));
        ++ctx->jobs_frame;
    }
    CK(cudaMemsetAsync(ctx->job_counter.p, 0, sizeof(unsigned int), ctx->stream));
    CK(cudaMemsetAsync(ctx->ray_counter.p, 0, RT_DEBUG_WORDS * sizeof(unsigned long long), ctx->stream));
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    // section thresholds: o.sched = th_s | th_l << 8 (0 = defaults, tuned on B200: see DESIGN.md)
    p.th_s = (o.sched & 0xff) ? (o.sched & 0xff) : (o.variant == RT_WAVEFRONT ? 22 : 28);       // wavefront: idle lanes before a pull (profiles/r1_sched_tune_final.txt)
    p.th_l = ((o.sched >> 8) & 0xff) ? ((o.sched >> 8) & 0xff) : 33;
    p.th_n = ((o.sched >> 20) & 0xf) ? ((o.sched >> 20) & 0xf) : 4;      // node loop runs while >= th_n/8 of its lanes remain
    if (o.variant == RT_WAVEFRONT) CK(launch_render_pool<T>(p, pshape, staged, ctx->sm_count, ctx->stream));
    else if (o.variant == RT_MEGA_LOCKSTEP) CK(launch_render_mega<T>(p, staged, smem, ctx->sm_count, ctx->stream));
    else CK(launch_render_sched<T>(p, sched_shape<T>(o.sched), staged, smem, ctx->sm_count, ctx->stream));
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    return RT_OK;
}

}  // namespace

extern "C" int rt_render(rt_ctx* ctx, const rt_camera* cam, const rt_render_opts* opts) {
    NvtxRange nvtx_("rt_render");
    if (!ctx || !cam || !opts) return RT_E_INVALID;
    if (opts->abi != RT_ABI_VERSION) return fail(ctx, RT_E_INVALID, "rt_render: ABI version mismatch");
    if (!ctx->has_scene) return fail(ctx, RT_E_STATE, "rt_render: no scene uploaded");
    if (cam->image_width < 1 || !(cam->aspect_ratio > 0)) return fail(ctx, RT_E_INVALID, "rt_render: bad image size");
    if (opts->precision != RT_F32 && opts->precision != RT_F64) return fail(ctx, RT_E_INVALID, "rt_render: bad precision");
    if (opts->variant != RT_MEGAKERNEL && opts->variant != RT_MEGA_LOCKSTEP && opts->variant != RT_WAVEFRONT) return fail(ctx, RT_E_INVALID, "rt_render: variant not built");
    if (opts->tile_world > 1 && (opts->tile_rank < 0 || opts->tile_rank >= opts->tile_world)) return fail(ctx, RT_E_INVALID, "rt_render: tile_rank out of range");
    if (opts->shared_jobs) {
        if (!ctx->jobs_ptr) return fail(ctx, RT_E_STATE, "rt_render: shared_jobs needs rt_jobs_export (root) or rt_jobs_open / rt_jobs_share (peers) first");
        if (opts->variant != RT_WAVEFRONT || opts->precision != RT_F32 || opts->tile_world > 1)
            return fail(ctx, RT_E_INVALID, "rt_render: shared_jobs is for the wavefront variant in fast mode (f32), without tile sharding");
    }
    // the wavefront pool keeps a path's sample index in 14 bits and its remaining depth in 16: beyond that the
    // warp-scheduled megakernel (same arithmetic, same image) takes the call
    rt_render_opts eff = *opts;
    // (and a job's tile origin in 15 + 16 bits)
    if (eff.variant == RT_WAVEFRONT && (cam->max_depth > 65535 || cam->image_width > 32000 || image_height_of(*cam) > 32000 ||
        (eff.sample_count > 0 ? eff.sample_count : cam->samples_per_pixel) > 16383)) { eff.variant = RT_MEGAKERNEL; eff.sched = 0; }
    if (eff.shared_jobs && eff.variant != RT_WAVEFRONT) return fail(ctx, RT_E_INVALID, "rt_render: shared_jobs needs the wavefront variant (and its 16-bit sample / depth limits)");
    if (eff.tile_world > 1 && eff.variant != RT_WAVEFRONT) return fail(ctx, RT_E_INVALID, "rt_render: tile-interleaved sharding needs the wavefront variant (and its 16-bit sample / depth limits)");
    opts = &eff;
    CK(cudaSetDevice(ctx->device));
    const int W = cam->image_width, H = image_height_of(*cam);
    if ((long long)W * H >= (1ll << 32)) return fail(ctx, RT_E_INVALID, "rt_render: image too large for the 32-bit pixel key");
    const int sample_begin = std::max(0, opts->sample_begin);
    const int sample_count = opts->sample_count > 0 ? opts->sample_count : std::max(0, cam->samples_per_pixel - sample_begin);

    // slices: enough warp jobs to keep every resident warp busy and the tail short
    long long tiles = (long long)((W + 7) / 8) * ((H + 3) / 4);
    int ctas_per_sm = 2;
    int block = opts->variant == RT_MEGA_LOCKSTEP ? (opts->precision == RT_F32 ? RT_MEGA_BLOCK(float) : RT_MEGA_BLOCK(double))
                                                        : sched_block(opts->precision == RT_F32 ? sched_shape<float>(opts->sched) : sched_shape<double>(opts->sched));
    if (opts->variant == RT_WAVEFRONT) {
        bool staged; int P;
        const int sh = opts->precision == RT_F32 ? pool_shape_for<float>(ctx, opts->sched, &staged) : pool_shape_for<double>(ctx, opts->sched, &staged);
        pool_shape(sh, &P, &block);
        tiles = (long long)((W + 15) / 16) * ((H + P / 16 - 1) / (P / 16));
        ctas_per_sm = 1;
    }
    // (dynamic sharding: the frame's jobs are shared by opts->shared_jobs GPUs -- the plan is made for all their warps)
    const long long resident_warps = (long long)ctx->sm_count * ctas_per_sm * (block / 32) * std::max(1, opts->shared_jobs);
    // Parity mode keeps ONE slice unless asked otherwise: the per-pixel sum is then formed sample by
    // sample in the reference's order (src/camera.h:41-44).  Any other order can differ in the last
    // bit, and this scene has sums that sit exactly on a byte boundary of write_color (albedo 0.5 x
    // sky blue 1.0: 0.5/8 spp -> sqrt = 0.25 -> 64.0), where one ulp decides the byte.
    int slices = opts->slices;
    if (slices <= 0 && opts->precision == RT_F64) slices = 1;
    int sizes[RT_MAX_SLICES], n_sized = -1, n_bulk = 0;
    if (slices <= 0 && opts->variant == RT_WAVEFRONT && sample_count > 0) {
        const long long my_tiles = opts->tile_world > 1 ? (tiles + opts->tile_world - 1) / opts->tile_world : tiles;
        slices = n_sized = plan_layers(sample_count, my_tiles, resident_warps, (opts->sched >> 24) & 0xff, sizes, &n_bulk);
    }
    if (slices <= 0) slices = int(std::min<long long>(64, (16 * resident_warps + tiles - 1) / tiles));
    slices = std::max(1, std::min(std::min(slices, RT_MAX_SLICES), std::max(1, sample_count)));
    const size_t elem = opts->precision == RT_F32 ? sizeof(float4) : sizeof(double4);
    // a progressive pass may use fewer slices than the buffer holds (it adds into the first ones; resolve sums all)
    if (opts->accumulate && opts->slices <= 0 && ctx->slices > 0) slices = std::min(slices, ctx->slices);
    const bool fits = ctx->W == W && ctx->H == H && ctx->precision == opts->precision && slices <= ctx->slices;
    const bool same = fits && (opts->accumulate || slices == ctx->slices);
    if (opts->accumulate && !fits) return fail(ctx, RT_E_STATE, "rt_render: accumulate needs the same image and precision as the buffer, and no more slices than it holds");
    if (!same) {
        // the frame changes shape: mappings of peers that rendered the old shape are dropped (the caller exchanges
        // handles again), and this context's own exported handle is stale if the buffer has to grow
        if (ctx->peers.n > 0 && (ctx->W != W || ctx->H != H || ctx->precision != opts->precision)) rt_ipc_close_peers(ctx);
        CK(ctx->accum.resize(size_t(slices) * W * H * elem));
        ctx->W = W; ctx->H = H; ctx->precision = opts->precision; ctx->slices = slices;
    }
    if (!opts->accumulate) ctx->samples_in_buffer = 0;
    ctx->folded = false;
    const bool sized = slices == n_sized;             // a progressive pass that got fewer layers than it wanted splits evenly
    ctx->bulk_layers = sized ? n_bulk : slices;
    ctx->slice_begin[0] = sample_begin;
    for (int k = 0; k < slices; ++k)
        ctx->slice_begin[k + 1] = sized ? ctx->slice_begin[k] + sizes[k] : sample_begin + int((long long)sample_count * (k + 1) / slices);
    int rc = opts->precision == RT_F32 ? render_impl<float>(ctx, *cam, *opts, sample_begin, sample_count, slices)
                                       : render_impl<double>(ctx, *cam, *opts, sample_begin, sample_count, slices);
    if (rc != RT_OK) return rc;
    ctx->samples_in_buffer += sample_count;
    ctx->render_timed = true;
    ctx->stats.samples = (uint64_t)W * H * sample_count;
    ctx->stats.image_width = W; ctx->stats.image_height = H; ctx->stats.slices = slices;
    ctx->stats.kernel_launches = 1;
    return RT_OK;
}

// a split read-back still copying out of rgb8 / text: whatever writes them next waits on the device (no host wait)
static cudaError_t after_pending_copies(rt_ctx* ctx) {
    return ctx->copy_pending ? cudaStreamWaitEvent(ctx->stream, ctx->copy_done, 0) : cudaSuccess;
}

extern "C" int rt_readback(rt_ctx* ctx, int spp_total, double* sum_rgb, uint8_t* rgb8, char* ppm, size_t* ppm_len) {
    NvtxRange nvtx_("rt_readback");
    if (!ctx) return RT_E_INVALID;
    if (ctx->W == 0 || !ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_readback: nothing rendered");
    if (ppm && !ppm_len) return fail(ctx, RT_E_INVALID, "rt_readback: ppm needs ppm_len");
    if (ctx->peers.n > 0 && !ctx->folded) return fail(ctx, RT_E_STATE, "rt_readback: peers are mapped; every rank must call rt_fold_slices before the fused resolve");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(ctx->W) * ctx->H;
    CK(ensure_rgb8(ctx, npix));
    if (sum_rgb) CK(ctx->sum.resize(npix * 3));
    const double scale = spp_total > 0 ? 1.0 / spp_total : 0.0;                 // src/color.h:20
    CK(after_pending_copies(ctx));
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    int launches = 1;
    const FrameViews no_views{nullptr, nullptr, 0, 0};
    const bool want_text = ppm || (ppm_len && !ppm);
    // whole frame on this GPU, text wanted, no per-pixel sums: ONE kernel from the accumulation layers to rgb8 + P3 text
    // ... unless the float fast path applies (few layers, no peers): then resolve + count in one kernel and the text in a second
    // one that reads the bytes back from L2 -- no look-back chain between tiles; measured faster at every size (profiles/)
    const bool two_pass = want_text && !sum_rgb && ctx->tiles.world <= 1 && ctx->precision == RT_F32 &&
                          fast_f32_eligible(live_slices(ctx), live_peers(ctx)) && !env_on("RT_NO_TWO_PASS_ENCODE");
    const bool fused = want_text && !sum_rgb && ctx->tiles.world <= 1 && !two_pass && !env_on("RT_NO_FUSED_ENCODE");
    unsigned long long text_len = 0;
    bool waited = false;
    if (two_pass) {
        EncodeHeader hdr;
        hdr.len = std::snprintf(hdr.text, sizeof hdr.text, "P3\n%d %d\n255\n", ctx->W, ctx->H);   // src/camera.h:35
        const int nb = encode_blocks(npix);
        CK(ctx->enc_sums.resize(nb + 1)); CK(ctx->enc_offsets.resize(nb + 3));
        CK(ctx->text.resize(npix * 12 + 64));
        CK(launch_resolve_count_f32(reinterpret_cast<const float4*>(ctx->accum.p), live_slices(ctx), npix, scale, ctx->rgb8.p, ctx->enc_sums.p, ctx->stream));
        CK(launch_encode_p3(ctx->rgb8.p, npix, hdr, ctx->enc_sums.p, ctx->enc_offsets.p, ctx->enc_offsets.p + nb + 1, ctx->text.p, ctx->stream, true));
        launches = 3;
        CK(cudaMemcpyAsync(&text_len, ctx->enc_offsets.p + nb + 1, sizeof text_len, cudaMemcpyDeviceToHost, ctx->stream));
    } else if (fused) {
        EncodeHeader hdr;
        hdr.len = std::snprintf(hdr.text, sizeof hdr.text, "P3\n%d %d\n255\n", ctx->W, ctx->H);   // src/camera.h:35
        const int nb = encode_blocks(npix);
        CK(ctx->enc_offsets.resize(nb + 3));
        CK(ctx->text.resize(npix * 12 + 64));
        if (ctx->precision == RT_F32)
            CK(launch_resolve_encode<float4>(reinterpret_cast<const float4*>(ctx->accum.p), live_slices(ctx), live_peers(ctx), npix, scale,
                                             ctx->rgb8.p, ctx->text.p, hdr, ctx->enc_offsets.p, ctx->enc_offsets.p + nb + 2, ctx->stream));
        else
            CK(launch_resolve_encode<double4>(reinterpret_cast<const double4*>(ctx->accum.p), live_slices(ctx), live_peers(ctx), npix, scale,
                                              ctx->rgb8.p, ctx->text.p, hdr, ctx->enc_offsets.p, ctx->enc_offsets.p + nb + 2, ctx->stream));
        CK(cudaMemcpyAsync(&text_len, ctx->enc_offsets.p + nb + 2, sizeof text_len, cudaMemcpyDeviceToHost, ctx->stream));
    } else {
        if (ctx->precision == RT_F32)
            CK(launch_resolve<float4>(reinterpret_cast<const float4*>(ctx->accum.p), live_slices(ctx), live_peers(ctx), npix, scale,
                                      sum_rgb ? ctx->sum.p : nullptr, ctx->rgb8.p, nullptr, ctx->stream, no_views, ctx->tiles));
        else
            CK(launch_resolve<double4>(reinterpret_cast<const double4*>(ctx->accum.p), live_slices(ctx), live_peers(ctx), npix, scale,
                                       sum_rgb ? ctx->sum.p : nullptr, ctx->rgb8.p, nullptr, ctx->stream, no_views, ctx->tiles));
        // tile-sharded frame on the root: the other ranks' resolve kernels store their tiles' bytes straight into this image over
        // NVLink and bump its counter; the stream waits for all of them here, before the image is encoded or copied
        if (ctx->tiles.world > 1 && ctx->frame_exported && ctx->tiles_pending) {
            ctx->frame_target += unsigned(ctx->tiles.world - 1);
            CK(launch_frame_wait(reinterpret_cast<const unsigned*>(ctx->rgb8.p + frame_counter_offset(npix)), ctx->frame_target, 10000,
                                 ctx->frame_timeout.p, ctx->stream));
            ctx->tiles_pending = false; waited = true; ++launches;
        }
        if (want_text) {
            EncodeHeader hdr;
            hdr.len = std::snprintf(hdr.text, sizeof hdr.text, "P3\n%d %d\n255\n", ctx->W, ctx->H);   // src/camera.h:35
            const int nb = encode_blocks(npix);
            CK(ctx->enc_sums.resize(nb + 1)); CK(ctx->enc_offsets.resize(nb + 3));
            CK(ctx->text.resize(npix * 12 + 64));
            CK(launch_encode_p3(ctx->rgb8.p, npix, hdr, ctx->enc_sums.p, ctx->enc_offsets.p, ctx->enc_offsets.p + nb + 1, ctx->text.p, ctx->stream));
            launches += 3;
            CK(cudaMemcpyAsync(&text_len, ctx->enc_offsets.p + nb + 1, sizeof text_len, cudaMemcpyDeviceToHost, ctx->stream));
        }
    }
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    if (sum_rgb) CK(cudaMemcpyAsync(sum_rgb, ctx->sum.p, npix * 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (rgb8) CK(cudaMemcpyAsync(rgb8, ctx->rgb8.p, npix * 3, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (waited) {
        int timed_out = 0;
        CK(cudaMemcpy(&timed_out, ctx->frame_timeout.p, sizeof timed_out, cudaMemcpyDeviceToHost));
        if (timed_out) {
            CK(cudaMemset(ctx->frame_timeout.p, 0, sizeof(int)));
            return fail(ctx, RT_E_STATE, "rt_readback: the other ranks' tiles did not arrive within 10 s (did every rank call rt_resolve_tiles for this frame?)");
        }
    }
    if (ppm_len) {
        if (ppm) {
            if (*ppm_len < text_len) { *ppm_len = size_t(text_len); return fail(ctx, RT_E_INVALID, "rt_readback: ppm buffer too small"); }
            CK(cudaMemcpy(ppm, ctx->text.p, size_t(text_len), cudaMemcpyDeviceToHost));
        }
        *ppm_len = size_t(text_len);
    }
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]) == cudaSuccess) ctx->stats.resolve_ms = ms;
    ctx->stats.kernel_launches = (ctx->render_timed ? 1 : 0) + launches;
    return RT_OK;
}

// Split read-back.  rt_readback_begin does what rt_readback does up to the frame's bytes being ready on the device (resolve, P3
// text, its length -- that is rt_readback itself, asked for no copies), then starts the two bulk copies on the context's copy
// stream and returns; rt_readback_end waits for them.  In between the caller may upload and rt_render the NEXT frame: a render
// writes the accumulation layers only, and whatever resolves next waits for copy_done on the device (after_pending_copies).
extern "C" int rt_readback_begin(rt_ctx* ctx, int spp_total, uint8_t* rgb8, char* ppm, size_t ppm_capacity) {
    NvtxRange nvtx_("rt_readback_begin");
    if (!ctx) return RT_E_INVALID;
    if (ctx->copy_pending) return fail(ctx, RT_E_STATE, "rt_readback_begin: a split read-back is in flight (rt_readback_end first)");
    if (!rgb8 && !ppm) return fail(ctx, RT_E_INVALID, "rt_readback_begin: neither rgb8 nor ppm wanted");
    size_t text_len = 0;
    const int rc = rt_readback(ctx, spp_total, nullptr, nullptr, nullptr, ppm ? &text_len : nullptr);   // synchronises the render stream
    if (rc != RT_OK) return rc;
    if (ppm && ppm_capacity < text_len) return fail(ctx, RT_E_INVALID, "rt_readback_begin: ppm buffer too small");
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->copy_done, cudaEventDisableTiming));
    }
    const size_t npix = size_t(ctx->W) * ctx->H;
    if (rgb8) CK(cudaMemcpyAsync(rgb8, ctx->rgb8.p, npix * 3, cudaMemcpyDeviceToHost, ctx->copy_stream));
    if (ppm) CK(cudaMemcpyAsync(ppm, ctx->text.p, text_len, cudaMemcpyDeviceToHost, ctx->copy_stream));
    CK(cudaEventRecord(ctx->copy_done, ctx->copy_stream));
    ctx->copy_pending = true;
    ctx->copy_text_len = ppm ? text_len : 0;
    return RT_OK;
}

extern "C" int rt_readback_end(rt_ctx* ctx, size_t* ppm_len) {
    NvtxRange nvtx_("rt_readback_end");
    if (!ctx) return RT_E_INVALID;
    if (!ctx->copy_pending) return fail(ctx, RT_E_STATE, "rt_readback_end: no split read-back in flight");
    CK(cudaSetDevice(ctx->device));
    ctx->copy_pending = false;                                  // whatever the copies' fate, the hand-over is finished
    CK(cudaEventSynchronize(ctx->copy_done));
    if (ppm_len) *ppm_len = ctx->copy_text_len;
    return RT_OK;
}

// ---------------------------------------------------------------- 8f-2: more output formats --
static int resolve_views(rt_ctx* ctx, int spp_total, const FrameViews& v) {
    if (ctx->peers.n > 0 && !ctx->folded) return fail(ctx, RT_E_STATE, "resolve: peers are mapped; every rank must call rt_fold_slices before the fused resolve");
    const size_t npix = size_t(ctx->W) * ctx->H;
    CK(ensure_rgb8(ctx, npix));
    const double scale = spp_total > 0 ? 1.0 / spp_total : 0.0;
    CK(after_pending_copies(ctx));
    if (ctx->precision == RT_F32)
        CK(launch_resolve<float4>(reinterpret_cast<const float4*>(ctx->accum.p), live_slices(ctx), live_peers(ctx), npix, scale,
                                  nullptr, ctx->rgb8.p, nullptr, ctx->stream, v, ctx->tiles));
    else
        CK(launch_resolve<double4>(reinterpret_cast<const double4*>(ctx->accum.p), live_slices(ctx), live_peers(ctx), npix, scale,
                                   nullptr, ctx->rgb8.p, nullptr, ctx->stream, v, ctx->tiles));
    return RT_OK;
}

extern "C" int rt_readback_image(rt_ctx* ctx, int spp_total, int format, void* buf, size_t* len) {
    if (!ctx || !len) return RT_E_INVALID;
    if (ctx->W == 0 || !ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_readback_image: nothing rendered");
    if (format == RT_IMG_P3) return rt_readback(ctx, spp_total, nullptr, nullptr, static_cast<char*>(buf), len);
    if (format != RT_IMG_P6 && format != RT_IMG_PFM) return fail(ctx, RT_E_INVALID, "rt_readback_image: unknown format");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(ctx->W) * ctx->H;
    char hdr[64];
    const int hl = format == RT_IMG_P6 ? std::snprintf(hdr, sizeof hdr, "P6\n%d %d\n255\n", ctx->W, ctx->H)
                                       : std::snprintf(hdr, sizeof hdr, "PF\n%d %d\n-1.0\n", ctx->W, ctx->H);
    const size_t body = format == RT_IMG_P6 ? npix * 3 : npix * 3 * sizeof(float);
    const size_t need = size_t(hl) + body;
    if (!buf) { *len = need; return RT_OK; }
    if (*len < need) { *len = need; return fail(ctx, RT_E_INVALID, "rt_readback_image: buffer too small"); }
    FrameViews v{nullptr, nullptr, ctx->W, ctx->H};
    if (format == RT_IMG_PFM) { CK(ctx->pfm.resize(npix * 3)); v.pfm = ctx->pfm.p; }
    int rc = resolve_views(ctx, spp_total, v);
    if (rc != RT_OK) return rc;
    std::memcpy(buf, hdr, size_t(hl));
    CK(cudaMemcpyAsync(static_cast<char*>(buf) + hl, format == RT_IMG_P6 ? static_cast<const void*>(ctx->rgb8.p) : static_cast<const void*>(ctx->pfm.p),
                       body, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *len = need;
    return RT_OK;
}

extern "C" int rt_preview(rt_ctx* ctx, int spp_total, void** dev_rgba8, uint8_t* host_rgba8) {
    if (!ctx) return RT_E_INVALID;
    if (ctx->W == 0 || !ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_preview: nothing rendered");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(ctx->W) * ctx->H;
    CK(ctx->rgba8.resize(npix));
    int rc = resolve_views(ctx, spp_total, FrameViews{nullptr, ctx->rgba8.p, ctx->W, ctx->H});
    if (rc != RT_OK) return rc;
    if (dev_rgba8) *dev_rgba8 = ctx->rgba8.p;
    if (host_rgba8) {
        CK(cudaMemcpyAsync(host_rgba8, ctx->rgba8.p, npix * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    return RT_OK;
}

// profiling aid: the TRACE-stage scheduler counters a -DRT_POOL_STATS build of rt_wavefront.cuh collects
// (all zero in a normal build); out[0] = rays.  Not part of the reference-facing surface.
extern "C" int rt_debug_counters(rt_ctx* ctx, uint64_t out[32]) {
    if (!ctx || !out) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->ray_counter.p, 32 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return RT_OK;
}

// profiling aid: the per-warp records a -DRT_WARP_TIMES build writes (8 words per warp: start, out-of-jobs and exit times in
// ns, rounds and rays before / after the frame ran out of jobs, valid flag); zeros otherwise.  n_words <= 8 * 32 * 256.
extern "C" int rt_debug_warp_times(rt_ctx* ctx, uint64_t* out, size_t n_words) {
    if (!ctx || !out || n_words > RT_DEBUG_WORDS - 32) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->ray_counter.p + 32, n_words * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return RT_OK;
}

extern "C" int rt_get_stats(rt_ctx* ctx, rt_stats* out) {
    if (!ctx || !out) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->render_timed) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]) == cudaSuccess) ctx->stats.render_ms = ms;
        unsigned long long rays = 0;
        CK(cudaMemcpy(&rays, ctx->ray_counter.p, sizeof rays, cudaMemcpyDeviceToHost));
        ctx->stats.rays = rays;
    }
    *out = ctx->stats;
    return RT_OK;
}

// ---------------------------------------------------------------- multi-GPU plumbing --
extern "C" int rt_accum_buffer(rt_ctx* ctx, void** dev_ptr, size_t* n_elems, int* elem_bytes) {
    if (!ctx) return RT_E_INVALID;
    if (!ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_accum_buffer: nothing rendered");
    if (dev_ptr) *dev_ptr = ctx->accum.p;
    if (n_elems) *n_elems = size_t(ctx->slices) * ctx->W * ctx->H * 4;
    if (elem_bytes) *elem_bytes = ctx->precision == RT_F32 ? 4 : 8;
    return RT_OK;
}

extern "C" int rt_plan_layers(int sample_count, long long tiles, long long resident_warps, int jobs_per_warp, int* sizes, int* n_bulk) {
    if (sample_count <= 0 || !sizes || !n_bulk) return 0;
    return plan_layers(sample_count, tiles, resident_warps, jobs_per_warp, sizes, n_bulk);
}

extern "C" int rt_fold_slices(rt_ctx* ctx) {
    NvtxRange nvtx_("rt_fold_slices");
    if (!ctx) return RT_E_INVALID;
    if (!ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_fold_slices: nothing rendered");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(ctx->W) * ctx->H;
    if (ctx->precision == RT_F32) CK(launch_fold<float4>(reinterpret_cast<float4*>(ctx->accum.p), ctx->slices, npix, ctx->stream));
    else CK(launch_fold<double4>(reinterpret_cast<double4*>(ctx->accum.p), ctx->slices, npix, ctx->stream));
    ctx->folded = true;
    return RT_OK;
}

// The blob a rank hands to the root: the CUDA IPC handle of its accumulation buffer + what the buffer is
// (image size, precision), so that the root can refuse a peer that rendered something else.
struct IpcBlob { cudaIpcMemHandle_t h; int32_t W, H, precision, magic; };
static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(IpcBlob) == RT_IPC_HANDLE_BYTES, "ipc blob size");

extern "C" int rt_ipc_export(rt_ctx* ctx, void* handle) {
    if (!ctx || !handle) return RT_E_INVALID;
    if (!ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_ipc_export: nothing rendered");
    CK(cudaSetDevice(ctx->device));
    IpcBlob b;
    CK(cudaIpcGetMemHandle(&b.h, ctx->accum.p));
    b.W = ctx->W; b.H = ctx->H; b.precision = ctx->precision; b.magic = 0x52544950;
    std::memcpy(handle, &b, sizeof b);
    return RT_OK;
}

extern "C" int rt_ipc_close_peers(rt_ctx* ctx) {
    if (!ctx) return RT_E_INVALID;
    for (int i = 0; i < RT_MAX_PEERS; ++i)
        if (ctx->peer_mapped[i]) { cudaIpcCloseMemHandle(ctx->peer_mapped[i]); ctx->peer_mapped[i] = nullptr; }
    ctx->peers.n = 0;
    return RT_OK;
}

extern "C" int rt_ipc_open_peers(rt_ctx* ctx, const void* handles, int n_peers) {
    if (!ctx || (n_peers && !handles)) return RT_E_INVALID;
    if (n_peers > RT_MAX_PEERS) return fail(ctx, RT_E_INVALID, "rt_ipc_open_peers: too many peers");
    if (n_peers && !ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_ipc_open_peers: render once first (the peers must match this context's frame)");
    CK(cudaSetDevice(ctx->device));
    rt_ipc_close_peers(ctx);
    for (int i = 0; i < n_peers; ++i) {
        IpcBlob b;
        std::memcpy(&b, static_cast<const char*>(handles) + sizeof(IpcBlob) * i, sizeof b);
        if (b.magic != 0x52544950 || b.W != ctx->W || b.H != ctx->H || b.precision != ctx->precision) {
            rt_ipc_close_peers(ctx);
            return fail(ctx, RT_E_INVALID, "rt_ipc_open_peers: a peer's buffer is not this context's image size / precision");
        }
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, b.h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { rt_ipc_close_peers(ctx); return cuda_fail(ctx, e, "cudaIpcOpenMemHandle"); }
        ctx->peer_mapped[i] = p;
        ctx->peers.ptr[i] = p;
        // only layer 0 of a peer is ever read: the resolve refuses to run with peers unless every rank has folded
        // (rt_fold_slices), so the peers' layer plans do not matter
        ctx->peers.slices[i] = 1;
        ctx->peers.n = i + 1;
    }
    return RT_OK;
}

// ---- tile-sharded frames: the image hand-over (rt_b200.h) ----
struct FrameBlob { cudaIpcMemHandle_t h; int32_t W, H, magic, pad; };
static_assert(sizeof(FrameBlob) == RT_IPC_HANDLE_BYTES, "frame blob size");

// ---- dynamic sharding: one job counter per frame for all GPUs of the node (include/rt_b200.h) ----
struct JobsBlob { cudaIpcMemHandle_t h; int magic, n, pad0, pad1; };
static_assert(sizeof(JobsBlob) == RT_IPC_HANDLE_BYTES, "jobs blob size");

extern "C" int rt_jobs_close(rt_ctx* ctx) {
    if (!ctx) return RT_E_INVALID;
    if (ctx->jobs_mapped) { cudaIpcCloseMemHandle(ctx->jobs_mapped); ctx->jobs_mapped = nullptr; }
    ctx->jobs_ptr = nullptr; ctx->jobs_root = false; ctx->jobs_frame = 0;
    return RT_OK;
}

extern "C" int rt_jobs_export(rt_ctx* ctx, void* handle) {
    if (!ctx || !handle) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    rt_jobs_close(ctx);
    // cudaMalloc'ed on purpose (not from the stream-ordered pool): the block is exported through CUDA IPC
    CK(ctx->jobs_own.resize(256));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemset(ctx->jobs_own.p, 0, 256 * sizeof(unsigned)));
    JobsBlob b;
    CK(cudaIpcGetMemHandle(&b.h, ctx->jobs_own.p));
    b.magic = 0x52544a42; b.n = RT_JOB_COUNTERS; b.pad0 = b.pad1 = 0;
    std::memcpy(handle, &b, sizeof b);
    ctx->jobs_ptr = ctx->jobs_own.p; ctx->jobs_root = true; ctx->jobs_frame = 0;
    return RT_OK;
}

extern "C" int rt_jobs_open(rt_ctx* ctx, const void* handle) {
    if (!ctx || !handle) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    JobsBlob b;
    std::memcpy(&b, handle, sizeof b);
    if (b.magic != 0x52544a42 || b.n != RT_JOB_COUNTERS) return fail(ctx, RT_E_INVALID, "rt_jobs_open: not a job-counter handle of this library version");
    rt_jobs_close(ctx);
    void* p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, b.h, cudaIpcMemLazyEnablePeerAccess));
    ctx->jobs_mapped = p; ctx->jobs_ptr = static_cast<unsigned*>(p); ctx->jobs_root = false; ctx->jobs_frame = 0;
    return RT_OK;
}

extern "C" int rt_jobs_share(rt_ctx* ctx, rt_ctx* root) {
    if (!ctx || !root) return RT_E_INVALID;
    if (!root->jobs_root || !root->jobs_own.p) return fail(ctx, RT_E_STATE, "rt_jobs_share: the root has not called rt_jobs_export");
    if (ctx == root) return RT_OK;
    CK(cudaSetDevice(ctx->device));
    if (ctx->device != root->device) {
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, ctx->device, root->device));
        if (!can) return fail(ctx, RT_E_INVALID, "rt_jobs_share: this device cannot reach the root's memory");
        cudaError_t e = cudaDeviceEnablePeerAccess(root->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(ctx, e, "cudaDeviceEnablePeerAccess");
        cudaGetLastError();
    }
    rt_jobs_close(ctx);
    ctx->jobs_ptr = root->jobs_own.p; ctx->jobs_root = false; ctx->jobs_frame = root->jobs_frame;
    return RT_OK;
}

extern "C" int rt_frame_export(rt_ctx* ctx, void* handle) {
    if (!ctx || !handle) return RT_E_INVALID;
    if (ctx->W == 0) return fail(ctx, RT_E_STATE, "rt_frame_export: render once first (the image size must be known)");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(ctx->W) * ctx->H;
    CK(ensure_rgb8(ctx, npix));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemset(ctx->rgb8.p + frame_counter_offset(npix), 0, 128));
    FrameBlob b;
    CK(cudaIpcGetMemHandle(&b.h, ctx->rgb8.p));
    b.W = ctx->W; b.H = ctx->H; b.magic = 0x52544652; b.pad = 0;
    std::memcpy(handle, &b, sizeof b);
    ctx->frame_exported = true; ctx->frame_target = 0;
    return RT_OK;
}

extern "C" int rt_frame_close(rt_ctx* ctx) {
    if (!ctx) return RT_E_INVALID;
    if (ctx->frame_remote) { cudaIpcCloseMemHandle(ctx->frame_remote); ctx->frame_remote = nullptr; }
    ctx->frame_exported = false; ctx->frame_target = 0;
    return RT_OK;
}

extern "C" int rt_frame_open(rt_ctx* ctx, const void* handle) {
    if (!ctx || !handle) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    FrameBlob b;
    std::memcpy(&b, handle, sizeof b);
    if (b.magic != 0x52544652) return fail(ctx, RT_E_INVALID, "rt_frame_open: not a frame handle");
    if (ctx->frame_remote) { cudaIpcCloseMemHandle(ctx->frame_remote); ctx->frame_remote = nullptr; }
    void* p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, b.h, cudaIpcMemLazyEnablePeerAccess));
    ctx->frame_remote = p; ctx->frame_remote_W = b.W; ctx->frame_remote_H = b.H;
    return RT_OK;
}

extern "C" int rt_resolve_tiles(rt_ctx* ctx, int spp_total) {
    NvtxRange nvtx_("rt_resolve_tiles");
    if (!ctx) return RT_E_INVALID;
    if (ctx->W == 0 || !ctx->accum.p) return fail(ctx, RT_E_STATE, "rt_resolve_tiles: nothing rendered");
    if (ctx->tiles.world <= 1) return fail(ctx, RT_E_STATE, "rt_resolve_tiles: the last render was not tile-sharded");
    if (!ctx->frame_remote) return fail(ctx, RT_E_STATE, "rt_resolve_tiles: no root image mapped (rt_frame_open)");
    if (ctx->frame_remote_W != ctx->W || ctx->frame_remote_H != ctx->H) return fail(ctx, RT_E_STATE, "rt_resolve_tiles: the root's image has another size");
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(ctx->W) * ctx->H;
    const double scale = spp_total > 0 ? 1.0 / spp_total : 0.0;
    unsigned char* image = static_cast<unsigned char*>(ctx->frame_remote);
    const FrameViews no_views{nullptr, nullptr, 0, 0};
    PeerList none = {};
    if (ctx->precision == RT_F32)
        CK(launch_resolve<float4>(reinterpret_cast<const float4*>(ctx->accum.p), live_slices(ctx), none, npix, scale, nullptr, image, nullptr, ctx->stream, no_views, ctx->tiles));
    else
        CK(launch_resolve<double4>(reinterpret_cast<const double4*>(ctx->accum.p), live_slices(ctx), none, npix, scale, nullptr, image, nullptr, ctx->stream, no_views, ctx->tiles));
    CK(launch_frame_signal(reinterpret_cast<unsigned*>(image + frame_counter_offset(npix)), ctx->stream));
    ctx->tiles_pending = false;
    ctx->stats.kernel_launches = (ctx->render_timed ? 1 : 0) + 2;
    return RT_OK;
}

// ------------------------------------------------------------------------ parity hooks --
extern "C" int rt_trace_closest(rt_ctx* ctx, const rt_ray* rays, size_t n, double tmin, double tmax,
                                int precision, int backend, rt_hit* out) {
    if (!ctx || (n && (!rays || !out))) return RT_E_INVALID;
    if (!ctx->has_scene) return fail(ctx, RT_E_STATE, "rt_trace_closest: no scene uploaded");
    if (n == 0) return RT_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf<double> d_rays; DevBuf<HitRec> d_out;
    CK(d_rays.resize(n * 7)); CK(d_out.resize(n));
    CK(cudaMemcpyAsync(d_rays.p, rays, n * sizeof(rt_ray), cudaMemcpyHostToDevice, ctx->stream));
    cudaError_t e = precision == RT_F32
        ? launch_trace_closest<float>(scene_view<float>(ctx), d_rays.p, n, tmin, tmax, backend == RT_TRACE_BRUTE, d_out.p, ctx->stream)
        : launch_trace_closest<double>(scene_view<double>(ctx), d_rays.p, n, tmin, tmax, backend == RT_TRACE_BRUTE, d_out.p, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out.p, n * sizeof(rt_hit), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    d_rays.release(); d_out.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_trace_closest");
    return RT_OK;
}

extern "C" int rt_scatter(rt_ctx* ctx, const rt_material* materials, const rt_ray* rays_in, const rt_hit* hits,
                          size_t n, uint64_t seed, int precision,
                          int32_t* scattered_flag, double* attenuation_rgb, rt_ray* scattered) {
    if (!ctx || (n && (!materials || !rays_in || !hits || !scattered_flag || !attenuation_rgb || !scattered))) return RT_E_INVALID;
    if (n == 0) return RT_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf<MatRec> d_m; DevBuf<double> d_in, d_att, d_out; DevBuf<HitRec> d_h; DevBuf<int> d_f;
    CK(d_m.resize(n)); CK(d_in.resize(7 * n)); CK(d_att.resize(3 * n)); CK(d_out.resize(7 * n)); CK(d_h.resize(n)); CK(d_f.resize(n));
    std::vector<rt_material> mats(materials, materials + n);
    for (auto& m : mats) if (m.type == RT_METAL) m.param = m.param < 1 ? m.param : 1;      // src/material.h:33
    CK(cudaMemcpyAsync(d_m.p, mats.data(), n * sizeof(rt_material), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_in.p, rays_in, n * sizeof(rt_ray), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_h.p, hits, n * sizeof(rt_hit), cudaMemcpyHostToDevice, ctx->stream));
    cudaError_t e = precision == RT_F32
        ? launch_scatter<float>(d_m.p, d_in.p, d_h.p, n, seed, d_f.p, d_att.p, d_out.p, ctx->stream)
        : launch_scatter<double>(d_m.p, d_in.p, d_h.p, n, seed, d_f.p, d_att.p, d_out.p, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(scattered_flag, d_f.p, n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(attenuation_rgb, d_att.p, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(scattered, d_out.p, n * sizeof(rt_ray), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    d_m.release(); d_in.release(); d_att.release(); d_out.release(); d_h.release(); d_f.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_scatter");
    return RT_OK;
}

extern "C" int rt_gen_rays(rt_ctx* ctx, const rt_camera* cam, const int32_t* px_i, const int32_t* px_j,
                           const uint32_t* sample, size_t n, uint64_t seed, int precision, rt_ray* out) {
    if (!ctx || !cam || (n && (!px_i || !px_j || !sample || !out))) return RT_E_INVALID;
    if (cam->image_width < 1 || !(cam->aspect_ratio > 0)) return fail(ctx, RT_E_INVALID, "rt_gen_rays: bad image size");
    if (n == 0) return RT_OK;
    CK(cudaSetDevice(ctx->device));
    DevBuf<int> d_i, d_j; DevBuf<uint32_t> d_s; DevBuf<double> d_out;
    CK(d_i.resize(n)); CK(d_j.resize(n)); CK(d_s.resize(n)); CK(d_out.resize(7 * n));
    CK(cudaMemcpyAsync(d_i.p, px_i, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_j.p, px_j, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_s.p, sample, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    const CameraFrame<double> f = camera_frame(*cam);
    cudaError_t e = precision == RT_F32
        ? launch_gen_rays<float>(narrow_frame<float>(f), d_i.p, d_j.p, d_s.p, n, seed, d_out.p, ctx->stream)
        : launch_gen_rays<double>(f, d_i.p, d_j.p, d_s.p, n, seed, d_out.p, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out.p, n * sizeof(rt_ray), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    d_i.release(); d_j.release(); d_s.release(); d_out.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_gen_rays");
    return RT_OK;
}

extern "C" int rt_bench_closest(rt_ctx* ctx, const rt_camera* cam, size_t n, uint64_t seed, int precision,
                                int backend, int repeats, double* ms, int32_t* prim_out, float* t_out,
                                rt_ray* rays_out, size_t rays_out_n) {
    if (!ctx || !cam) return RT_E_INVALID;
    if (!ctx->has_scene) return fail(ctx, RT_E_STATE, "rt_bench_closest: no scene uploaded");
    if (n == 0) return RT_OK;
    CK(cudaSetDevice(ctx->device));
    rt_camera c = *cam;
    c.image_width = 8192; c.aspect_ratio = 1.0;                 // the virtual 8192 x 8192 image of SURVEY.md 8d
    const CameraFrame<double> f = camera_frame(c);
    rays_out_n = rays_out ? std::min(rays_out_n, n) : 0;
    DevBuf<int> d_prim; DevBuf<float> d_t; DevBuf<double> d_rays;
    CK(d_prim.resize(n)); CK(d_t.resize(n)); CK(d_rays.resize(std::max<size_t>(1, 7 * rays_out_n)));
    size_t smem = 0;
    cudaError_t e = cudaSuccess;
    repeats = std::max(1, repeats);
    float best = 1e30f;
    for (int it = 0; it < repeats && e == cudaSuccess; ++it) {
        cudaEventRecord(ctx->ev[2], ctx->stream);
        if (precision == RT_F32) {
            const bool staged = can_stage<float>(ctx, &smem);
            e = launch_bench_closest<float>(scene_view<float>(ctx), narrow_frame<float>(f), n, seed, backend == RT_TRACE_BRUTE, staged, smem,
                                            ctx->sm_count, d_prim.p, d_t.p, d_rays.p, rays_out_n, ctx->stream);
        } else {
            const bool staged = can_stage<double>(ctx, &smem);
            e = launch_bench_closest<double>(scene_view<double>(ctx), f, n, seed, backend == RT_TRACE_BRUTE, staged, smem,
                                             ctx->sm_count, d_prim.p, d_t.p, d_rays.p, rays_out_n, ctx->stream);
        }
        cudaEventRecord(ctx->ev[3], ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        float t = 0;
        if (e == cudaSuccess && cudaEventElapsedTime(&t, ctx->ev[2], ctx->ev[3]) == cudaSuccess) best = std::min(best, t);
    }
    if (ms) *ms = best;
    if (e == cudaSuccess && prim_out) e = cudaMemcpy(prim_out, d_prim.p, n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && t_out) e = cudaMemcpy(t_out, d_t.p, n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && rays_out_n) e = cudaMemcpy(rays_out, d_rays.p, rays_out_n * sizeof(rt_ray), cudaMemcpyDeviceToHost);
    d_prim.release(); d_t.release(); d_rays.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_bench_closest");
    return RT_OK;
}

// -------------------------------------------------------------------- scene generators --
// src/main.cpp:11-53.  The reference draws from one default-seeded std::mt19937 through
// uniform_real_distribution<double>(0,1) (src/rtweekend.h:25-29); g++ evaluates the arguments
// of vec3(...) right to left, so the draws are sequenced explicitly here.
extern "C" size_t rt_scene_rtiow(int g, double p_diffuse, double p_metal, uint32_t mt_seed,
                                 rt_sphere* spheres, rt_material* materials, size_t cap) {
    std::mt19937 gen(mt_seed);
    std::uniform_real_distribution<double> dist(0.0, 1.0);
    auto rnd = [&]() { return dist(gen); };
    auto rnd_range = [&](double lo, double hi) { return lo + (hi - lo) * rnd(); };
    size_t n = 0;
    auto push = [&](double cx, double cy, double cz, double vy, double r, bool moving, int type, double ar, double ag, double ab, double param) {
        if (n < cap && spheres && materials) {
            rt_sphere& s = spheres[n];
            s.center1[0] = cx; s.center1[1] = cy; s.center1[2] = cz;
            s.center_vec[0] = 0; s.center_vec[1] = vy; s.center_vec[2] = 0;
            s.radius = r; s.is_moving = moving; s.material = int32_t(n);
            rt_material& m = materials[n];
            m.type = type; m.pad = 0; m.albedo[0] = ar; m.albedo[1] = ag; m.albedo[2] = ab; m.param = param;
        }
        ++n;
    };
    push(0, -1000, 0, 0, 1000, false, RT_LAMBERTIAN, 0.5, 0.5, 0.5, 0);
    for (int a = -g; a < g; a++) {
        for (int b = -g; b < g; b++) {
            const double choose_mat = rnd();
            const double cz = b + 0.9 * rnd();               // first draw feeds z (:20)
            const double cx = a + 0.9 * rnd();
            const double dx = cx - 4, dy = 0.2 - 0.2, dz = cz - 0;
            if (std::sqrt(dx * dx + dy * dy + dz * dz) > 0.9) {
                if (choose_mat < p_diffuse) {
                    double rz = rnd(), ry = rnd(), rx = rnd();   // right operand of color::random()*color::random() first (:27)
                    double lz = rnd(), ly = rnd(), lx = rnd();
                    const double c2y = 0.2 + rnd_range(0, .5);   // center + vec3(0, random_double(0,.5), 0) (:29)
                    push(cx, 0.2, cz, c2y - 0.2, 0.2, true, RT_LAMBERTIAN, lx * rx, ly * ry, lz * rz, 0);
                } else if (choose_mat < p_metal) {
                    double az = rnd_range(0.5, 1), ay = rnd_range(0.5, 1), ax = rnd_range(0.5, 1);
                    const double fuzz = rnd_range(0, 0.5);
                    push(cx, 0.2, cz, 0, 0.2, false, RT_METAL, ax, ay, az, fuzz < 1 ? fuzz : 1);
                } else {
                    push(cx, 0.2, cz, 0, 0.2, false, RT_DIELECTRIC, 0, 0, 0, 1.5);
                }
            }
        }
    }
    push(0, 1, 0, 0, 1.0, false, RT_DIELECTRIC, 0, 0, 0, 1.5);
    push(-4, 1, 0, 0, 1.0, false, RT_LAMBERTIAN, 0.4, 0.2, 0.1, 0);
    push(4, 1, 0, 0, 1.0, false, RT_METAL, 0.7, 0.6, 0.5, 0.0);
    return n;
}

// src/main.cpp:57-68; scale = g/11 moves the camera back for the wider grids of config 5
extern "C" int rt_camera_rtiow(rt_camera* cam, int image_width, int spp, int max_depth, double scale) {
    if (!cam) return RT_E_INVALID;
    if (!(scale > 0)) scale = 1.0;
    std::memset(cam, 0, sizeof *cam);
    cam->aspect_ratio = 16.0 / 9.0;
    cam->image_width = image_width;
    cam->samples_per_pixel = spp;
    cam->max_depth = max_depth;
    cam->vfov = 20;
    cam->lookfrom[0] = 13 * scale; cam->lookfrom[1] = 2 * scale; cam->lookfrom[2] = 3 * scale;
    cam->lookat[0] = cam->lookat[1] = cam->lookat[2] = 0;
    cam->vup[0] = 0; cam->vup[1] = 1; cam->vup[2] = 0;
    cam->defocus_angle = 0.6;
    cam->focus_dist = 10.0 * scale;
    return RT_OK;
}

// ------------------------------------------------------- framebuffer / builder hooks --
extern "C" int rt_write_color(rt_ctx* ctx, const double* sums_rgb, size_t n, int samples_per_pixel, uint8_t* rgb8) {
    if (!ctx || (n && (!sums_rgb || !rgb8)) || samples_per_pixel < 1) return RT_E_INVALID;
    if (n == 0) return RT_OK;
    CK(cudaSetDevice(ctx->device));
    std::vector<double4> q(n);
    for (size_t i = 0; i < n; ++i) q[i] = make_double4(sums_rgb[3 * i], sums_rgb[3 * i + 1], sums_rgb[3 * i + 2], double(samples_per_pixel));
    DevBuf<double4> d_q; DevBuf<unsigned char> d_b;
    CK(d_q.resize(n)); CK(d_b.resize(3 * n + 16));
    CK(cudaMemcpyAsync(d_q.p, q.data(), n * sizeof(double4), cudaMemcpyHostToDevice, ctx->stream));
    PeerList none = {};
    cudaError_t e = launch_resolve<double4>(d_q.p, 1, none, n, 1.0 / samples_per_pixel, nullptr, d_b.p, nullptr, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb8, d_b.p, 3 * n, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    d_q.release(); d_b.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_write_color");
    return RT_OK;
}

extern "C" int rt_encode_p3(rt_ctx* ctx, const uint8_t* rgb8, int width, int height, char* ppm, size_t* ppm_len) {
    if (!ctx || !rgb8 || !ppm_len || width < 1 || height < 1) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    const size_t npix = size_t(width) * height;
    EncodeHeader hdr;
    hdr.len = std::snprintf(hdr.text, sizeof hdr.text, "P3\n%d %d\n255\n", width, height);
    const int nb = encode_blocks(npix);
    DevBuf<unsigned char> d_b; DevBuf<unsigned> d_s; DevBuf<unsigned long long> d_o; DevBuf<char> d_t;
    CK(d_b.resize(3 * npix + 16)); CK(d_s.resize(nb + 1)); CK(d_o.resize(nb + 2)); CK(d_t.resize(npix * 12 + 64));
    CK(cudaMemcpyAsync(d_b.p, rgb8, 3 * npix, cudaMemcpyHostToDevice, ctx->stream));
    unsigned long long len = 0;
    cudaError_t e = launch_encode_p3(d_b.p, npix, hdr, d_s.p, d_o.p, d_o.p + nb + 1, d_t.p, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&len, d_o.p + nb + 1, sizeof len, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    int rc = RT_OK;
    if (e == cudaSuccess) {
        if (ppm && *ppm_len >= len) e = cudaMemcpy(ppm, d_t.p, size_t(len), cudaMemcpyDeviceToHost);
        else if (ppm) rc = fail(ctx, RT_E_INVALID, "rt_encode_p3: ppm buffer too small");
        *ppm_len = size_t(len);
    }
    d_b.release(); d_s.release(); d_o.release(); d_t.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_encode_p3");
    return rc;
}

// No counterpart in the reference: the encoder for callers that may not allocate (rt_static.cu holds the scratch in a
// file-scope __device__ array).  Nothing here or below it calls cudaMalloc / cudaFree; asynchronous on the context's stream.
extern "C" int rt_encode_p3_static(rt_ctx* ctx, const uint8_t* dev_rgb8, int width, int height, char* dev_ppm, size_t ppm_capacity,
                                   unsigned long long* dev_ppm_len) {
    if (!ctx) return RT_E_INVALID;
    char msg[256] = "";
    const int rc = encode_p3_static(ctx->device, ctx->stream, dev_rgb8, width, height, dev_ppm, ppm_capacity, dev_ppm_len, msg, sizeof msg);
    if (rc != 0) return fail(ctx, rc == -2 ? RT_E_CUDA : RT_E_INVALID, msg);
    return RT_OK;
}

extern "C" int rt_demo_compute(rt_ctx* ctx, const rt_demo_sphere* spheres, size_t n, int width, int height, uint8_t* rgba8) {
    static_assert(sizeof(rt_demo_sphere) == sizeof(DemoSphere), "rt_demo_sphere layout");
    if (!ctx || !rgba8 || width < 1 || height < 1 || (long long)width * height > (1ll << 30)) return RT_E_INVALID;
    CK(cudaSetDevice(ctx->device));
    const rt_demo_sphere hard_coded = { { 3.0f, 0.0f, 0.0f }, 1.0f, { 1.0f, 0.3f, 0.7f, 1.0f } };   // ray_tracing.comp:49-52
    if (!spheres) { spheres = &hard_coded; n = 1; }
    const size_t bytes = size_t(width) * height * 4;
    DevBuf<DemoSphere> d_s; DevBuf<unsigned char> d_img;
    CK(d_s.resize(std::max<size_t>(1, n))); CK(d_img.resize(bytes));
    cudaError_t e = n ? cudaMemcpyAsync(d_s.p, spheres, n * sizeof(DemoSphere), cudaMemcpyHostToDevice, ctx->stream) : cudaSuccess;
    if (e == cudaSuccess) e = launch_demo_compute(d_s.p, int(n), width, height, d_img.p, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgba8, d_img.p, bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    d_s.release(); d_img.release();
    if (e != cudaSuccess) return cuda_fail(ctx, e, "rt_demo_compute");
    return RT_OK;
}

extern "C" int rt_build_bvh_host(const rt_sphere* spheres, size_t n, int max_leaf, float* nodes_out, int32_t* order_out,
                                 int32_t* root_ref, int32_t* depth, int32_t* n_inner) {
    if (n && !spheres) return RT_E_INVALID;
    std::vector<HostBox> boxes(n);
    for (size_t k = 0; k < n; ++k) boxes[k] = sphere_box(spheres[k].center1, spheres[k].center_vec, spheres[k].radius, spheres[k].is_moving != 0);
    HostBVH bvh;
    build_bvh_sah(boxes, max_leaf > 0 ? max_leaf : 4, bvh);
    if (nodes_out && !bvh.nodes.empty()) std::memcpy(nodes_out, bvh.nodes.data(), bvh.nodes.size() * sizeof(float));
    if (order_out && n) std::memcpy(order_out, bvh.order.data(), n * sizeof(int));
    if (root_ref) *root_ref = bvh.root_ref;
    if (depth) *depth = bvh.depth;
    if (n_inner) *n_inner = bvh.n_inner;
    return RT_OK;
}
