// viso_ba.cpp — host side of the C ABI declared in include/viso_ba.h.
//
// Replaces the header-only g2o wrapper of the reference (include/bundle_adjuster.h:108-187):
// the pools the reference pre-allocates in its constructor (:122-166) become one device arena
// that is kept across Clear()/SetProblem(); graph construction (addVertex/addEdge) becomes the
// structure build in upload(); optimizer_.optimize() — never reached in the reference —
// becomes the LM driver in run_lm().  No CPU compute path exists here: every numeric result
// comes from the kernels in ba_kernels.cu (there is deliberately no fallback).
#include "../../include/viso_ba.h"
#include "ba_device.h"
#include "ba_setup.h"
#include "ba_solve_plan.h"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

using namespace visoba;

namespace {

// ------------------------------------------------------------------------------------------
// NCCL through dlopen: the library must load (and every single-GPU path must work) on a box
// without NCCL; only viso_ba_shard_init needs it.
// ------------------------------------------------------------------------------------------
struct Id128 { // ncclUniqueId, passed by value
    char b[128];
};
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, Id128, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclDouble = 8; // ncclFloat64
constexpr int kNcclSum = 0, kNcclMax = 2;

void nccl_load(NcclApi& api)
{
    const char* names[] = { "libnccl.so.2", "libnccl.so" };
    for (const char* n : names) {
        api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.lib) break;
    }
    if (!api.lib) return;
#define VISO_SYM(field, name)                                                   \
    *(void**)(&api.field) = dlsym(api.lib, name);                               \
    if (!api.field) {                                                           \
        api.lib = nullptr;                                                      \
        return;                                                                 \
    }
    VISO_SYM(GetUniqueId, "ncclGetUniqueId")
    VISO_SYM(CommInitRank, "ncclCommInitRank")
    VISO_SYM(AllReduce, "ncclAllReduce")
    VISO_SYM(AllGather, "ncclAllGather")
    VISO_SYM(Broadcast, "ncclBroadcast")
    VISO_SYM(GroupStart, "ncclGroupStart")
    VISO_SYM(GroupEnd, "ncclGroupEnd")
    VISO_SYM(CommDestroy, "ncclCommDestroy")
    VISO_SYM(GetErrorString, "ncclGetErrorString")
#undef VISO_SYM
}

NcclApi* nccl_api()
{ // loaded once per process, whichever thread asks first
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] { nccl_load(api); });
    return api.lib ? &api : nullptr;
}

// ------------------------------------------------------------------------------------------
// device arena: one allocation, bump-pointer sub-allocation, kept while it is large enough
// (the reference's "allocate the pools once in the constructor", bundle_adjuster.h:122-166)
// ------------------------------------------------------------------------------------------
struct Arena {
    char* base = nullptr;
    size_t cap = 0, used = 0;
    template <typename T> T* take(size_t count)
    {
        const size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
        T* p = reinterpret_cast<T*>(base ? base + used : nullptr);
        used += bytes;
        return p;
    }
};

struct DevBuf { // grow-only device buffer
    void* p = nullptr;
    size_t cap = 0;
};

// std::vector<double>-like host array whose storage is page-locked when a CUDA device is usable:
// the refined points then travel device -> this array in one DMA, without a bounce buffer
struct PinVec {
    double* p = nullptr;
    size_t n = 0, cap = 0;
    bool pinned = false;
    PinVec() = default;
    PinVec(const PinVec&) = delete;
    PinVec& operator=(const PinVec&) = delete;
    ~PinVec() { release(); }
    size_t size() const { return n; }
    double* data() { return p; }
    const double* data() const { return p; }
    void clear() { n = 0; }
    void release()
    {
        if (p) pinned ? (void)cudaFreeHost(p) : std::free(p);
        p = nullptr, cap = 0, n = 0;
    }
    bool reserve(size_t m, bool try_pinned)
    {
        if (m <= cap) return true;
        const size_t c = std::max(m, cap + cap / 2);
        double* q = nullptr;
        bool pin = false;
        if (try_pinned && c * sizeof(double) >= (1u << 16) && cudaHostAlloc(reinterpret_cast<void**>(&q), c * sizeof(double), cudaHostAllocDefault) == cudaSuccess) pin = true;
        else {
            if (try_pinned) (void)cudaGetLastError();
            q = static_cast<double*>(std::malloc(c * sizeof(double)));
        }
        if (!q) return false;
        if (n) std::memcpy(q, p, n * sizeof(double));
        const size_t keep = n;
        release();
        p = q, cap = c, n = keep, pinned = pin;
        return true;
    }
    bool append(const double* a, size_t k, bool try_pinned)
    {
        if (!reserve(n + k, try_pinned)) return false;
        std::memcpy(p + n, a, k * sizeof(double));
        n += k;
        return true;
    }
};

// dst <- src (n doubles) on up to 8 threads; returns the index of a non-finite entry or n
size_t copy_and_check(double* dst, const double* src, size_t n)
{
    auto work = [&](size_t a, size_t b, size_t* bad) {
        std::memcpy(dst + a, src + a, (b - a) * sizeof(double));
        for (size_t i = a; i < b; ++i)
            if (!std::isfinite(src[i])) {
                *bad = i;
                return;
            }
    };
    const size_t nt = n >= (1u << 20) ? std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;
    std::vector<size_t> bad(nt, n);
    if (nt == 1) work(0, n, &bad[0]);
    else {
        std::vector<std::thread> th;
        for (size_t t = 0; t < nt; ++t) th.emplace_back(work, n * t / nt, n * (t + 1) / nt, &bad[t]);
        for (auto& t : th) t.join();
    }
    return *std::min_element(bad.begin(), bad.end());
}

struct KernelAcc {
    int64_t launches = 0;
    double ms = 0.0;
};

struct PendingEvent {
    std::string name;
    cudaEvent_t a, b;
    int n;
};

double now_ms()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

} // namespace

constexpr int64_t kDevGeneralMinEdges = 200000; // general path, one problem: from here on the structure is built on the device (below, the host build is cheaper than a dozen launches and four syncs)

struct viso_ba {
    viso_ba_config cfg;
    std::string err;
    // ---- host staging (what AddPoint / AddPose / AddEdge append to) ----
    std::vector<double> poses;
    PinVec points;
    std::vector<int32_t> e_pose, e_point;
    std::vector<float> e_uv;
    std::vector<int8_t> fixed; // -1 default policy, 0 free, 1 fixed
    std::vector<int64_t> pose_off, point_off, edge_off; // batched: B+1 entries, else empty
    int64_t n_problems = 1;
    // ---- device ----
    bool cuda_ok = false;
    bool uploaded = false;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    Arena arena;
    DevProblem P;
    LargeSolve L;
    bool use_large = false;
    std::vector<int32_t> blk_ij_host; // (I, J) of every stored S block when the storage is compact, else empty
    bool window = false; // tile fast path (ba_window.cu) instead of the pair-list kernels
    int max_n = 0; // largest reduced system over the problems
    double *d_pose_init = nullptr, *d_pts_init = nullptr, *d_pose_out = nullptr, *d_pts_out = nullptr;
    int32_t* d_tile_map = nullptr; // window path: tiles of the problems still active (straggler trials)
    double* d_allreduce = nullptr; // S_blk | bs   (contiguous)
    size_t allreduce_count = 0;
    // pinned host mirrors
    void* pin = nullptr;
    size_t pin_cap = 0;
    std::vector<LmState> lm_host;
    float setup_ms = 0.f;
    // ---- bulk problems whose edges went straight to the device (window fast path) ----
    bool no_window = false;       // a duplicate edge was seen: this problem stays on the general path
    bool edges_on_device = false; // e_pose / e_point / e_uv are empty; the raw arrays live in `raw`
    bool edges_local = false;     // raw indices are local to the edge's problem
    int64_t n_edges_dev = 0;
    std::vector<int64_t> edge_off_all; // B+1 (also for B == 1)
    DevBuf raw, scratch, pairbuf, pts_stage;
    DevBuf gen_obs, gen_tabs, gen_work; // device-side build of the general path (one large problem)
    bool pts_staged = false;            // pts_stage holds the points exactly as `points` does (bulk set_problem)
    std::vector<int32_t> stage_c0, stage_p0; // source of an H2D copy that may still be in flight
    // ---- photometric edges (ba_photo.cu): the handle holds either reprojection or photometric edges ----
    bool photo = false;      // the staged edges are (point, source keyframe, pose vertex); e_uv carries the source index
    bool photo_next = false; // a photometric bulk problem is being staged (keeps it off the window path)
    std::vector<std::vector<uint8_t>> ph_img; // [pose] level-0 image of the pose's keyframe, rows * cols bytes (empty: none)
    std::vector<int32_t> ph_rows, ph_cols;
    DevBuf ph_img_dev, ph_frames_dev, ph_srcrt_dev, ph_sum_dev, ph_mid_dev;
    int host_threads_override = 0; // > 0: host threads of the general structure build (debug / tests)
    // ---- sharding ----
    int rank = 0, n_ranks = 1;
    void* comm = nullptr;
    std::vector<int32_t> shard_pt; // n_ranks+1 point boundaries
    // ---- accounting ----
    int64_t launches = 0;
    bool timing = false;
    std::map<std::string, KernelAcc> ktimes;
    std::vector<PendingEvent> pending;
    // ---- CUDA graph of one LM trial ----
    cudaGraphExec_t trial_graph = nullptr;
    int trial_graph_launches = 0;

    viso_ba() { std::memset(&P, 0, sizeof(P)), std::memset(&L, 0, sizeof(L)); }
};

namespace {

int32_t fail(viso_ba* h, int32_t code, const char* fmt, ...)
{
    if (h) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        h->err = buf;
        if (h->cfg.verbose) fprintf(stderr, "[viso_ba] %s\n", buf);
    }
    return code;
}

#define CU(call)                                                                                \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess)                                                                  \
            return fail(h, VISO_BA_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                __FILE__, __LINE__);                                                            \
    } while (0)

#define NC(call)                                                                                \
    do {                                                                                        \
        int e_ = (call);                                                                        \
        if (e_ != 0)                                                                            \
            return fail(h, VISO_BA_ERR_NCCL, "%s: %s", #call, nccl_api()->GetErrorString(e_));  \
    } while (0)

// ---- host threads of the general structure build ----
// The pair-list build of a large problem is tens of millions of simple operations (config 4: 55 M pairs); it runs on a few
// host threads, each on a contiguous range of points, and produces exactly the tables the one-thread build does (the
// lists are ordered by point, ranges are taken in order).  VISO_BA_HOST_THREADS overrides the count (1 = serial).
int host_threads(const viso_ba* h)
{
    if (h->host_threads_override > 0) return h->host_threads_override;
    if (const char* e = std::getenv("VISO_BA_HOST_THREADS")) {
        const int v = std::atoi(e);
        if (v > 0) return std::min(v, 64);
    }
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 1;
    const int per_rank = (int)hw / std::max(1, h->n_ranks); // the ranks of a sharded problem share one box
    return std::max(1, std::min(16, per_rank));
}

template <typename F> void parallel_ranges(int nt, int64_t lo, int64_t hi, F&& f)
{ // f(t, range_lo, range_hi) on nt contiguous ranges of [lo, hi)
    if (nt <= 1 || hi - lo < 4096) {
        f(0, lo, hi);
        return;
    }
    std::vector<std::thread> th;
    th.reserve((size_t)nt);
    for (int t = 0; t < nt; ++t) {
        const int64_t a = lo + (hi - lo) * t / nt, b = lo + (hi - lo) * (t + 1) / nt;
        th.emplace_back([&f, t, a, b] { f(t, a, b); });
    }
    for (std::thread& x : th) x.join();
}

void drop_graph(viso_ba* h)
{
    if (h->trial_graph) {
        cudaGraphExecDestroy(h->trial_graph);
        h->trial_graph = nullptr;
    }
}

int32_t ensure_cuda(viso_ba* h)
{
    if (h->cuda_ok) {
        CU(cudaSetDevice(h->cfg.device));
        return VISO_BA_OK;
    }
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0)
        return fail(h, VISO_BA_ERR_CUDA, "no CUDA device usable (%s); this backend has no CPU fallback",
            e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    if (h->cfg.device < 0 || h->cfg.device >= n)
        return fail(h, VISO_BA_ERR_CUDA, "device %d out of range (%d devices)", h->cfg.device, n);
    CU(cudaSetDevice(h->cfg.device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, h->cfg.device));
    if (prop.major != 10)
        return fail(h, VISO_BA_ERR_CUDA, "device %d is sm_%d%d; the kernels are built for sm_100a only",
            h->cfg.device, prop.major, prop.minor);
    CU(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    if (!h->stream) h->stream = h->own_stream;
    CU(cudaEventCreate(&h->ev0));
    CU(cudaEventCreate(&h->ev1));
    h->cuda_ok = true;
    return VISO_BA_OK;
}

// run one launcher, count its kernels and (optionally) time it with CUDA events
template <typename F> inline void run_k(viso_ba* h, const char* name, F&& f)
{
    if (h->timing) {
        PendingEvent pe;
        pe.name = name;
        cudaEventCreate(&pe.a);
        cudaEventCreate(&pe.b);
        cudaEventRecord(pe.a, h->stream);
        pe.n = f();
        cudaEventRecord(pe.b, h->stream);
        h->launches += pe.n;
        h->pending.push_back(pe);
    } else {
        h->launches += f();
    }
}

void resolve_timing(viso_ba* h)
{
    for (auto& pe : h->pending) {
        float ms = 0.f;
        cudaEventSynchronize(pe.b);
        cudaEventElapsedTime(&ms, pe.a, pe.b);
        KernelAcc& k = h->ktimes[pe.name];
        k.launches += pe.n;
        k.ms += ms;
        cudaEventDestroy(pe.a);
        cudaEventDestroy(pe.b);
    }
    h->pending.clear();
}

} // namespace

// ==========================================================================================
// C ABI: lifetime, staging
// ==========================================================================================
extern "C" int32_t viso_ba_version(void) { return VISO_BA_VERSION; }

extern "C" const char* viso_ba_status_string(int32_t s)
{
    switch (s) {
    case VISO_BA_OK: return "ok";
    case VISO_BA_ERR_CAPACITY: return "capacity exceeded";
    case VISO_BA_ERR_INDEX: return "index out of range";
    case VISO_BA_ERR_CUDA: return "CUDA error";
    case VISO_BA_ERR_NCCL: return "NCCL error";
    case VISO_BA_ERR_NONFINITE: return "non-finite input";
    case VISO_BA_ERR_ARG: return "bad argument";
    case VISO_BA_ERR_STATE: return "invalid state";
    case VISO_BA_ERR_UNSUPPORTED: return "unsupported";
    default: return "unknown status";
    }
}

extern "C" void viso_ba_default_config(viso_ba_config* c)
{
    if (!c) return;
    std::memset(c, 0, sizeof(*c));
    c->struct_size = (int32_t)sizeof(*c);
    c->device = 0;
    c->max_points = 1000; // bundle_adjuster.h:182-184
    c->max_poses = 1000;
    c->max_edges = 1000;
    c->fx = 517.3; // src/main.cpp:14-17
    c->fy = 516.5;
    c->cx = 325.1;
    c->cy = 249.7;
    c->huber_delta = 0.0; // the reference sets no robust kernel (bundle_adjuster.h:160)
    c->lm_tau = 1e-5;
    c->lm_lambda_init = 0.0;
    c->lm_max_trials = 10;
    c->verbose = 0;
    c->fix_first_pose = 1;
    c->use_cuda_graph = 1;
    c->path = 0;
    c->solver_segments = 0;
    c->host_structure_build = 0;
}

extern "C" int32_t viso_ba_create(const viso_ba_config* cfg, viso_ba** out)
{
    if (!out) return VISO_BA_ERR_ARG;
    *out = nullptr;
    viso_ba_config c;
    viso_ba_default_config(&c);
    if (cfg) {
        if (cfg->struct_size != (int32_t)sizeof(viso_ba_config)) return VISO_BA_ERR_ARG;
        c = *cfg;
    }
    if (c.max_points < 0 || c.max_poses < 0 || c.max_edges < 0) return VISO_BA_ERR_ARG;
    if (c.max_points > 0x7fffffff || c.max_poses > 0x7fffffff || c.max_edges > 0x7fffffff)
        return VISO_BA_ERR_UNSUPPORTED;
    if (c.lm_max_trials < 1) c.lm_max_trials = 1;
    viso_ba* h = new (std::nothrow) viso_ba();
    if (!h) return VISO_BA_ERR_CAPACITY;
    h->cfg = c;
    // CUDA is initialised lazily (first upload) so that staging-only use and the
    // "library loads and exports its symbols" check need no GPU.
    *out = h;
    return VISO_BA_OK;
}

extern "C" void viso_ba_destroy(viso_ba* h)
{
    if (!h) return;
    if (h->cuda_ok) {
        cudaSetDevice(h->cfg.device);
        cudaStreamSynchronize(h->stream);
        drop_graph(h);
        resolve_timing(h);
        if (h->comm && nccl_api()) nccl_api()->CommDestroy(h->comm);
        if (h->arena.base) cudaFree(h->arena.base);
        for (DevBuf* b : { &h->raw, &h->scratch, &h->pairbuf, &h->pts_stage, &h->gen_obs, &h->gen_tabs, &h->gen_work, &h->ph_img_dev, &h->ph_frames_dev, &h->ph_srcrt_dev, &h->ph_sum_dev, &h->ph_mid_dev })
            if (b->p) cudaFree(b->p);
        if (h->pin) cudaFreeHost(h->pin);
        if (h->ev0) cudaEventDestroy(h->ev0);
        if (h->ev1) cudaEventDestroy(h->ev1);
        if (h->own_stream) cudaStreamDestroy(h->own_stream);
    }
    delete h;
}

extern "C" const char* viso_ba_last_error(const viso_ba* h) { return h ? h->err.c_str() : "null handle"; }

extern "C" int32_t viso_ba_clear(viso_ba* h)
{
    if (!h) return VISO_BA_ERR_ARG;
    h->ph_img.clear(), h->ph_rows.clear(), h->ph_cols.clear(); // the keyframe images go with the poses
    h->photo = false;
    h->poses.clear(), h->points.clear(), h->e_pose.clear(), h->e_point.clear(), h->e_uv.clear();
    h->fixed.clear();
    h->pose_off.clear(), h->point_off.clear(), h->edge_off.clear(), h->edge_off_all.clear();
    h->edges_on_device = false, h->edges_local = false, h->n_edges_dev = 0, h->no_window = false;
    h->pts_staged = false;
    h->n_problems = 1;
    h->uploaded = false;
    drop_graph(h);
    return VISO_BA_OK;
}

extern "C" int64_t viso_ba_add_point(viso_ba* h, const double xyz[3])
{
    if (!h || !xyz) return -(int64_t)VISO_BA_ERR_ARG;
    if (!h->pose_off.empty()) return -(int64_t)fail(h, VISO_BA_ERR_STATE, "handle holds a batched problem; clear it first");
    if (h->edges_on_device) return -(int64_t)fail(h, VISO_BA_ERR_STATE, "handle holds a bulk problem; clear it first");
    const int64_t n = (int64_t)h->points.size() / 3;
    if (n >= h->cfg.max_points) return -(int64_t)fail(h, VISO_BA_ERR_CAPACITY, "point pool full (%lld)", (long long)h->cfg.max_points);
    for (int i = 0; i < 3; ++i)
        if (!std::isfinite(xyz[i])) return -(int64_t)fail(h, VISO_BA_ERR_NONFINITE, "point %lld is not finite", (long long)n);
    if (!h->points.append(xyz, 3, false)) return -(int64_t)fail(h, VISO_BA_ERR_CAPACITY, "out of host memory");
    h->pts_staged = false;
    h->uploaded = false;
    return n;
}

extern "C" int64_t viso_ba_add_pose(viso_ba* h, const double q[7])
{
    if (!h || !q) return -(int64_t)VISO_BA_ERR_ARG;
    if (!h->pose_off.empty()) return -(int64_t)fail(h, VISO_BA_ERR_STATE, "handle holds a batched problem; clear it first");
    if (h->edges_on_device) return -(int64_t)fail(h, VISO_BA_ERR_STATE, "handle holds a bulk problem; clear it first");
    const int64_t n = (int64_t)h->poses.size() / 7;
    if (n >= h->cfg.max_poses) return -(int64_t)fail(h, VISO_BA_ERR_CAPACITY, "pose pool full (%lld)", (long long)h->cfg.max_poses);
    for (int i = 0; i < 7; ++i)
        if (!std::isfinite(q[i])) return -(int64_t)fail(h, VISO_BA_ERR_NONFINITE, "pose %lld is not finite", (long long)n);
    const double nn = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    if (!(nn > 1e-12)) return -(int64_t)fail(h, VISO_BA_ERR_ARG, "pose %lld has a zero quaternion", (long long)n);
    h->poses.insert(h->poses.end(), q, q + 7);
    h->fixed.push_back(-1);
    h->uploaded = false;
    return n;
}

extern "C" int32_t viso_ba_add_edge(viso_ba* h, int64_t point, int64_t pose, float u, float v)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (!h->pose_off.empty()) return fail(h, VISO_BA_ERR_STATE, "handle holds a batched problem; clear it first");
    if (h->edges_on_device) return fail(h, VISO_BA_ERR_STATE, "handle holds a bulk problem; clear it first");
    if (h->photo) return fail(h, VISO_BA_ERR_UNSUPPORTED, "handle holds photometric edges: reprojection and photometric edges do not mix");
    if ((int64_t)h->e_pose.size() >= h->cfg.max_edges) return fail(h, VISO_BA_ERR_CAPACITY, "edge pool full (%lld)", (long long)h->cfg.max_edges);
    if (point < 0 || point >= (int64_t)h->points.size() / 3) return fail(h, VISO_BA_ERR_INDEX, "edge refers to point %lld", (long long)point);
    if (pose < 0 || pose >= (int64_t)h->poses.size() / 7) return fail(h, VISO_BA_ERR_INDEX, "edge refers to pose %lld", (long long)pose);
    if (!std::isfinite(u) || !std::isfinite(v)) return fail(h, VISO_BA_ERR_NONFINITE, "edge measurement is not finite");
    h->e_pose.push_back((int32_t)pose);
    h->e_point.push_back((int32_t)point);
    h->e_uv.push_back(u);
    h->e_uv.push_back(v);
    h->uploaded = false;
    return VISO_BA_OK;
}

namespace {
int32_t check_finite(viso_ba* h, const double* a, size_t n, const char* what)
{
    for (size_t i = 0; i < n; ++i)
        if (!std::isfinite(a[i])) return fail(h, VISO_BA_ERR_NONFINITE, "%s[%zu] is not finite", what, i);
    return VISO_BA_OK;
}
} // namespace

// ------------------------------------------------------------------------------------------
// bulk problems of the window fast path: the edge arrays go from the caller's memory straight
// to the device (no host copy, no host pass over the edges); validation is a kernel
// ------------------------------------------------------------------------------------------
namespace {

int32_t pin_reserve(viso_ba* h, size_t bytes);

int32_t devbuf_reserve(viso_ba* h, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return VISO_BA_OK;
    if (b.p) {
        CU(cudaStreamSynchronize(h->stream));
        CU(cudaFree(b.p));
        b.p = nullptr, b.cap = 0;
    }
    const size_t want = bytes + bytes / 8;
    CU(cudaMalloc(&b.p, want));
    b.cap = want;
    return VISO_BA_OK;
}

struct Carver { // bump allocation inside one buffer; base == nullptr: dry run for the size
    char* base;
    size_t used = 0;
    explicit Carver(void* b) : base(static_cast<char*>(b)) {}
    template <typename T> T* take(size_t count)
    {
        T* p = reinterpret_cast<T*>(base ? base + used : nullptr);
        used += (count * sizeof(T) + 255) & ~(size_t)255;
        return p;
    }
};

struct RawEdges {
    int32_t *e_pose, *e_point;
    float2* e_uv;
    int64_t* edge_off;
    int32_t *prob_cam0, *prob_pt0;
};

size_t carve_raw(void* base, RawEdges& r, int64_t Ne, int64_t B)
{
    Carver c(base);
    r.e_pose = c.take<int32_t>((size_t)Ne);
    r.e_point = c.take<int32_t>((size_t)Ne);
    r.e_uv = c.take<float2>((size_t)Ne);
    r.edge_off = c.take<int64_t>((size_t)B + 1);
    r.prob_cam0 = c.take<int32_t>((size_t)B + 1);
    r.prob_pt0 = c.take<int32_t>((size_t)B + 1);
    return c.used;
}

size_t carve_scratch(void* base, SetupWork& w, int64_t Ne, int64_t B, int64_t tiles_max)
{
    Carver c(base);
    w.keys = c.take<unsigned long long>((size_t)Ne);
    w.keys_sorted = c.take<unsigned long long>((size_t)Ne);
    w.vals = c.take<unsigned int>((size_t)Ne);
    w.vals_sorted = c.take<unsigned int>((size_t)Ne);
    w.temp_bytes = std::max(setup_sort_temp_bytes(Ne), setup_sort_temp_bytes(tiles_max + 1));
    w.temp = c.take<char>(w.temp_bytes);
    w.prob_ntiles = c.take<int32_t>((size_t)B);
    w.prob_blkptr_off = c.take<int64_t>((size_t)B + 1);
    w.tile_pairs = c.take<int64_t>((size_t)tiles_max + 1);
    w.tile_list_off = c.take<int64_t>((size_t)tiles_max + 1);
    w.totals = c.take<int64_t>(4);
    w.err = c.take<int>(4);
    return c.used;
}

// a tile is closed when the next point would break a cap, so every tile but a problem's last
// holds kWinPtCap points or more than kWinObsCap - kWinMaxCams observations
int64_t tiles_upper_bound(int64_t B, const int64_t* point_off, const int64_t* edge_off)
{
    int64_t n = 0;
    for (int64_t b = 0; b < B; ++b)
        n += (point_off[b + 1] - point_off[b]) / kWinPtCap + (edge_off[b + 1] - edge_off[b]) / (kWinObsCap - kWinMaxCams + 1) + 2;
    return n;
}

// can this problem take the window path whatever the gauge turns out to be?
bool window_shape_ok(const viso_ba* h, int64_t B, const int64_t* pose_off, int64_t Ne, int max_free)
{
    if (h->cfg.path == 2 || h->n_ranks != 1 || h->no_window || h->photo || h->photo_next) return false;
    for (int64_t b = 0; b < B; ++b)
        if (pose_off[b + 1] - pose_off[b] > std::min(max_free, kWinMaxCams)) return false;
    return true;
}

// H2D of the caller's edge arrays + validation kernel.  `local`: indices are local to the problem.
int32_t stage_finish(viso_ba* h);

// `defer`: return with the copies and the validation kernel still in flight; the caller overlaps its
// host work with them and calls stage_finish before it returns to ITS caller
int32_t stage_edges_on_device(viso_ba* h, int64_t B, const int64_t* pose_off, const int64_t* point_off, const int64_t* edge_off,
    const int32_t* edge_pose, const int32_t* edge_point, const float* edge_uv, bool local, SetupIn* in_out, SetupWork* w_out,
    bool defer = false)
{
    const int64_t Ne = edge_off[B];
    int32_t rc;
    RawEdges r;
    if ((rc = devbuf_reserve(h, h->raw, carve_raw(nullptr, r, Ne, B)))) return rc;
    carve_raw(h->raw.p, r, Ne, B);
    SetupWork w;
    const int64_t tiles_max = tiles_upper_bound(B, point_off, edge_off);
    if ((rc = devbuf_reserve(h, h->scratch, carve_scratch(nullptr, w, Ne, B, tiles_max)))) return rc;
    carve_scratch(h->scratch.p, w, Ne, B, tiles_max);
    cudaStream_t st = h->stream;
    if (Ne) {
        CU(cudaMemcpyAsync(r.e_pose, edge_pose, sizeof(int32_t) * (size_t)Ne, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(r.e_point, edge_point, sizeof(int32_t) * (size_t)Ne, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(r.e_uv, edge_uv, sizeof(float) * 2 * (size_t)Ne, cudaMemcpyHostToDevice, st));
    }
    std::vector<int32_t>&c0 = h->stage_c0, &p0 = h->stage_p0;
    c0.resize((size_t)B + 1), p0.resize((size_t)B + 1);
    for (int64_t b = 0; b <= B; ++b) c0[(size_t)b] = (int32_t)pose_off[b], p0[(size_t)b] = (int32_t)point_off[b];
    CU(cudaMemcpyAsync(r.edge_off, edge_off, sizeof(int64_t) * (size_t)(B + 1), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(r.prob_cam0, c0.data(), sizeof(int32_t) * (size_t)(B + 1), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(r.prob_pt0, p0.data(), sizeof(int32_t) * (size_t)(B + 1), cudaMemcpyHostToDevice, st));
    SetupIn in;
    std::memset(&in, 0, sizeof(in));
    in.n_edges = Ne, in.n_points = (int32_t)point_off[B], in.n_cams = (int32_t)pose_off[B], in.n_problems = (int32_t)B;
    in.local_indices = local ? 1 : 0;
    in.e_pose = r.e_pose, in.e_point = r.e_point, in.e_uv = r.e_uv, in.edge_off = r.edge_off;
    in.prob_cam0 = r.prob_cam0, in.prob_pt0 = r.prob_pt0;
    run_k(h, "setup", [&] { return setup_stage0(in, w, st); });
    if ((rc = pin_reserve(h, 64))) return rc;
    CU(cudaMemcpyAsync(h->pin, w.err, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (in_out) *in_out = in;
    if (w_out) *w_out = w;
    return defer ? VISO_BA_OK : stage_finish(h);
}

int32_t stage_finish(viso_ba* h)
{
    CU(cudaStreamSynchronize(h->stream)); // the caller's arrays are free again
    CU(cudaGetLastError());
    const int* err = static_cast<const int*>(h->pin);
    if (err[0]) {
        long long e = 0;
        std::memcpy(&e, err + 2, sizeof(e));
        return fail(h, err[0], err[0] == VISO_BA_ERR_INDEX ? "edge %lld refers to a point or pose outside its problem" : "edge_uv of edge %lld is not finite", e);
    }
    return VISO_BA_OK;
}

} // namespace

extern "C" int32_t viso_ba_set_problem_batched(viso_ba* h, int64_t B, const int64_t* pose_off,
    const int64_t* point_off, const int64_t* edge_off, const double* poses7, const double* points3,
    const int32_t* edge_pose, const int32_t* edge_point, const float* edge_uv)
{
    if (!h || B < 1 || !pose_off || !point_off || !edge_off) return VISO_BA_ERR_ARG;
    const int64_t Nc = pose_off[B], Np = point_off[B], Ne = edge_off[B];
    if (pose_off[0] != 0 || point_off[0] != 0 || edge_off[0] != 0) return fail(h, VISO_BA_ERR_ARG, "offset arrays must start at 0");
    if (Nc < 0 || Np < 0 || Ne < 0) return fail(h, VISO_BA_ERR_ARG, "negative sizes");
    if ((Nc && !poses7) || (Np && !points3) || (Ne && (!edge_pose || !edge_point || !edge_uv))) return fail(h, VISO_BA_ERR_ARG, "null array");
    if (Nc > h->cfg.max_poses || Np > h->cfg.max_points || Ne > h->cfg.max_edges)
        return fail(h, VISO_BA_ERR_CAPACITY, "problem (%lld poses, %lld points, %lld edges) exceeds the pools (%lld, %lld, %lld)",
            (long long)Nc, (long long)Np, (long long)Ne, (long long)h->cfg.max_poses, (long long)h->cfg.max_points, (long long)h->cfg.max_edges);
    for (int64_t b = 0; b < B; ++b)
        if (pose_off[b + 1] < pose_off[b] || point_off[b + 1] < point_off[b] || edge_off[b + 1] < edge_off[b])
            return fail(h, VISO_BA_ERR_ARG, "offsets of problem %lld decrease", (long long)b);
    int32_t rc;
    if ((rc = check_finite(h, poses7, (size_t)Nc * 7, "poses"))) return rc;
    for (int64_t c = 0; c < Nc; ++c) {
        const double* q = poses7 + 7 * c;
        if (!(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3] > 1e-12)) return fail(h, VISO_BA_ERR_ARG, "pose %lld has a zero quaternion", (long long)c);
    }
    // window fast path for sure (<= 27 keyframes per problem: eligible whatever gets fixed later) and a
    // usable device: the edges never touch host memory of this library
    h->no_window = false; // a duplicate edge of the PREVIOUS problem says nothing about this one
    // ... and one LARGE problem whose structure the device builds (build_general_device) is staged the same way
    const bool big_general = B == 1 && Ne >= kDevGeneralMinEdges && h->cfg.host_structure_build == 0 && !h->photo_next;
    const bool direct = Nc > 0 && Np > 0 && (window_shape_ok(h, B, pose_off, Ne, kSmallMaxN / 6) || big_general) && ensure_cuda(h) == VISO_BA_OK;
    const int nt_host = host_threads(h);
    if (!direct) { // validation of the edges on a few host threads: the first offending edge (lowest index) is reported
        for (int64_t b = 0; b < B; ++b) {
            const int64_t nc = pose_off[b + 1] - pose_off[b], np = point_off[b + 1] - point_off[b];
            std::vector<int64_t> bad((size_t)nt_host, -1);
            parallel_ranges(nt_host, edge_off[b], edge_off[b + 1], [&](int t, int64_t lo, int64_t hi) {
                for (int64_t e = lo; e < hi; ++e)
                    if (edge_pose[e] < 0 || edge_pose[e] >= nc || edge_point[e] < 0 || edge_point[e] >= np || !std::isfinite(edge_uv[2 * e])
                        || !std::isfinite(edge_uv[2 * e + 1])) {
                        bad[(size_t)t] = e;
                        return;
                    }
            });
            for (int64_t e : bad) {
                if (e < 0) continue;
                if (edge_pose[e] < 0 || edge_pose[e] >= nc) return fail(h, VISO_BA_ERR_INDEX, "edge %lld refers to pose %d", (long long)e, edge_pose[e]);
                if (edge_point[e] < 0 || edge_point[e] >= np) return fail(h, VISO_BA_ERR_INDEX, "edge %lld refers to point %d", (long long)e, edge_point[e]);
                return fail(h, VISO_BA_ERR_NONFINITE, "edge_uv[%lld] is not finite", (long long)(std::isfinite(edge_uv[2 * e]) ? 2 * e + 1 : 2 * e));
            }
        }
    }
    viso_ba_clear(h);
    if (direct) {
        // edges and points leave for the device first (asynchronously when the caller's arrays are
        // page-locked); the host-side copy and checks of the points below run while the DMA does
        if ((rc = stage_edges_on_device(h, B, pose_off, point_off, edge_off, edge_pose, edge_point, edge_uv, true, nullptr, nullptr, true))) return rc;
        if ((rc = devbuf_reserve(h, h->pts_stage, sizeof(double) * 3 * (size_t)Np))) {
            cudaStreamSynchronize(h->stream);
            return rc;
        }
        if (cudaMemcpyAsync(h->pts_stage.p, points3, sizeof(double) * 3 * (size_t)Np, cudaMemcpyHostToDevice, h->stream) != cudaSuccess) {
            cudaStreamSynchronize(h->stream);
            return fail(h, VISO_BA_ERR_CUDA, "H2D of the points failed");
        }
    }
    { // host copy of the points (+ finiteness) — what GetPoint returns before Optimize
        size_t bad = 3 * (size_t)Np;
        const bool ok = h->points.reserve(3 * (size_t)Np, h->cuda_ok);
        if (ok) bad = copy_and_check(h->points.data(), points3, 3 * (size_t)Np), h->points.n = 3 * (size_t)Np;
        if (direct) {
            rc = stage_finish(h);
            if (!ok || bad < 3 * (size_t)Np || rc) h->points.clear();
            if (rc) return rc;
        }
        if (!ok) return fail(h, VISO_BA_ERR_CAPACITY, "out of host memory");
        if (bad < 3 * (size_t)Np) {
            h->points.clear();
            return fail(h, VISO_BA_ERR_NONFINITE, "points[%zu] is not finite", bad);
        }
    }
    if (direct) {
        h->edges_on_device = true, h->edges_local = true, h->n_edges_dev = Ne;
        h->pts_staged = true;
    } else {
        h->e_uv.resize(2 * (size_t)Ne);
        h->e_pose.resize((size_t)Ne);
        h->e_point.resize((size_t)Ne);
        for (int64_t b = 0; b < B; ++b) // store GLOBAL indices
            parallel_ranges(B == 1 ? nt_host : 1, edge_off[b], edge_off[b + 1], [&](int, int64_t lo, int64_t hi) {
                for (int64_t e = lo; e < hi; ++e) {
                    h->e_pose[(size_t)e] = (int32_t)(edge_pose[e] + pose_off[b]);
                    h->e_point[(size_t)e] = (int32_t)(edge_point[e] + point_off[b]);
                    h->e_uv[2 * (size_t)e] = edge_uv[2 * e], h->e_uv[2 * (size_t)e + 1] = edge_uv[2 * e + 1];
                }
            });
    }
    h->poses.assign(poses7, poses7 + 7 * Nc);
    h->fixed.assign((size_t)Nc, -1);
    h->n_problems = B;
    h->edge_off_all.assign(edge_off, edge_off + B + 1);
    if (B > 1) {
        h->pose_off.assign(pose_off, pose_off + B + 1);
        h->point_off.assign(point_off, point_off + B + 1);
        h->edge_off.assign(edge_off, edge_off + B + 1);
    }
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_problem(viso_ba* h, int64_t n_poses, const double* poses7, int64_t n_points,
    const double* points3, int64_t n_edges, const int32_t* edge_pose, const int32_t* edge_point, const float* edge_uv)
{
    const int64_t po[2] = { 0, n_poses }, pt[2] = { 0, n_points }, eo[2] = { 0, n_edges };
    return viso_ba_set_problem_batched(h, 1, po, pt, eo, poses7, points3, edge_pose, edge_point, edge_uv);
}

// ------------------------------------------------------------------------------------------
// photometric edges: EdgeDirectProjection(srcFrame, targetFrame) between VertexPoint and VertexPose
// (reference bundle_adjuster.h:58-106, :154-160).  The keyframe of a pose vertex is its level-0 image; an edge's
// source keyframe is another pose's keyframe and projects with the pose that keyframe had at upload (:79).
// ------------------------------------------------------------------------------------------
namespace {
inline float src_bits(int32_t src)
{
    float f;
    std::memcpy(&f, &src, sizeof(f));
    return f;
}
} // namespace

extern "C" int32_t viso_ba_set_pose_image(viso_ba* h, int64_t pose, const uint8_t* data, int32_t rows, int32_t cols, int32_t stride)
{
    if (!h || !data) return VISO_BA_ERR_ARG;
    if (pose < 0 || pose >= (int64_t)h->poses.size() / 7) return fail(h, VISO_BA_ERR_INDEX, "no pose %lld", (long long)pose);
    if (rows < 8 || cols < 8 || stride < cols || rows > 16384 || cols > 16384) return fail(h, VISO_BA_ERR_ARG, "bad image size %d x %d (stride %d)", cols, rows, stride);
    const size_t n = h->poses.size() / 7;
    if (h->ph_img.size() < n) h->ph_img.resize(n), h->ph_rows.resize(n, 0), h->ph_cols.resize(n, 0);
    std::vector<uint8_t>& im = h->ph_img[(size_t)pose];
    im.resize((size_t)rows * (size_t)cols);
    for (int32_t r = 0; r < rows; ++r) std::memcpy(im.data() + (size_t)r * cols, data + (size_t)r * stride, (size_t)cols);
    h->ph_rows[(size_t)pose] = rows, h->ph_cols[(size_t)pose] = cols;
    h->uploaded = false;
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_add_photo_edge(viso_ba* h, int64_t point, int64_t src_pose, int64_t dst_pose)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (!h->pose_off.empty()) return fail(h, VISO_BA_ERR_STATE, "handle holds a batched problem; clear it first");
    if (h->edges_on_device) return fail(h, VISO_BA_ERR_STATE, "handle holds a bulk problem; clear it first");
    if (!h->photo && !h->e_pose.empty()) return fail(h, VISO_BA_ERR_UNSUPPORTED, "handle holds reprojection edges: reprojection and photometric edges do not mix");
    if ((int64_t)h->e_pose.size() >= h->cfg.max_edges) return fail(h, VISO_BA_ERR_CAPACITY, "edge pool full (%lld)", (long long)h->cfg.max_edges);
    const int64_t nc = (int64_t)h->poses.size() / 7;
    if (point < 0 || point >= (int64_t)h->points.size() / 3) return fail(h, VISO_BA_ERR_INDEX, "edge refers to point %lld", (long long)point);
    if (dst_pose < 0 || dst_pose >= nc) return fail(h, VISO_BA_ERR_INDEX, "edge refers to pose %lld", (long long)dst_pose);
    if (src_pose < 0 || src_pose >= nc) return fail(h, VISO_BA_ERR_INDEX, "edge refers to source keyframe %lld", (long long)src_pose);
    h->e_pose.push_back((int32_t)dst_pose);
    h->e_point.push_back((int32_t)point);
    h->e_uv.push_back(src_bits((int32_t)src_pose));
    h->e_uv.push_back(0.f);
    h->photo = true;
    h->uploaded = false;
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_photo_problem(viso_ba* h, int64_t n_poses, const double* poses7, int64_t n_points, const double* points3,
    int64_t n_edges, const int32_t* edge_point, const int32_t* edge_src, const int32_t* edge_dst)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (n_edges < 0 || (n_edges && (!edge_point || !edge_src || !edge_dst))) return fail(h, VISO_BA_ERR_ARG, "null array");
    std::vector<float> uv(2 * (size_t)n_edges, 0.f);
    for (int64_t e = 0; e < n_edges; ++e) {
        if (edge_src[e] < 0 || edge_src[e] >= n_poses) return fail(h, VISO_BA_ERR_INDEX, "edge %lld refers to source keyframe %d", (long long)e, edge_src[e]);
        uv[2 * (size_t)e] = src_bits(edge_src[e]);
    }
    h->photo_next = true; // keeps the edges off the window path
    const int32_t rc = viso_ba_set_problem(h, n_poses, poses7, n_points, points3, n_edges, edge_dst, edge_point, uv.data());
    h->photo_next = false;
    if (rc) return rc;
    h->photo = n_edges > 0;
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_pose_fixed(viso_ba* h, int64_t pose, int32_t fixed)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (pose < 0 || pose >= (int64_t)h->fixed.size()) return fail(h, VISO_BA_ERR_INDEX, "no pose %lld", (long long)pose);
    h->fixed[(size_t)pose] = fixed ? 1 : 0;
    h->uploaded = false;
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_intrinsics(viso_ba* h, double fx, double fy, double cx, double cy)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (!(fx > 0.0) || !(fy > 0.0) || !std::isfinite(fx) || !std::isfinite(fy) || !std::isfinite(cx) || !std::isfinite(cy))
        return fail(h, VISO_BA_ERR_ARG, "bad intrinsics");
    h->cfg.fx = fx, h->cfg.fy = fy, h->cfg.cx = cx, h->cfg.cy = cy;
    h->P.fx = fx, h->P.fy = fy, h->P.cx = cx, h->P.cy = cy;
    drop_graph(h);
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_huber_delta(viso_ba* h, double delta)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (std::isnan(delta)) return fail(h, VISO_BA_ERR_ARG, "bad huber delta");
    h->cfg.huber_delta = delta;
    h->P.huber_delta = delta;
    drop_graph(h);
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_stream(viso_ba* h, void* s)
{
    if (!h) return VISO_BA_ERR_ARG;
    int32_t rc = ensure_cuda(h);
    if (rc) return rc;
    h->stream = s ? (cudaStream_t)s : h->own_stream;
    drop_graph(h);
    return VISO_BA_OK;
}

extern "C" int64_t viso_ba_num_points(const viso_ba* h) { return h ? (int64_t)h->points.size() / 3 : -1; }
extern "C" int64_t viso_ba_num_poses(const viso_ba* h) { return h ? (int64_t)h->poses.size() / 7 : -1; }
extern "C" int64_t viso_ba_num_edges(const viso_ba* h) { return h ? (h->edges_on_device ? h->n_edges_dev : (int64_t)h->e_pose.size()) : -1; }
extern "C" int64_t viso_ba_num_problems(const viso_ba* h) { return h ? h->n_problems : -1; }
extern "C" int64_t viso_ba_kernel_launches(const viso_ba* h) { return h ? h->launches : -1; }

// ==========================================================================================
// structure build: what g2o does in addVertex/addEdge/initializeOptimization
// (bundle_adjuster.h:122-166) — sort the edges by (point, pose), index the block structure
// of the reduced camera system and list, for every S block, the observation pairs feeding it.
// ==========================================================================================
namespace {

struct Built {
    std::vector<int32_t> pt_ptr, pt_prob, obs_cam, cam_free, cam_prob, prob_cam0, prob_tri0, prob_nfree;
    std::vector<float> obs_uv;
    std::vector<int32_t> pairs;      // 2 per pair
    std::vector<int32_t> items;      // 4 per item
    std::vector<int32_t> diag_items, block_item_ptr, block_cam;
    std::vector<int32_t> diag_blk;   // [Nc] S block of the pose's diagonal, -1 for fixed poses
    std::vector<int32_t> blk_ij;     // 2 per S block: (I, J) free indices (compact storage, large solver only)
    SolvePlan plan;                  // segments / separators / elimination steps of the large solver (large solver only)
    bool compact = false;            // S blocks stored for co-visible pose pairs only
    bool dev_general = false;        // general path built on the device: obs arrays and pairs live in the handle's DevBufs
    std::vector<int32_t> chunks;     // 4 per chunk
    std::vector<int32_t> prob_chunk_ptr;
    // window fast path: the structure is built on the device; the host only bounds its size
    bool window = false;
    int64_t tiles_max = 0, part_max = 0, blkptr_max = 0;
    std::vector<int64_t> pt0; // first point of every problem, B+1
    int64_t n_pairs = 0;
    int32_t n_blocks = 0, max_n = 0;
};

// rank r owns points [out[r], out[r+1]): the first point whose CSR offset reaches r/n of the edges
template <typename T> void shard_bounds(const T* pt_ptr, int64_t Np, int64_t Ne, int n_ranks, int64_t* out)
{
    out[0] = 0;
    out[n_ranks] = Np;
    for (int r = 1; r < n_ranks; ++r) {
        const int64_t target = Ne * r / n_ranks;
        const T* it = std::lower_bound(pt_ptr, pt_ptr + Np + 1, (T)target);
        out[r] = std::min<int64_t>(Np, it - pt_ptr);
    }
}

// problems, gauge, free indices, S block offsets: O(poses + problems); decides the path
int32_t build_tables(viso_ba* h, Built& b)
{
    const int64_t Nc = (int64_t)h->poses.size() / 7, Np = (int64_t)h->points.size() / 3;
    const int64_t Ne = h->edges_on_device ? h->n_edges_dev : (int64_t)h->e_pose.size();
    const int64_t B = h->n_problems;
    // ---- problems, gauge, free indices, block offsets ----
    b.cam_prob.assign((size_t)Nc, 0);
    b.prob_cam0.assign((size_t)B + 1, 0);
    std::vector<int64_t>& pt0 = b.pt0;
    pt0.assign((size_t)B + 1, 0);
    if (B > 1) {
        for (int64_t p = 0; p <= B; ++p) b.prob_cam0[(size_t)p] = (int32_t)h->pose_off[(size_t)p], pt0[(size_t)p] = h->point_off[(size_t)p];
        for (int64_t p = 0; p < B; ++p) {
            for (int64_t c = h->pose_off[(size_t)p]; c < h->pose_off[(size_t)p + 1]; ++c) b.cam_prob[(size_t)c] = (int32_t)p;
        }
    } else {
        b.prob_cam0[1] = (int32_t)Nc;
        pt0[1] = Np;
    }
    b.cam_free.assign((size_t)Nc, -1);
    b.prob_nfree.assign((size_t)B, 0);
    b.prob_tri0.assign((size_t)B + 1, 0);
    int64_t tri = 0;
    b.max_n = 0;
    for (int64_t p = 0; p < B; ++p) {
        int32_t nf = 0;
        for (int32_t c = b.prob_cam0[(size_t)p]; c < b.prob_cam0[(size_t)p + 1]; ++c) {
            int8_t f = h->fixed[(size_t)c];
            if (f < 0) f = (h->cfg.fix_first_pose && c == b.prob_cam0[(size_t)p]) ? 1 : 0;
            b.cam_free[(size_t)c] = f ? -1 : nf++;
        }
        b.prob_nfree[(size_t)p] = nf;
        b.prob_tri0[(size_t)p] = (int32_t)tri;
        tri += (int64_t)nf * (nf + 1) / 2;
        b.max_n = std::max(b.max_n, 6 * nf);
        if (tri > 0x7fffffff / 36) return fail(h, VISO_BA_ERR_UNSUPPORTED, "reduced camera system too large (%lld blocks)", (long long)tri);
    }
    b.prob_tri0[(size_t)B] = (int32_t)tri;
    b.n_blocks = (int32_t)tri;

    b.block_cam.assign((size_t)b.n_blocks, -1);
    b.diag_blk.assign((size_t)Nc, -1);
    for (int64_t p = 0; p < B; ++p)
        for (int32_t c = b.prob_cam0[(size_t)p]; c < b.prob_cam0[(size_t)p + 1]; ++c) {
            const int32_t f = b.cam_free[(size_t)c];
            if (f >= 0) {
                b.diag_blk[(size_t)c] = b.prob_tri0[(size_t)p] + f * (f + 1) / 2 + f;
                b.block_cam[(size_t)b.diag_blk[(size_t)c]] = c;
            }
        }
    // ---- window fast path? (every problem <= 32 keyframes and <= 27 free poses, one GPU) ----
    std::vector<int64_t> cam_off((size_t)B + 1);
    for (int64_t p = 0; p <= B; ++p) cam_off[(size_t)p] = b.prob_cam0[(size_t)p];
    b.window = Np > 0 && b.max_n <= kSmallMaxN && window_shape_ok(h, B, cam_off.data(), Ne, kWinMaxCams);
    // device-resident edges off the window path: one large problem whose structure build_general_device makes (upload
    // brings them back to the host if it has to fall back to the host build)
    if (b.window) {
        if (B == 1) h->edge_off_all.assign({ 0, Ne }); // AddEdge may have grown a single problem since
        b.tiles_max = tiles_upper_bound(B, pt0.data(), h->edge_off_all.data());
        for (int64_t p = 0; p < B; ++p) {
            const int64_t nf = b.prob_nfree[(size_t)p], nblk = nf * (nf + 1) / 2;
            const int64_t tp = (pt0[(size_t)p + 1] - pt0[(size_t)p]) / kWinPtCap +
                (h->edge_off_all[(size_t)p + 1] - h->edge_off_all[(size_t)p]) / (kWinObsCap - kWinMaxCams + 1) + 2;
            b.part_max += tp * nblk * kItemPart;
            b.blkptr_max += tp * win_blkptr_entries(nblk);
        }
    }
    return VISO_BA_OK;
}

// general path: sort the edges, shard the points, list the observation pairs of every S block (host)
// compact storage of S (one large problem): blocks of co-visible pose pairs only.  `present` is indexed by the dense
// triangular index I (I + 1) / 2 + J of FREE poses (the diagonal is added here); fills remap (dense -> stored block or -1),
// blk_ij, the solve plan, block_cam and diag_blk, and shrinks n_blocks.
void compact_structure(viso_ba* h, Built& b, int32_t nf, std::vector<uint8_t>& present, std::vector<int32_t>& remap)
{
    const int64_t Nc = (int64_t)h->poses.size() / 7;
    for (int32_t f = 0; f < nf; ++f) present[(size_t)f * (f + 1) / 2 + f] = 1;
    remap.assign((size_t)b.n_blocks, -1);
    int32_t k = 0;
    for (int32_t I = 0; I < nf; ++I)
        for (int32_t J = 0; J <= I; ++J)
            if (present[(size_t)I * (I + 1) / 2 + J]) {
                remap[(size_t)I * (I + 1) / 2 + J] = k++;
                b.blk_ij.push_back(I);
                b.blk_ij.push_back(J);
            }
    // how the solver cuts the pose chain (ba_solve_plan.h): config.solver_segments 0 = auto (up to 8 interior segments
    // when the co-visibility span is small against the chain), 1 = never, k >= 2 = up to k whenever the structure allows
    {
        const int seg = h->cfg.solver_segments;
        b.plan = make_solve_plan(nf, b.blk_ij, large_solve_cta_count(), seg == 0 ? 8 : seg, seg >= 2);
    }
    b.n_blocks = k;
    b.prob_tri0[1] = k;
    b.block_cam.assign((size_t)k, -1);
    for (int64_t c = 0; c < Nc; ++c) {
        const int32_t f = b.cam_free[(size_t)c];
        if (f >= 0) {
            b.diag_blk[(size_t)c] = remap[(size_t)f * (f + 1) / 2 + f];
            b.block_cam[(size_t)b.diag_blk[(size_t)c]] = (int32_t)c;
        }
    }
}

// work items (slices of one block's pairs, one warp each) from the block offsets of the pair list, and the chunks of points
// of the deterministic per-problem sums
void general_items_and_chunks(viso_ba* h, Built& b, const std::vector<int64_t>& blk_ptr, int32_t pt_begin, int32_t pt_end)
{
    const int64_t B = h->n_problems;
    const std::vector<int64_t>& pt0 = b.pt0;
    int64_t item_pairs = (b.n_pairs / 16384 + 31) / 32 * 32;
    item_pairs = std::max<int64_t>(64, std::min<int64_t>(kItemMaxPairs, item_pairs));
    b.block_item_ptr.assign((size_t)b.n_blocks + 1, 0);
    {
        int32_t prob = 0;
        for (int32_t k = 0; k < b.n_blocks; ++k) {
            while (k >= b.prob_tri0[(size_t)prob + 1]) ++prob;
            const int32_t is_diag = b.block_cam[(size_t)k] >= 0 ? 1 : 0;
            for (int64_t q = blk_ptr[(size_t)k]; q < blk_ptr[(size_t)k + 1]; q += item_pairs) {
                const int64_t q1 = std::min(q + item_pairs, blk_ptr[(size_t)k + 1]);
                if (is_diag) b.diag_items.push_back((int32_t)(b.items.size() / 4));
                b.items.push_back((int32_t)q);
                b.items.push_back((int32_t)q1);
                b.items.push_back(k);
                b.items.push_back((prob << 1) | is_diag);
            }
            b.block_item_ptr[(size_t)k + 1] = (int32_t)(b.items.size() / 4);
        }
    }
    b.prob_chunk_ptr.assign((size_t)B + 1, 0);
    for (int64_t p = 0; p < B; ++p) {
        const int32_t lo = std::max<int64_t>(pt0[(size_t)p], pt_begin), hi = (int32_t)std::min<int64_t>(pt0[(size_t)p + 1], pt_end);
        for (int32_t q = lo; q < hi; q += kChunkPoints) {
            b.chunks.push_back((int32_t)p);
            b.chunks.push_back(q);
            b.chunks.push_back(std::min(q + kChunkPoints, hi));
            b.chunks.push_back(0);
        }
        b.prob_chunk_ptr[(size_t)p + 1] = (int32_t)(b.chunks.size() / 4);
    }
}

int32_t build_general(viso_ba* h, Built& b)
{
    const int64_t Nc = (int64_t)h->poses.size() / 7, Np = (int64_t)h->points.size() / 3;
    const int64_t Ne = (int64_t)h->e_pose.size();
    const int64_t B = h->n_problems;
    const double t_begin = now_ms();
    const std::vector<int64_t>& pt0 = b.pt0;
    b.pt_prob.assign((size_t)Np, 0);
    for (int64_t p = 0; p < B; ++p)
        for (int64_t q = pt0[(size_t)p]; q < pt0[(size_t)p + 1]; ++q) b.pt_prob[(size_t)q] = (int32_t)p;
    // ---- stable counting sort by point, then by pose inside each point ----
    // every thread owns a contiguous range of POINTS and scans all edges for them: no shared counters, and the order of a
    // point's edges stays the order they were added in (stable), whatever the thread count
    const int nt = host_threads(h);
    b.pt_ptr.assign((size_t)Np + 1, 0);
    parallel_ranges(nt, 0, Np, [&](int, int64_t lo, int64_t hi) {
        for (int64_t e = 0; e < Ne; ++e) {
            const int32_t p = h->e_point[(size_t)e];
            if (p >= lo && p < hi) b.pt_ptr[(size_t)p + 1]++;
        }
    });
    for (int64_t p = 0; p < Np; ++p) b.pt_ptr[(size_t)p + 1] += b.pt_ptr[(size_t)p];
    b.obs_cam.resize((size_t)Ne);
    b.obs_uv.resize((size_t)Ne * 2);
    {
        std::vector<int32_t> cur(b.pt_ptr.begin(), b.pt_ptr.end() - 1);
        parallel_ranges(nt, 0, Np, [&](int, int64_t lo, int64_t hi) {
            for (int64_t e = 0; e < Ne; ++e) {
                const int32_t p = h->e_point[(size_t)e];
                if (p < lo || p >= hi) continue;
                const int32_t s = cur[(size_t)p]++;
                b.obs_cam[(size_t)s] = h->e_pose[(size_t)e];
                b.obs_uv[2 * (size_t)s] = h->e_uv[2 * (size_t)e];
                b.obs_uv[2 * (size_t)s + 1] = h->e_uv[2 * (size_t)e + 1];
            }
            for (int64_t p = lo; p < hi; ++p) { // insertion sort by pose: tracks are short
                const int32_t s0 = b.pt_ptr[(size_t)p], s1 = b.pt_ptr[(size_t)p + 1];
                for (int32_t i = s0 + 1; i < s1; ++i) {
                    const int32_t c = b.obs_cam[(size_t)i];
                    const float u = b.obs_uv[2 * (size_t)i], v = b.obs_uv[2 * (size_t)i + 1];
                    int32_t j = i - 1;
                    while (j >= s0 && b.obs_cam[(size_t)j] > c) {
                        b.obs_cam[(size_t)j + 1] = b.obs_cam[(size_t)j];
                        b.obs_uv[2 * (size_t)j + 2] = b.obs_uv[2 * (size_t)j];
                        b.obs_uv[2 * (size_t)j + 3] = b.obs_uv[2 * (size_t)j + 1];
                        --j;
                    }
                    b.obs_cam[(size_t)j + 1] = c;
                    b.obs_uv[2 * (size_t)j + 2] = u;
                    b.obs_uv[2 * (size_t)j + 3] = v;
                }
            }
        });
    }

    const double t_sort = now_ms();
    // ---- this rank's point range (sharding: contiguous points, balanced on edge count) ----
    {
        std::vector<int64_t> bounds((size_t)h->n_ranks + 1);
        shard_bounds(b.pt_ptr.data(), Np, Ne, h->n_ranks, bounds.data());
        h->shard_pt.assign(bounds.begin(), bounds.end());
    }
    const int32_t pt_begin = h->shard_pt[(size_t)h->rank], pt_end = h->shard_pt[(size_t)h->rank + 1];

    // ---- large single problem: keep S blocks for co-visible pose pairs only (the structure is
    // global: every rank of a sharded solve derives it from ALL points), and the envelope of the
    // dense matrix the solver works in.  Small / batched problems keep the dense triangle. ----
    std::vector<int32_t> remap;
    b.compact = B == 1 && b.max_n > kSmallMaxN;
    if (b.compact) {
        const int32_t nf = b.prob_nfree[0];
        std::vector<uint8_t> present((size_t)b.n_blocks, 0);
        // every thread writes the same value (1) and nothing reads until the threads have joined
        uint8_t* const pres = present.data();
        parallel_ranges(nt, 0, Np, [&](int, int64_t lo, int64_t hi) {
            for (int64_t p = lo; p < hi; ++p) {
                const int32_t s0 = b.pt_ptr[(size_t)p], s1 = b.pt_ptr[(size_t)p + 1];
                for (int32_t a = s0; a < s1; ++a) {
                    const int32_t fa = b.cam_free[(size_t)b.obs_cam[(size_t)a]];
                    if (fa < 0) continue;
                    const size_t row = (size_t)fa * (fa + 1) / 2;
                    for (int32_t c = s0; c < a; ++c) {
                        const int32_t fc = b.cam_free[(size_t)b.obs_cam[(size_t)c]];
                        if (fc >= 0 && !__atomic_load_n(&pres[row + (size_t)fc], __ATOMIC_RELAXED)) __atomic_store_n(&pres[row + (size_t)fc], (uint8_t)1, __ATOMIC_RELAXED);
                    }
                }
            }
        });
        compact_structure(h, b, nf, present, remap);
    }

    const double t_struct = now_ms();
    // ---- pair list sorted by S block (two passes: count, fill) ----
    // thread t takes the t-th contiguous range of this rank's points; inside a block the pairs of range 0 come first, then
    // range 1, ...: the list is ordered by point exactly as a single pass over the points leaves it
    std::vector<int64_t> blk_ptr((size_t)b.n_blocks + 1, 0);
    auto for_pairs = [&](int64_t lo, int64_t hi, auto&& emit0) {
        auto emit = [&](int32_t blk, int32_t a, int32_t c) { emit0(b.compact ? remap[(size_t)blk] : blk, a, c); };
        for (int64_t p = lo; p < hi; ++p) {
            const int32_t tri0 = b.prob_tri0[(size_t)b.pt_prob[(size_t)p]];
            const int32_t s0 = b.pt_ptr[(size_t)p], s1 = b.pt_ptr[(size_t)p + 1];
            for (int32_t a = s0; a < s1; ++a) {
                const int32_t fa = b.cam_free[(size_t)b.obs_cam[(size_t)a]];
                if (fa < 0) continue;
                for (int32_t c = s0; c <= a; ++c) { // sorted by pose: fc <= fa
                    const int32_t fc = b.cam_free[(size_t)b.obs_cam[(size_t)c]];
                    if (fc < 0) continue;
                    emit(tri0 + fa * (fa + 1) / 2 + fc, a, c);
                }
            }
        }
    };
    const int ntp = (pt_end - pt_begin) >= 4096 ? nt : 1;
    std::vector<std::vector<int64_t>> cnt((size_t)ntp, std::vector<int64_t>((size_t)b.n_blocks, 0));
    parallel_ranges(ntp, pt_begin, pt_end, [&](int t, int64_t lo, int64_t hi) {
        std::vector<int64_t>& c = cnt[(size_t)t];
        for_pairs(lo, hi, [&](int32_t blk, int32_t, int32_t) { c[(size_t)blk]++; });
    });
    for (int32_t k = 0; k < b.n_blocks; ++k) { // cnt[t][k] becomes the first slot of thread t in block k
        int64_t run = blk_ptr[(size_t)k];
        for (int t = 0; t < ntp; ++t) {
            const int64_t c = cnt[(size_t)t][(size_t)k];
            cnt[(size_t)t][(size_t)k] = run;
            run += c;
        }
        blk_ptr[(size_t)k + 1] = run;
    }
    b.n_pairs = blk_ptr[(size_t)b.n_blocks];
    if (b.n_pairs > 0x7fffffff) return fail(h, VISO_BA_ERR_UNSUPPORTED, "pair list too long (%lld)", (long long)b.n_pairs);
    b.pairs.resize((size_t)b.n_pairs * 2);
    parallel_ranges(ntp, pt_begin, pt_end, [&](int t, int64_t lo, int64_t hi) {
        std::vector<int64_t>& cur = cnt[(size_t)t];
        for_pairs(lo, hi, [&](int32_t blk, int32_t a, int32_t c) {
            const int64_t q = cur[(size_t)blk]++;
            b.pairs[2 * (size_t)q] = a;
            b.pairs[2 * (size_t)q + 1] = c;
        });
    });

    const double t_pairs = now_ms();
    general_items_and_chunks(h, b, blk_ptr, pt_begin, pt_end);
    if (h->cfg.verbose)
        fprintf(stderr, "[viso_ba] host structure build, %d threads: sort %.1f ms, block structure + solve plan %.1f ms, pair lists %.1f ms (%lld pairs), items %.1f ms\n",
            nt, t_sort - t_begin, t_struct - t_sort, t_pairs - t_struct, (long long)b.n_pairs, now_ms() - t_pairs);
    return VISO_BA_OK;
}

// The general path's structure build ON THE DEVICE (one large problem, B == 1): the edges are staged and validated like the
// window path's, sorted by (point, pose) with the same radix sort, the structure of S is marked by a kernel, the pairs are
// generated point by point and ordered by S block with a stable radix sort.  The host keeps what is small: shard bounds
// (from pt_ptr), the compact block numbering and the solve plan (from the presence bytes), the work items (from the block
// offsets).  The tables are the ones build_general makes (same order of observations, same order of pairs inside a block).
// Observations and pairs stay in the handle's own buffers; the arena holds everything else.

int32_t build_general_device(viso_ba* h, Built& b)
{
    const int64_t Nc = (int64_t)h->poses.size() / 7, Np = (int64_t)h->points.size() / 3;
    const int64_t Ne = h->edges_on_device ? h->n_edges_dev : (int64_t)h->e_pose.size();
    const double t_begin = now_ms();
    cudaStream_t st = h->stream;
    int32_t rc;
    SetupIn in;
    SetupWork w;
    const int64_t cam_off[2] = { 0, Nc }, edge_off[2] = { 0, Ne };
    if (h->edges_on_device) { // staged (and validated) by set_problem straight from the caller's arrays; only the views are rebuilt
        RawEdges r;
        carve_raw(h->raw.p, r, Ne, 1);
        carve_scratch(h->scratch.p, w, Ne, 1, tiles_upper_bound(1, b.pt0.data(), edge_off));
        std::memset(&in, 0, sizeof(in));
        in.n_edges = Ne, in.n_points = (int32_t)Np, in.n_cams = (int32_t)Nc, in.n_problems = 1;
        in.local_indices = h->edges_local ? 1 : 0;
        in.e_pose = r.e_pose, in.e_point = r.e_point, in.e_uv = r.e_uv, in.edge_off = r.edge_off;
        in.prob_cam0 = r.prob_cam0, in.prob_pt0 = r.prob_pt0;
    } else if ((rc = stage_edges_on_device(h, 1, cam_off, b.pt0.data(), edge_off, h->e_pose.data(), h->e_point.data(), h->e_uv.data(), false, &in, &w))) {
        return rc;
    }
    // small tables the kernels read: cam_free, prob_nfree; later the compact block numbering
    const size_t dense_blocks = (size_t)b.n_blocks;
    Carver ct(nullptr);
    ct.take<int32_t>((size_t)Nc), ct.take<int32_t>(1), ct.take<int32_t>(dense_blocks), ct.take<uint8_t>(dense_blocks);
    if ((rc = devbuf_reserve(h, h->gen_tabs, ct.used))) return rc;
    Carver c2(h->gen_tabs.p);
    int32_t* d_cam_free = c2.take<int32_t>((size_t)Nc);
    int32_t* d_nfree = c2.take<int32_t>(1);
    int32_t* d_remap = c2.take<int32_t>(dense_blocks);
    uint8_t* d_present = c2.take<uint8_t>(dense_blocks);
    CU(cudaMemcpyAsync(d_cam_free, b.cam_free.data(), sizeof(int32_t) * (size_t)Nc, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_nfree, b.prob_nfree.data(), sizeof(int32_t), cudaMemcpyHostToDevice, st));
    in.cam_free = d_cam_free, in.prob_nfree = d_nfree;
    // observations in (point, pose) order
    Carver co(nullptr);
    co.take<int32_t>((size_t)Np + 1), co.take<int32_t>((size_t)Np), co.take<int32_t>((size_t)Ne), co.take<float2>((size_t)Ne);
    if ((rc = devbuf_reserve(h, h->gen_obs, co.used))) return rc;
    Carver c3(h->gen_obs.p);
    SetupOut o;
    std::memset(&o, 0, sizeof(o));
    o.pt_ptr = c3.take<int32_t>((size_t)Np + 1), o.pt_prob = c3.take<int32_t>((size_t)Np);
    o.obs_cam = c3.take<int32_t>((size_t)Ne), o.obs_uv = c3.take<float2>((size_t)Ne);
    run_k(h, "setup", [&] { return setup_sort_only(in, w, o, st); });
    const bool compact = b.max_n > kSmallMaxN;
    if (compact) {
        CU(cudaMemsetAsync(d_present, 0, dense_blocks, st));
        run_k(h, "setup", [&] { return gen_present((int32_t)Np, o.pt_ptr, o.obs_cam, d_cam_free, d_present, st); });
    }
    b.pt_ptr.resize((size_t)Np + 1);
    CU(cudaMemcpyAsync(b.pt_ptr.data(), o.pt_ptr, sizeof(int32_t) * ((size_t)Np + 1), cudaMemcpyDeviceToHost, st));
    std::vector<uint8_t> present;
    if (compact) {
        present.resize(dense_blocks);
        CU(cudaMemcpyAsync(present.data(), d_present, dense_blocks, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    const double t_sort = now_ms();
    { // this rank's point range (sharding: contiguous points, balanced on edge count)
        std::vector<int64_t> bounds((size_t)h->n_ranks + 1);
        shard_bounds(b.pt_ptr.data(), Np, Ne, h->n_ranks, bounds.data());
        h->shard_pt.assign(bounds.begin(), bounds.end());
    }
    const int32_t pt_begin = h->shard_pt[(size_t)h->rank], pt_end = h->shard_pt[(size_t)h->rank + 1];
    std::vector<int32_t> remap;
    b.compact = compact;
    if (compact) {
        compact_structure(h, b, b.prob_nfree[0], present, remap);
        CU(cudaMemcpyAsync(d_remap, remap.data(), sizeof(int32_t) * dense_blocks, cudaMemcpyHostToDevice, st));
    }
    const double t_struct = now_ms();
    // pairs of this rank's points: count, scan, generate, sort by block, block offsets
    const int64_t npl = (int64_t)pt_end - pt_begin;
    size_t temp_bytes = gen_temp_bytes(npl, 0);
    Carver cw0(nullptr);
    cw0.take<int64_t>((size_t)npl + 1), cw0.take<char>(temp_bytes);
    if ((rc = devbuf_reserve(h, h->gen_work, cw0.used))) return rc;
    {
        Carver cw(h->gen_work.p);
        int64_t* d_counts = cw.take<int64_t>((size_t)npl + 1);
        void* d_temp = cw.take<char>(temp_bytes);
        run_k(h, "setup", [&] { return gen_pair_offsets(pt_begin, pt_end, o.pt_ptr, o.obs_cam, d_cam_free, d_counts, d_temp, temp_bytes, st); });
        if ((rc = pin_reserve(h, 64))) return rc;
        CU(cudaMemcpyAsync(h->pin, d_counts + npl, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    b.n_pairs = *static_cast<const int64_t*>(h->pin);
    if (b.n_pairs > 0x7fffffff) return fail(h, VISO_BA_ERR_UNSUPPORTED, "pair list too long (%lld)", (long long)b.n_pairs);
    const size_t npairs = (size_t)std::max<int64_t>(b.n_pairs, 1);
    temp_bytes = gen_temp_bytes(npl, b.n_pairs);
    Carver cw1(nullptr); // the offsets stay at the front: the buffer may move when it grows, so they are recomputed below
    cw1.take<int64_t>((size_t)npl + 1), cw1.take<char>(temp_bytes), cw1.take<unsigned int>(npairs), cw1.take<unsigned int>(npairs);
    cw1.take<unsigned long long>(npairs), cw1.take<int64_t>((size_t)b.n_blocks + 1);
    const bool moved = cw1.used > h->gen_work.cap;
    if ((rc = devbuf_reserve(h, h->gen_work, cw1.used))) return rc;
    if ((rc = devbuf_reserve(h, h->pairbuf, sizeof(unsigned long long) * npairs))) return rc;
    std::vector<int64_t> blk_ptr((size_t)b.n_blocks + 1, 0);
    {
        Carver cw(h->gen_work.p);
        int64_t* d_counts = cw.take<int64_t>((size_t)npl + 1);
        void* d_temp = cw.take<char>(temp_bytes);
        unsigned int* d_keys = cw.take<unsigned int>(npairs);
        unsigned int* d_keys_sorted = cw.take<unsigned int>(npairs);
        unsigned long long* d_vals = cw.take<unsigned long long>(npairs);
        int64_t* d_blk_ptr = cw.take<int64_t>((size_t)b.n_blocks + 1);
        if (moved) run_k(h, "setup", [&] { return gen_pair_offsets(pt_begin, pt_end, o.pt_ptr, o.obs_cam, d_cam_free, d_counts, d_temp, temp_bytes, st); });
        run_k(h, "setup", [&] {
            return gen_pairs(pt_begin, pt_end, o.pt_ptr, o.obs_cam, d_cam_free, compact ? d_remap : nullptr, d_counts, b.n_pairs, b.n_blocks, d_keys,
                d_keys_sorted, d_vals, static_cast<unsigned long long*>(h->pairbuf.p), d_blk_ptr, d_temp, temp_bytes, st);
        });
        CU(cudaMemcpyAsync(blk_ptr.data(), d_blk_ptr, sizeof(int64_t) * ((size_t)b.n_blocks + 1), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        CU(cudaGetLastError());
    }
    const double t_pairs = now_ms();
    general_items_and_chunks(h, b, blk_ptr, pt_begin, pt_end);
    b.dev_general = true;
    if (h->cfg.verbose)
        fprintf(stderr, "[viso_ba] device structure build: stage + sort %.1f ms, block structure + solve plan %.1f ms, pair lists %.1f ms (%lld pairs), items %.1f ms\n",
            t_sort - t_begin, t_struct - t_sort, t_pairs - t_struct, (long long)b.n_pairs, now_ms() - t_pairs);
    return VISO_BA_OK;
}

template <typename T> int32_t h2d(viso_ba* h, T* dst, const std::vector<T>& src)
{
    if (src.empty()) return VISO_BA_OK;
    CU(cudaMemcpyAsync(dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice, h->stream));
    return VISO_BA_OK;
}

} // namespace

// ==========================================================================================
// upload: H2D + device layout
// ==========================================================================================
namespace {

struct DevPtrs { // everything carved out of the arena, in one place so that the dry run and the
                 // real pass cannot diverge
    int32_t *pt_ptr, *pt_prob, *obs_cam, *cam_free, *cam_prob, *prob_cam0, *prob_tri0, *prob_nfree;
    int32_t* diag_blk;
    int2* blk_ij;
    float2* obs_uv;
    int2* pairs;
    int4 *items, *chunks;
    int32_t *diag_items, *block_item_ptr, *block_cam, *prob_chunk_ptr;
    WinTile* tiles;
    uint32_t *pt_mask, *win_meta, *win_blkptr;
    float2* win_uv;
    uint16_t* win_slot;
    int32_t *prob_tile_ptr, *prob_pt0;
    int64_t* prob_part_off;
    double *tile_part, *tile_sums;
    int32_t* tile_map;
    double *pts[2], *pose[2], *cam_rt[2], *pose_init, *pts_init, *pose_out, *pts_out;
    LmState* lm;
    IterStat* iter_stats;
    double *rec, *ptinfo, *cost_pt, *cost_new_pt, *scale_pt, *item_part, *S_blk, *bs, *dp, *chunk_part, *prob_sums;
    unsigned long long* maxdiag_bits;
    int32_t* n_active;
    // large solver (ba_dense.cu): task matrices / diagonal factors / solutions in one block of doubles, tables in one of ints
    double *Ldoubles, *Lx;
    int32_t *Lints, *Lfail;
    int4* Lsteps;
    CholTask* Ltasks;
    unsigned int* Lbars;
};

constexpr int kMaxIterStats = 4096;

void carve(Arena& A, DevPtrs& d, int64_t Nc, int64_t Np, int64_t Ne, int64_t B, const Built& b, bool large, int ldL)
{
    A.used = 0;
    const bool own = !b.dev_general; // the device-side general build keeps observations and pairs in the handle's own buffers
    d.pt_ptr = A.take<int32_t>(own ? (size_t)Np + 1 : 0);
    d.pt_prob = A.take<int32_t>(own ? (size_t)Np : 0);
    d.obs_cam = A.take<int32_t>(own ? (size_t)Ne : 0);
    d.obs_uv = A.take<float2>(own ? (size_t)Ne : 0);
    d.cam_free = A.take<int32_t>((size_t)Nc);
    d.cam_prob = A.take<int32_t>((size_t)Nc);
    d.prob_cam0 = A.take<int32_t>((size_t)B + 1);
    d.prob_tri0 = A.take<int32_t>((size_t)B + 1);
    d.prob_nfree = A.take<int32_t>((size_t)B);
    d.pairs = A.take<int2>(own ? (size_t)b.n_pairs : 0);
    const size_t NeW = b.window ? (size_t)Ne : 0, NpW = b.window ? (size_t)Np : 0, BW = b.window ? (size_t)B + 1 : 0;
    d.items = A.take<int4>(b.items.size() / 4);
    d.diag_items = A.take<int32_t>(b.diag_items.size());
    d.block_item_ptr = A.take<int32_t>((size_t)b.n_blocks + 1);
    d.block_cam = A.take<int32_t>((size_t)b.n_blocks);
    d.diag_blk = A.take<int32_t>((size_t)Nc);
    d.blk_ij = A.take<int2>(b.blk_ij.size() / 2);
    d.chunks = A.take<int4>(b.chunks.size() / 4);
    d.prob_chunk_ptr = A.take<int32_t>((size_t)B + 1);
    d.tiles = A.take<WinTile>((size_t)b.tiles_max);
    d.pt_mask = A.take<uint32_t>(NpW);
    d.win_meta = A.take<uint32_t>(NeW);
    d.win_blkptr = A.take<uint32_t>((size_t)b.blkptr_max);
    d.win_uv = A.take<float2>(NeW);
    d.win_slot = A.take<uint16_t>(NeW);
    d.prob_tile_ptr = A.take<int32_t>(BW);
    d.prob_part_off = A.take<int64_t>(BW);
    d.prob_pt0 = A.take<int32_t>(BW);
    d.tile_part = A.take<double>((size_t)b.part_max);
    d.tile_sums = A.take<double>((size_t)b.tiles_max * 2);
    d.tile_map = A.take<int32_t>((size_t)b.tiles_max);
    for (int k = 0; k < 2; ++k) {
        d.pts[k] = A.take<double>((size_t)Np * 3);
        d.pose[k] = A.take<double>((size_t)Nc * 7);
        d.cam_rt[k] = A.take<double>((size_t)Nc * 12);
    }
    d.pose_init = A.take<double>((size_t)Nc * 7);
    d.pts_init = A.take<double>((size_t)Np * 3);
    d.pose_out = A.take<double>((size_t)Nc * 7);
    d.pts_out = A.take<double>((size_t)Np * 3);
    d.lm = A.take<LmState>((size_t)B);
    d.iter_stats = A.take<IterStat>(kMaxIterStats);
    d.rec = A.take<double>(b.window ? 0 : (size_t)Ne * kRecDoubles);
    d.ptinfo = A.take<double>((size_t)Np * kPtInfoDoubles);
    d.cost_pt = A.take<double>((size_t)Np);
    d.cost_new_pt = A.take<double>((size_t)Np);
    d.scale_pt = A.take<double>((size_t)Np);
    d.item_part = A.take<double>(b.items.size() / 4 * kItemPart);
    // S_blk | bs | prob_sums must stay contiguous: they are what a sharded handle all-reduces
    d.S_blk = A.take<double>((size_t)b.n_blocks * 36 + (size_t)Nc * 6 + (size_t)B * 2);
    d.bs = d.S_blk + (size_t)b.n_blocks * 36;
    d.prob_sums = d.bs + (size_t)Nc * 6;
    d.dp = A.take<double>((size_t)Nc * 6);
    d.chunk_part = A.take<double>(b.chunks.size() / 4 * 2);
    d.maxdiag_bits = A.take<unsigned long long>((size_t)B);
    d.n_active = A.take<int32_t>(2); // [1]: tiles of the active problems (k_win_active_tiles)
    if (large) {
        size_t nd = 0, ni = 0, ns = 0;
        for (const PlanTask& t : b.plan.tasks) {
            nd += t.a_doubles() + t.dd_doubles() + (size_t)t.nrows;
            ni += 2 * (size_t)(t.nrows / 64); // flags, first64
            ns += (size_t)(t.n_elim / 64);
        }
        ni += 3 * (size_t)b.plan.nf + b.plan.tasks.size() + 1; // cam_task, cam_off, cam_sep, row_ptr
        d.Ldoubles = A.take<double>(nd);
        d.Lx = A.take<double>((size_t)b.max_n);
        d.Lints = A.take<int32_t>(ni);
        d.Lsteps = A.take<int4>(ns);
        d.Ltasks = A.take<CholTask>(b.plan.tasks.size());
        d.Lbars = A.take<unsigned int>(kBarsPerTask * b.plan.tasks.size());
        d.Lfail = A.take<int32_t>(1);
    } else {
        d.Ldoubles = d.Lx = nullptr;
        d.Lints = d.Lfail = nullptr;
        d.Lsteps = nullptr, d.Ltasks = nullptr, d.Lbars = nullptr;
    }
    (void)ldL;
}

} // namespace

namespace {

// large solver: the plan's tables and the task descriptors (device pointers into the arena) -> device
int32_t upload_solve_plan(viso_ba* h, const SolvePlan& plan, const DevPtrs& d)
{
    const size_t nt = plan.tasks.size();
    std::vector<CholTask> tasks(nt);
    std::vector<int32_t> ints;
    std::vector<int32_t> steps;
    double* pd = d.Ldoubles;
    int4* ps = d.Lsteps;
    // ints layout: per task {flags, first64}, then cam_task, cam_off, cam_sep, row_ptr
    size_t io = 0;
    for (size_t t = 0; t < nt; ++t) {
        const PlanTask& q = plan.tasks[t];
        CholTask& c = tasks[t];
        std::memset(&c, 0, sizeof(c));
        c.A = pd, pd += q.a_doubles();
        c.Dd = pd, pd += q.dd_doubles();
        c.x = pd, pd += (size_t)q.nrows;
        const size_t nb = (size_t)(q.nrows / 64);
        c.flags = d.Lints + io, io += nb;
        ints.insert(ints.end(), nb, 0);
        c.first64 = d.Lints + io, io += nb;
        ints.insert(ints.end(), q.first64.begin(), q.first64.end());
        c.steps = ps, ps += q.n_elim / 64;
        steps.insert(steps.end(), q.steps.begin(), q.steps.end());
        c.ld = q.ld, c.n_elim = q.n_elim, c.nrows = q.nrows, c.n_real = q.n_real;
        c.cta0 = q.cta0, c.ncta = q.ncta, c.sL = q.sL, c.sR = q.sR;
    }
    LargeSolve& L = h->L;
    std::memset(&L, 0, sizeof(L));
    L.n = 6 * plan.nf, L.n_seg = plan.n_seg, L.sep_cols = plan.sep_cols, L.total_rows = plan.total_rows;
    L.tasks = d.Ltasks;
    L.cam_task = d.Lints + io, io += (size_t)plan.nf, ints.insert(ints.end(), plan.cam_task.begin(), plan.cam_task.end());
    L.cam_off = d.Lints + io, io += (size_t)plan.nf, ints.insert(ints.end(), plan.cam_off.begin(), plan.cam_off.end());
    L.cam_sep = d.Lints + io, io += (size_t)plan.nf, ints.insert(ints.end(), plan.cam_sep.begin(), plan.cam_sep.end());
    L.row_ptr = d.Lints + io, io += nt + 1, ints.insert(ints.end(), plan.row_ptr.begin(), plan.row_ptr.end());
    L.x = d.Lx, L.fail = d.Lfail, L.bars = d.Lbars;
    CU(cudaMemcpyAsync(d.Lints, ints.data(), ints.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(d.Lsteps, steps.data(), steps.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(d.Ltasks, tasks.data(), nt * sizeof(CholTask), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemsetAsync(d.Lx, 0, sizeof(double) * (size_t)L.n, h->stream));
    CU(cudaStreamSynchronize(h->stream)); // the host vectors die here
    return VISO_BA_OK;
}

// window fast path: sort, CSR, tiles, slots and pair lists are built by kernels (ba_setup.cu).
// Returns VISO_BA_OK with *fallback = true when the problem has duplicate (point, pose) edges.
int32_t build_window_on_device(viso_ba* h, const Built& b, const DevPtrs& d, bool* fallback)
{
    const int64_t Np = (int64_t)h->points.size() / 3, B = h->n_problems;
    int32_t rc;
    *fallback = false;
    SetupIn in;
    SetupWork w;
    if (h->edges_on_device) { // staged (and validated) by set_problem; only the views are rebuilt
        RawEdges r;
        carve_raw(h->raw.p, r, h->n_edges_dev, B);
        carve_scratch(h->scratch.p, w, h->n_edges_dev, B, b.tiles_max);
        std::memset(&in, 0, sizeof(in));
        in.n_edges = h->n_edges_dev, in.n_points = (int32_t)Np, in.n_cams = (int32_t)(h->poses.size() / 7), in.n_problems = (int32_t)B;
        in.local_indices = h->edges_local ? 1 : 0;
        in.e_pose = r.e_pose, in.e_point = r.e_point, in.e_uv = r.e_uv, in.edge_off = r.edge_off;
        in.prob_cam0 = r.prob_cam0, in.prob_pt0 = r.prob_pt0;
    } else { // incremental API or a small bulk problem: the host vectors hold global indices
        std::vector<int64_t> cam_off((size_t)B + 1);
        for (int64_t p = 0; p <= B; ++p) cam_off[(size_t)p] = b.prob_cam0[(size_t)p];
        if ((rc = stage_edges_on_device(h, B, cam_off.data(), b.pt0.data(), h->edge_off_all.data(), h->e_pose.data(), h->e_point.data(),
                 h->e_uv.data(), false, &in, &w)))
            return rc;
    }
    in.prob_nfree = d.prob_nfree, in.cam_free = d.cam_free; // uploaded by the caller on the same stream
    SetupOut o;
    o.pt_ptr = d.pt_ptr, o.pt_prob = d.pt_prob, o.obs_cam = d.obs_cam, o.obs_uv = d.obs_uv;
    o.prob_tile_ptr = d.prob_tile_ptr, o.prob_part_off = d.prob_part_off, o.tiles = d.tiles, o.pt_mask = d.pt_mask;
    o.win_meta = d.win_meta, o.win_uv = d.win_uv, o.win_slot = d.win_slot, o.win_blkptr = d.win_blkptr, o.win_pairs = nullptr;
    cudaStream_t st = h->stream;
    run_k(h, "setup", [&] { return setup_stage1(in, w, o, st); });
    if ((rc = pin_reserve(h, 128))) return rc;
    char* pin = static_cast<char*>(h->pin);
    CU(cudaMemcpyAsync(pin, w.err, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(pin + 32, w.totals, 4 * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    if (reinterpret_cast<const int*>(pin)[1]) { // duplicate edges: the visibility-mask logic does not apply
        *fallback = true;
        return VISO_BA_OK;
    }
    const int64_t* totals = reinterpret_cast<const int64_t*>(pin + 32);
    const int64_t n_tiles = totals[0];
    if (n_tiles > b.tiles_max || totals[1] > b.part_max || totals[2] > b.blkptr_max || n_tiles > 0x7fffffff)
        return fail(h, VISO_BA_ERR_STATE, "internal: tile bound exceeded (%lld tiles, bound %lld)", (long long)n_tiles, (long long)b.tiles_max);
    run_k(h, "setup", [&] { return setup_stage2(in, w, o, (int32_t)n_tiles, st); });
    CU(cudaMemcpyAsync(pin + 64, w.tile_list_off + n_tiles, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    const int64_t n_pairs = n_tiles > 0 ? *reinterpret_cast<const int64_t*>(pin + 64) : 0;
    if ((rc = devbuf_reserve(h, h->pairbuf, sizeof(uint32_t) * (size_t)std::max<int64_t>(n_pairs, 1)))) return rc;
    o.win_pairs = static_cast<uint32_t*>(h->pairbuf.p);
    run_k(h, "setup", [&] { return setup_stage3(in, w, o, (int32_t)n_tiles, st); });
    CU(cudaGetLastError());
    DevProblem& P = h->P;
    P.n_tiles = (int32_t)n_tiles, P.n_pairs = n_pairs;
    P.win_pairs = o.win_pairs;
    return VISO_BA_OK;
}

// after a duplicate edge sent a device-staged problem to the general path: bring the edges back
int32_t edges_back_to_host(viso_ba* h)
{
    if (!h->edges_on_device) return VISO_BA_OK;
    const int64_t Ne = h->n_edges_dev, B = h->n_problems;
    RawEdges r;
    carve_raw(h->raw.p, r, Ne, B);
    h->e_pose.resize((size_t)Ne), h->e_point.resize((size_t)Ne), h->e_uv.resize(2 * (size_t)Ne);
    CU(cudaMemcpyAsync(h->e_pose.data(), r.e_pose, sizeof(int32_t) * (size_t)Ne, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(h->e_point.data(), r.e_point, sizeof(int32_t) * (size_t)Ne, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(h->e_uv.data(), r.e_uv, sizeof(float) * 2 * (size_t)Ne, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (h->edges_local && B > 1)
        for (int64_t b = 0; b < B; ++b)
            for (int64_t e = h->edge_off[(size_t)b]; e < h->edge_off[(size_t)b + 1]; ++e) {
                h->e_pose[(size_t)e] += (int32_t)h->pose_off[(size_t)b];
                h->e_point[(size_t)e] += (int32_t)h->point_off[(size_t)b];
            }
    h->edges_on_device = false, h->edges_local = false, h->n_edges_dev = 0;
    return VISO_BA_OK;
}

} // namespace

namespace {
// device side of a photometric problem: the keyframe images, their descriptors, the per-edge sums
int32_t upload_photo(viso_ba* h, int64_t Nc, int64_t Ne, int64_t B)
{
    if (B > 1 || h->n_ranks > 1) return fail(h, VISO_BA_ERR_UNSUPPORTED, "photometric edges: one problem on one GPU");
    if (h->window || h->edges_on_device) return fail(h, VISO_BA_ERR_STATE, "internal: photometric edges on the window path");
    std::vector<uint8_t> used((size_t)Nc, 0);
    for (size_t e = 0; e < h->e_pose.size(); ++e) {
        int32_t src;
        std::memcpy(&src, &h->e_uv[2 * e], sizeof(src));
        if (src < 0 || src >= Nc) return fail(h, VISO_BA_ERR_INDEX, "edge %zu refers to source keyframe %d", e, src);
        used[(size_t)src] = 1, used[(size_t)h->e_pose[e]] = 1;
    }
    struct Img { const uint8_t* data; int32_t cols, rows, step, pad; }; // = visoimg::DevImage
    std::vector<Img> frames((size_t)Nc);
    size_t bytes = 0;
    for (int64_t c = 0; c < Nc; ++c) {
        const bool have = (size_t)c < h->ph_img.size() && !h->ph_img[(size_t)c].empty();
        if (used[(size_t)c] && !have) return fail(h, VISO_BA_ERR_STATE, "pose %lld takes part in a photometric edge but has no image (viso_ba_set_pose_image)", (long long)c);
        if (have) bytes += (h->ph_img[(size_t)c].size() + 255) & ~(size_t)255;
    }
    int32_t rc;
    if ((rc = devbuf_reserve(h, h->ph_img_dev, bytes + 256)) || (rc = devbuf_reserve(h, h->ph_frames_dev, sizeof(Img) * (size_t)Nc))
        || (rc = devbuf_reserve(h, h->ph_srcrt_dev, sizeof(double) * 12 * (size_t)Nc))
        || (rc = devbuf_reserve(h, h->ph_sum_dev, sizeof(double) * photo_sum_doubles() * (size_t)std::max<int64_t>(Ne, 1)))
        || (rc = devbuf_reserve(h, h->ph_mid_dev, sizeof(double) * 4 * (size_t)std::max<int64_t>(Ne, 1))))
        return rc;
    size_t off = 0;
    for (int64_t c = 0; c < Nc; ++c) {
        Img& f = frames[(size_t)c];
        f.data = nullptr, f.cols = f.rows = f.step = f.pad = 0;
        if ((size_t)c >= h->ph_img.size() || h->ph_img[(size_t)c].empty()) continue;
        const std::vector<uint8_t>& im = h->ph_img[(size_t)c];
        f.data = static_cast<const uint8_t*>(h->ph_img_dev.p) + off;
        f.cols = h->ph_cols[(size_t)c], f.rows = h->ph_rows[(size_t)c], f.step = f.cols;
        CU(cudaMemcpyAsync(const_cast<uint8_t*>(f.data), im.data(), im.size(), cudaMemcpyHostToDevice, h->stream));
        off += (im.size() + 255) & ~(size_t)255;
    }
    CU(cudaMemcpyAsync(h->ph_frames_dev.p, frames.data(), sizeof(Img) * (size_t)Nc, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemsetAsync(h->ph_sum_dev.p, 0, sizeof(double) * photo_sum_doubles() * (size_t)std::max<int64_t>(Ne, 1), h->stream));
    CU(cudaStreamSynchronize(h->stream)); // `frames` dies here
    h->P.ph_frames = h->ph_frames_dev.p;
    h->P.ph_src_rt = static_cast<const double*>(h->ph_srcrt_dev.p);
    h->P.ph_sum = static_cast<double*>(h->ph_sum_dev.p);
    h->P.ph_mid = static_cast<double*>(h->ph_mid_dev.p);
    return VISO_BA_OK;
}
} // namespace

extern "C" int32_t viso_ba_upload(viso_ba* h)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (h->uploaded) return VISO_BA_OK;
    int32_t rc = ensure_cuda(h);
    if (rc) return rc;
    const double t0 = now_ms();
    const int64_t Nc = (int64_t)h->poses.size() / 7, Np = (int64_t)h->points.size() / 3;
    const int64_t B = h->n_problems;
    if (Nc < 1 || Np < 1) return fail(h, VISO_BA_ERR_STATE, "nothing to optimise: %lld poses, %lld points", (long long)Nc, (long long)Np);
    drop_graph(h);
    for (int attempt = 0;; ++attempt) {
        const int64_t Ne = h->edges_on_device ? h->n_edges_dev : (int64_t)h->e_pose.size();
        Built b;
        if ((rc = build_tables(h, b))) break;
        if (!b.window) {
            const bool on_device = B == 1 && !h->photo && Ne >= kDevGeneralMinEdges && h->cfg.host_structure_build == 0;
            if (!on_device && (rc = edges_back_to_host(h))) break; // the host build reads the host vectors
            if ((rc = on_device ? build_general_device(h, b) : build_general(h, b))) break;
        }
        h->max_n = b.max_n;
        h->use_large = b.max_n > kSmallMaxN;
        if (h->use_large && B > 1) {
            rc = fail(h, VISO_BA_ERR_UNSUPPORTED, "batched problems need <= %d free poses each (got a reduced system of %d)", kSmallMaxN / 6, b.max_n);
            break;
        }
        if (h->use_large && b.max_n > large_solve_max_n()) {
            rc = fail(h, VISO_BA_ERR_UNSUPPORTED, "the dense reduced-system solver takes at most %d free poses (got %d)", large_solve_max_n() / 6, b.max_n / 6);
            break;
        }
        if (h->n_ranks > 1 && B > 1) {
            rc = fail(h, VISO_BA_ERR_UNSUPPORTED, "batched problems are not sharded: run replicas instead");
            break;
        }
        const int ldL = (b.max_n + 8) & ~7;

        DevPtrs d;
        Arena dry;
        carve(dry, d, Nc, Np, Ne, B, b, h->use_large, ldL);
        if (dry.used > h->arena.cap) {
            if (h->arena.base) {
                CU(cudaStreamSynchronize(h->stream));
                CU(cudaFree(h->arena.base));
                h->arena.base = nullptr, h->arena.cap = 0;
            }
            const size_t want = dry.used + dry.used / 8; // head-room so that a slightly larger window re-uses it
            CU(cudaMalloc((void**)&h->arena.base, want));
            h->arena.cap = want;
        }
        carve(h->arena, d, Nc, Np, Ne, B, b, h->use_large, ldL);

        DevProblem& P = h->P;
        std::memset(&P, 0, sizeof(P));
        // tables every path needs
        if ((rc = h2d(h, d.cam_free, b.cam_free))) break;
        if ((rc = h2d(h, d.cam_prob, b.cam_prob))) break;
        if ((rc = h2d(h, d.prob_cam0, b.prob_cam0))) break;
        if ((rc = h2d(h, d.prob_tri0, b.prob_tri0))) break;
        if ((rc = h2d(h, d.prob_nfree, b.prob_nfree))) break;
        if ((rc = h2d(h, d.block_cam, b.block_cam))) break;
        if ((rc = h2d(h, d.diag_blk, b.diag_blk))) break;
        if ((rc = h2d(h, reinterpret_cast<int32_t*>(d.blk_ij), b.blk_ij))) break;
        if (b.window) {
            bool fallback = false;
            if ((rc = build_window_on_device(h, b, d, &fallback))) break;
            if (fallback) { // duplicate (point, pose) edges: the general path sums them correctly
                if (attempt > 0) {
                    rc = fail(h, VISO_BA_ERR_STATE, "internal: window fallback loop");
                    break;
                }
                if ((rc = edges_back_to_host(h))) break;
                h->no_window = true;
                continue;
            }
        } else {
            if (b.dev_general) { // observations and pairs were built on the device and live in the handle's own buffers
                Carver c3(h->gen_obs.p);
                d.pt_ptr = c3.take<int32_t>((size_t)Np + 1), d.pt_prob = c3.take<int32_t>((size_t)Np);
                d.obs_cam = c3.take<int32_t>((size_t)Ne), d.obs_uv = c3.take<float2>((size_t)Ne);
                d.pairs = static_cast<int2*>(h->pairbuf.p);
            } else {
                if ((rc = h2d(h, d.pt_ptr, b.pt_ptr))) break;
                if ((rc = h2d(h, d.pt_prob, b.pt_prob))) break;
                if ((rc = h2d(h, d.obs_cam, b.obs_cam))) break;
                if ((rc = h2d(h, reinterpret_cast<float*>(d.obs_uv), b.obs_uv))) break;
                if ((rc = h2d(h, reinterpret_cast<int32_t*>(d.pairs), b.pairs))) break;
            }
            if ((rc = h2d(h, reinterpret_cast<int32_t*>(d.items), b.items))) break;
            if ((rc = h2d(h, d.diag_items, b.diag_items))) break;
            if ((rc = h2d(h, d.block_item_ptr, b.block_item_ptr))) break;
            if ((rc = h2d(h, reinterpret_cast<int32_t*>(d.chunks), b.chunks))) break;
            if ((rc = h2d(h, d.prob_chunk_ptr, b.prob_chunk_ptr))) break;
        }
        // state: normalised quaternions (Sophus keeps unit quaternions), buffer 0 current
        std::vector<double> poses = h->poses;
        for (int64_t c = 0; c < Nc; ++c) {
            double* q = &poses[(size_t)c * 7];
            const double n = 1.0 / std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
            for (int i = 0; i < 4; ++i) q[i] *= n;
        }
        if ((rc = h2d(h, d.pose_init, poses))) break;
        if (h->pts_staged) { // the points went to the device when the problem was set
            CU(cudaMemcpyAsync(d.pts_init, h->pts_stage.p, sizeof(double) * 3 * (size_t)Np, cudaMemcpyDeviceToDevice, h->stream));
        } else if (Np) {
            CU(cudaMemcpyAsync(d.pts_init, h->points.data(), sizeof(double) * 3 * (size_t)Np, cudaMemcpyHostToDevice, h->stream));
        }
        CU(cudaStreamSynchronize(h->stream)); // `poses` and `b` die at the end of this scope

        P.n_cams = (int32_t)Nc, P.n_points = (int32_t)Np, P.n_problems = (int32_t)B;
        if (b.window) P.pt_begin = 0, P.pt_end = (int32_t)Np;
        else P.pt_begin = h->shard_pt[(size_t)h->rank], P.pt_end = h->shard_pt[(size_t)h->rank + 1];
        P.n_obs = Ne;
        if (!b.window) P.n_pairs = b.n_pairs;
        P.n_items = (int32_t)(b.items.size() / 4), P.n_diag_items = (int32_t)b.diag_items.size();
        P.n_blocks = b.n_blocks, P.n_chunks = (int32_t)(b.chunks.size() / 4);
        P.fx = h->cfg.fx, P.fy = h->cfg.fy, P.cx = h->cfg.cx, P.cy = h->cfg.cy, P.huber_delta = h->cfg.huber_delta;
        P.pt_ptr = d.pt_ptr, P.pt_prob = d.pt_prob, P.obs_cam = d.obs_cam, P.obs_uv = d.obs_uv;
        P.cam_free = d.cam_free, P.cam_prob = d.cam_prob, P.prob_cam0 = d.prob_cam0, P.prob_tri0 = d.prob_tri0;
        P.prob_nfree = d.prob_nfree, P.pairs = d.pairs, P.items = d.items, P.diag_items = d.diag_items;
        P.block_item_ptr = d.block_item_ptr, P.block_cam = d.block_cam, P.chunks = d.chunks, P.prob_chunk_ptr = d.prob_chunk_ptr;
        P.diag_blk = d.diag_blk, P.blk_ij = b.compact ? d.blk_ij : nullptr;
        for (int k = 0; k < 2; ++k) P.pts[k] = d.pts[k], P.pose[k] = d.pose[k], P.cam_rt[k] = d.cam_rt[k];
        P.lm = d.lm, P.iter_stats = (B == 1) ? d.iter_stats : nullptr;
        P.rec = d.rec, P.ptinfo = d.ptinfo, P.cost_pt = d.cost_pt, P.cost_new_pt = d.cost_new_pt, P.scale_pt = d.scale_pt;
        P.item_part = d.item_part, P.S_blk = d.S_blk, P.bs = d.bs, P.dp = d.dp, P.chunk_part = d.chunk_part;
        P.prob_sums = d.prob_sums, P.maxdiag_bits = d.maxdiag_bits, P.n_active = d.n_active;
        P.lambda_scale = h->rank == 0 ? 1.0 : 0.0;
        h->window = b.window;
        P.tiles = d.tiles, P.pt_mask = d.pt_mask; // n_tiles, n_pairs, win_pairs: set by build_window_on_device
        P.win_meta = d.win_meta, P.win_blkptr = d.win_blkptr, P.win_uv = d.win_uv, P.win_slot = d.win_slot;
        P.prob_tile_ptr = d.prob_tile_ptr, P.prob_part_off = d.prob_part_off, P.tile_part = d.tile_part, P.tile_sums = d.tile_sums;
        P.tile_map = nullptr, P.n_map_tiles = 0, h->d_tile_map = d.tile_map;
        h->d_pose_init = d.pose_init, h->d_pts_init = d.pts_init, h->d_pose_out = d.pose_out, h->d_pts_out = d.pts_out;
        h->d_allreduce = d.S_blk;
        h->allreduce_count = (size_t)b.n_blocks * 36 + (size_t)Nc * 6;
        std::memset(&h->L, 0, sizeof(h->L));
        if (h->use_large) {
            if ((rc = upload_solve_plan(h, b.plan, d))) break;
        }
        h->blk_ij_host = b.compact ? b.blk_ij : std::vector<int32_t>();
        CU(cudaMemsetAsync(d.S_blk, 0, sizeof(double) * (h->allreduce_count + (size_t)B * 2), h->stream));
        CU(cudaMemsetAsync(d.bs, 0, sizeof(double) * (size_t)Nc * 6, h->stream));
        CU(cudaMemsetAsync(d.dp, 0, sizeof(double) * (size_t)Nc * 6, h->stream));
        if (h->photo && (rc = upload_photo(h, Nc, Ne, B))) break;
        rc = VISO_BA_OK;
        break;
    }
    if (rc) return rc;
    h->uploaded = true;
    if ((rc = viso_ba_reset_state(h))) {
        h->uploaded = false;
        return rc;
    }
    CU(cudaStreamSynchronize(h->stream));
    h->setup_ms = (float)(now_ms() - t0);
    return VISO_BA_OK;
}

// debug / test hook: the HOST part of the general structure build (tables, sort by point and pose, pair lists by S block,
// work items) with `threads` host threads, no GPU needed: elapsed time and an FNV-1a checksum of everything it produced
extern "C" int32_t viso_ba_debug_host_build(viso_ba* h, int32_t threads, double* ms_out, uint64_t* checksum_out)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (h->edges_on_device) return fail(h, VISO_BA_ERR_STATE, "the edges of this problem live on the device (window path)");
    if (h->poses.size() < 7 || h->points.size() < 3) return fail(h, VISO_BA_ERR_STATE, "nothing staged");
    const int keep = h->host_threads_override;
    h->host_threads_override = threads > 0 ? threads : 0;
    const double t0 = now_ms();
    Built b;
    int32_t rc = build_tables(h, b);
    if (!rc) {
        b.window = false;
        rc = build_general(h, b);
    }
    const double t1 = now_ms();
    h->host_threads_override = keep;
    if (rc) return rc;
    uint64_t x = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t n) {
        const unsigned char* c = static_cast<const unsigned char*>(p);
        for (size_t i = 0; i < n; ++i) x = (x ^ c[i]) * 1099511628211ull;
    };
    auto mixv = [&](const std::vector<int32_t>& v) { mix(v.data(), v.size() * sizeof(int32_t)); };
    mixv(b.pt_ptr), mixv(b.obs_cam), mix(b.obs_uv.data(), b.obs_uv.size() * sizeof(float)), mixv(b.pairs), mixv(b.items);
    mixv(b.diag_items), mixv(b.block_item_ptr), mixv(b.blk_ij), mixv(b.chunks), mix(&b.n_pairs, sizeof(b.n_pairs));
    if (ms_out) *ms_out = t1 - t0;
    if (checksum_out) *checksum_out = x;
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_reset_state(viso_ba* h)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (!h->uploaded) return viso_ba_upload(h); // upload ends with a reset
    int32_t rc = ensure_cuda(h);
    if (rc) return rc;
    const DevProblem& P = h->P;
    CU(cudaMemcpyAsync(P.pose[0], h->d_pose_init, sizeof(double) * 7 * (size_t)P.n_cams, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaMemcpyAsync(P.pts[0], h->d_pts_init, sizeof(double) * 3 * (size_t)P.n_points, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaMemsetAsync(P.lm, 0, sizeof(LmState) * (size_t)P.n_problems, h->stream)); // cur = 0
    run_k(h, "state_init", [&] { return launch_state_init(P, 0, h->cfg.lm_max_trials, 0.0, h->stream); });
    run_k(h, "cam_rt", [&] { return launch_cam_rt(P, 0, h->stream); });
    if (P.ph_sum) // srcFrame->Project uses the pose the keyframe had when the edges were made (bundle_adjuster.h:79): frozen here
        CU(cudaMemcpyAsync(const_cast<double*>(P.ph_src_rt), P.cam_rt[0], sizeof(double) * 12 * (size_t)P.n_cams, cudaMemcpyDeviceToDevice, h->stream));
    CU(cudaGetLastError());
    return VISO_BA_OK;
}

// ==========================================================================================
// the LM driver: optimizer_.initializeOptimization() + optimize(n), which the reference never
// calls.  All decisions (gain ratio, accept / reject, lambda) are taken on the
// device by k_decide; the host only enqueues trials and polls one counter.
// ==========================================================================================
namespace {

int32_t pin_reserve(viso_ba* h, size_t bytes)
{
    if (bytes <= h->pin_cap) return VISO_BA_OK;
    if (h->pin) cudaFreeHost(h->pin);
    h->pin = nullptr, h->pin_cap = 0;
    CU(cudaHostAlloc(&h->pin, bytes, cudaHostAllocDefault));
    h->pin_cap = bytes;
    return VISO_BA_OK;
}

int32_t allreduce(viso_ba* h, double* buf, size_t count, int op)
{
    if (h->n_ranks <= 1) return VISO_BA_OK;
    NcclApi* api = nccl_api();
    if (!api || !h->comm) return fail(h, VISO_BA_ERR_NCCL, "sharded handle without a communicator");
    int e = 0; // timed like a kernel (name "allreduce") when kernel timing is on; it launches none of OUR kernels
    run_k(h, "allreduce", [&] {
        e = api->AllReduce(buf, buf, count, kNcclDouble, op, h->comm, h->stream);
        return 0;
    });
    NC(e);
    return VISO_BA_OK;
}

// linearise the current state and form the reduced camera system (S_blk, bs); init_pass: only
// what g2o's computeLambdaInit needs (cost, diagonal blocks at lambda = 0)
void enqueue_linearize(viso_ba* h, int init_pass, double lambda_override)
{
    const DevProblem& P = h->P;
    cudaStream_t st = h->stream;
    if (h->window) {
        run_k(h, "win_linearize_schur", [&] { return launch_win_linearize_schur(P, init_pass, lambda_override, st); });
        run_k(h, "win_finalize", [&] { return launch_win_finalize(P, init_pass, lambda_override, st); });
    } else {
        if (P.ph_sum) { // photometric edges: sample the patches, then the photometric per-point pass
            run_k(h, "photo_sample", [&] { return launch_photo_sample(P, 0, 1, lambda_override >= 0.0, st); });
            run_k(h, "photo_linearize", [&] { return launch_photo_linearize(P, init_pass, lambda_override, st); });
        } else {
            run_k(h, "linearize", [&] { return launch_linearize(P, init_pass, lambda_override, st); });
        }
        run_k(h, "schur_blocks", [&] { return launch_schur_blocks(P, init_pass, st); });
        run_k(h, "block_finalize", [&] { return launch_block_finalize(P, init_pass, lambda_override, st); });
    }
}

int32_t enqueue_trial(viso_ba* h)
{
    const DevProblem& P = h->P;
    cudaStream_t st = h->stream;
    int32_t rc;
    if (h->window) { // every problem has <= 32 keyframes: tile kernels, small solve
        enqueue_linearize(h, 0, -1.0);
        run_k(h, "solve_small", [&] { return launch_solve_small(P, h->max_n, st); });
        run_k(h, "win_update", [&] { return launch_win_update(P, st); });
        run_k(h, "decide", [&] { return launch_win_sums_decide(P, st); }); // per-problem totals + the LM decision
        return VISO_BA_OK;
    }
    enqueue_linearize(h, 0, -1.0);
    if ((rc = allreduce(h, h->d_allreduce, h->allreduce_count, kNcclSum))) return rc;
    if (h->use_large) run_k(h, "solve_large", [&] { return launch_solve_large(P, h->L, st); });
    else run_k(h, "solve_small", [&] { return launch_solve_small(P, h->max_n, st); });
    run_k(h, "update", [&] { return launch_update(P, st); });
    if (P.ph_sum) { // the candidate's cost needs the images again
        run_k(h, "photo_sample", [&] { return launch_photo_sample(P, 1, 0, 0, st); });
        run_k(h, "photo_cost", [&] { return launch_photo_cost(P, 1, 0, st); });
    }
    run_k(h, "reduce", [&] { return launch_reduce(P, 1, st); });
    if ((rc = allreduce(h, P.prob_sums, 2 * (size_t)P.n_problems, kNcclSum))) return rc;
    run_k(h, "decide", [&] { return launch_decide(P, st); });
    return VISO_BA_OK;
}

int32_t run_trials(viso_ba* h, int count)
{
    // a compacted grid changes from batch to batch: straggler trials are launched directly
    const bool graph_ok = h->cfg.use_cuda_graph && !h->timing && h->n_ranks == 1 && !h->P.tile_map;
    if (!graph_ok) {
        for (int i = 0; i < count; ++i) {
            int32_t rc = enqueue_trial(h);
            if (rc) return rc;
        }
        CU(cudaGetLastError());
        return VISO_BA_OK;
    }
    if (!h->trial_graph) {
        cudaGraph_t g = nullptr;
        const int64_t before = h->launches;
        CU(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
        int32_t rc = enqueue_trial(h);
        cudaError_t e = cudaStreamEndCapture(h->stream, &g);
        h->trial_graph_launches = (int)(h->launches - before);
        h->launches = before;
        if (rc) return rc;
        if (e != cudaSuccess) return fail(h, VISO_BA_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
        e = cudaGraphInstantiate(&h->trial_graph, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) return fail(h, VISO_BA_ERR_CUDA, "graph instantiate: %s", cudaGetErrorString(e));
    }
    for (int i = 0; i < count; ++i) {
        CU(cudaGraphLaunch(h->trial_graph, h->stream));
        h->launches += h->trial_graph_launches;
    }
    return VISO_BA_OK;
}

int32_t run_lm(viso_ba* h, int iterations, viso_ba_summary* summary)
{
    const DevProblem& P = h->P;
    cudaStream_t st = h->stream;
    int32_t rc;
    if (iterations > kMaxIterStats) return fail(h, VISO_BA_ERR_ARG, "at most %d iterations per call", kMaxIterStats);
    if (iterations < 0) iterations = 0;
    h->P.tile_map = nullptr;
    CU(cudaEventRecord(h->ev0, st));
    run_k(h, "state_init", [&] { return launch_state_init(P, iterations > 0 ? iterations : 1, h->cfg.lm_max_trials, h->cfg.lm_lambda_init, st); });
    // first linearisation: cost of the start state, lambda_0 = tau * max diag(H)   (g2o computeLambdaInit)
    enqueue_linearize(h, 1, -1.0);
    if (h->window) run_k(h, "reduce", [&] { return launch_win_prob_sums(P, 0, st); });
    else run_k(h, "reduce", [&] { return launch_reduce(P, 0, st); });
    if (h->n_ranks > 1) {
        if ((rc = allreduce(h, h->d_allreduce, h->allreduce_count + 2 * (size_t)P.n_problems, kNcclSum))) return rc;
        if ((rc = allreduce(h, reinterpret_cast<double*>(P.maxdiag_bits), (size_t)P.n_problems, kNcclMax))) return rc;
    }
    run_k(h, "lm_init", [&] { return launch_lm_init(P, h->cfg.lm_tau, h->cfg.lm_lambda_init, st); });
    CU(cudaGetLastError());

    if ((rc = pin_reserve(h, 64 + sizeof(LmState) * (size_t)P.n_problems))) return rc;
    int32_t* n_active = static_cast<int32_t*>(h->pin);
    int done = 0;
    const long long max_rounds = (long long)iterations * h->cfg.lm_max_trials;
    int batch = iterations;
    while (iterations > 0 && done < max_rounds) {
        batch = (int)std::min<long long>(batch, max_rounds - done);
        rc = run_trials(h, batch);
        h->P.tile_map = nullptr;
        if (rc) return rc;
        done += batch;
        // window path: the tiles of the problems still active, so that the straggler trials launch only those
        if (h->window) run_k(h, "active_tiles", [&] { return launch_win_active_tiles(P, h->d_tile_map, st); });
        CU(cudaMemcpyAsync(n_active, P.n_active, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (*n_active <= 0) break;
        batch = std::max(1, std::min(iterations, 4)); // stragglers: a few rejected trials to make up
        if (h->window && n_active[1] > 0 && n_active[1] < P.n_tiles) h->P.tile_map = h->d_tile_map, h->P.n_map_tiles = n_active[1];
    }
    h->P.tile_map = nullptr;
    CU(cudaEventRecord(h->ev1, st));
    LmState* lm = reinterpret_cast<LmState*>(static_cast<char*>(h->pin) + 64);
    CU(cudaMemcpyAsync(lm, P.lm, sizeof(LmState) * (size_t)P.n_problems, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    h->lm_host.assign(lm, lm + P.n_problems);
    resolve_timing(h);
    if (summary) {
        std::memset(summary, 0, sizeof(*summary));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->ev0, h->ev1);
        summary->device_ms = ms;
        summary->setup_ms = h->setup_ms;
        for (const LmState& s : h->lm_host) { // batched: totals over the problems
            summary->initial_cost += s.chi_initial;
            summary->final_cost += s.chi;
            summary->iterations = std::max(summary->iterations, s.it);
            summary->trials += s.trials;
            summary->terminated += s.terminated;
        }
        summary->final_lambda = h->lm_host[0].lambda;
    }
    return VISO_BA_OK;
}

// current state -> contiguous device buffers -> host staging vectors
int32_t read_back(viso_ba* h)
{
    const DevProblem& P = h->P;
    run_k(h, "gather_current", [&] { return launch_gather_current(P, h->d_pose_out, h->d_pts_out, h->stream); });
    if (h->n_ranks > 1) { // every rank moved only its own points
        NcclApi* api = nccl_api();
        NC(api->GroupStart());
        int bad = 0; // a failed call inside the group must not leave the group open
        for (int r = 0; r < h->n_ranks && !bad; ++r) {
            const size_t a = 3 * (size_t)h->shard_pt[(size_t)r], n = 3 * (size_t)h->shard_pt[(size_t)r + 1] - a;
            if (n) bad = api->Broadcast(h->d_pts_out + a, h->d_pts_out + a, n, kNcclDouble, r, h->comm, h->stream);
        }
        const int end = api->GroupEnd();
        NC(bad ? bad : end);
    }
    const size_t nb_pose = sizeof(double) * 7 * (size_t)P.n_cams, nb_pts = sizeof(double) * 3 * (size_t)P.n_points;
    int32_t rc = pin_reserve(h, nb_pose + nb_pts + 64 + sizeof(LmState) * (size_t)P.n_problems);
    if (rc) return rc;
    char* pin = static_cast<char*>(h->pin);
    CU(cudaMemcpyAsync(pin, h->d_pose_out, nb_pose, cudaMemcpyDeviceToHost, h->stream));
    const bool direct = h->points.pinned; // page-locked storage: the DMA lands in place
    CU(cudaMemcpyAsync(direct ? reinterpret_cast<char*>(h->points.data()) : pin + nb_pose, h->d_pts_out, nb_pts, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    std::memcpy(h->poses.data(), pin, nb_pose);
    if (!direct) std::memcpy(h->points.data(), pin + nb_pose, nb_pts);
    h->pts_staged = false; // `points` now holds the refined state, the staging buffer the initial one
    return VISO_BA_OK;
}

} // namespace

extern "C" int32_t viso_ba_optimize_resident(viso_ba* h, int32_t iterations, viso_ba_summary* summary)
{
    if (!h) return VISO_BA_ERR_ARG;
    int32_t rc = viso_ba_upload(h);
    if (rc) return rc;
    if ((rc = ensure_cuda(h))) return rc;
    return run_lm(h, iterations, summary);
}

extern "C" int32_t viso_ba_download(viso_ba* h)
{
    if (!h) return VISO_BA_ERR_ARG;
    if (!h->uploaded) return fail(h, VISO_BA_ERR_STATE, "nothing on the device to read back");
    int32_t rc = ensure_cuda(h);
    if (rc) return rc;
    return read_back(h);
}

extern "C" int32_t viso_ba_optimize(viso_ba* h, int32_t iterations, viso_ba_iter_stats* stats, viso_ba_summary* summary)
{
    if (!h) return VISO_BA_ERR_ARG;
    int32_t rc = viso_ba_upload(h);
    if (rc) return rc;
    if ((rc = ensure_cuda(h))) return rc;
    if (stats && h->n_problems != 1) return fail(h, VISO_BA_ERR_ARG, "per-iteration stats exist for single-problem handles only");
    viso_ba_summary local;
    if ((rc = run_lm(h, iterations, &local))) return rc;
    if (summary) *summary = local;
    if (stats && iterations > 0) {
        std::vector<IterStat> tmp((size_t)iterations);
        CU(cudaMemcpyAsync(tmp.data(), h->P.iter_stats, sizeof(IterStat) * (size_t)local.iterations, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        for (int i = 0; i < iterations; ++i) {
            if (i < local.iterations) {
                stats[i].cost_before = tmp[(size_t)i].cost_before, stats[i].cost_after = tmp[(size_t)i].cost_after;
                stats[i].lambda = tmp[(size_t)i].lambda, stats[i].trials = tmp[(size_t)i].trials, stats[i].accepted = tmp[(size_t)i].accepted;
            } else {
                std::memset(&stats[i], 0, sizeof(stats[i]));
            }
        }
    }
    // the host copy now holds the refined state; the device keeps it too (uploaded stays true)
    return read_back(h);
}

extern "C" int32_t viso_ba_get_poses(viso_ba* h, int64_t first, int64_t count, double* out7)
{
    if (!h || !out7 || first < 0 || count < 0) return VISO_BA_ERR_ARG;
    if (first + count > (int64_t)h->poses.size() / 7) return fail(h, VISO_BA_ERR_INDEX, "poses [%lld, %lld) out of range", (long long)first, (long long)(first + count));
    std::memcpy(out7, h->poses.data() + 7 * first, sizeof(double) * 7 * (size_t)count);
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_get_points(viso_ba* h, int64_t first, int64_t count, double* out3)
{
    if (!h || !out3 || first < 0 || count < 0) return VISO_BA_ERR_ARG;
    if (first + count > (int64_t)h->points.size() / 3) return fail(h, VISO_BA_ERR_INDEX, "points [%lld, %lld) out of range", (long long)first, (long long)(first + count));
    std::memcpy(out3, h->points.data() + 3 * first, sizeof(double) * 3 * (size_t)count);
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_get_problem_summaries(viso_ba* h, viso_ba_summary* out)
{
    if (!h || !out) return VISO_BA_ERR_ARG;
    if (h->lm_host.size() != (size_t)h->n_problems) return fail(h, VISO_BA_ERR_STATE, "no optimize call yet");
    for (size_t b = 0; b < h->lm_host.size(); ++b) {
        const LmState& s = h->lm_host[b];
        std::memset(&out[b], 0, sizeof(out[b]));
        out[b].initial_cost = s.chi_initial, out[b].final_cost = s.chi, out[b].final_lambda = s.lambda;
        out[b].iterations = s.it, out[b].trials = s.trials, out[b].terminated = s.terminated;
    }
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_eval_cost(viso_ba* h, double* cost)
{
    if (!h || !cost) return VISO_BA_ERR_ARG;
    int32_t rc = viso_ba_upload(h);
    if (rc) return rc;
    if ((rc = ensure_cuda(h))) return rc;
    const DevProblem& P = h->P;
    if (h->window) { // the init pass of the tile kernel leaves the cost of every tile in tile_sums
        run_k(h, "win_linearize_schur", [&] { return launch_win_linearize_schur(P, 1, 0.0, h->stream); });
        run_k(h, "reduce", [&] { return launch_win_prob_sums(P, 1, h->stream); });
    } else {
        if (P.ph_sum) {
            run_k(h, "photo_sample", [&] { return launch_photo_sample(P, 0, 0, 1, h->stream); });
            run_k(h, "photo_cost", [&] { return launch_photo_cost(P, 0, 1, h->stream); });
        } else {
            run_k(h, "eval_cost", [&] { return launch_eval_cost(P, h->stream); });
        }
        run_k(h, "reduce", [&] { return launch_reduce(P, 2, h->stream); });
    }
    if ((rc = allreduce(h, P.prob_sums, 2 * (size_t)P.n_problems, kNcclSum))) return rc;
    if ((rc = pin_reserve(h, sizeof(double) * 2 * (size_t)P.n_problems))) return rc;
    CU(cudaMemcpyAsync(h->pin, P.prob_sums, sizeof(double) * 2 * (size_t)P.n_problems, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    CU(cudaGetLastError());
    resolve_timing(h);
    double s = 0.0;
    for (int b = 0; b < P.n_problems; ++b) s += static_cast<double*>(h->pin)[2 * b];
    *cost = s;
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_debug_reduced_system(viso_ba* h, double lambda, int32_t n_cap, double* S, double* bs, int32_t* n_out)
{
    if (!h || !n_out) return VISO_BA_ERR_ARG;
    int32_t rc = viso_ba_upload(h);
    if (rc) return rc;
    if ((rc = ensure_cuda(h))) return rc;
    if (h->n_problems != 1) return fail(h, VISO_BA_ERR_UNSUPPORTED, "debug hook works on single-problem handles");
    const DevProblem& P = h->P;
    const int n = h->max_n, nf = n / 6;
    *n_out = n;
    if (!S && !bs) return VISO_BA_OK;
    if (n_cap < n) return fail(h, VISO_BA_ERR_CAPACITY, "reduced system is %d x %d", n, n);
    cudaStream_t st = h->stream;
    run_k(h, "state_init", [&] { return launch_state_init(P, 1, h->cfg.lm_max_trials, lambda, st); });
    enqueue_linearize(h, 0, lambda);
    if ((rc = allreduce(h, h->d_allreduce, h->allreduce_count, kNcclSum))) return rc;
    run_k(h, "state_init", [&] { return launch_state_init(P, 0, h->cfg.lm_max_trials, 0.0, st); });
    std::vector<double> blk((size_t)P.n_blocks * 36), b6((size_t)P.n_cams * 6);
    std::vector<int32_t> cam_free((size_t)P.n_cams);
    CU(cudaMemcpyAsync(blk.data(), P.S_blk, sizeof(double) * blk.size(), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(b6.data(), P.bs, sizeof(double) * b6.size(), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(cam_free.data(), P.cam_free, sizeof(int32_t) * cam_free.size(), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    CU(cudaGetLastError());
    resolve_timing(h);
    if (S) {
        const bool compact = !h->blk_ij_host.empty();
        if (compact) std::fill(S, S + (size_t)n * n, 0.0);
        const int64_t nblk = compact ? (int64_t)h->blk_ij_host.size() / 2 : (int64_t)nf * (nf + 1) / 2;
        int I = 0, J = -1;
        for (int64_t k = 0; k < nblk; ++k) {
            if (compact) I = h->blk_ij_host[2 * (size_t)k], J = h->blk_ij_host[2 * (size_t)k + 1];
            else if (++J > I) ++I, J = 0;
            const double* x = &blk[36 * (size_t)k];
            for (int r = 0; r < 6; ++r)
                for (int c = 0; c < 6; ++c) {
                    S[(size_t)(6 * I + r) * n + 6 * J + c] = x[6 * r + c];
                    S[(size_t)(6 * J + c) * n + 6 * I + r] = x[6 * r + c];
                }
        }
    }
    if (bs)
        for (int c = 0; c < P.n_cams; ++c)
            if (cam_free[(size_t)c] >= 0)
                for (int r = 0; r < 6; ++r) bs[6 * cam_free[(size_t)c] + r] = b6[6 * (size_t)c + r];
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_set_kernel_timing(viso_ba* h, int32_t on)
{
    if (!h) return VISO_BA_ERR_ARG;
    h->timing = on != 0;
    h->ktimes.clear();
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_get_kernel_times(viso_ba* h, int32_t cap, viso_ba_kernel_time* out, int32_t* n_out)
{
    if (!h || !n_out) return VISO_BA_ERR_ARG;
    resolve_timing(h);
    *n_out = (int32_t)h->ktimes.size();
    if (!out) return VISO_BA_OK;
    int i = 0;
    for (const auto& kv : h->ktimes) {
        if (i >= cap) break;
        std::memset(&out[i], 0, sizeof(out[i]));
        std::strncpy(out[i].name, kv.first.c_str(), sizeof(out[i].name) - 1);
        out[i].launches = kv.second.launches;
        out[i].device_ms = kv.second.ms;
        ++i;
    }
    return VISO_BA_OK;
}

// ==========================================================================================
// sharding over the GPUs of one box
// ==========================================================================================
extern "C" int32_t viso_ba_nccl_unique_id(void* out128)
{
    if (!out128) return VISO_BA_ERR_ARG;
    NcclApi* api = nccl_api();
    if (!api) return VISO_BA_ERR_NCCL;
    return api->GetUniqueId(out128) == 0 ? VISO_BA_OK : VISO_BA_ERR_NCCL;
}

extern "C" int32_t viso_ba_shard_point_ranges(int64_t n_points, int64_t n_edges, const int32_t* edge_point,
    int32_t n_ranks, int64_t* out)
{
    if (n_points < 0 || n_edges < 0 || n_ranks < 1 || !out || (n_edges && !edge_point)) return VISO_BA_ERR_ARG;
    std::vector<int64_t> ptr((size_t)n_points + 1, 0);
    for (int64_t e = 0; e < n_edges; ++e) {
        if (edge_point[e] < 0 || edge_point[e] >= n_points) return VISO_BA_ERR_INDEX;
        ptr[(size_t)edge_point[e] + 1]++;
    }
    for (int64_t p = 0; p < n_points; ++p) ptr[(size_t)p + 1] += ptr[(size_t)p];
    shard_bounds(ptr.data(), n_points, n_edges, n_ranks, out);
    return VISO_BA_OK;
}

extern "C" int32_t viso_ba_shard_init(viso_ba* h, int32_t rank, int32_t n_ranks, const void* id128)
{
    if (!h || n_ranks < 1 || rank < 0 || rank >= n_ranks) return VISO_BA_ERR_ARG;
    if (n_ranks == 1) {
        h->rank = 0, h->n_ranks = 1;
        return VISO_BA_OK;
    }
    if (!id128) return VISO_BA_ERR_ARG;
    int32_t rc = ensure_cuda(h);
    if (rc) return rc;
    NcclApi* api = nccl_api();
    if (!api) return fail(h, VISO_BA_ERR_NCCL, "libnccl.so.2 not found: %s", dlerror());
    if (h->comm) {
        api->CommDestroy(h->comm);
        h->comm = nullptr;
    }
    Id128 id;
    std::memcpy(id.b, id128, 128);
    NC(api->CommInitRank(&h->comm, n_ranks, id, rank));
    h->rank = rank, h->n_ranks = n_ranks;
    h->uploaded = false;
    drop_graph(h);
    return VISO_BA_OK;
}
