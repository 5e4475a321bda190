// ba_engine.cu — the CUDA engine behind include/ba_engine.h: host driver of the Levenberg-Marquardt loop.
//
// Replaces ceres::Solve at src/adjuster/BundleAdjuster.h:126 (trust-region LM as configured at :167-174; rules restated
// in SURVEY.md §8 a-LM) with hand-written sm_100a kernels (ba_kernels.cuh).  Map state lives on the device as FP64
// SoA buffers for the whole solve; the host only sequences kernels and reads one small struct of scalars per decision.
// There is NO CPU fallback: without a CUDA device every entry point that needs one returns BA_ERR_CUDA.
//
// Multi-GPU (one process per GPU): points and their observations are sharded by contiguous range; poses, intrinsics
// and all PCG vectors are replicated; per-rank partial sums of camera-side quantities are combined with NCCL
// all-reduces on the compute stream.  NCCL is loaded with dlopen only when world_size > 1.
#include "../../include/ba_engine.h"
#include "../../include/ba_merge.hpp"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <exception>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ba_kernels.cuh"
#include "ba_rows.cuh"
#ifdef BA_WITH_STREAM          // the persistent cp.async.bulk streaming kernel (scatter_mode 6) is an experiment kept out of the default build
#include "ba_stream.cuh"
#endif
#include "ba_p2p.cuh"
#include "ba_dense.cuh"
#include "ba_pcg_cluster.cuh"

namespace {
using std::vector;
thread_local std::string g_create_error;

#define CU_TRY(expr)                                                                                         \
    do {                                                                                                     \
        cudaError_t _e = (expr);                                                                             \
        if (_e != cudaSuccess) {                                                                             \
            char _b[512];                                                                                    \
            std::snprintf(_b, sizeof _b, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
            throw BaFail{BA_ERR_CUDA, _b};                                                                   \
        }                                                                                                    \
    } while (0)

struct BaFail { int code; std::string msg; };
// Every kernel launch is counted (ba_summary.kernel_launches) and every launch site first checks that no earlier launch on this host
// thread failed (a <<<>>> launch reports configuration errors, a missing sm_100a image or a missing attribute only through
// cudaGetLastError): a failed launch can never be skipped silently.  The last launch of a sequence is checked by read_scalars() / finish().
inline void launch_check(const char* file, int line) {
    const cudaError_t x = cudaPeekAtLastError();
    if (x != cudaSuccess) {
        cudaGetLastError();
        char b[256];
        std::snprintf(b, sizeof b, "a kernel launch before %s:%d failed: %s", file, line, cudaGetErrorString(x));
        throw BaFail{BA_ERR_CUDA, b};
    }
}
#define KL(eng) (launch_check(__FILE__, __LINE__), ++(eng)->kernel_launches)
#ifndef BA_L2_PERSIST_DEFAULT
#define BA_L2_PERSIST_DEFAULT 0
#endif
#ifndef BA_JPACK_DEFAULT
#define BA_JPACK_DEFAULT 1
#endif
#ifndef BA_PDL_DEFAULT
#define BA_PDL_DEFAULT 1
#endif

// ---- minimal NCCL binding, resolved at run time (torch bundles its own libnccl.so.2; whichever is already loaded wins) ----
struct Nccl {
    typedef struct { char internal[128]; } UniqueId;
    typedef void* Comm;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(Comm*, int, UniqueId, int) = nullptr;
    int (*CommDestroy)(Comm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, Comm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    void* handle = nullptr;
    enum { kInt8 = 0, kInt32 = 2, kFloat64 = 8, kSum = 0, kMax = 2 };   // ncclDataType_t / ncclRedOp_t values (stable since NCCL 2.0)
    bool load(std::string& err) {
        if (handle) return true;
        for (const char* n : {"libnccl.so.2", "libnccl.so"}) { handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (handle) break; }
        if (!handle) { err = std::string("dlopen libnccl.so.2 failed: ") + dlerror(); return false; }
        auto sym = [&](const char* n) { void* p = dlsym(handle, n); if (!p) err = std::string("missing NCCL symbol ") + n; return p; };
        GetUniqueId = reinterpret_cast<decltype(GetUniqueId)>(sym("ncclGetUniqueId"));
        CommInitRank = reinterpret_cast<decltype(CommInitRank)>(sym("ncclCommInitRank"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
        AllReduce = reinterpret_cast<decltype(AllReduce)>(sym("ncclAllReduce"));
        Broadcast = reinterpret_cast<decltype(Broadcast)>(sym("ncclBroadcast"));
        AllGather = reinterpret_cast<decltype(AllGather)>(sym("ncclAllGather"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
        return GetUniqueId && CommInitRank && CommDestroy && AllReduce && Broadcast && AllGather && GroupStart && GroupEnd && GetErrorString;
    }
};
Nccl g_nccl;

#define NCCL_TRY(expr)                                                                                       \
    do {                                                                                                     \
        int _r = (expr);                                                                                     \
        if (_r != 0) {                                                                                       \
            char _b[512];                                                                                    \
            std::snprintf(_b, sizeof _b, "%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, g_nccl.GetErrorString(_r)); \
            throw BaFail{BA_ERR_NCCL, _b};                                                                   \
        }                                                                                                    \
    } while (0)

int64_t g_h2d_bytes = 0, g_d2h_bytes = 0;   // cumulative host<->device traffic of this library (ba_get_io_bytes)

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    void alloc(size_t count) {
        if (count <= n && p) return;
        release();
        CU_TRY(cudaMalloc(reinterpret_cast<void**>(&p), std::max<size_t>(count, 1) * sizeof(T)));
        n = count;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    void upload(const T* h, size_t count, cudaStream_t s) { alloc(count); if (count) { CU_TRY(cudaMemcpyAsync(p, h, count * sizeof(T), cudaMemcpyHostToDevice, s)); g_h2d_bytes += int64_t(count * sizeof(T)); } }
    void zero(cudaStream_t s) { if (p && n) CU_TRY(cudaMemsetAsync(p, 0, n * sizeof(T), s)); }
    ~DevBuf() { release(); }
};

// Page-locked std::vector storage for the host arrays ba_set_problem builds and uploads: kept in the engine between calls, so a call
// neither page-faults its way through 150 MB of fresh allocations nor stages its uploads through the driver's bounce buffer.
template <class T>
struct PinnedAlloc {
    using value_type = T;
    PinnedAlloc() = default;
    template <class U> PinnedAlloc(const PinnedAlloc<U>&) {}
    T* allocate(size_t n) { void* p = nullptr; if (cudaMallocHost(&p, std::max<size_t>(n, 1) * sizeof(T)) != cudaSuccess) throw std::bad_alloc(); return static_cast<T*>(p); }
    void deallocate(T* p, size_t) { cudaFreeHost(p); }
    template <class U> bool operator==(const PinnedAlloc<U>&) const { return true; }
    template <class U> bool operator!=(const PinnedAlloc<U>&) const { return false; }
};
template <class T> using pvec = std::vector<T, PinnedAlloc<T>>;

struct HostScratch {   // ba_set_problem's work arrays (capacity survives between calls)
    pvec<int4> tmeta;
    pvec<int32_t> perm, obs_pt, crows;
    std::vector<unsigned char> slot, lvl1, hrank, hpos;
    std::vector<int32_t> spack, row_cam, hbase, tpts;
};

enum KernelId { KID_K0 = 0, KID_K1, KID_K1C, KID_K2C, KID_K3P, KID_K3C, KID_K4, KID_K5, KID_K7, KID_K8, KID_PCGV, KID_MISC, KID_DFORM, KID_DCHOL, KID_GATHER, KID_TAIL,
                KID_COUNT };
const char* const kKernelNames[KID_COUNT] = {"K0_camera_precompute", "K1_residual_jacobian", "K1c_candidate_cost", "K2c_camera_blocks",
                                             "K3p_point_blocks",      "K3c_camera_blocks",    "K4_reduced_rhs",     "K5_schur_matvec",
                                             "K7_backsub_model",      "K8_candidate",         "PCG_vector_update",  "misc_reductions",
                                             "D1_dense_schur_form",   "D2_dense_cholesky_solve", "K5b_camera_gather",  "PCG_cluster_tail"};

}  // namespace

// ------------------------------------------------------------------------------------------------------------
struct ba_engine {
    ba_options opt{};
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool has_problem = false;
    std::string err;
    // global sizes and this rank's shard
    int Nc = 0, Ng = 0, NpG = 0, NoG = 0;   // global
    int pb = 0, pe = 0, ob = 0, oe = 0;     // local point / observation ranges [pb,pe) [ob,oe)
    int Np = 0, No = 0, ldo = 0, nTiles = 0, nf = 0, fi = 0;
    // host copies kept for ba_get_structure and ba_reset_state
    vector<int32_t> h_obs_cam, h_obs_pt;
    vector<int32_t> h_cam_group;
    vector<double> h_obs_uv;           // host mirror of the observation pixels (ba_append merges into it)
    double last_radius = 0.0;          // final trust-region radius of the previous solve (warm_start_radius)
    vector<uint8_t> h_cam_fixed, h_cam_free;
    vector<int32_t> h_shard_pt;   // [world+1] point range boundaries of every rank
    // device: static problem data
    DevBuf<int> d_tile_off, d_obs_cam, d_obs_pt, d_pt_off, d_cam_group, d_perm, d_cam_off;
    DevBuf<int> d_pt_by_cam;          // local point of the i-th observation in camera order (pose-factored K2c)
    bool lean = false;                // K1 writes neither the pose columns of J2 nor the by-observation records: K2c/K4/K5/K7 all run pose-factored (PCG, block-Jacobi)
    DevBuf<double2> d_obs_uv;
    DevBuf<unsigned char> d_cam_free, d_grp_free;
    // device: parameters (current, candidate, initial)
    DevBuf<double> d_pose, d_intr, d_pt, d_cpose, d_cintr, d_cpt, d_pose0, d_intr0, d_pt0, d_ptfull;
    // device: work arrays
    DevBuf<double> d_camrec, d_ccamrec;
    DevBuf<double2> d_J2, d_r2, d_JcA;
    int k1_grid = 1;
    DevBuf<double> d_Hpp, d_gp, d_Hinv, d_vp, d_yp, d_sc_pt;
    DevBuf<double> d_G, d_vt;         // pose-factored matvec: per-camera 6x6 factors, vt = [G_c p_c | intrinsics entries of p]
    DevBuf<double> d_red;        // [Hcc 36Nc | gc 6Nc | Hii 16Ng | gi 4Ng] — one all-reduce per evaluation
    DevBuf<double> d_sc_f, d_Df2, d_Minv, d_b, d_x, d_r, d_z, d_p, d_y;
    DevBuf<double> d_sdiag;           // Schur-Jacobi: per-pose correction blocks (k_schur_diag)
    DevBuf<double> d_part0, d_part1;   // per-tile / per-block partials
    // direct solve of the reduced camera system (ba_dense.cuh); dense == true selects it for this problem
    DevBuf<double> d_ipart;            // [Nc][14] per-camera shares of the group blocks (free intrinsics)
    DevBuf<int> d_grp_off, d_grp_cams;  // cameras of every group, CSR
    bool dense = false;
    DevBuf<double> d_S, d_WU, d_GH, d_corner;
    DevBuf<int> d_blk_off;
    DevBuf<int2> d_blk_key, d_pairs;
    DevBuf<int4> d_items;
    DevBuf<int> d_blk_item_off, d_blk_ticket;
    DevBuf<double> d_item_part, d_Zinv;
    DevBuf<int2> d_chol_tasks;
    DevBuf<int> d_chol_ready;           // tile flags + abort word (last)
    bak::CholFlow CF{};
    int chol_mode = 1;                  // 1: dataflow (k_dense_chol_df), 0: barrier per phase (k_dense_chol); BA_CHOL_MODE overrides
    bak::DenseView DV{};
    int chol_grid = 1;
    // per-(tile, camera) partial rows + by-camera gather (ba_rows.cuh, ba_stream.cuh)
    DevBuf<int4> d_tmeta, d_tile_info;
    HostScratch hs;
    DevBuf<int> d_cam_row_off, d_cam_rows;
    DevBuf<double> d_rows, d_tile_intr;
    bak::RowSched RS{};
    int scatter_mode = 6;              // 1 = red.global atomics, 4 = partial rows, one CTA per tile (default), 6 = partial rows, persistent streaming kernel (opt-in)
    int scatter_eff = 4;               // what scatter_mode resolves to for the current problem (default 4)
    bool stream_attr_set = false;      // max dynamic shared memory of the k_schur_stream instantiations (per engine: per device)
    int num_sms = 148;
    DevBuf<bak::Scalars> d_scal;
    bak::Scalars* h_scal = nullptr;    // pinned
    bak::View V{};
    // multi-GPU
    Nccl::Comm comm = nullptr;
    struct P2P {   // one-shot peer-memory all-reduce of the PCG matvec (ba_p2p.cuh)
        bool on = false;
        size_t cap = 0;                       // doubles in the local receive buffer
        double* local = nullptr;
        unsigned int* lflags = nullptr;
            void* opened[2 * bak::P2P_MAX_RANKS] = {};   // IPC mappings to close
        int nopened = 0;
        bak::PeerView view{};
        unsigned int seq = 0;
    } p2p;
    // profiling
    int64_t kernel_launches = 0;
    int64_t host_syncs = 0;           // Scalars read-backs (stream synchronisations) of the current / last ba_solve
    unsigned profiling = 0;           // bit k set: time kernel id k with CUDA events
    struct Stat { int64_t launches = 0; int64_t timed = 0; double ms = 0.0; double bytes = 0.0; } stats[KID_COUNT];
    int dbg_tail1 = 0;                // BA_TAIL_MODE1
    int tail_wide_min_nf = 32768;     // BA_TAIL_WIDE_MIN_NF: reduced unknowns from which the 16-CTA cluster is used
    int tail_cl = 8;                  // CTAs of the PCG tail's cluster for reduced systems of >= 32768 unknowns: 16 where the device can place such a cluster (BA_TAIL_CL=8 forces 8)
    int factored_opt = 1;             // BA_POSE_FACTORED (default 1): K5a streams Jp only (ba_rows.cuh: k_pose_factors); 0 = explicit Jc and Jp (bisecting / A-B)
    int factored = 0;                 // in effect for the current problem (partial rows, one CTA per tile, PCG)
    int p2p_wide = -1;                // peer-memory all-reduce: pull on many SMs in a kernel of its own (1) or inside the PCG tail's cluster (0); -1 (default): the
                                      // kernel of its own from 4 ranks and 32768 reduced unknowns on.  Measured at 8 GPUs (profiles/r2i_pull_ab_n8.txt): Final-13682
                                      // (82 092 unknowns) 274.7 -> 259.6 us per PCG iteration, Venice (10 668) 60.8 -> 64.2 us (the extra launch costs more than the
                                      // wider pull saves); at 2 GPUs no difference.  BA_P2P_WIDE overrides.
    int pcg_tag = -1;                 // ordinal of the PCG iteration being launched (LaunchTimer tags; -1 outside the PCG loop)
    int pcg_pred = 0;                 // PCG iterations of the previous linear solve on this handle (first batch size)
    int k5_sample_stride = 1;         // time every n-th K5 launch only (events between two launches serialise them; BA_K5_SAMPLE)
    int pdl = 0;                      // programmatic dependent launch along the PCG chain (BA_PDL)
    int jpack = 0;                    // packed translation columns of J (pinhole model)
    bool l2_window_ever = false;
    bool l2_window = false;           // a persisting-L2 access-policy window for the partial rows is set on the stream
    struct Pending { int kid; cudaEvent_t a, b; int tag; };   // tag: PCG iteration ordinal within its linear solve (-1: not a PCG-loop launch)
    vector<Pending> pending;
    vector<cudaEvent_t> ev_pool;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
    // results
    vector<ba_iter_record> traj;
    ba_summary summary{};
    bool scales_set = false;

    double* Hcc() { return d_red.p; }
    double* gc() { return d_red.p + size_t(36) * Nc; }
    double* Hii() { return d_red.p + size_t(42) * Nc; }
    double* gi() { return d_red.p + size_t(42) * Nc + size_t(16) * Ng; }
    size_t red_count() const { return size_t(42) * Nc + size_t(20) * Ng; }
};

namespace {

int fail(ba_engine* e, int code, const std::string& msg) { if (e) e->err = msg; else g_create_error = msg; return code; }

// Profiling hook around one launch: CUDA events on the engine's stream, resolved after the solve's final sync.
struct LaunchTimer {
    ba_engine* e; int kid; cudaEvent_t a = nullptr, b = nullptr; int tag;
    LaunchTimer(ba_engine* e_, int kid_, bool sample = true, int tag_ = -1) : e(e_), kid(kid_), tag(tag_) {
        ++e->stats[kid].launches;
        if (!((e->profiling >> kid) & 1u) || !sample) return;
        ++e->stats[kid].timed;
        auto get = [&]() { cudaEvent_t ev; if (!e->ev_pool.empty()) { ev = e->ev_pool.back(); e->ev_pool.pop_back(); } else cudaEventCreate(&ev); return ev; };
        a = get(); b = get();
        cudaEventRecord(a, e->stream);
    }
    ~LaunchTimer() { if (a) { cudaEventRecord(b, e->stream); e->pending.push_back({kid, a, b, tag}); } }
};

void resolve_timers(ba_engine* e) {
    for (auto& p : e->pending) {
        float ms = 0.f;
        if (p.kid >= 0 && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) e->stats[p.kid].ms += ms;   // kid < 0: a voided no-op launch
        e->ev_pool.push_back(p.a); e->ev_pool.push_back(p.b);
    }
    e->pending.clear();
}

void allreduce_sum(ba_engine* e, double* buf, size_t n) {
    if (e->opt.world_size <= 1) return;
    NCCL_TRY(g_nccl.AllReduce(buf, buf, n, Nccl::kFloat64, Nccl::kSum, e->comm, e->stream));
}
void allreduce_max(ba_engine* e, double* buf, size_t n) {
    if (e->opt.world_size <= 1) return;
    NCCL_TRY(g_nccl.AllReduce(buf, buf, n, Nccl::kFloat64, Nccl::kMax, e->comm, e->stream));
}
void allreduce_flags(ba_engine* e, int* buf, size_t n) {
    if (e->opt.world_size <= 1) return;
    NCCL_TRY(g_nccl.AllReduce(buf, buf, n, Nccl::kInt32, Nccl::kMax, e->comm, e->stream));
}

// Cut the point-sorted observation list [0,No) into tiles of whole points with <= TILE_OBS observations; a point with
// more observations than that becomes a tile of its own (the kernels walk it in chunks).
// Host threads over [0, n) in contiguous ranges (the per-tile schedule construction of ba_set_problem; ranges write disjoint slices).
int host_threads() {
    int nt = int(std::thread::hardware_concurrency());
    if (const char* v = std::getenv("BA_HOST_THREADS")) nt = std::atoi(v);
    return std::max(1, std::min(nt, 16));
}
// f(k, a, b) on nt host threads: chunk k of nt covers [a, b) of [0, n)
template <class F>
void parallel_chunks(int n, int nt, F&& f) {
    if (nt <= 1) { f(0, 0, n); return; }
    std::vector<std::thread> th;
    std::exception_ptr err;
    std::mutex mu;
    for (int k = 0; k < nt; ++k) {
        const int a = int(int64_t(n) * k / nt), b = int(int64_t(n) * (k + 1) / nt);
        th.emplace_back([&, k, a, b] { try { f(k, a, b); } catch (...) { std::lock_guard<std::mutex> g(mu); err = std::current_exception(); } });
    }
    for (auto& t : th) t.join();
    if (err) std::rethrow_exception(err);
}

template <class F>
void parallel_ranges(int n, F&& f) {
    int nt = int(std::thread::hardware_concurrency());
    if (const char* v = std::getenv("BA_HOST_THREADS")) nt = std::atoi(v);
    nt = std::max(1, std::min(nt, 16));
    if (n < 2048 || nt == 1) { f(0, n); return; }
    std::vector<std::thread> th;
    std::exception_ptr err;
    std::mutex mu;
    for (int k = 0; k < nt; ++k) {
        const int a = int(int64_t(n) * k / nt), b = int(int64_t(n) * (k + 1) / nt);
        th.emplace_back([&, a, b] { try { f(a, b); } catch (...) { std::lock_guard<std::mutex> g(mu); err = std::current_exception(); } });
    }
    for (auto& t : th) t.join();
    if (err) std::rethrow_exception(err);
}

// A tile also holds at most TILE_MAX_PTS points, consecutive ids all of them observed (the streaming kernel copies the Hinv blocks of a
// tile's points as one piece, and point j of the tile is handled by thread j < observations of the tile).
void build_tiles(const vector<int32_t>& pt_off, int Np, vector<int32_t>& tile_off) {
    tile_off.clear();
    tile_off.push_back(0);
    int start = 0;      // first observation of the open tile
    int pstart = -1;    // first point of the open tile (-1: the tile is empty)
    for (int p = 0; p < Np; ++p) {
        const int a = pt_off[p], b = pt_off[p + 1];
        if (b == a) {   // an unobserved point closes the open tile: a tile's points are consecutive ids, one point thread each
            if (pstart >= 0) { tile_off.push_back(a); start = a; pstart = -1; }
            continue;
        }
        if (pstart >= 0 && (b - start > bak::TILE_OBS || p - pstart + 1 > bak::TILE_MAX_PTS)) {   // point p does not fit into the open tile
            tile_off.push_back(a); start = a; pstart = -1;
        }
        if (b - a > bak::TILE_OBS) { tile_off.push_back(b); start = b; pstart = -1; continue; }     // long point: a tile of its own
        if (pstart < 0) pstart = p;
    }
    if (tile_off.back() != pt_off[Np]) tile_off.push_back(pt_off[Np]);
}

// Two phases (CUDA IPC: freeing an exported allocation while a peer still has it mapped is undefined): every rank first closes the
// mappings it imported, then all ranks meet in a collective (one-int NCCL all-reduce + stream sync), and only then does each rank free
// the buffers it exported.  `collective` = every rank is known to be in this call (setup_p2p's reallocation and failure paths,
// ba_destroy of a healthy handle); a handle that is destroyed after an error skips the rendezvous rather than risk a hang.
void teardown_p2p(ba_engine* e, bool collective) {
    for (int i = 0; i < e->p2p.nopened; ++i) cudaIpcCloseMemHandle(e->p2p.opened[i]);
    e->p2p.nopened = 0;
    if (collective && e->comm && e->p2p.local && e->stream) {
        int* d = nullptr;
        if (cudaMalloc(reinterpret_cast<void**>(&d), sizeof(int)) == cudaSuccess) {
            cudaMemsetAsync(d, 0, sizeof(int), e->stream);
            g_nccl.AllReduce(d, d, 1, Nccl::kInt32, Nccl::kMax, e->comm, e->stream);
            cudaStreamSynchronize(e->stream);
            cudaFree(d);
        }
    }
    if (e->p2p.local) cudaFree(e->p2p.local);
    if (e->p2p.lflags) cudaFree(e->p2p.lflags);
    e->p2p = ba_engine::P2P{};
}

// Map every rank's receive buffer and flag array into this process (CUDA IPC; handles travel through one ncclAllGather).
// Collective: every rank calls it with the same nf.  On any failure anywhere all ranks fall back to the NCCL all-reduce
// (or fail, when p2p_allreduce = 2).
void setup_p2p(ba_engine* e) {
    const int W = e->opt.world_size, R = e->opt.rank;
    if (W <= 1 || e->opt.p2p_allreduce == 1) return;
    if (W > bak::P2P_MAX_RANKS) { if (e->opt.p2p_allreduce == 2) throw BaFail{BA_ERR_UNSUPPORTED, "peer-memory all-reduce supports at most 8 ranks"}; return; }
    const int ldf = (e->nf + 15) / 16 * 16;
    const size_t need = size_t(2) * W * ldf;
    cudaStream_t s = e->stream;
    if (e->p2p.on && e->p2p.cap >= need) { e->p2p.view.ldf = ldf; return; }
    teardown_p2p(e, true);
    ba_engine::P2P& Q = e->p2p;
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&Q.local), need * sizeof(double)));
    CU_TRY(cudaMalloc(reinterpret_cast<void**>(&Q.lflags), 64 * sizeof(unsigned int)));
    CU_TRY(cudaMemsetAsync(Q.local, 0, need * sizeof(double), s));
    CU_TRY(cudaMemsetAsync(Q.lflags, 0, 64 * sizeof(unsigned int), s));
    Q.cap = need;
    struct Rec { cudaIpcMemHandle_t data, flags; };
    static_assert(sizeof(Rec) == 128, "two 64-byte IPC handles");
    vector<Rec> all(W);
    int fail_local = 0;
    if (cudaIpcGetMemHandle(&all[R].data, Q.local) != cudaSuccess || cudaIpcGetMemHandle(&all[R].flags, Q.lflags) != cudaSuccess) { fail_local = 1; cudaGetLastError(); }
    DevBuf<char> d_all; d_all.alloc(size_t(W) * sizeof(Rec));
    CU_TRY(cudaMemcpyAsync(d_all.p + size_t(R) * sizeof(Rec), &all[R], sizeof(Rec), cudaMemcpyHostToDevice, s));
    NCCL_TRY(g_nccl.AllGather(d_all.p + size_t(R) * sizeof(Rec), d_all.p, sizeof(Rec), Nccl::kInt8, e->comm, s));
    CU_TRY(cudaMemcpyAsync(all.data(), d_all.p, size_t(W) * sizeof(Rec), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaStreamSynchronize(s));
    bak::PeerView& V = Q.view;
    V = bak::PeerView{};
    V.W = W; V.rank = R; V.ldf = ldf; V.timeout_flag = &e->d_scal.p->bad_lin;
    V.data[R] = Q.local; V.flags[R] = Q.lflags;
    for (int r = 0; r < W && !fail_local; ++r) {
        if (r == R) continue;
        void *pd = nullptr, *pf = nullptr;
        if (cudaIpcOpenMemHandle(&pd, all[r].data, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { fail_local = 1; cudaGetLastError(); break; }
        Q.opened[Q.nopened++] = pd;
        if (cudaIpcOpenMemHandle(&pf, all[r].flags, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { fail_local = 1; cudaGetLastError(); break; }
        Q.opened[Q.nopened++] = pf;
        V.data[r] = static_cast<double*>(pd); V.flags[r] = static_cast<unsigned int*>(pf);
    }
    // agree: one failure anywhere disables the path everywhere
    DevBuf<int> d_fail; d_fail.alloc(1);
    CU_TRY(cudaMemcpyAsync(d_fail.p, &fail_local, sizeof(int), cudaMemcpyHostToDevice, s));
    NCCL_TRY(g_nccl.AllReduce(d_fail.p, d_fail.p, 1, Nccl::kInt32, Nccl::kMax, e->comm, s));
    int fail_any = 0;
    CU_TRY(cudaMemcpyAsync(&fail_any, d_fail.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    CU_TRY(cudaStreamSynchronize(s));
    if (fail_any) {
        teardown_p2p(e, true);
        if (e->opt.p2p_allreduce == 2) throw BaFail{BA_ERR_NCCL, "peer-memory all-reduce required (p2p_allreduce = 2) but CUDA IPC mapping of a peer buffer failed"};
        return;
    }
    Q.on = true;
    Q.seq = 0;
}

// Point ranges of the W shards from the global CSR-by-point offsets: rank r starts at the first point whose first
// observation is >= r*No/W (ba_shard_points).
void shard_points(const vector<int32_t>& gpt_off, int NpG, int NoG, int W, vector<int32_t>& shard_pt) {
    shard_pt.assign(size_t(W) + 1, 0);
    shard_pt[W] = NpG;
    for (int r = 1; r < W; ++r) {
        const int64_t target = int64_t(NoG) * r / W;
        const int p = int(std::lower_bound(gpt_off.begin(), gpt_off.end(), int32_t(target)) - gpt_off.begin());
        shard_pt[r] = std::min(std::max(p, shard_pt[r - 1]), NpG);
    }
}

constexpr int kDenseAutoMaxDim = 2048;                  // BA_LINEAR_AUTO: direct solve up to this many reduced unknowns (341 poses)
constexpr int kDenseMaxDim = 8192;                      // BA_LINEAR_EXACT refuses beyond (S would exceed 512 MB)
constexpr int64_t kDenseMaxPairs = int64_t(1) << 27;    // co-observation pair records (8 B each)

// Direct-solve schedule (ba_dense.cuh): every ordered pair (i, j) of observations of one point with cam(i) >= cam(j), grouped by
// (cam(i), cam(j)) with two stable counting sorts, so that block (ci, cj) of S is the sum over its list in point order.
void build_dense(ba_engine* e, const ba_problem* P, const vector<int32_t>& pt_off) {
    e->dense = false;
    const int mode = e->opt.linear_solver;
    if (mode == BA_LINEAR_PCG) return;
    const bool force = mode == BA_LINEAR_EXACT;
    const int Nc = e->Nc, Np = e->Np, No = e->No;
    const char* why = nullptr;
    if (e->opt.world_size > 1) why = "the direct solve runs on one GPU (world_size = 1)";
    else if (e->nf > (force ? kDenseMaxDim : kDenseAutoMaxDim)) why = "too many reduced unknowns for a dense Schur complement";
    else if (e->fi && e->Ng != 1) why = "free intrinsics with more than one Camera group";
    int64_t bound = 0;
    if (!why) {
        for (int p = 0; p < Np; ++p) { const int64_t k = pt_off[p + 1] - pt_off[p]; bound += k * k; }
        if (bound > kDenseMaxPairs) why = "too many co-observation pairs";
    }
    if (why) {
        if (force) throw BaFail{BA_ERR_UNSUPPORTED, std::string("linear_solver = BA_LINEAR_EXACT: ") + why + " (use BA_LINEAR_PCG or BA_LINEAR_AUTO)"};
        return;
    }
    cudaStream_t s = e->stream;
    const int32_t* cam = P->obs_cam + e->ob;
    vector<int2> A, B;
    A.reserve(size_t(bound / 2 + No));
    for (int p = 0; p < Np; ++p)
        for (int i = pt_off[p]; i < pt_off[p + 1]; ++i)
            for (int j = pt_off[p]; j < pt_off[p + 1]; ++j)
                if (cam[i] >= cam[j]) A.push_back(make_int2(i, j));
    const size_t np = A.size();
    B.resize(np);
    vector<int32_t> cnt(size_t(Nc) + 1);
    std::fill(cnt.begin(), cnt.end(), 0);
    for (size_t q = 0; q < np; ++q) ++cnt[cam[A[q].y] + 1];
    for (int c = 0; c < Nc; ++c) cnt[c + 1] += cnt[c];
    for (size_t q = 0; q < np; ++q) B[cnt[cam[A[q].y]]++] = A[q];
    std::fill(cnt.begin(), cnt.end(), 0);
    for (size_t q = 0; q < np; ++q) ++cnt[cam[B[q].x] + 1];
    for (int c = 0; c < Nc; ++c) cnt[c + 1] += cnt[c];
    for (size_t q = 0; q < np; ++q) A[cnt[cam[B[q].x]]++] = B[q];
    vector<int32_t> blk_off;
    vector<int2> blk_key;
    for (size_t q = 0; q < np; ++q) {
        const int ci = cam[A[q].x], cj = cam[A[q].y];
        if (blk_key.empty() || blk_key.back().x != ci || blk_key.back().y != cj) { blk_key.push_back(make_int2(ci, cj)); blk_off.push_back(int32_t(q)); }
    }
    blk_off.push_back(int32_t(np));
    vector<int4> items;
    vector<int32_t> blk_item_off(blk_key.size() + 1, 0);
    for (size_t b = 0; b < blk_key.size(); ++b) {
        const int first = int(items.size());
        for (int q = blk_off[b]; q < blk_off[b + 1]; q += bak::PAIR_CHUNK) items.push_back(make_int4(int(b), q, std::min(q + bak::PAIR_CHUNK, blk_off[b + 1]), first));
        blk_item_off[b + 1] = int32_t(items.size());
    }
    bak::DenseView& D = e->DV;
    D.n = e->nf; D.T = (e->nf + bak::DT - 1) / bak::DT; D.ld = bak::DT * D.T;
    D.nblk = int(blk_key.size());
    D.corner_blocks = e->fi ? std::max(1, (Np + 127) / 128) : 0;
    e->d_S.alloc(size_t(bak::DT) * (D.T + 1) * D.ld);
    e->d_WU.alloc(size_t(36) * std::max(No, 1));
    if (e->fi) { e->d_GH.alloc(size_t(24) * std::max(Np, 1)); e->d_corner.alloc(size_t(16) * D.corner_blocks); }
    e->d_blk_off.upload(blk_off.data(), blk_off.size(), s);
    e->d_blk_key.upload(blk_key.data(), blk_key.size(), s);
    e->d_pairs.upload(A.data(), np, s);
    e->d_items.upload(items.data(), items.size(), s);
    e->d_blk_item_off.upload(blk_item_off.data(), blk_item_off.size(), s);
    e->d_item_part.alloc(size_t(36) * std::max<size_t>(items.size(), 1));
    e->d_blk_ticket.alloc(std::max<size_t>(blk_key.size(), 1)); e->d_blk_ticket.zero(s);
    e->d_Zinv.alloc(size_t(bak::DT) * bak::DT * D.T);
    CU_TRY(cudaStreamSynchronize(s));
    D.nitems = int(items.size()); D.items = e->d_items.p; D.blk_item_off = e->d_blk_item_off.p; D.item_part = e->d_item_part.p;
    D.blk_ticket = e->d_blk_ticket.p; D.Zinv = e->d_Zinv.p;
    D.S = e->d_S.p; D.WU = e->d_WU.p; D.GH = e->d_GH.p; D.corner_part = e->d_corner.p;
    D.blk_off = e->d_blk_off.p; D.blk_key = e->d_blk_key.p; D.pairs = e->d_pairs.p;
    {   // dataflow task list: tiles column by column, diagonal first
        vector<int2> tasks;
        for (int j = 0; j < D.T; ++j) for (int i = j; i <= D.T; ++i) tasks.push_back(make_int2(i, j));
        e->d_chol_tasks.upload(tasks.data(), tasks.size(), s);
        const size_t nflags = size_t(D.T + 1) * (D.T + 1) + 1;
        e->d_chol_ready.alloc(nflags);
        CU_TRY(cudaMemsetAsync(e->d_chol_ready.p, 0, nflags * sizeof(int), s));
        CU_TRY(cudaStreamSynchronize(s));
        e->CF.ntasks = int(tasks.size()); e->CF.tasks = e->d_chol_tasks.p; e->CF.ready = e->d_chol_ready.p; e->CF.abort = e->d_chol_ready.p + (nflags - 1); e->CF.epoch = 0;
        if (const char* m = std::getenv("BA_CHOL_MODE")) e->chol_mode = std::atoi(m);
    }
    int occ = 1;
    if (e->chol_mode == 1) CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, bak::k_dense_chol_df, 256, 0));
    else CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, bak::k_dense_chol, 256, 0));
    e->chol_grid = std::max(1, std::min(std::max(occ, 1) * e->num_sms, D.T * (D.T + 1) / 2));
    e->dense = true;
    // algorithmic bytes: formation reads the Jacobian columns it needs once and every pair's two 144-byte records,
    // writes the records and S; the factorisation reads and writes the lower triangle once per elimination step on average T/3 times
    const double jb = (e->fi ? 208.0 : 144.0) - 16.0 * e->jpack;
    e->stats[KID_DFORM].bytes = double(No) * (jb + 288 + 4) + double(np) * (8 + 288) + 8.0 * double(D.ld) * D.ld;
    e->stats[KID_DCHOL].bytes = 8.0 * double(D.ld) * D.ld * (D.T + 2) / 6.0;
}

struct DevState { const double* pose; const double* intr; const double* pt; };   // device-resident parameters of the whole (merged) problem

// BA_TIMING=1: host-side phase times of ba_set_problem on stderr (what separates bench.py's e2e from its HBM-resident value)
struct PhaseTimer {
    bool on = std::getenv("BA_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void tick(const char* what) {
        if (!on) return;
        const auto n = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[ba_set_problem] %-28s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};

void do_set_problem(ba_engine* e, const ba_problem* P, const DevState* dev = nullptr) {
    PhaseTimer pt_;
    const int Nc = P->num_cameras, Ng = P->num_groups, NpG = P->num_points, NoG = P->num_observations;
    const int W = e->opt.world_size, R = e->opt.rank;
    CU_TRY(cudaSetDevice(e->device));
    cudaStream_t s = e->stream;
    e->Nc = Nc; e->Ng = Ng; e->NpG = NpG; e->NoG = NoG;
    e->fi = e->opt.intrinsics_free ? 1 : 0;
    e->nf = 6 * Nc + (e->fi ? 4 * Ng : 0);
    // the engine's own copy of the index structure and pixels (ba_append merges into it): 120 MB on the Venice shape, copied by a
    // helper thread while this one builds the schedule; joined before returning (the caller's arrays are only borrowed for the call)
    struct Joiner { std::thread t; ~Joiner() { if (t.joinable()) t.join(); } } copier;
    copier.t = std::thread([e, P, NoG, Nc] {
        e->h_obs_cam.assign(P->obs_cam, P->obs_cam + NoG);
        e->h_obs_pt.assign(P->obs_pt, P->obs_pt + NoG);
        e->h_cam_fixed.assign(P->cam_fixed, P->cam_fixed + Nc);
        e->h_cam_group.assign(P->cam_group, P->cam_group + Nc);
        e->h_obs_uv.assign(P->obs_uv, P->obs_uv + size_t(2) * NoG);
    });

    pt_.tick("host copies (append)");
    // global CSR by point, global free masks.  The observations are sorted by point, so the offsets are the positions where the point id
    // changes; the by-camera permutation is a stable counting sort with one histogram per host thread (chunks of observations in order).
    const int nth = NoG >= (1 << 16) ? host_threads() : 1;
    vector<int32_t> gpt_off(size_t(NpG) + 1, 0);
    vector<int32_t> cam_cnt(Nc, 0);
    parallel_chunks(NoG, nth, [&](int, int a, int b) {
        for (int o = a; o < b; ++o) {
            const int p0 = o ? P->obs_pt[o - 1] : -1, p1 = P->obs_pt[o];
            for (int p = p0 + 1; p <= p1; ++p) gpt_off[p] = o;          // first observation of every point in (p0, p1]
        }
    });
    for (int p = (NoG ? P->obs_pt[NoG - 1] : -1) + 1; p <= NpG; ++p) gpt_off[p] = NoG;
    {
        vector<int32_t> hist(size_t(nth) * std::max(Nc, 1), 0);
        parallel_chunks(NoG, nth, [&](int k, int a, int b) { int32_t* h = hist.data() + size_t(k) * Nc; for (int o = a; o < b; ++o) ++h[P->obs_cam[o]]; });
        for (int k = 0; k < nth; ++k) for (int c = 0; c < Nc; ++c) cam_cnt[c] += hist[size_t(k) * Nc + c];
    }
    e->h_cam_free.assign(Nc, 0);
    vector<unsigned char> grp_free(std::max(Ng, 1), 0);
    for (int c = 0; c < Nc; ++c) {
        e->h_cam_free[c] = (!P->cam_fixed[c] && cam_cnt[c] > 0) ? 1 : 0;
        if (e->fi && cam_cnt[c] > 0) grp_free[P->cam_group[c]] = 1;
    }

    // shard: contiguous point ranges balanced by observation count, cut at point boundaries
    shard_points(gpt_off, NpG, NoG, W, e->h_shard_pt);
    e->pb = e->h_shard_pt[R]; e->pe = e->h_shard_pt[R + 1];
    e->ob = gpt_off[e->pb]; e->oe = gpt_off[e->pe];
    const int Np = e->pe - e->pb, No = e->oe - e->ob;
    e->Np = Np; e->No = No;
    e->ldo = (No + 31) / 32 * 32 + 32;

    // local index arrays
    HostScratch& hs = e->hs;
    vector<int32_t> pt_off(size_t(Np) + 1), tile_off, cam_off(size_t(Nc) + 1, 0);
    pvec<int32_t>& obs_pt = hs.obs_pt; obs_pt.resize(std::max(No, 1));
    pvec<int32_t>& perm = hs.perm; perm.resize(std::max(No, 1));
    for (int p = 0; p <= Np; ++p) pt_off[p] = gpt_off[e->pb + p] - e->ob;
    {
        const int nt2 = No >= (1 << 16) ? nth : 1;
        vector<int32_t> hist(size_t(nt2) * std::max(Nc, 1), 0);
        const int32_t* lcam = P->obs_cam + e->ob;
        const int32_t* lpt = P->obs_pt + e->ob;
        const int pb = e->pb;
        parallel_chunks(No, nt2, [&](int k, int a, int b) {
            int32_t* h = hist.data() + size_t(k) * Nc;
            for (int o = a; o < b; ++o) { obs_pt[o] = lpt[o] - pb; ++h[lcam[o]]; }
        });
        for (int c = 0; c < Nc; ++c) {           // exclusive scan over (camera, chunk): chunk k's first slot of camera c
            int32_t run = cam_off[c];
            for (int k = 0; k < nt2; ++k) { const int32_t n = hist[size_t(k) * Nc + c]; hist[size_t(k) * Nc + c] = run; run += n; }
            cam_off[c + 1] = run;
        }
        parallel_chunks(No, nt2, [&](int k, int a, int b) {
            int32_t* cur = hist.data() + size_t(k) * Nc;
            for (int o = a; o < b; ++o) perm[cur[lcam[o]]++] = o;
        });
    }
    pt_.tick("csr, shard, perm");
    build_tiles(pt_off, Np, tile_off);
    e->nTiles = int(tile_off.size()) - 1;
    // default: partial rows with one CTA per tile (mode 4).  The persistent streaming kernel (mode 6) is an explicit opt-in: measured on the
    // Venice shape its K5a is 3 % faster (181 vs 186 us) but the PCG iteration is 20 % slower (its CTAs hold every SM's shared memory, so
    // the gather's CTAs cannot start early and the chain loses ~45 us per iteration), and on the Final-13682 shape its result is not
    // reproducible from run to run (an unresolved race; tests/test_engine_gpu.py holds it to bit-identity on the small shapes only).
    e->scatter_eff = e->scatter_mode ? e->scatter_mode : 4;
    pt_.tick("tiles");
    // static scatter schedule: within each tile, the rank of every observation in camera-sorted (stable) order, and per
    // sorted position the camera + segment length at segment heads (ba_scatter.cuh)
    std::vector<unsigned char>& slot = hs.slot; slot.resize(std::max(No, 1));      // every entry is (re)written per tile below
    std::vector<int32_t>& spack = hs.spack; spack.resize(std::max(No, 1));
    if (Nc >= (1 << 22)) throw BaFail{BA_ERR_UNSUPPORTED, "more than 4M cameras"};
    static_assert(bak::TILE_OBS <= 256, "tile positions are packed into 8 bits");
    const int32_t* ocam = P->obs_cam + e->ob;
    // mode-4 schedule: level-1 chunk counts, head ranks, compact row numbering and the by-camera row lists.  Tiles are independent:
    // pass A (host threads over tile ranges) sorts each tile by camera and counts its rows, a prefix sum numbers the rows, pass B fills.
    {
        std::vector<unsigned char>&lvl1 = hs.lvl1, &hrank = hs.hrank, &hpos = hs.hpos;
        lvl1.resize(std::max(No, 1)); hrank.resize(std::max(No, 1)); hpos.resize(std::max(No, 1));
        std::vector<int32_t>&hbase = hs.hbase, &row_cam = hs.row_cam;
        hbase.assign(size_t(e->nTiles) + 1, 0);
        parallel_ranges(e->nTiles, [&](int t0, int t1) {
            uint32_t key[bak::TILE_OBS];
            for (int t = t0; t < t1; ++t) {
                const int o0 = tile_off[t], n = tile_off[t + 1] - o0;
                for (int i = 0; i < n; ++i) { slot[o0 + i] = 0; spack[o0 + i] = -1; lvl1[o0 + i] = 0; hrank[o0 + i] = 0; hpos[o0 + i] = 0; }   // the arrays are reused between calls
                if (n > bak::TILE_OBS) { hbase[t + 1] = n; continue; }   // long-point tile: one row per observation, no schedule
                for (int i = 0; i < n; ++i) key[i] = uint32_t(ocam[o0 + i]) << 9 | uint32_t(i);   // unique keys: std::sort == stable sort by camera
                std::sort(key, key + n);
                // Rows: one per maximal run of equal cameras INSIDE a warp's 32 sorted positions (a camera segment that crosses a
                // 32-boundary gives one row per warp it touches: the fold is a warp-segmented shuffle reduction, ba_rows.cuh).  For the
                // head of every run: camera | run length << 22 and its row number in the tile; for every position: how many positions
                // of its run follow it (`lvl1` field of the record).
                int rank = 0;
                for (int j = 0; j < n;) {
                    const int cam = int(key[j] >> 9);
                    int k = j;
                    while (k < n && int(key[k] >> 9) == cam) { slot[o0 + (key[k] & 511u)] = static_cast<unsigned char>(k); hpos[o0 + (key[k] & 511u)] = static_cast<unsigned char>(j); ++k; }
                    for (int q = j; q < k;) {
                        const int nxt = std::min((q / 32 + 1) * 32, k);     // end of this run: the segment's end or the warp boundary
                        spack[o0 + q] = cam | ((nxt - q) << 22);
                        hrank[o0 + q] = static_cast<unsigned char>(rank++);
                        for (int r = q; r < nxt; ++r) lvl1[o0 + r] = static_cast<unsigned char>(nxt - 1 - r);
                        q = nxt;
                    }
                    j = k;
                }
                hbase[t + 1] = rank;
            }
        });
        for (int t = 0; t < e->nTiles; ++t) hbase[t + 1] += hbase[t];
        row_cam.resize(hbase[e->nTiles]);
        parallel_ranges(e->nTiles, [&](int t0, int t1) {
            for (int t = t0; t < t1; ++t) {
                const int o0 = tile_off[t], n = tile_off[t + 1] - o0;
                int32_t* rc = row_cam.data() + hbase[t];
                if (n > bak::TILE_OBS) { for (int i = 0; i < n; ++i) rc[i] = ocam[o0 + i]; continue; }
                for (int j = 0; j < n;) { const int pk = spack[o0 + j]; *rc++ = pk & 0x3fffff; j += pk >> 22; }
            }
        });
        const int nRows = int(row_cam.size());
        vector<int32_t> cro(size_t(Nc) + 1, 0);
        pvec<int32_t>& crows = hs.crows; crows.resize(std::max(nRows, 1));
        {   // rows by camera: stable counting sort, one histogram per host thread
            const int nt3 = nRows >= (1 << 16) ? nth : 1;
            vector<int32_t> hist(size_t(nt3) * std::max(Nc, 1), 0);
            parallel_chunks(nRows, nt3, [&](int k, int a, int b) { int32_t* h = hist.data() + size_t(k) * Nc; for (int r = a; r < b; ++r) ++h[row_cam[r]]; });
            for (int c = 0; c < Nc; ++c) {
                int32_t run = cro[c];
                for (int k = 0; k < nt3; ++k) { const int32_t n = hist[size_t(k) * Nc + c]; hist[size_t(k) * Nc + c] = run; run += n; }
                cro[c + 1] = run;
            }
            parallel_chunks(nRows, nt3, [&](int k, int a, int b) { int32_t* cur = hist.data() + size_t(k) * Nc; for (int r = a; r < b; ++r) crows[cur[row_cam[r]]++] = r; });
        }
        // pack the per-position records (RowSched::tmeta) and per-tile records (RowSched::tile_info)
        pvec<int4>& tmeta = hs.tmeta; tmeta.resize(std::max(No, 1));
        vector<int4> tinfo(std::max(e->nTiles, 1));
        parallel_ranges(e->nTiles, [&](int t0, int t1) {
          for (int t = t0; t < t1; ++t) {
            const int o0 = tile_off[t], n = tile_off[t + 1] - o0;
            const int pfirst = obs_pt[o0];
            tinfo[t] = make_int4(o0, n, hbase[t], pfirst);
            if (n > bak::TILE_OBS) { for (int i = 0; i < n; ++i) tmeta[o0 + i] = make_int4(P->obs_cam[e->ob + o0 + i], 0, -1, 0); continue; }
            const int plast = obs_pt[o0 + n - 1];
            tinfo[t].y = int(unsigned(n) | unsigned(plast - pfirst + 1) << 24);   // points of the tile's range (<= TILE_MAX_PTS)
            for (int i = 0; i < n; ++i) {
                const int lpt = obs_pt[o0 + i] - pfirst;
                const unsigned packA = unsigned(slot[o0 + i]) | unsigned(lvl1[o0 + i]) << 8 | unsigned(hrank[o0 + i]) << 16 | unsigned(lpt) << 24;
                int pw = 0;
                if (pfirst + i <= plast) { const int a = pt_off[pfirst + i] - o0, cnt = pt_off[pfirst + i + 1] - pt_off[pfirst + i]; pw = cnt > 0 ? (a | cnt << 16) : 0; }
                tmeta[o0 + i] = make_int4(int(unsigned(P->obs_cam[e->ob + o0 + i]) | unsigned(hpos[o0 + i]) << 24), int(packA), spack[o0 + i], pw);
            }
          }
        });
        pt_.tick("row schedule (host threads)");
        e->d_tmeta.upload(tmeta.data(), No, s); e->d_tile_info.upload(tinfo.data(), tinfo.size(), s);
        e->d_cam_row_off.upload(cro.data(), cro.size(), s); e->d_cam_rows.upload(crows.data(), nRows, s);
        e->d_rows.alloc(size_t(nRows) * bak::ROW_DOUBLES); e->d_rows.zero(s);
        // The rows are written by K5a and read back by the gather a few microseconds later, every PCG iteration: ask L2 to keep them
        // (persisting access-policy window on the engine's stream) so that they need not travel to HBM and back.  BA_L2_PERSIST=0 disables.
        {
            int on = BA_L2_PERSIST_DEFAULT;
            if (const char* v = std::getenv("BA_L2_PERSIST")) on = std::atoi(v);
            cudaDeviceProp prop{};
            if (on) CU_TRY(cudaGetDeviceProperties(&prop, e->device));
            const size_t bytes = size_t(nRows) * bak::ROW_DOUBLES * sizeof(double);
            cudaStreamAttrValue attr{};
            if (on && bytes > 0 && prop.persistingL2CacheMaxSize > 0) {
                const size_t setaside = std::min<size_t>(prop.persistingL2CacheMaxSize, on > 1 ? size_t(on) << 20 : bytes);
                CU_TRY(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, setaside));
                const size_t win = std::min<size_t>(bytes, size_t(prop.accessPolicyMaxWindowSize));
                attr.accessPolicyWindow.base_ptr = e->d_rows.p;
                attr.accessPolicyWindow.num_bytes = win;
                attr.accessPolicyWindow.hitRatio = float(std::min(1.0, double(setaside) / double(win)));
                attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                e->l2_window = true;
            } else {
                attr.accessPolicyWindow.num_bytes = 0;
                e->l2_window = false;
            }
            if (on || e->l2_window_ever) { CU_TRY(cudaStreamSetAttribute(e->stream, cudaStreamAttributeAccessPolicyWindow, &attr)); e->l2_window_ever = true; }
        }
        e->d_tile_intr.alloc(size_t(4) * std::max(e->nTiles, 1)); e->d_tile_intr.zero(s);
        CU_TRY(cudaStreamSynchronize(s));
        bak::RowSched& RS = e->RS;
        RS.tmeta = e->d_tmeta.p; RS.tile_info = e->d_tile_info.p;
        RS.rows = e->d_rows.p; RS.tile_intr = e->d_tile_intr.p; RS.cam_row_off = e->d_cam_row_off.p; RS.cam_rows = e->d_cam_rows.p; RS.nRows = nRows;
    }
    pt_.tick("schedule uploads, rows alloc");
    e->d_tile_off.upload(tile_off.data(), tile_off.size(), s);
    e->d_obs_cam.upload(P->obs_cam + e->ob, No, s);
    e->d_obs_pt.upload(obs_pt.data(), No, s);
    e->d_obs_uv.upload(reinterpret_cast<const double2*>(P->obs_uv) + e->ob, No, s);
    e->d_pt_off.upload(pt_off.data(), pt_off.size(), s);
    e->d_cam_group.upload(P->cam_group, Nc, s);
    e->d_perm.upload(perm.data(), No, s);
    e->d_cam_off.upload(cam_off.data(), cam_off.size(), s);
    e->d_cam_free.upload(e->h_cam_free.data(), Nc, s);
    e->d_grp_free.upload(grp_free.data(), std::max(Ng, 1), s);
    {   // cameras of every intrinsics group, in camera order (k_group_blocks)
        vector<int32_t> goff(size_t(std::max(Ng, 1)) + 1, 0), gcams(std::max(Nc, 1));
        for (int c = 0; c < Nc; ++c) ++goff[P->cam_group[c] + 1];
        for (int g = 0; g < Ng; ++g) goff[g + 1] += goff[g];
        vector<int32_t> cur(goff.begin(), goff.end() - 1);
        for (int c = 0; c < Nc; ++c) gcams[cur[P->cam_group[c]]++] = c;
        e->d_grp_off.upload(goff.data(), goff.size(), s); e->d_grp_cams.upload(gcams.data(), gcams.size(), s);
        e->d_ipart.alloc(size_t(14) * std::max(Nc, 1));
    }
    if (dev) {   // ba_append: the parameters are already on the device
        e->d_pose0.alloc(size_t(6) * Nc); e->d_intr0.alloc(size_t(4) * Ng); e->d_pt0.alloc(size_t(3) * Np);
        CU_TRY(cudaMemcpyAsync(e->d_pose0.p, dev->pose, size_t(6) * Nc * sizeof(double), cudaMemcpyDeviceToDevice, s));
        CU_TRY(cudaMemcpyAsync(e->d_intr0.p, dev->intr, size_t(4) * Ng * sizeof(double), cudaMemcpyDeviceToDevice, s));
        CU_TRY(cudaMemcpyAsync(e->d_pt0.p, dev->pt, size_t(3) * Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
    } else {
        e->d_pose0.upload(P->pose, size_t(6) * Nc, s);
        e->d_intr0.upload(P->intr, size_t(4) * Ng, s);
        e->d_pt0.upload(P->pt + size_t(3) * e->pb, size_t(3) * Np, s);
    }
    for (auto* b : {&e->d_pose, &e->d_cpose}) b->alloc(size_t(6) * Nc);
    for (auto* b : {&e->d_intr, &e->d_cintr}) b->alloc(size_t(4) * Ng);
    for (auto* b : {&e->d_pt, &e->d_cpt}) b->alloc(size_t(3) * Np);
    e->d_camrec.alloc(size_t(bak::CAMREC) * Nc); e->d_ccamrec.alloc(size_t(bak::CAMREC) * Nc);
    // pinhole model: translation columns 3 and 4 share one stored column (ba_kernels.cuh: load_jc)
    e->jpack = (e->opt.model == BA_MODEL_PINHOLE && BA_JPACK_DEFAULT) ? 1 : 0;
    if (const char* v = std::getenv("BA_JPACK")) e->jpack = (std::atoi(v) && e->opt.model == BA_MODEL_PINHOLE) ? 1 : 0;
    e->d_J2.alloc(size_t((e->fi ? 13 : 9) - e->jpack) * e->ldo); e->d_r2.alloc(e->ldo);
    e->d_JcA.alloc(size_t(e->fi ? 12 : 8) * std::max(No, 1));
    e->k1_grid = std::max(1, std::min(e->nTiles, e->num_sms * K1_CTAS_PER_SM));
    e->d_Hpp.alloc(size_t(6) * Np); e->d_gp.alloc(size_t(3) * Np); e->d_Hinv.alloc(size_t(6) * Np);
    e->d_vp.alloc(size_t(3) * Np); e->d_yp.alloc(size_t(3) * Np); e->d_sc_pt.alloc(size_t(3) * Np);
    e->d_red.alloc(e->red_count());
    for (auto* b : {&e->d_sc_f, &e->d_Df2, &e->d_b, &e->d_x, &e->d_r, &e->d_z, &e->d_p, &e->d_y}) b->alloc(e->nf);
    e->d_Minv.alloc(size_t(36) * Nc + size_t(16) * Ng);
    e->factored = (e->factored_opt && e->scatter_eff == 4 && Nc > 0) ? 1 : 0;
    if (e->factored) { e->d_G.alloc(size_t(36) * Nc); e->d_vt.alloc(size_t(e->nf) + size_t(bak::VT_STRIDE - 6) * Nc); e->d_vt.zero(s); }
    const size_t nparts = std::max<size_t>(std::max(e->nTiles, Nc + 8), 2048);
    e->d_part0.alloc(nparts); e->d_part1.alloc(nparts);
    e->d_scal.alloc(1);
    e->d_Hpp.zero(s); e->d_gp.zero(s); e->d_yp.zero(s); e->d_Hinv.zero(s); e->d_vp.zero(s); e->d_scal.zero(s);
    CU_TRY(cudaMemcpyAsync(e->d_pose.p, e->d_pose0.p, size_t(6) * Nc * sizeof(double), cudaMemcpyDeviceToDevice, s));
    CU_TRY(cudaMemcpyAsync(e->d_intr.p, e->d_intr0.p, size_t(4) * Ng * sizeof(double), cudaMemcpyDeviceToDevice, s));
    CU_TRY(cudaMemcpyAsync(e->d_pt.p, e->d_pt0.p, size_t(3) * Np * sizeof(double), cudaMemcpyDeviceToDevice, s));

    pt_.tick("index/parameter uploads, allocs");
    setup_p2p(e);

    bak::View& V = e->V;
    V.Nc = Nc; V.Ng = Ng; V.Np = Np; V.No = No; V.ldo = e->ldo; V.nTiles = e->nTiles; V.fi = e->fi; V.left_twist = e->opt.lm_policy == BA_POLICY_G2O ? 1 : 0; V.skip_jc = 0;
    V.tile_off = e->d_tile_off.p; V.obs_cam = e->d_obs_cam.p; V.obs_pt = e->d_obs_pt.p; V.obs_uv = e->d_obs_uv.p;
    V.pt_off = e->d_pt_off.p; V.cam_group = e->d_cam_group.p; V.cam_free = e->d_cam_free.p; V.grp_free = e->d_grp_free.p;
    V.perm_by_cam = e->d_perm.p; V.cam_off = e->d_cam_off.p;
    V.jpack = e->jpack;
    V.J2 = e->d_J2.p; V.r2 = e->d_r2.p; V.JcA = e->d_JcA.p; V.jca_stride = e->fi ? 12 : 8; V.Hpp = e->d_Hpp.p; V.gp = e->d_gp.p; V.Hinv = e->d_Hinv.p; V.vp = e->d_vp.p; V.yp = e->d_yp.p;
    V.sc_pt = e->d_sc_pt.p;
    CU_TRY(cudaStreamSynchronize(s));   // host staging vectors go out of scope
    e->traj.clear();
    e->scales_set = false;
    e->pcg_pred = 0;
    e->has_problem = true;

    pt_.tick("p2p, view");
    build_dense(e, P, pt_off);
    e->lean = e->factored && !e->dense && e->opt.preconditioner != BA_PRECOND_SCHUR_JACOBI;
    if (e->lean && No > 0) {
        e->d_pt_by_cam.alloc(No);
        KL(e), bak::k_points_by_camera<<<(No + 255) / 256, 256, 0, s>>>(No, e->d_perm.p, e->d_obs_pt.p, e->d_pt_by_cam.p);
    }
    pt_.tick("dense setup");
    CU_TRY(cudaStreamSynchronize(s));   // the uploads read page-locked arrays this engine reuses in its next call
    pt_.tick("upload drain");

    // algorithmic bytes per launch of each kernel in THIS formulation, for roofline.achieved
    const double no = No, np = Np, nc = Nc, jb = (e->fi ? 208.0 : 144.0) - 16.0 * e->jpack;
    e->stats[KID_K0].bytes = (48 + 160) * nc;
    const double jca = e->fi ? 192.0 : 128.0;   // by-observation record for K2c
    e->stats[KID_K1].bytes = no * (24 + jb + 16 + jca) + np * (24 + 72) + 160 * nc;   // (pose columns skipped: adjusted below, once `dense` and `factored` are known)
    e->stats[KID_K1C].bytes = no * 24 + np * 24 + 160 * nc;
    e->stats[KID_K2C].bytes = no * (jca + 4) + (288 + 48) * nc;
    e->stats[KID_K3P].bytes = np * (48 + 24 + 24 + 48 + 24);
    e->stats[KID_K3C].bytes = nc * (288 + 48 + 48 + 288);
    e->stats[KID_K4].bytes = no * (jb + 16 + 8 + 5) + np * 24 + 48 * nc;
    e->stats[KID_K5].bytes = no * (jb + 8 + 5) + np * (48 + 4) + 96 * nc;
    e->stats[KID_K7].bytes = no * (jb + 8 + 16) + np * (48 + 24 + 24 + 4) + 48 * nc;
    e->stats[KID_K8].bytes = np * 72 + nc * 144;
    e->stats[KID_PCGV].bytes = double(e->nf) * 8 * 14 + 288 * nc;
    e->stats[KID_MISC].bytes = 0;
    e->stats[KID_GATHER].bytes = 0; e->stats[KID_TAIL].bytes = double(e->nf) * 8 * 12 + 288 * nc;
    if (e->scatter_eff >= 4) {   // partial-row formulation: 16 B/obs of schedule, one packed row written per (tile, camera) pair, read back by the gather
        const double nr = e->RS.nRows;
        const double rb = 8.0 * bak::ROW_DOUBLES;   // bytes per partial row as stored
        e->stats[KID_K4].bytes = no * (jb + 16 + 16) + np * 24 + rb * nr + 16.0 * e->nTiles;
        if (e->factored) e->stats[KID_K7].bytes = no * (48.0 + (e->fi ? 64.0 : 0.0) + 8 + 16) + np * (48 + 24 + 24 + 4 + 24) + (48 + 288 + 48) * nc;
        if (e->lean) {   // neither the pose columns nor the by-observation records are written; K2c gathers Jp, r (4 [8] sectors per observation), two indices and the point
            e->stats[KID_K1].bytes -= no * (jb - 48.0 - (e->fi ? 64.0 : 0.0) + jca) + 64 * nc;   // (and 96 of the record's 160 bytes)
            e->stats[KID_K2C].bytes = no * (128.0 + (e->fi ? 128.0 : 0.0) + 8 + 24) + (288 + 288 + 48) * nc;
        }
        if (e->factored) e->stats[KID_K4].bytes = no * (48.0 + (e->fi ? 64.0 : 0.0) + 16 + 16) + np * (24 + 24) + rb * nr + 16.0 * e->nTiles;
        e->stats[KID_K5].bytes = no * (jb + 16) + np * 48 + 48 * nc + rb * nr + 16.0 * e->nTiles;
        if (e->factored)   // pose-factored: Jp (48 B) [+ Ji 64 B] and the schedule record per observation, Hinv and the point per point, vt per camera
            e->stats[KID_K5].bytes = no * (48.0 + (e->fi ? 64.0 : 0.0) + 16) + np * (48 + 24) + 48 * nc + rb * nr + 16.0 * e->nTiles;
        e->stats[KID_GATHER].bytes = (rb + 4) * nr + 48 * nc;
        e->stats[KID_PCGV].bytes = e->stats[KID_GATHER].bytes + e->stats[KID_TAIL].bytes;
    }
}

}  // namespace

namespace {

// ---- kernel sequences ----------------------------------------------------------------------------------------
inline int cdiv(int a, int b) { return std::max(1, (a + b - 1) / b); }   // never a grid of 0 (every kernel bounds-checks its index)
constexpr int kNormBlocks = 592;

void read_scalars(ba_engine* e) {
    launch_check(__FILE__, __LINE__);
    ++e->host_syncs;
    CU_TRY(cudaMemcpyAsync(e->h_scal, e->d_scal.p, sizeof(bak::Scalars), cudaMemcpyDeviceToHost, e->stream));
    g_d2h_bytes += int64_t(sizeof(bak::Scalars));
    CU_TRY(cudaStreamSynchronize(e->stream));
}

// Evaluate residuals, Jacobian, cost and the block Hessian / gradient at the CURRENT parameters (K0, K1, K2c).
void evaluate_full(ba_engine* e, bool full_jacobian = false) {
    cudaStream_t s = e->stream;
    bak::Scalars* sc = e->d_scal.p;
    const bak::RobustParams rp{e->opt.robust_kind, e->opt.robust_a};
    CU_TRY(cudaMemsetAsync(&sc->bad_eval, 0, sizeof(int), s));
    { LaunchTimer t(e, KID_K0);
      KL(e), bak::k0_camera_precompute<<<cdiv(e->Nc, 128), 128, 0, s>>>(e->Nc, e->d_pose.p, e->d_camrec.p, e->opt.taylor_rule);
      if (e->factored) KL(e), bak::k_pose_factors<<<cdiv(e->Nc, 128), 128, 0, s>>>(e->Nc, e->d_camrec.p, e->d_cam_free.p, e->V.left_twist, e->d_G.p); }
    int k1_parts = 0;   // cost partials K1 writes (one per CTA of its persistent grid)
    if (e->nTiles > 0) {
        LaunchTimer t(e, KID_K1);
        // neither the pose columns of J2 nor the by-observation records have a reader when K2c/K4/K5/K7 all run pose-factored (PCG with the
        // block-Jacobi preconditioner): K1 then skips those 208 / 288 bytes per observation -- except for ba_evaluate, which returns the columns
        bak::View V1 = e->V;
        V1.skip_jc = (!full_jacobian && e->lean) ? 1 : 0;
        const int grid = V1.skip_jc ? std::max(1, std::min(e->nTiles, e->num_sms * K1_LEAN_CTAS_PER_SM)) : e->k1_grid;   // persistent: CTAs per SM x SMs
        k1_parts = grid;
        auto k1 = [&](auto kern) { KL(e); kern<<<grid, bak::TILE_OBS, 0, s>>>(V1, e->d_camrec.p, e->d_intr.p, e->d_pt.p, rp, e->d_part0.p, &sc->bad_eval); };
        if (e->opt.model == BA_MODEL_SNAVELY) { if (V1.skip_jc) k1(bak::k1_residual_jacobian<1, true>); else k1(bak::k1_residual_jacobian<1, false>); }
        else { if (V1.skip_jc) k1(bak::k1_residual_jacobian<0, true>); else k1(bak::k1_residual_jacobian<0, false>); }
    }
    { LaunchTimer t(e, KID_MISC);
      KL(e), bak::k_reduce_partials<<<1, 1024, 0, s>>>(e->d_part0.p, k1_parts, &sc->cost); }
    if (e->Nc > 0) {
        LaunchTimer t(e, KID_K2C);
        if (e->lean && e->No > 0) {   // pose-factored: from the SoA point columns, the residuals and the points (ba_rows.cuh)
            if (e->fi) { if (e->jpack) KL(e), bak::k2c_camera_blocks_f<true, 1><<<e->Nc, 128, 0, s>>>(e->V, e->d_G.p, e->d_pt.p, e->d_pt_by_cam.p, e->Hcc(), e->gc(), e->d_ipart.p);
                         else KL(e), bak::k2c_camera_blocks_f<true, 0><<<e->Nc, 128, 0, s>>>(e->V, e->d_G.p, e->d_pt.p, e->d_pt_by_cam.p, e->Hcc(), e->gc(), e->d_ipart.p); }
            else { if (e->jpack) KL(e), bak::k2c_camera_blocks_f<false, 1><<<e->Nc, 128, 0, s>>>(e->V, e->d_G.p, e->d_pt.p, e->d_pt_by_cam.p, e->Hcc(), e->gc(), nullptr);
                   else KL(e), bak::k2c_camera_blocks_f<false, 0><<<e->Nc, 128, 0, s>>>(e->V, e->d_G.p, e->d_pt.p, e->d_pt_by_cam.p, e->Hcc(), e->gc(), nullptr); }
        } else if (e->fi) KL(e), bak::k2c_camera_blocks<true><<<e->Nc, 128, 0, s>>>(e->V, e->Hcc(), e->gc(), e->d_ipart.p);
        else KL(e), bak::k2c_camera_blocks<false><<<e->Nc, 128, 0, s>>>(e->V, e->Hcc(), e->gc(), nullptr);
        if (e->fi) KL(e), bak::k_group_blocks<<<cdiv(e->Ng * 32, 128), 128, 0, s>>>(e->Ng, e->d_grp_off.p, e->d_grp_cams.p, e->d_ipart.p, e->Hii(), e->gi());
    }
    if (e->opt.world_size > 1) {
        NCCL_TRY(g_nccl.GroupStart());
        allreduce_sum(e, e->d_red.p, e->red_count());
        allreduce_sum(e, &sc->cost, 1);
        allreduce_flags(e, &sc->bad_eval, 1);
        NCCL_TRY(g_nccl.GroupEnd());
    }
}

void gradient_max(ba_engine* e) {
    cudaStream_t s = e->stream;
    LaunchTimer t(e, KID_MISC);
    KL(e), bak::k_gradmax_points<<<kNormBlocks, 256, 0, s>>>(e->V, e->d_pt.p, e->d_part1.p);
    KL(e), bak::k_gradmax_final<<<1, 1024, 0, s>>>(e->Nc, e->Ng, e->fi, e->d_cam_free.p, e->d_grp_free.p, e->d_pose.p, e->d_intr.p, e->gc(), e->gi(),
                                            e->d_part1.p, kNormBlocks, e->d_scal.p);
    allreduce_max(e, &e->d_scal.p->gmax_pt, 1);
}

// x+ = x - y and the norms of K8; yf = PCG solution (pose | intrinsics part), yp in V.yp.
void make_candidate(ba_engine* e, const double* yf, double lambda = 0.0) {
    cudaStream_t s = e->stream;
    bak::Scalars* sc = e->d_scal.p;
    LaunchTimer t(e, KID_K8);
    double* scale_part = e->d_part1.p + 1024;   // third partial array (d_part1 holds >= 2048 doubles)
    KL(e), bak::k8_candidate_points<<<kNormBlocks, 256, 0, s>>>(e->V, e->d_pt.p, e->d_cpt.p, e->d_part0.p, e->d_part1.p, scale_part, lambda);
    KL(e), bak::k8_candidate_cameras<<<1, 1024, 0, s>>>(e->Nc, e->Ng, e->fi, e->d_cam_free.p, e->d_grp_free.p, e->d_pose.p, e->d_intr.p, yf,
                                                 e->d_cpose.p, e->d_cintr.p, e->d_part0.p, e->d_part1.p, kNormBlocks, sc, scale_part, e->gc(), lambda,
                                                 e->V.left_twist);
}

void candidate_cost(ba_engine* e) {
    cudaStream_t s = e->stream;
    bak::Scalars* sc = e->d_scal.p;
    const bak::RobustParams rp{e->opt.robust_kind, e->opt.robust_a};
    CU_TRY(cudaMemsetAsync(&sc->bad_cand, 0, sizeof(int), s));
    { LaunchTimer t(e, KID_K0);
      KL(e), bak::k0_camera_precompute<<<cdiv(e->Nc, 128), 128, 0, s>>>(e->Nc, e->d_cpose.p, e->d_ccamrec.p, e->opt.taylor_rule); }
    const int cblocks = e->No > 0 ? std::min(std::min(cdiv(e->No, bak::K1C_THREADS), e->num_sms * bak::K1C_CTAS_PER_SM), 2048) : 0;   // <= 2048 partials (d_part0)
    if (cblocks > 0) {
        LaunchTimer t(e, KID_K1C);
        if (e->opt.model == BA_MODEL_SNAVELY) KL(e), bak::k1c_candidate_cost<1><<<cblocks, bak::K1C_THREADS, 0, s>>>(e->V, e->d_ccamrec.p, e->d_cintr.p, e->d_cpt.p, rp, e->d_part0.p, &sc->bad_cand);
        else KL(e), bak::k1c_candidate_cost<0><<<cblocks, bak::K1C_THREADS, 0, s>>>(e->V, e->d_ccamrec.p, e->d_cintr.p, e->d_cpt.p, rp, e->d_part0.p, &sc->bad_cand);
    }
    { LaunchTimer t(e, KID_MISC);
      KL(e), bak::k_reduce_partials<<<1, 1024, 0, s>>>(e->d_part0.p, cblocks, &sc->ccost); }
}

// Launch along the PCG chain (K5a -> K5b -> tail -> K5a ...) with programmatic dependent launch: the next kernel's CTAs are scheduled
// once every CTA of its predecessor has executed griddepcontrol.launch_dependents (all three do, first thing), run their prologue
// (K5a: the tile's J columns and schedule record, none of which the predecessor writes) and block in griddepcontrol.wait until the
// predecessor grid has completed and flushed.  Every kernel of the chain executes the wait before touching anything a predecessor
// writes, so completion stays transitive.  Without the attribute (e->pdl == 0, or any other launch site) both instructions are no-ops.
template <class... KArgs, class... Args>
void launch_chain(ba_engine* e, void (*kern)(KArgs...), dim3 grid, dim3 block, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = e->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = e->pdl ? 1 : 0;
    KL(e);
    CU_TRY(cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...));
}

template <int MODE>
void launch_sweep(ba_engine* e, int kid, const double* xin, double* yout, const int* done, double* partial) {
    if (e->nTiles == 0) return;
    LaunchTimer t(e, kid);
    KL(e);
    if (e->fi) { if (e->jpack) bak::k_schur_sweep<MODE, true, 1><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, xin, yout, done, partial);
                 else bak::k_schur_sweep<MODE, true, 0><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, xin, yout, done, partial); }
    else { if (e->jpack) bak::k_schur_sweep<MODE, false, 1><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, xin, yout, done, partial);
           else bak::k_schur_sweep<MODE, false, 0><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, xin, yout, done, partial); }
}

// K7: back-substitution of the point steps and the model cost change, from the reduced solution in d_x.  Pose-factored form (streams Jp, r and
// the index arrays; F y = Jp (v x X + tau) with (v, tau) = G_c y_c from k_pose_apply) whenever the factors exist, else the explicit sweep.
void launch_backsub(ba_engine* e) {
    if (!e->factored || e->nTiles == 0) { launch_sweep<bak::SWEEP_BACKSUB>(e, KID_K7, e->d_x.p, nullptr, nullptr, e->d_part0.p); return; }
    LaunchTimer t(e, KID_K7);
    KL(e), bak::k_pose_apply<<<cdiv(e->nf, 256), 256, 0, e->stream>>>(e->Nc, e->nf, e->d_G.p, e->d_x.p, e->d_vt.p);
    KL(e);
    if (e->fi) { if (e->jpack) bak::k_backsub_f<true, 1><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, e->d_vt.p, e->d_pt.p, e->d_part0.p);
                 else bak::k_backsub_f<true, 0><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, e->d_vt.p, e->d_pt.p, e->d_part0.p); }
    else { if (e->jpack) bak::k_backsub_f<false, 1><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, e->d_vt.p, e->d_pt.p, e->d_part0.p);
           else bak::k_backsub_f<false, 0><<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, e->d_vt.p, e->d_pt.p, e->d_part0.p); }
}

// K4 / K5a by the configured scatter: 1 = L2 atomics straight into `yout` (k_schur_sweep); 4 = partial rows, one CTA per
// tile (k_schur_rows); 6 = partial rows, persistent streaming kernel (k_schur_stream).  4 and 6 leave the rows for the by-camera gather.
#ifdef BA_WITH_STREAM
template <int MODE, bool FI, int PK>
void launch_stream(ba_engine* e, const double* xin, double* yout, const int* done, bool chain) {
    using Cfg = bak::StreamCfg<MODE, FI, PK>;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(unsigned(std::max(1, std::min(e->num_sms, e->nTiles)))); cfg.blockDim = dim3(Cfg::THREADS);
    cfg.dynamicSmemBytes = Cfg::SMEM; cfg.stream = e->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = (chain && e->pdl) ? 1 : 0;
    KL(e);
    CU_TRY(cudaLaunchKernelEx(&cfg, bak::k_schur_stream<MODE, FI, PK>, e->V, e->RS, xin, yout, done));
}

template <int MODE, bool FI, int PK>
void set_stream_attr() {
    CU_TRY(cudaFuncSetAttribute(bak::k_schur_stream<MODE, FI, PK>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bak::StreamCfg<MODE, FI, PK>::SMEM)));
}
// Once per engine (i.e. per device): the streaming kernels use more dynamic shared memory than the default limit.
void set_stream_attrs(ba_engine* e) {
    if (e->stream_attr_set) return;
    set_stream_attr<bak::SWEEP_MATVEC, false, 0>(); set_stream_attr<bak::SWEEP_MATVEC, false, 1>();
    set_stream_attr<bak::SWEEP_MATVEC, true, 0>();  set_stream_attr<bak::SWEEP_MATVEC, true, 1>();
    set_stream_attr<bak::SWEEP_RHS, false, 0>();    set_stream_attr<bak::SWEEP_RHS, false, 1>();
    set_stream_attr<bak::SWEEP_RHS, true, 0>();     set_stream_attr<bak::SWEEP_RHS, true, 1>();
    e->stream_attr_set = true;
}

#endif

template <int MODE>
void launch_scatter(ba_engine* e, int kid, const double* xin, double* yout, const int* done) {
    if (e->nTiles == 0) return;
    if (e->scatter_eff == 1) { launch_sweep<MODE>(e, kid, xin, yout, done, nullptr); return; }
    const bool sample = kid != KID_K5 || e->stats[KID_K5].launches % e->k5_sample_stride == 0;
    LaunchTimer t(e, kid, sample, kid == KID_K5 ? e->pcg_tag : -1);
    if (e->fi && e->Ng > 1) CU_TRY(cudaMemsetAsync(yout + 6 * e->Nc, 0, size_t(4) * e->Ng * sizeof(double), e->stream));
    // K5a goes down the PDL chain; K4 is an ordinary launch (both griddepcontrol instructions are no-ops there)
    const bool chain = MODE == bak::SWEEP_MATVEC;
#ifdef BA_WITH_STREAM
    if (e->scatter_eff == 6) {
        if (e->fi) { if (e->jpack) launch_stream<MODE, true, 1>(e, xin, yout, done, chain); else launch_stream<MODE, true, 0>(e, xin, yout, done, chain); }
        else { if (e->jpack) launch_stream<MODE, false, 1>(e, xin, yout, done, chain); else launch_stream<MODE, false, 0>(e, xin, yout, done, chain); }
        return;
    }
#endif
    if (e->factored) {   // pose-factored rows.  K5a: gathers vt = G p (written by the PCG tail) instead of p; K4: no pose input.  Both read the points of the linearisation
        auto gof = [&](auto kern) {
            if (chain) launch_chain(e, kern, dim3(e->nTiles), dim3(bak::TILE_OBS), e->V, e->RS, e->d_vt.p, e->d_pt.p, yout, done);
            else { KL(e); kern<<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, e->RS, nullptr, e->d_pt.p, yout, done); }
        };
        if (e->fi) { if (e->jpack) gof(bak::k_schur_rows_f<MODE, true, 1>); else gof(bak::k_schur_rows_f<MODE, true, 0>); }
        else { if (e->jpack) gof(bak::k_schur_rows_f<MODE, false, 1>); else gof(bak::k_schur_rows_f<MODE, false, 0>); }
        return;
    }
    auto go = [&](auto kern) {
        if (chain) launch_chain(e, kern, dim3(e->nTiles), dim3(bak::TILE_OBS), e->V, e->RS, xin, yout, done);
        else { KL(e); kern<<<e->nTiles, bak::TILE_OBS, 0, e->stream>>>(e->V, e->RS, xin, yout, done); }
    };
    if (e->fi) { if (e->jpack) go(bak::k_schur_rows<MODE, true, 1>); else go(bak::k_schur_rows<MODE, true, 0>); }
    else { if (e->jpack) go(bak::k_schur_rows<MODE, false, 1>); else go(bak::k_schur_rows<MODE, false, 0>); }
}

// One linear solve of the LM step at trust-region radius `radius` (K3, K4, PCG with K5, K7).  Leaves the step in
// d_x (pose|intrinsics part) and V.yp (points), the model cost change in scal.mcc.  Returns PCG iterations, or -1.
int solve_linear(ba_engine* e, double radius, double* rel_out, double lambda = 0.0) {
    cudaStream_t s = e->stream;
    bak::Scalars* sc = e->d_scal.p;
    const bak::LmParams lm{radius, e->opt.min_lm_diagonal, e->opt.max_lm_diagonal, e->opt.lm_policy == BA_POLICY_G2O ? 1 : 0, lambda};
    const bool rowsm = e->scatter_eff >= 4;
    const int gblocks = e->Nc + 1;
    const dim3 gthreads(bak::GATHER_WARPS * 32);
    CU_TRY(cudaMemsetAsync(&sc->bad_lin, 0, sizeof(int), s));
    if (e->Np > 0) { LaunchTimer t(e, KID_K3P); KL(e), bak::k3_point_blocks<<<cdiv(e->Np, 256), 256, 0, s>>>(e->V, lm, &sc->bad_lin); }
    { LaunchTimer t(e, KID_K3C);
      KL(e), bak::k3_camera_blocks<<<cdiv(e->Nc + e->Ng, 128), 128, 0, s>>>(e->Nc, e->Ng, e->fi, e->Hcc(), e->Hii(), e->d_sc_f.p, lm, e->d_Df2.p, e->dense ? nullptr : e->d_Minv.p, &sc->bad_lin);
      if (e->opt.preconditioner == BA_PRECOND_SCHUR_JACOBI && !e->dense && e->Nc > 0) {   // pose blocks of S instead of B (needs K3p's Hinv of this damping)
          e->d_sdiag.alloc(size_t(36) * e->Nc);
          KL(e), bak::k_schur_diag<<<e->Nc, 128, 0, s>>>(e->V, e->d_sdiag.p);
          allreduce_sum(e, e->d_sdiag.p, size_t(36) * e->Nc);
          KL(e), bak::k3_schur_jacobi_blocks<<<cdiv(e->Nc, 128), 128, 0, s>>>(e->Nc, e->Hcc(), e->d_Df2.p, e->d_sdiag.p, e->d_Minv.p);
      } }
    // reduced right-hand side b = F^T (r - E vp) = gc - F^T E vp
    if (!rowsm || e->nTiles == 0) e->d_b.zero(s);
    launch_scatter<bak::SWEEP_RHS>(e, KID_K4, nullptr, e->d_b.p, nullptr);
    if (rowsm && e->nTiles > 0) { LaunchTimer t(e, KID_GATHER);
        KL(e), bak::k_camera_gather<<<gblocks, gthreads, 0, s>>>(e->Nc, e->Ng, e->fi, e->nTiles, e->RS, nullptr, nullptr, e->d_b.p, nullptr, nullptr, 0, e->factored ? e->d_G.p : nullptr); }
    allreduce_sum(e, e->d_b.p, e->nf);
    if (e->dense) {   // direct solve (ba_dense.cuh): S formed densely, Cholesky with the right-hand side as an extra row, no host round trip
        const bak::DenseView& D = e->DV;
        {
            LaunchTimer t(e, KID_DFORM);
            CU_TRY(cudaMemsetAsync(D.S, 0, size_t(bak::DT) * (D.T + 1) * D.ld * sizeof(double), s));
            if (e->No > 0) KL(e), bak::k_dense_wu<<<cdiv(e->No, 128), 128, 0, s>>>(e->V, D);
            if (e->fi) KL(e), bak::k_dense_point_g<<<D.corner_blocks, 128, 0, s>>>(e->V, D);
            KL(e), bak::k_dense_init<<<cdiv(36 * e->Nc + D.ld + 32 + 16 * 32, 256), 256, 0, s>>>(e->Nc, e->fi, D, e->Hcc(), e->Hii(), e->d_Df2.p, e->d_b.p);
            if (D.nitems > 0) KL(e), bak::k_dense_pairs<<<std::min(cdiv(D.nitems, 4), 16 * e->num_sms), 128, 0, s>>>(D);
            if (e->fi && e->Nc > 0) KL(e), bak::k_dense_border<<<e->Nc, 128, 0, s>>>(e->V, D);
        }
        {
            LaunchTimer t(e, KID_DCHOL);
            bak::DenseView dv = D;
            double* xp = e->d_x.p; bak::PcgState2* pcg = &sc->pcg; int* bad = &sc->bad_lin;
            KL(e);
            if (e->chol_mode == 1) {
                bak::CholFlow cf = e->CF;
                cf.epoch = ++e->CF.epoch;
                void* args[] = {&dv, &cf, &xp, &pcg, &bad};
                CU_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(bak::k_dense_chol_df), dim3(e->chol_grid), dim3(256), args, 0, s));
            } else {
                void* args[] = {&dv, &xp, &pcg, &bad};
                CU_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(bak::k_dense_chol), dim3(e->chol_grid), dim3(256), args, 0, s));
            }
        }
        launch_backsub(e);
        { LaunchTimer t(e, KID_MISC);
          KL(e), bak::k_reduce_partials<<<1, 1024, 0, s>>>(e->d_part0.p, e->nTiles, &sc->mcc); }
        *rel_out = 0.0;
        return 0;   // failure (bad_lin) is seen by the caller's next read-back
    }
    const bak::PcgParams pp{e->opt.pcg_max_iterations, e->opt.pcg_forcing, e->opt.pcg_rel_tolerance, e->opt.eta};
    const double* Gf = e->factored ? e->d_G.p : nullptr;   // pose-factored K5a: its rows are multiplied by G_c^T in the gather, the tail hands vt = G p back
    { LaunchTimer t(e, KID_PCGV);
      KL(e), bak::k_pcg_init2<<<1, 1024, 0, s>>>(e->nf, e->Nc, e->d_Minv.p, e->d_b.p, e->d_x.p, e->d_r.p, e->d_z.p, e->d_p.p, &sc->pcg, Gf, e->factored ? e->d_vt.p : nullptr); }
    // PCG iterations (K5a -> K5b -> cluster tail, a chain of programmatic dependent launches) go out in batches without host sync; a
    // device-side `done` flag turns the rest of a batch into no-ops once the stop test fires, and the host reads the PCG state once
    // per batch.  Fixed tolerance: iteration counts of consecutive linear solves differ by a few, so the first batch is the previous
    // solve's count minus a margin and the remainder goes in eights (2-3 read-backs per linear solve).  Forcing sequence (a handful
    // of iterations per solve): 2, 4, 8, 8, ...
    int batch = e->opt.pcg_forcing ? 2 : (e->pcg_pred > 16 ? e->pcg_pred - 4 : 8);
    const size_t pend0 = e->pending.size();
    int launched = 0;
    bak::TailArgs ta{};
    ta.Nc = e->Nc; ta.Ng = e->Ng; ta.fi = e->fi; ta.Minv = e->d_Minv.p; ta.b = e->d_b.p; ta.Df2 = e->d_Df2.p;
    ta.y = e->d_y.p; ta.p = e->d_p.p; ta.x = e->d_x.p; ta.r = e->d_r.p; ta.z = e->d_z.p; ta.st = &sc->pcg; ta.pp = pp;
    ta.dotp = e->d_part1.p; ta.ndot = gblocks;
    ta.G = Gf; ta.vt = e->factored ? e->d_vt.p : nullptr;
    const bool multi = e->opt.world_size > 1;
    const bool no_rows = !rowsm || e->nTiles == 0;   // y is accumulated in d_y by atomics (or stays zero on a rank without observations)
    for (;;) {
        for (int i = 0; i < batch; ++i) {
            ++launched;
            e->pcg_tag = launched;
            const bool sample = e->stats[KID_K5].launches % e->k5_sample_stride == 0;
            if (no_rows) e->d_y.zero(s);
            launch_scatter<bak::SWEEP_MATVEC>(e, KID_K5, e->d_p.p, e->d_y.p, &sc->pcg.done);
            LaunchTimer t(e, KID_PCGV, sample, launched);
            // y from the rows (or already in d_y); what is left for the tail: mode 0 = q and the shares of p.q are there (one GPU),
            // 1 = y is complete in local memory, 2 = pull the ranks' slots from peer memory (ba_p2p.cuh)
            ta.mode = 1;
            if (multi && e->p2p.on) {   // one-shot NVLink all-reduce: local gather here, flags + pull + sum inside the tail
                ta.seq = ++e->p2p.seq; ta.P = e->p2p.view; ta.mode = 2;
                LaunchTimer tg(e, KID_GATHER, sample, launched);
                if (no_rows) {   // this rank's slot <- d_y
                    double* mine = e->p2p.view.data[e->p2p.view.rank] + size_t(ta.seq & 1u) * e->p2p.view.ldf;
                    CU_TRY(cudaMemcpyAsync(mine, e->d_y.p, size_t(e->nf) * sizeof(double), cudaMemcpyDeviceToDevice, s));
                } else launch_chain(e, bak::k_camera_gather_push, dim3(gblocks), gthreads, e->Nc, e->Ng, e->fi, e->nTiles, e->RS, e->d_y.p, e->p2p.view, ta.seq, &sc->pcg, 1, Gf);
                if (e->p2p_wide == 1 || (e->p2p_wide < 0 && e->opt.world_size >= 4 && e->nf >= 32768)) {   // pull and sum the W slots on many SMs (ba_p2p.cuh: k_allreduce_pull); the tail then finds y complete in local memory
                    const int pblocks = std::max(1, std::min(2 * e->num_sms, cdiv((e->nf + 1) / 2, bak::PULL_THREADS)));
                    launch_chain(e, bak::k_allreduce_pull, dim3(pblocks), dim3(bak::PULL_THREADS), e->nf, e->p2p.view, ta.seq, e->d_y.p, &sc->pcg, 1);
                    ta.mode = 1;
                }
            } else if (multi) {
                { LaunchTimer tg(e, KID_GATHER, sample, launched);
                  if (!no_rows) KL(e), bak::k_camera_gather<<<gblocks, gthreads, 0, s>>>(e->Nc, e->Ng, e->fi, e->nTiles, e->RS, nullptr, nullptr, e->d_y.p, nullptr, &sc->pcg, 0, Gf); }
                allreduce_sum(e, e->d_y.p, e->nf);
            } else if (!no_rows && e->dbg_tail1) {   // BA_TAIL_MODE1=1 (bisecting aid): plain gather, q and p.q by the tail's cluster reduction
                LaunchTimer tg(e, KID_GATHER, sample, launched);
                launch_chain(e, bak::k_camera_gather, dim3(gblocks), gthreads, e->Nc, e->Ng, e->fi, e->nTiles, e->RS, nullptr, nullptr, e->d_y.p, nullptr, &sc->pcg, 0, Gf);
            } else if (!no_rows) {   // one GPU: the gather forms q and the shares of p.q as well
                LaunchTimer tg(e, KID_GATHER, sample, launched);
                launch_chain(e, bak::k_camera_gather, dim3(gblocks), gthreads, e->Nc, e->Ng, e->fi, e->nTiles, e->RS, e->d_Df2.p, e->d_p.p, e->d_y.p, e->d_part1.p, &sc->pcg, 1, Gf);
                ta.mode = 0;
            }
            // q, alpha, x, r, z, rho, beta, stop test and the new search direction: one thread-block cluster (ba_pcg_cluster.cuh)
            LaunchTimer tt(e, KID_TAIL, sample, launched);
            if (e->tail_cl == bak::TAIL_CL_WIDE && e->nf >= e->tail_wide_min_nf) launch_chain(e, bak::k_pcg_tail<bak::TAIL_CL_WIDE>, dim3(bak::TAIL_CL_WIDE), dim3(bak::TAIL_THREADS), ta);
            else launch_chain(e, bak::k_pcg_tail<bak::TAIL_CL>, dim3(bak::TAIL_CL), dim3(bak::TAIL_THREADS), ta);
        }
        read_scalars(e);
        if (e->h_scal->bad_lin & 8) throw BaFail{BA_ERR_NCCL, "peer-memory all-reduce timed out waiting for another rank"};
        if (e->h_scal->pcg.done) break;
        batch = e->opt.pcg_forcing ? std::min(2 * batch, 8) : 8;
    }
    e->pcg_tag = -1;
    {   // launches after the stop test fired did no work: they are neither counted nor timed (ba_get_kernel_stats)
        const int real = e->h_scal->pcg.k;
        if (real > 0) e->pcg_pred = real;
        if (launched > real) {
            for (int kid : {KID_K5, KID_PCGV, KID_GATHER, KID_TAIL}) e->stats[kid].launches = std::max<int64_t>(0, e->stats[kid].launches - (launched - real));
            for (size_t i = pend0; i < e->pending.size(); ++i) {
                ba_engine::Pending& q = e->pending[i];
                if (q.tag > real && q.kid >= 0) { --e->stats[q.kid].timed; q.kid = -1; }
            }
        }
    }
    *rel_out = e->h_scal->pcg.rel;
    if (e->h_scal->bad_lin || e->h_scal->pcg.done == 3) return -1;
    // back-substitution and model cost change
    launch_backsub(e);
    { LaunchTimer t(e, KID_MISC);
      KL(e), bak::k_reduce_partials<<<1, 1024, 0, s>>>(e->d_part0.p, e->nTiles, &sc->mcc); }
    return e->h_scal->pcg.k;
}

void swap_params(ba_engine* e) {
    std::swap(e->d_pose.p, e->d_cpose.p); std::swap(e->d_intr.p, e->d_cintr.p); std::swap(e->d_pt.p, e->d_cpt.p);
}

}  // namespace

namespace {

// ---- the Levenberg-Marquardt loop: Ceres 2.2 TrustRegionMinimizer + LevenbergMarquardtStrategy (SURVEY §8 a-LM) ----
void lm_solve(ba_engine* e) {
    const ba_options& O = e->opt;
    cudaStream_t s = e->stream;
    bak::Scalars* sc = e->d_scal.p;
    bak::Scalars* H = e->h_scal;
    ba_summary& S = e->summary;
    std::memset(&S, 0, sizeof S);
    S.num_residuals = 2 * e->NoG;
    e->traj.clear();
    e->kernel_launches = 0;
    e->host_syncs = 0;
    CU_TRY(cudaEventRecord(e->ev_t0, s));
    auto finish = [&](int term, const char* msg, double cost) {
        S.termination = term; S.usable = (term != BA_TERM_FAILURE); S.final_cost = cost; S.num_iterations = int(e->traj.size());
        if (!e->traj.empty() && e->traj.back().radius > 0.0) e->last_radius = e->traj.back().radius;
        std::snprintf(S.message, sizeof S.message, "%s", msg);
        launch_check(__FILE__, __LINE__);
        CU_TRY(cudaEventRecord(e->ev_t1, s));
        CU_TRY(cudaEventSynchronize(e->ev_t1));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, e->ev_t0, e->ev_t1));
        S.solve_seconds = ms * 1e-3;
        S.kernel_launches = e->kernel_launches;
        resolve_timers(e);
    };

    // ---- iteration 0: evaluate, Jacobi scales, gradient norm, ||x|| ----
    e->d_yp.zero(s); e->d_x.zero(s);
    evaluate_full(e);
    {   // s_j = 1/(1 + sqrt(H_jj)) from the iteration-0 diagonals, frozen for the solve
        LaunchTimer t(e, KID_MISC);
        if (e->Np > 0) KL(e), bak::k_scales_points<<<cdiv(3 * e->Np, 256), 256, 0, s>>>(e->V, O.jacobi_scaling);
        KL(e), bak::k_scales_cameras<<<cdiv(e->nf, 256), 256, 0, s>>>(e->Nc, e->Ng, e->fi, e->Hcc(), e->Hii(), e->d_sc_f.p, O.jacobi_scaling);
    }
    gradient_max(e);
    make_candidate(e, e->d_x.p);   // y = 0: gives ||x|| of the free parameters
    allreduce_sum(e, &sc->ccost, 4);
    read_scalars(e);
    double cost = H->cost;
    S.initial_cost = cost;
    if (H->bad_eval || !std::isfinite(cost)) { finish(BA_TERM_FAILURE, "initial residual/Jacobian evaluation failed (non-finite)", cost); return; }
    double gmax = std::max(H->gmax_pt, H->gmax_f);
    double x_norm = std::sqrt(H->xn2_pt + H->xn2_f);
    double radius = (O.warm_start_radius && e->last_radius > 0.0) ? e->last_radius : O.initial_trust_region_radius, decrease_factor = 2.0;
    int num_invalid = 0;
    {
        ba_iter_record R{}; R.step_is_valid = 1; R.step_is_successful = 1; R.cost = cost; R.radius = radius; R.radius_used = radius; R.gradient_max_norm = gmax;
        e->traj.push_back(R); S.num_successful_steps = 1;
    }
    int iteration = 0;
    for (;;) {
        const ba_iter_record last = e->traj.back();
        if (iteration >= O.max_num_iterations) { finish(BA_TERM_NO_CONVERGENCE, "maximum number of iterations reached", cost); return; }
        if (last.step_is_successful && gmax <= O.gradient_tolerance) { finish(BA_TERM_CONVERGENCE, "gradient tolerance reached", cost); return; }
        if (radius <= O.min_trust_region_radius) { finish(BA_TERM_CONVERGENCE, "minimum trust region radius reached", cost); return; }
        if (O.target_chi2 > 0.0 && 2.0 * cost <= O.target_chi2) { finish(BA_TERM_TARGET_REACHED, "target chi2 reached", cost); return; }
        ++iteration;
        ba_iter_record R{}; R.iteration = iteration; R.radius_used = radius; R.gradient_max_norm = gmax; R.cost = cost;

        double rel = 0.0;
        const int its = solve_linear(e, radius, &rel);
        R.pcg_iterations = std::max(its, 0); R.pcg_rel_residual = rel; S.total_pcg_iterations += std::max(its, 0);
        bool valid = its >= 0;
        if (valid) {   // launch candidate + its cost right away; one read-back serves model change, norms and candidate cost
            make_candidate(e, e->d_x.p);
            candidate_cost(e);
            if (O.world_size > 1) {
                NCCL_TRY(g_nccl.GroupStart());
                allreduce_sum(e, &sc->ccost, 4);
                allreduce_flags(e, &sc->bad_cand, 1);
                NCCL_TRY(g_nccl.GroupEnd());
            }
            read_scalars(e);
            R.model_cost_change = -H->mcc;
            valid = !H->bad_lin && std::isfinite(R.model_cost_change) && R.model_cost_change > 0.0;
        }
        R.step_is_valid = valid;
        if (!valid) {   // HandleInvalidStep
            if (++num_invalid >= O.max_num_consecutive_invalid_steps) { e->traj.push_back(R); ++S.num_unsuccessful_steps; finish(BA_TERM_FAILURE, "too many consecutive invalid steps", cost); return; }
            radius /= decrease_factor; decrease_factor *= 2.0;
            R.radius = radius; e->traj.push_back(R); ++S.num_unsuccessful_steps;
            continue;
        }
        num_invalid = 0;
        double ccost = H->ccost;
        if (H->bad_cand || !std::isfinite(ccost)) ccost = std::numeric_limits<double>::max();
        R.candidate_cost = ccost;
        R.step_norm = std::sqrt(H->step2_pt + H->step2_f);
        // TrustRegionMinimizer::ParameterToleranceReached / FunctionToleranceReached, in this order, before the accept test, on every valid
        // step: Ceres 2.2 has NO "at least one successful step" guard on these two tests (SURVEY 8 a-LM lists one; that guard is not in the
        // 2.2 minimizer and is deliberately not implemented here) — restated from the published source, not pinned by a Ceres run.
        if (R.step_norm <= O.parameter_tolerance * (x_norm + O.parameter_tolerance)) { finish(BA_TERM_CONVERGENCE, "parameter tolerance reached", cost); return; }
        if (std::fabs(cost - ccost) <= O.function_tolerance * cost) { finish(BA_TERM_CONVERGENCE, "function tolerance reached", cost); return; }
        R.relative_decrease = (ccost >= std::numeric_limits<double>::max()) ? std::numeric_limits<double>::lowest() : (cost - ccost) / R.model_cost_change;
        if (R.relative_decrease > O.min_relative_decrease) {   // HandleSuccessfulStep
            swap_params(e);
            x_norm = std::sqrt(H->xn2_pt + H->xn2_f);
            evaluate_full(e);
            gradient_max(e);
            read_scalars(e);
            cost = H->cost;
            if (H->bad_eval || !std::isfinite(cost)) { finish(BA_TERM_FAILURE, "residual/Jacobian evaluation failed after an accepted step", cost); return; }
            gmax = std::max(H->gmax_pt, H->gmax_f);
            radius = std::min(O.max_trust_region_radius, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * R.relative_decrease - 1.0, 3)));
            decrease_factor = 2.0;
            R.step_is_successful = 1; R.cost = cost; R.gradient_max_norm = gmax; ++S.num_successful_steps;
        } else {
            radius /= decrease_factor; decrease_factor *= 2.0; ++S.num_unsuccessful_steps;
        }
        R.radius = radius;
        e->traj.push_back(R);
    }
}

// ---- g2o policy: OptimizationAlgorithmLevenberg::solve + VertexSE3Expmap (same rules as oracle/ba_oracle.cpp::lm_solve_g2o) ----
void lm_solve_g2o(ba_engine* e) {
    const ba_options& O = e->opt;
    cudaStream_t s = e->stream;
    bak::Scalars* sc = e->d_scal.p;
    bak::Scalars* H = e->h_scal;
    ba_summary& S = e->summary;
    std::memset(&S, 0, sizeof S);
    S.num_residuals = 2 * e->NoG;
    e->traj.clear();
    e->kernel_launches = 0;
    e->host_syncs = 0;
    CU_TRY(cudaEventRecord(e->ev_t0, s));
    auto finish = [&](int term, const char* msg, double cost) {
        S.termination = term; S.usable = (term != BA_TERM_FAILURE); S.final_cost = cost; S.num_iterations = int(e->traj.size());
        std::snprintf(S.message, sizeof S.message, "%s", msg);
        launch_check(__FILE__, __LINE__);
        CU_TRY(cudaEventRecord(e->ev_t1, s));
        CU_TRY(cudaEventSynchronize(e->ev_t1));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, e->ev_t0, e->ev_t1));
        S.solve_seconds = ms * 1e-3;
        S.kernel_launches = e->kernel_launches;
        resolve_timers(e);
    };
    e->d_yp.zero(s); e->d_x.zero(s);
    evaluate_full(e);
    {   // no Jacobi scaling in g2o: unit scales; lambda0 = 1e-5 * max diag(H)
        LaunchTimer t(e, KID_MISC);
        if (e->Np > 0) KL(e), bak::k_scales_points<<<cdiv(3 * e->Np, 256), 256, 0, s>>>(e->V, 0);
        KL(e), bak::k_scales_cameras<<<cdiv(e->nf, 256), 256, 0, s>>>(e->Nc, e->Ng, e->fi, e->Hcc(), e->Hii(), e->d_sc_f.p, 0);
        KL(e), bak::k_maxdiag_points<<<kNormBlocks, 256, 0, s>>>(e->V, e->d_part1.p);
        KL(e), bak::k_maxdiag_final<<<1, 1024, 0, s>>>(e->Nc, e->Hcc(), e->d_part1.p, kNormBlocks, sc);
        allreduce_max(e, &sc->maxdiag_pt, 1);
    }
    read_scalars(e);
    double cost = H->cost;
    S.initial_cost = cost;
    if (H->bad_eval || !std::isfinite(cost)) { finish(BA_TERM_FAILURE, "initial residual/Jacobian evaluation failed (non-finite)", cost); return; }
    double lambda = 1e-5 * std::max(H->maxdiag_pt, H->maxdiag_f), nu = 2.0;
    { ba_iter_record R{}; R.step_is_valid = 1; R.step_is_successful = 1; R.cost = cost; e->traj.push_back(R); S.num_successful_steps = 1; }
    int record = 0;
    for (int iteration = 0; iteration < O.max_num_iterations; ++iteration) {
        if (O.target_chi2 > 0.0 && 2.0 * cost <= O.target_chi2) { finish(BA_TERM_TARGET_REACHED, "target chi2 reached", cost); return; }
        double rho = 0.0;
        int qmax = 0;
        do {
            ba_iter_record R{}; R.iteration = ++record; R.radius_used = 1.0 / lambda; R.cost = cost;
            double rel = 0.0;
            const int its = solve_linear(e, 1.0 / lambda, &rel, lambda);
            R.pcg_iterations = std::max(its, 0); R.pcg_rel_residual = rel; S.total_pcg_iterations += std::max(its, 0);
            double tempChi = std::numeric_limits<double>::max(), scale = 0.0;
            if (its >= 0) {
                make_candidate(e, e->d_x.p, lambda);
                candidate_cost(e);
                if (O.world_size > 1) {
                    NCCL_TRY(g_nccl.GroupStart());
                    allreduce_sum(e, &sc->ccost, 1);
                    allreduce_sum(e, &sc->scale_pt, 1);
                    allreduce_flags(e, &sc->bad_cand, 1);
                    NCCL_TRY(g_nccl.GroupEnd());
                }
                read_scalars(e);
                if (!H->bad_cand && !H->bad_lin && std::isfinite(H->ccost)) tempChi = 2.0 * H->ccost;
                scale = H->scale_pt + H->scale_f;
            }
            R.step_is_valid = its >= 0 && !H->bad_lin;
            R.candidate_cost = tempChi >= std::numeric_limits<double>::max() ? tempChi : 0.5 * tempChi;
            R.model_cost_change = 0.5 * scale;
            rho = (2.0 * cost - tempChi) / (scale + 1e-3);
            R.relative_decrease = rho;
            if (rho > 0.0 && std::isfinite(tempChi)) {
                lambda *= std::max(1.0 / 3.0, std::min(1.0 - std::pow(2.0 * rho - 1.0, 3), 2.0 / 3.0)); nu = 2.0;
                swap_params(e);
                evaluate_full(e);
                read_scalars(e);
                cost = H->cost;
                if (H->bad_eval || !std::isfinite(cost)) { finish(BA_TERM_FAILURE, "residual/Jacobian evaluation failed after an accepted step", cost); return; }
                R.step_is_successful = 1; R.cost = cost; ++S.num_successful_steps;
            } else {
                lambda *= nu; nu *= 2.0; ++S.num_unsuccessful_steps;
            }
            R.radius = 1.0 / lambda;
            e->traj.push_back(R);
            if (!std::isfinite(lambda)) break;
            ++qmax;
        } while (rho < 0.0 && qmax < 10);
        if (qmax == 10 || rho == 0.0 || !std::isfinite(lambda)) { finish(BA_TERM_CONVERGENCE, "g2o: no further improvement (trial limit, zero gain or lambda overflow)", cost); return; }
    }
    finish(BA_TERM_NO_CONVERGENCE, "maximum number of iterations reached", cost);
}

template <class F>
int guarded(ba_engine* e, F&& f) {
    try { f(); return BA_OK; }
    catch (const BaFail& x) { return fail(e, x.code, x.msg); }
    catch (const std::exception& x) { return fail(e, BA_ERR_INTERNAL, x.what()); }
    catch (...) { return fail(e, BA_ERR_INTERNAL, "unknown exception"); }
}

}  // namespace

// ------------------------------------------------------------------------------------------------------------
// C-ABI
extern "C" {

int ba_options_default(ba_options* o) {
    if (!o) return BA_ERR_INVALID_ARGUMENT;
    std::memset(o, 0, sizeof *o);
    o->abi_version = BA_ABI_VERSION;
    o->model = BA_MODEL_PINHOLE;
    o->robust_kind = BA_ROBUST_HUBER; o->robust_a = 4.0;          // BundleAdjuster.h:109
    o->intrinsics_free = 1;                                        // :106-107
    o->taylor_rule = BA_TAYLOR_THETA2_LE_EPS;
    o->max_num_iterations = 50; o->initial_trust_region_radius = 1e4; o->max_trust_region_radius = 1e16; o->min_trust_region_radius = 1e-32;
    o->min_relative_decrease = 1e-3; o->min_lm_diagonal = 1e-6; o->max_lm_diagonal = 1e32; o->max_num_consecutive_invalid_steps = 5;
    o->function_tolerance = 1e-6; o->gradient_tolerance = 1e-10; o->parameter_tolerance = 1e-8; o->jacobi_scaling = 1;
    o->linear_solver = BA_LINEAR_AUTO;                             // direct solve for small camera counts, PCG beyond
    o->preconditioner = BA_PRECOND_JACOBI;                         // :171
    o->pcg_max_iterations = 2000; o->pcg_rel_tolerance = 1e-12; o->pcg_forcing = 0; o->eta = 0.1;
    o->target_chi2 = 0.0; o->device = 0; o->num_threads = 1;
    o->rank = 0; o->world_size = 1; o->nccl_unique_id = nullptr; o->stream = nullptr; o->scatter_mode = 0; o->p2p_allreduce = 0; o->lm_policy = BA_POLICY_CERES;
    return BA_OK;
}

int ba_create(const ba_options* opt, ba_engine** out) {
    if (!opt || !out) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "null argument");
    if (opt->abi_version != BA_ABI_VERSION) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "abi_version mismatch");
    if (opt->linear_solver < BA_LINEAR_EXACT || opt->linear_solver > BA_LINEAR_AUTO) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "unknown linear_solver");
    if (opt->model != BA_MODEL_PINHOLE && opt->model != BA_MODEL_SNAVELY) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "unknown residual model");
    if (opt->lm_policy == BA_POLICY_G2O && opt->model != BA_MODEL_PINHOLE) return fail(nullptr, BA_ERR_UNSUPPORTED, "g2o policy: EdgeSE3ProjectXYZ is the pinhole model");
    if (opt->lm_policy == BA_POLICY_G2O && opt->intrinsics_free) return fail(nullptr, BA_ERR_UNSUPPORTED, "g2o policy: EdgeSE3ProjectXYZ has constant pinhole intrinsics (set intrinsics_free = 0)");
    if (opt->preconditioner != BA_PRECOND_JACOBI && opt->preconditioner != BA_PRECOND_SCHUR_JACOBI) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "unknown preconditioner");
    if (opt->world_size < 1 || opt->rank < 0 || opt->rank >= opt->world_size) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "bad rank/world_size");
    if (opt->world_size > 1 && !opt->nccl_unique_id) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "world_size > 1 needs nccl_unique_id");
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(nullptr, BA_ERR_CUDA, std::string("no usable CUDA device (") + cudaGetErrorString(ce) + ") — this engine has no CPU fallback");
    if (opt->device < 0 || opt->device >= ndev) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "device ordinal out of range");
    ba_engine* e = new (std::nothrow) ba_engine();
    if (!e) return fail(nullptr, BA_ERR_INTERNAL, "out of memory");
    e->opt = *opt; e->opt.nccl_unique_id = nullptr;
    e->device = opt->device;
    const int rc = guarded(e, [&] {
        CU_TRY(cudaSetDevice(e->device));
        cudaDeviceProp prop{};
        CU_TRY(cudaGetDeviceProperties(&prop, e->device));
        e->num_sms = prop.multiProcessorCount;
        auto known = [](int m) { return m == 1 || m == 4 || m == 6; };
        e->scatter_mode = known(opt->scatter_mode) ? opt->scatter_mode : 0;   // 0: default
        if (e->scatter_mode == 0) if (const char* v = std::getenv("BA_SCATTER_MODE")) { const int m = std::atoi(v); if (known(m)) e->scatter_mode = m; }   // A/B knob for the default
#ifdef BA_WITH_STREAM
        set_stream_attrs(e);
#else
        if (e->scatter_mode == 6) throw BaFail{BA_ERR_UNSUPPORTED, "scatter_mode 6 (the cp.async.bulk streaming kernel) is not compiled into this library (build with -DBA_WITH_STREAM)"};
#endif
        if (const char* g = std::getenv("BA_L2_FETCH")) CU_TRY(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, size_t(std::atoi(g))));   // experiment knob: 32 / 64 / 128
        if (const char* v = std::getenv("BA_K5_SAMPLE")) e->k5_sample_stride = std::max(1, std::atoi(v));
        e->pdl = BA_PDL_DEFAULT;
        if (const char* v = std::getenv("BA_PDL")) e->pdl = std::atoi(v) ? 1 : 0;
        if (const char* v = std::getenv("BA_TAIL_MODE1")) e->dbg_tail1 = std::atoi(v) ? 1 : 0;
        if (const char* v = std::getenv("BA_POSE_FACTORED")) e->factored_opt = std::atoi(v) ? 1 : 0;
        {   // a 16-CTA cluster is a non-portable size: opt in, then ask whether this device can co-schedule one
            int want = bak::TAIL_CL_WIDE;
            if (const char* v = std::getenv("BA_TAIL_CL")) want = std::atoi(v);
            if (const char* v = std::getenv("BA_TAIL_WIDE_MIN_NF")) e->tail_wide_min_nf = std::max(0, std::atoi(v));
            e->tail_cl = bak::TAIL_CL;
            if (want == bak::TAIL_CL_WIDE && cudaFuncSetAttribute(bak::k_pcg_tail<bak::TAIL_CL_WIDE>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
                cudaLaunchConfig_t cfg{};
                cfg.gridDim = dim3(bak::TAIL_CL_WIDE); cfg.blockDim = dim3(bak::TAIL_THREADS); cfg.dynamicSmemBytes = 0;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = bak::TAIL_CL_WIDE; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                cfg.attrs = at; cfg.numAttrs = 1;
                int nclusters = 0;
                if (cudaOccupancyMaxActiveClusters(&nclusters, bak::k_pcg_tail<bak::TAIL_CL_WIDE>, &cfg) != cudaSuccess) {   // the kernel carries __cluster_dims__: ask again without the attribute
                    (void)cudaGetLastError();
                    cfg.numAttrs = 0; nclusters = 0;
                    if (cudaOccupancyMaxActiveClusters(&nclusters, bak::k_pcg_tail<bak::TAIL_CL_WIDE>, &cfg) != cudaSuccess) nclusters = 0;
                }
                if (nclusters >= 1) e->tail_cl = bak::TAIL_CL_WIDE;
            }
            (void)cudaGetLastError();   // a refusal above is not an error of this engine
        }
        if (const char* v = std::getenv("BA_P2P_WIDE")) e->p2p_wide = std::atoi(v) ? 1 : 0;
        // sm_100a images are architecture-specific: they run on compute capability 10.0 only (not on 10.3 / 12.x); probe one kernel as well
        if (prop.major != 10 || prop.minor != 0)
            throw BaFail{BA_ERR_CUDA, std::string("device ") + prop.name + " has compute capability " + std::to_string(prop.major) + "." + std::to_string(prop.minor) +
                                          "; this library holds sm_100a code only (compute capability 10.0) and has no CPU fallback"};
        {
            cudaFuncAttributes fa{};
            const cudaError_t pe = cudaFuncGetAttributes(&fa, bak::k0_camera_precompute);
            if (pe != cudaSuccess) { cudaGetLastError(); throw BaFail{BA_ERR_CUDA, std::string("no sm_100a kernel image usable on ") + prop.name + ": " + cudaGetErrorString(pe)}; }
        }
        if (opt->stream) e->stream = static_cast<cudaStream_t>(opt->stream);
        else { CU_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking)); e->own_stream = true; }
        CU_TRY(cudaMallocHost(reinterpret_cast<void**>(&e->h_scal), sizeof(bak::Scalars)));
        CU_TRY(cudaEventCreate(&e->ev_t0)); CU_TRY(cudaEventCreate(&e->ev_t1));
        if (opt->world_size > 1) {
            std::string err;
            if (!g_nccl.load(err)) throw BaFail{BA_ERR_NCCL, err};
            Nccl::UniqueId id; std::memcpy(&id, opt->nccl_unique_id, sizeof id);
            NCCL_TRY(g_nccl.CommInitRank(&e->comm, opt->world_size, id, opt->rank));
        }
    });
    if (rc != BA_OK) { g_create_error = e->err; ba_destroy(e); return rc; }
    *out = e;
    return BA_OK;
}

int ba_set_problem(ba_engine* e, const ba_problem* p) {
    if (!e || !p) return fail(e, BA_ERR_INVALID_ARGUMENT, "null argument");
    if (p->num_cameras < 0 || p->num_groups < 0 || p->num_points < 0 || p->num_observations < 0) return fail(e, BA_ERR_INVALID_ARGUMENT, "negative size");
    const int Nc = p->num_cameras, Ng = p->num_groups, Np = p->num_points, No = p->num_observations;
    if ((Nc && (!p->pose || !p->cam_group || !p->cam_fixed)) || (Ng && !p->intr) || (Np && !p->pt) || (No && (!p->obs_cam || !p->obs_pt || !p->obs_uv)))
        return fail(e, BA_ERR_INVALID_ARGUMENT, "null array");
    for (int c = 0; c < Nc; ++c) if (p->cam_group[c] < 0 || p->cam_group[c] >= Ng) return fail(e, BA_ERR_INVALID_ARGUMENT, "cam_group out of range");
    for (int o = 0; o < No; ++o) {
        if (p->obs_cam[o] < 0 || p->obs_cam[o] >= Nc || p->obs_pt[o] < 0 || p->obs_pt[o] >= Np) return fail(e, BA_ERR_INVALID_ARGUMENT, "observation index out of range");
        if (o && p->obs_pt[o] < p->obs_pt[o - 1]) return fail(e, BA_ERR_INVALID_ARGUMENT, "observations must be sorted by point id");
    }
    e->has_problem = false;
    return guarded(e, [&] { do_set_problem(e, p); });
}

int ba_append(ba_engine* e, const ba_delta* d) {
    if (!e || !d) return fail(e, BA_ERR_INVALID_ARGUMENT, "null argument");
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (e->opt.world_size > 1) return fail(e, BA_ERR_UNSUPPORTED, "ba_append runs on one GPU (world_size = 1)");
    ba_merge::Merged M;
    const std::string why = ba_merge::merge(e->Nc, e->Ng, e->NpG, e->NoG, e->h_cam_group.data(), e->h_cam_fixed.data(), e->h_obs_cam.data(), e->h_obs_pt.data(),
                                            e->h_obs_uv.data(), *d, M);
    if (!why.empty()) return fail(e, BA_ERR_INVALID_ARGUMENT, "ba_append: " + why);
    return guarded(e, [&] {
        CU_TRY(cudaSetDevice(e->device));
        cudaStream_t s = e->stream;
        // merged parameters, assembled on the device: [current state | new entries]; only the new entries cross PCIe
        DevBuf<double> t_pose, t_intr, t_pt;
        t_pose.alloc(size_t(6) * M.Nc); t_intr.alloc(size_t(4) * M.Ng); t_pt.alloc(size_t(3) * M.Np);
        auto join = [&](DevBuf<double>& t, const double* cur, size_t n_old, const double* add, size_t n_new) {
            if (n_old) CU_TRY(cudaMemcpyAsync(t.p, cur, n_old * sizeof(double), cudaMemcpyDeviceToDevice, s));
            if (n_new) { CU_TRY(cudaMemcpyAsync(t.p + n_old, add, n_new * sizeof(double), cudaMemcpyHostToDevice, s)); g_h2d_bytes += int64_t(n_new * sizeof(double)); }
        };
        join(t_pose, e->d_pose.p, size_t(6) * e->Nc, d->pose, size_t(6) * d->num_new_cameras);
        join(t_intr, e->d_intr.p, size_t(4) * e->Ng, d->intr, size_t(4) * d->num_new_groups);
        join(t_pt, e->d_pt.p, size_t(3) * e->Np, d->pt, size_t(3) * d->num_new_points);
        ba_problem P{};
        P.num_cameras = M.Nc; P.num_groups = M.Ng; P.num_points = M.Np; P.num_observations = M.No;
        P.cam_group = M.cam_group.data(); P.cam_fixed = M.cam_fixed.data();
        P.obs_cam = M.obs_cam.data(); P.obs_pt = M.obs_pt.data(); P.obs_uv = M.obs_uv.data();
        const DevState dev{t_pose.p, t_intr.p, t_pt.p};
        e->has_problem = false;
        do_set_problem(e, &P, &dev);
        CU_TRY(cudaStreamSynchronize(s));   // the temporaries go out of scope
    });
}

int ba_update_parameters(ba_engine* e, int32_t kind, int32_t first, int32_t count, const double* values) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (e->opt.world_size > 1) return fail(e, BA_ERR_UNSUPPORTED, "ba_update_parameters runs on one GPU (world_size = 1)");
    if (kind < 0 || kind > 2) return fail(e, BA_ERR_INVALID_ARGUMENT, "ba_update_parameters: kind must be 0 (poses), 1 (intrinsics) or 2 (points)");
    const int w = kind == 0 ? 6 : (kind == 1 ? 4 : 3), n = kind == 0 ? e->Nc : (kind == 1 ? e->Ng : e->Np);
    if (first < 0 || count < 0 || int64_t(first) + count > n || (count && !values)) return fail(e, BA_ERR_INVALID_ARGUMENT, "ba_update_parameters: range out of bounds");
    if (!count) return BA_OK;
    return guarded(e, [&] {
        CU_TRY(cudaSetDevice(e->device));
        double* cur = kind == 0 ? e->d_pose.p : (kind == 1 ? e->d_intr.p : e->d_pt.p);
        double* ini = kind == 0 ? e->d_pose0.p : (kind == 1 ? e->d_intr0.p : e->d_pt0.p);
        const size_t bytes = size_t(w) * count * sizeof(double), off = size_t(w) * first;
        CU_TRY(cudaMemcpyAsync(cur + off, values, bytes, cudaMemcpyHostToDevice, e->stream));
        CU_TRY(cudaMemcpyAsync(ini + off, cur + off, bytes, cudaMemcpyDeviceToDevice, e->stream));
        CU_TRY(cudaStreamSynchronize(e->stream));   // `values` may be pageable and reused by the caller
        g_h2d_bytes += int64_t(bytes);
    });
}

int ba_get_observations(ba_engine* e, int32_t* obs_cam, int32_t* obs_pt, double* obs_uv, int32_t cap, int32_t* n) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (n) *n = e->NoG;
    const int m = std::max(0, std::min<int>(cap, e->NoG));
    if (obs_cam) std::copy(e->h_obs_cam.begin(), e->h_obs_cam.begin() + m, obs_cam);
    if (obs_pt) std::copy(e->h_obs_pt.begin(), e->h_obs_pt.begin() + m, obs_pt);
    if (obs_uv) std::copy(e->h_obs_uv.begin(), e->h_obs_uv.begin() + size_t(2) * m, obs_uv);
    return BA_OK;
}

int ba_reset_state(ba_engine* e) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    return guarded(e, [&] {
        CU_TRY(cudaSetDevice(e->device));
        cudaStream_t s = e->stream;
        CU_TRY(cudaMemcpyAsync(e->d_pose.p, e->d_pose0.p, size_t(6) * e->Nc * sizeof(double), cudaMemcpyDeviceToDevice, s));
        CU_TRY(cudaMemcpyAsync(e->d_intr.p, e->d_intr0.p, size_t(4) * e->Ng * sizeof(double), cudaMemcpyDeviceToDevice, s));
        CU_TRY(cudaMemcpyAsync(e->d_pt.p, e->d_pt0.p, size_t(3) * e->Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
        e->traj.clear();
    });
}

int ba_solve(ba_engine* e, ba_summary* out) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    const int rc = guarded(e, [&] { CU_TRY(cudaSetDevice(e->device)); if (e->opt.lm_policy == BA_POLICY_G2O) lm_solve_g2o(e); else lm_solve(e); });
    if (out) *out = e->summary;
    return rc;
}

int ba_get_state(ba_engine* e, double* pose, double* intr, double* pt) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    return guarded(e, [&] {
        CU_TRY(cudaSetDevice(e->device));
        cudaStream_t s = e->stream;
        if (pose) { CU_TRY(cudaMemcpyAsync(pose, e->d_pose.p, size_t(6) * e->Nc * sizeof(double), cudaMemcpyDeviceToHost, s)); g_d2h_bytes += int64_t(48) * e->Nc; }
        if (intr) { CU_TRY(cudaMemcpyAsync(intr, e->d_intr.p, size_t(4) * e->Ng * sizeof(double), cudaMemcpyDeviceToHost, s)); g_d2h_bytes += int64_t(32) * e->Ng; }
        if (pt) g_d2h_bytes += int64_t(24) * e->NpG;
        if (pt) {
            if (e->opt.world_size == 1) CU_TRY(cudaMemcpyAsync(pt, e->d_pt.p, size_t(3) * e->Np * sizeof(double), cudaMemcpyDeviceToHost, s));
            else {   // every rank ends with the full point array: one broadcast per shard (collective: all ranks must call)
                e->d_ptfull.alloc(size_t(3) * e->NpG);
                CU_TRY(cudaMemcpyAsync(e->d_ptfull.p + size_t(3) * e->pb, e->d_pt.p, size_t(3) * e->Np * sizeof(double), cudaMemcpyDeviceToDevice, s));
                NCCL_TRY(g_nccl.GroupStart());
                for (int r = 0; r < e->opt.world_size; ++r) {
                    double* seg = e->d_ptfull.p + size_t(3) * e->h_shard_pt[r];
                    const size_t cnt = size_t(3) * (e->h_shard_pt[r + 1] - e->h_shard_pt[r]);
                    if (cnt) NCCL_TRY(g_nccl.Broadcast(seg, seg, cnt, Nccl::kFloat64, r, e->comm, s));
                }
                NCCL_TRY(g_nccl.GroupEnd());
                CU_TRY(cudaMemcpyAsync(pt, e->d_ptfull.p, size_t(3) * e->NpG * sizeof(double), cudaMemcpyDeviceToHost, s));
            }
        }
        CU_TRY(cudaStreamSynchronize(s));
    });
}

int ba_get_trajectory(ba_engine* e, ba_iter_record* out, int32_t cap, int32_t* n) {
    if (!e || !n) return BA_ERR_INVALID_ARGUMENT;
    *n = int32_t(e->traj.size());
    if (out) for (int i = 0; i < std::min<int>(cap, *n); ++i) out[i] = e->traj[i];
    return BA_OK;
}

int ba_get_structure(ba_engine* e, int32_t* perm_by_cam, int32_t* cam_off, int32_t* pt_off, int32_t* schur_pairs, int64_t pairs_cap, int64_t* n_pairs) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    // One GPU: the arrays the kernels consume, read back from the device (d_perm / d_cam_off / d_pt_off as the threaded counting sorts
    // of ba_set_problem built them).  Several GPUs: every rank holds the schedule of its own shard only, so the global canonical
    // structure (independent of sharding) is rebuilt on the host from the index arrays kept at ba_set_problem.
    const int Nc = e->Nc, Np = e->NpG, No = e->NoG;
    vector<int32_t> po(size_t(Np) + 1, 0), co(size_t(Nc) + 1, 0);
    if (e->opt.world_size == 1) {
        const int rc = guarded(e, [&] {
            CU_TRY(cudaSetDevice(e->device));
            CU_TRY(cudaStreamSynchronize(e->stream));
            CU_TRY(cudaMemcpy(po.data(), e->d_pt_off.p, po.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
            CU_TRY(cudaMemcpy(co.data(), e->d_cam_off.p, co.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
            if (perm_by_cam && No) CU_TRY(cudaMemcpy(perm_by_cam, e->d_perm.p, size_t(No) * sizeof(int32_t), cudaMemcpyDeviceToHost));
            g_d2h_bytes += int64_t(po.size() + co.size() + (perm_by_cam ? No : 0)) * 4;
        });
        if (rc != BA_OK) return rc;
    } else {
        for (int o = 0; o < No; ++o) { ++po[e->h_obs_pt[o] + 1]; ++co[e->h_obs_cam[o] + 1]; }
        for (int p = 0; p < Np; ++p) po[p + 1] += po[p];
        for (int c = 0; c < Nc; ++c) co[c + 1] += co[c];
        if (perm_by_cam) { vector<int32_t> cur(co.begin(), co.end() - 1); for (int o = 0; o < No; ++o) perm_by_cam[cur[e->h_obs_cam[o]]++] = o; }
    }
    if (pt_off) std::copy(po.begin(), po.end(), pt_off);
    if (cam_off) std::copy(co.begin(), co.end(), cam_off);
    if (schur_pairs || n_pairs) {
        // co-visibility bitset, one row per pose: bit j of row i set iff i<j are both free and share a point
        const size_t words = (size_t(Nc) + 63) / 64;
        vector<uint64_t> bits(size_t(Nc) * words, 0);
        for (int p = 0; p < Np; ++p)
            for (int a = po[p]; a < po[p + 1]; ++a) {
                const int i = e->h_obs_cam[a];
                if (!e->h_cam_free[i]) continue;
                for (int b = po[p]; b < po[p + 1]; ++b) { const int j = e->h_obs_cam[b]; if (j > i && e->h_cam_free[j]) bits[size_t(i) * words + (j >> 6)] |= uint64_t(1) << (j & 63); }
            }
        int64_t n = 0;
        for (int i = 0; i < Nc; ++i)
            for (size_t w = 0; w < words; ++w) {
                uint64_t v = bits[size_t(i) * words + w];
                while (v) { const int j = int(w * 64) + __builtin_ctzll(v); v &= v - 1; if (schur_pairs && n < pairs_cap) { schur_pairs[2 * n] = i; schur_pairs[2 * n + 1] = j; } ++n; }
            }
        if (n_pairs) *n_pairs = n;
    }
    return BA_OK;
}

int ba_evaluate(ba_engine* e, double* cost, double* r, double* Jc, double* Jp, double* Ji, double* g) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    if (e->opt.world_size != 1) return fail(e, BA_ERR_UNSUPPORTED, "ba_evaluate is a single-GPU debugging entry point");
    int bad = 0;
    const int rc = guarded(e, [&] {
        CU_TRY(cudaSetDevice(e->device));
        evaluate_full(e, true);
        read_scalars(e);
        bad = e->h_scal->bad_eval;
        if (cost) *cost = e->h_scal->cost;
        const int No = e->No, ld = e->ldo, pk = e->jpack, ncol = (e->fi ? 13 : 9) - pk;
        vector<double2> hJ(size_t(ncol) * ld), hr(ld);
        CU_TRY(cudaMemcpy(hJ.data(), e->d_J2.p, hJ.size() * sizeof(double2), cudaMemcpyDeviceToHost));
        CU_TRY(cudaMemcpy(hr.data(), e->d_r2.p, hr.size() * sizeof(double2), cudaMemcpyDeviceToHost));
        for (int o = 0; o < No; ++o) {
            if (r) { r[2 * o] = hr[o].x; r[2 * o + 1] = hr[o].y; }
            if (Jc) for (int k = 0; k < 6; ++k) {
                double2 v = hJ[size_t(k - (pk && k > 3 ? 1 : 0)) * ld + o];
                if (pk && k == 3) v.y = 0.0;      // packed translation columns (load_jc): stored (d r0/d tx, d r1/d ty)
                if (pk && k == 4) v.x = 0.0;
                Jc[(2 * o) * 6 + k] = v.x; Jc[(2 * o + 1) * 6 + k] = v.y;
            }
            if (Jp) for (int k = 0; k < 3; ++k) { Jp[(2 * o) * 3 + k] = hJ[size_t(6 - pk + k) * ld + o].x; Jp[(2 * o + 1) * 3 + k] = hJ[size_t(6 - pk + k) * ld + o].y; }
            if (Ji) for (int k = 0; k < 4; ++k) { Ji[(2 * o) * 4 + k] = e->fi ? hJ[size_t(9 - pk + k) * ld + o].x : 0.0; Ji[(2 * o + 1) * 4 + k] = e->fi ? hJ[size_t(9 - pk + k) * ld + o].y : 0.0; }
        }
        if (g) {
            CU_TRY(cudaMemcpy(g, e->gc(), size_t(6) * e->Nc * sizeof(double), cudaMemcpyDeviceToHost));
            if (e->fi) CU_TRY(cudaMemcpy(g + size_t(6) * e->Nc, e->gi(), size_t(4) * e->Ng * sizeof(double), cudaMemcpyDeviceToHost));
            else std::fill(g + size_t(6) * e->Nc, g + size_t(6) * e->Nc + size_t(4) * e->Ng, 0.0);
            CU_TRY(cudaMemcpy(g + size_t(6) * e->Nc + size_t(4) * e->Ng, e->d_gp.p, size_t(3) * e->Np * sizeof(double), cudaMemcpyDeviceToHost));
        }
    });
    if (rc != BA_OK) return rc;
    return bad ? fail(e, BA_ERR_EVAL_FAILED, "non-finite residual or Jacobian") : BA_OK;
}

int ba_set_profiling(ba_engine* e, int32_t enabled) {
    if (!e) return BA_ERR_INVALID_ARGUMENT;
    e->profiling = enabled < 0 ? 0xffffffffu : unsigned(enabled);   // bit mask of kernel ids (ba_kernel_stat order); 1 keeps meaning "all" for convenience
    if (enabled == 1) e->profiling = 0xffffffffu;
    for (auto& st : e->stats) { st.launches = 0; st.timed = 0; st.ms = 0.0; }
    return BA_OK;
}

int ba_set_profiling_stride(ba_engine* e, int32_t stride) {
    if (!e || stride < 1) return BA_ERR_INVALID_ARGUMENT;
    e->k5_sample_stride = stride;
    return BA_OK;
}

int ba_linear_solver_in_use(ba_engine* e, int32_t* solver) {
    if (!e || !solver) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_INVALID_ARGUMENT, "no problem set");
    *solver = e->dense ? BA_LINEAR_EXACT : BA_LINEAR_PCG;
    return BA_OK;
}

int ba_get_kernel_stats(ba_engine* e, ba_kernel_stat* out, int32_t cap, int32_t* n) {
    if (!e || !n) return BA_ERR_INVALID_ARGUMENT;
    *n = KID_COUNT;
    if (out) for (int i = 0; i < std::min<int>(cap, KID_COUNT); ++i) {
        std::memset(&out[i], 0, sizeof out[i]);
        std::snprintf(out[i].name, sizeof out[i].name, "%s", kKernelNames[i]);
        out[i].launches = e->stats[i].launches; out[i].bytes_per_launch = e->stats[i].bytes;
        // launches that were not bracketed by events (sampled timing of K5) count at the mean of the timed ones
        out[i].total_ms = e->stats[i].timed > 0 ? e->stats[i].ms * double(e->stats[i].launches) / double(e->stats[i].timed) : 0.0;
    }
    return BA_OK;
}

int ba_shard_points(int32_t num_points, int32_t num_observations, const int32_t* obs_pt, int32_t world_size, int32_t* shard_pt) {
    if (num_points < 0 || num_observations < 0 || world_size < 1 || !shard_pt || (num_observations && !obs_pt)) return BA_ERR_INVALID_ARGUMENT;
    vector<int32_t> off(size_t(num_points) + 1, 0), sp;
    for (int o = 0; o < num_observations; ++o) {
        if (obs_pt[o] < 0 || obs_pt[o] >= num_points || (o && obs_pt[o] < obs_pt[o - 1])) return BA_ERR_INVALID_ARGUMENT;
        ++off[obs_pt[o] + 1];
    }
    for (int p = 0; p < num_points; ++p) off[p + 1] += off[p];
    shard_points(off, num_points, num_observations, world_size, sp);
    std::copy(sp.begin(), sp.end(), shard_pt);
    return BA_OK;
}

int ba_nccl_unique_id(void* out128) {
    if (!out128) return BA_ERR_INVALID_ARGUMENT;
    std::string err;
    if (!g_nccl.load(err)) return fail(nullptr, BA_ERR_NCCL, err);
    Nccl::UniqueId id;
    const int r = g_nccl.GetUniqueId(&id);
    if (r != 0) return fail(nullptr, BA_ERR_NCCL, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
    std::memcpy(out128, &id, sizeof id);
    return BA_OK;
}

int ba_get_execution_info(ba_engine* e, int64_t* info, int32_t cap, int32_t* n) {
    if (!e || !n) return BA_ERR_INVALID_ARGUMENT;
    if (!e->has_problem) return fail(e, BA_ERR_NO_PROBLEM, "no problem set");
    const int64_t v[8] = {e->dense ? BA_LINEAR_EXACT : BA_LINEAR_PCG, e->scatter_eff, e->opt.world_size > 1 ? (e->p2p.on ? 2 : 1) : 0, e->host_syncs, e->nTiles, e->RS.nRows,
                          (e->factored && !e->dense) ? 1 : 0, (e->tail_cl == bak::TAIL_CL_WIDE && e->nf >= e->tail_wide_min_nf) ? bak::TAIL_CL_WIDE : bak::TAIL_CL};
    *n = 8;
    if (info) for (int i = 0; i < std::min<int>(cap, 8); ++i) info[i] = v[i];
    return BA_OK;
}

int ba_get_io_bytes(int64_t* h2d, int64_t* d2h) { if (h2d) *h2d = g_h2d_bytes; if (d2h) *d2h = g_d2h_bytes; return BA_OK; }
const char* ba_last_error(const ba_engine* e) { return e ? e->err.c_str() : g_create_error.c_str(); }
const char* ba_backend(void) { return "cuda-sm100a"; }

void ba_destroy(ba_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->l2_window && e->stream) {   // a borrowed stream must not keep a window onto memory that is about to be freed
        cudaStreamAttrValue attr{};
        attr.accessPolicyWindow.num_bytes = 0;
        cudaStreamSetAttribute(e->stream, cudaStreamAttributeAccessPolicyWindow, &attr);
        cudaCtxResetPersistingL2Cache();
    }
    teardown_p2p(e, e->err.empty());   // collective: every rank destroys its handle (a handle that failed does not wait for the others)
    if (e->comm) g_nccl.CommDestroy(e->comm);
    for (auto& p : e->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto ev : e->ev_pool) cudaEventDestroy(ev);
    if (e->ev_t0) cudaEventDestroy(e->ev_t0);
    if (e->ev_t1) cudaEventDestroy(e->ev_t1);
    if (e->h_scal) cudaFreeHost(e->h_scal);
    if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------
// ba_cost_noalloc — an entry point that allocates nothing (include/ba_engine.h).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// It is the candidate-cost evaluation (K0 camera precompute + a K1'-style residual / robust-loss sweep, the same device functions the
// LM loop uses) with its scratch memory — the per-camera records and the per-block partial sums — in ONE file-scope device array of a
// fixed, modest size instead of engine-owned cudaMalloc buffers.  Every GPU that loads this library holds the array until exit.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: its address is looked up once per device
// with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
namespace {
constexpr int kNoallocBlocks = 1184;                                             // 148 SMs x 8 CTAs of 256 threads, grid-stride
constexpr size_t kNoallocCamDoubles = size_t(BA_NOALLOC_MAX_CAMERAS) * bak::CAMREC;
constexpr size_t kNoallocDoubles = kNoallocCamDoubles + 2048;                    // + block partials (1184) + ticket word, padded
}  // namespace
static __device__ __align__(256) double g_noalloc_scratch[kNoallocDoubles];     // 672 KB per device

namespace {
// Residual + robust loss of every observation (grid-stride), one partial per block; the last block to finish (ticket) adds the partials in
// index order, so the result does not depend on block scheduling.  `ticket` must be zero on entry (cleared by the host per call).
template <int MODEL>
__global__ void __launch_bounds__(256)
k_cost_static(int No, int Ng, const double* __restrict__ camrec, const double* __restrict__ intr, const int* __restrict__ cam_group,
              const double* __restrict__ pt, const int* __restrict__ obs_cam, const int* __restrict__ obs_pt, const double2* __restrict__ obs_uv,
              bak::RobustParams rp, double* __restrict__ partial, unsigned int* __restrict__ ticket, double* __restrict__ cost_out,
              int* __restrict__ bad_out) {
    __shared__ double sred[32];
    __shared__ bool last;
    __shared__ int sbad;
    if (threadIdx.x == 0) sbad = 0;
    __syncthreads();
    double cost = 0.0;
    int bad = 0;
    for (int o = blockIdx.x * blockDim.x + threadIdx.x; o < No; o += gridDim.x * blockDim.x) {
        const int c = obs_cam[o], p = obs_pt[o];
        bak::Proj q;
        double cr[12];   // R, t
        bak::load_camrec<6>(camrec, c, cr);
        bak::project_obs<MODEL>(cr, pt + 3 * size_t(p), intr + (Ng == 1 ? 0 : 4 * cam_group[c]), __ldg(obs_uv + o), q);
        double w;
        const double rho = bak::robust(rp, q.r0 * q.r0 + q.r1 * q.r1, w);
        if (!isfinite(rho)) bad = 1;
        cost += 0.5 * rho;
    }
    if (bad) atomicOr(&sbad, 1);
    const double cs = bak::block_sum(cost, sred);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = cs;
        if (sbad) atomicOr(bad_out, 1);
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double s = 0.0;
    for (int i = threadIdx.x; i < int(gridDim.x); i += blockDim.x) s += __ldcg(partial + i);
    s = bak::block_sum(s, sred);
    if (threadIdx.x == 0) *cost_out = s;
}

std::mutex g_noalloc_mu;
double* g_noalloc_ptr[64] = {};   // device address of g_noalloc_scratch, per device ordinal
}  // namespace

extern "C" int ba_cost_noalloc(int32_t device, void* stream, int32_t model, int32_t robust_kind, double robust_a, int32_t taylor_rule,
                               int32_t num_cameras, int32_t num_groups, int32_t num_observations, const double* d_pose, const double* d_intr,
                               const int32_t* d_cam_group, const double* d_pt, const int32_t* d_obs_cam, const int32_t* d_obs_pt,
                               const double* d_obs_uv, double* d_cost, int32_t* d_bad) {
    if (num_cameras < 0 || num_groups < 1 || num_observations < 0) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "ba_cost_noalloc: negative size or no intrinsics group");
    if (num_cameras > BA_NOALLOC_MAX_CAMERAS) {
        char b[200];
        std::snprintf(b, sizeof b, "ba_cost_noalloc: %d cameras exceed the static scratch of this entry point (BA_NOALLOC_MAX_CAMERAS = %d); use a ba_engine handle",
                      int(num_cameras), int(BA_NOALLOC_MAX_CAMERAS));
        return fail(nullptr, BA_ERR_INVALID_ARGUMENT, b);
    }
    if (model != BA_MODEL_PINHOLE && model != BA_MODEL_SNAVELY) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "ba_cost_noalloc: unknown residual model");
    if (!d_cost || !d_bad || (num_cameras && (!d_pose || !d_cam_group)) || !d_intr || (num_observations && (!d_pt || !d_obs_cam || !d_obs_pt || !d_obs_uv)))
        return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "ba_cost_noalloc: null device pointer");
    if (device < 0 || device >= 64) return fail(nullptr, BA_ERR_INVALID_ARGUMENT, "ba_cost_noalloc: device ordinal out of range");
    return guarded(nullptr, [&] {
        int prev = -1;
        CU_TRY(cudaGetDevice(&prev));
        struct Restore { int d; ~Restore() { if (d >= 0) cudaSetDevice(d); } } restore{prev == device ? -1 : prev};
        CU_TRY(cudaSetDevice(device));
        double* scratch = nullptr;
        {   // once per device, with that device current: the device address of the symbol
            std::lock_guard<std::mutex> g(g_noalloc_mu);
            if (!g_noalloc_ptr[device]) {
                void* p = nullptr;
                CU_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
                g_noalloc_ptr[device] = static_cast<double*>(p);
            }
            scratch = g_noalloc_ptr[device];
        }
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        double* camrec = scratch;
        double* partial = scratch + kNoallocCamDoubles;
        unsigned int* ticket = reinterpret_cast<unsigned int*>(partial + kNoallocBlocks);
        // the array keeps the last call's contents: the ticket and the outputs must start at zero; records and partials are overwritten before use
        CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s));
        CU_TRY(cudaMemsetAsync(d_bad, 0, sizeof(int32_t), s));
        if (num_cameras > 0) bak::k0_camera_precompute<<<cdiv(num_cameras, 128), 128, 0, s>>>(num_cameras, d_pose, camrec, taylor_rule);
        CU_TRY(cudaGetLastError());
        const int blocks = std::max(1, std::min(kNoallocBlocks, cdiv(num_observations, 256)));
        const bak::RobustParams rp{robust_kind, robust_a};
        const double2* uv = reinterpret_cast<const double2*>(d_obs_uv);
        if (model == BA_MODEL_SNAVELY) k_cost_static<1><<<blocks, 256, 0, s>>>(num_observations, num_groups, camrec, d_intr, d_cam_group, d_pt, d_obs_cam, d_obs_pt, uv, rp, partial, ticket, d_cost, d_bad);
        else k_cost_static<0><<<blocks, 256, 0, s>>>(num_observations, num_groups, camrec, d_intr, d_cam_group, d_pt, d_obs_cam, d_obs_pt, uv, rp, partial, ticket, d_cost, d_bad);
        CU_TRY(cudaGetLastError());
    });
}
