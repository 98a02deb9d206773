// libheat3d: C-ABI implementation (include/heat3d.h).  Host side of the heat3D hot path on B200:
// slab bookkeeping, pitched double-buffered fields, TMA descriptors, streams/events, the step loop with
// ghost-plane exchange overlapped with interior compute, fused residual, error norm, snapshots.
//
// There is no CPU path in this file: every entry point needs a CUDA device and says so when there is none.
#include "../../include/heat3d.h"
#include "kernels.cuh"
#include "kernels_t2.cuh"
#include "internal.h"

#include <algorithm>
#include <array>
#include <cfloat>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <map>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace {

constexpr uint32_t HALO_RING = 256;
constexpr uint32_t GRAPH_BLOCKS = 32;   // even: the current-buffer index is the same before and after a graph launch
thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CU_TRY(expr)                                                                                    \
    do {                                                                                                \
        cudaError_t e__ = (expr);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return fail(HEAT3D_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

// ---- NCCL, bound at run time so the library loads (and single-process runs work) without it ------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclFloat64 = 8 };
enum { ncclMax = 2, ncclSum = 0 };
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
} g_nccl;

int load_nccl()
{
    if (g_nccl.handle) return HEAT3D_OK;
    // If the host process already mapped a libnccl.so.2 (e.g. torch's bundled one) this returns that copy.
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(HEAT3D_ERR_NCCL, "cannot load libnccl.so.2: %s", dlerror());
#define H3_SYM(field, name)                                                                  \
    *(void**)(&g_nccl.field) = dlsym(h, name);                                               \
    if (!g_nccl.field) return fail(HEAT3D_ERR_NCCL, "libnccl lacks symbol %s", name);
    H3_SYM(GetUniqueId, "ncclGetUniqueId")
    H3_SYM(CommInitRank, "ncclCommInitRank")
    H3_SYM(CommDestroy, "ncclCommDestroy")
    H3_SYM(Send, "ncclSend")
    H3_SYM(Recv, "ncclRecv")
    H3_SYM(AllReduce, "ncclAllReduce")
    H3_SYM(AllGather, "ncclAllGather")
    H3_SYM(GroupStart, "ncclGroupStart")
    H3_SYM(GroupEnd, "ncclGroupEnd")
    H3_SYM(GetErrorString, "ncclGetErrorString")
#undef H3_SYM
    g_nccl.handle = h;
    return HEAT3D_OK;
}

#define NCCL_TRY(expr)                                                                                  \
    do {                                                                                                \
        int r__ = (expr);                                                                               \
        if (r__ != ncclSuccess)                                                                         \
            return fail(HEAT3D_ERR_NCCL, "%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(r__), __FILE__, __LINE__); \
    } while (0)

// ---- cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda) -----------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

int load_encode()
{
    if (g_encode) return HEAT3D_OK;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    CU_TRY(cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &fn, 12000, cudaEnableDefault, &q));
    if (q != cudaDriverEntryPointSuccess || !fn) return fail(HEAT3D_ERR_CUDA, "driver lacks cuTensorMapEncodeTiled");
    g_encode = (EncodeTiledFn)fn;
    return HEAT3D_OK;
}

// ---- stream memory operations (flag writes / waits ordered in a stream, no kernel involved) ---------------
typedef CUresult (*StreamValue64Fn)(CUstream, CUdeviceptr, cuuint64_t, unsigned int);
StreamValue64Fn g_stream_write64 = nullptr, g_stream_wait64 = nullptr;

int load_stream_memops()
{
    if (g_stream_write64 && g_stream_wait64) return HEAT3D_OK;
    void* fw = nullptr;
    void* fa = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    CU_TRY(cudaGetDriverEntryPointByVersion("cuStreamWriteValue64", &fw, 12000, cudaEnableDefault, &q1));
    CU_TRY(cudaGetDriverEntryPointByVersion("cuStreamWaitValue64", &fa, 12000, cudaEnableDefault, &q2));
    if (q1 != cudaDriverEntryPointSuccess || q2 != cudaDriverEntryPointSuccess || !fw || !fa)
        return fail(HEAT3D_ERR_CUDA, "driver lacks cuStreamWriteValue64/cuStreamWaitValue64");
    g_stream_write64 = (StreamValue64Fn)fw;
    g_stream_wait64 = (StreamValue64Fn)fa;
    return HEAT3D_OK;
}

// ---- TMA kernel variants --------------------------------------------------------------------------------
struct TmaVariant {
    const char* name;
    int tx, ty, ry, stages, minb;
    size_t smem;
    int threads;
    const void* fn[2][2][2];   // [arith][res][mirror]
};

template <int TX, int TY, int RY, int ST, int MINB>
TmaVariant make_variant(const char* name)
{
    TmaVariant v;
    v.name = name;
    v.tx = TX; v.ty = TY; v.ry = RY; v.stages = ST; v.minb = MINB;
    v.smem = h3::TmaCfg<TX, TY, ST>::SMEM_BYTES;
    v.threads = (TX / 2) * (TY / RY);
    v.fn[0][0][0] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_STRICT, false, MINB, false>;
    v.fn[0][1][0] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_STRICT, true, MINB, false>;
    v.fn[1][0][0] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_FMA, false, MINB, false>;
    v.fn[1][1][0] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_FMA, true, MINB, false>;
    v.fn[0][0][1] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_STRICT, false, MINB, true>;
    v.fn[0][1][1] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_STRICT, true, MINB, true>;
    v.fn[1][0][1] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_FMA, false, MINB, true>;
    v.fn[1][1][1] = (const void*)h3::stencil_tma_kernel<TX, TY, RY, ST, h3::ARITH_FMA, true, MINB, true>;
    return v;
}

// tile_x, tile_y, rows/thread, stages, CTAs/SM.  First entry = default; others are selectable via params.tune.
const std::vector<TmaVariant>& variants()
{
    static const std::vector<TmaVariant> v = {
        make_variant<128, 8, 2, 4, 3>("tma_128x8_r2_s4_c3"),
        make_variant<128, 8, 4, 4, 4>("tma_128x8_r4_s4_c4"),
        make_variant<128, 16, 2, 4, 1>("tma_128x16_r2_s4_c1"),
        make_variant<128, 16, 4, 4, 2>("tma_128x16_r4_s4_c2"),
        make_variant<128, 32, 4, 4, 1>("tma_128x32_r4_s4_c1"),
        make_variant<64, 16, 4, 4, 4>("tma_64x16_r4_s4_c4"),
    };
    return v;
}

// ---- depth-2 temporal-blocking kernel variants --------------------------------------------------------
struct T2Variant {
    const char* name;
    int cx, cy, ry, stages, minb, xmode;
    size_t smem;
    int threads;
    const void* fn[2][2][2];   // [arith][res][mirror]
};

template <int CX, int CY, int RY, int ST, int MINB, int XSW>
T2Variant make_t2(const char* name)
{
    T2Variant v;
    v.name = name;
    v.cx = CX; v.cy = CY; v.ry = RY; v.stages = ST; v.minb = MINB; v.xmode = XSW;
    v.smem = h3::Tma2Cfg<CX, CY, ST>::SMEM_BYTES;
    v.threads = (CX / 2) * (CY / RY);
    v.fn[0][0][0] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_STRICT, false, MINB, XSW, false>;
    v.fn[0][1][0] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_STRICT, true, MINB, XSW, false>;
    v.fn[1][0][0] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_FMA, false, MINB, XSW, false>;
    v.fn[1][1][0] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_FMA, true, MINB, XSW, false>;
    v.fn[0][0][1] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_STRICT, false, MINB, XSW, true>;
    v.fn[0][1][1] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_STRICT, true, MINB, XSW, true>;
    v.fn[1][0][1] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_FMA, false, MINB, XSW, true>;
    v.fn[1][1][1] = (const void*)h3::stencil_tma_t2_kernel<CX, CY, RY, ST, h3::ARITH_FMA, true, MINB, XSW, true>;
    return v;
}

// compute region cx x cy (output tile (cx-4) x (cy-2)), rows/thread, stages, CTAs/SM.  First = default;
// HEAT3D_T2_TUNE="cx,cy,ry,stages,ctas[,chunk_len[,xmode]]" selects another (experiments; xmode defaults to the first
// entry's).  Shared-memory access mode (kernels_t2.cuh): "_L" 0 = plain loads (round 1), "_S" 1 = swapped x-neighbour
// loads at both levels, "_P" 2 = plain loads from the stage + parity-split mid plane, "_X" 3 = swapped loads from the
// stage + parity-split mid plane (no bank conflicts left at either level).
const std::vector<T2Variant>& t2_variants()
{
    static const std::vector<T2Variant> v = {
        make_t2<64, 32, 4, 4, 2, 2>("t2_64x32_r4_s4_c2_P"),
        make_t2<64, 32, 4, 4, 2, 0>("t2_64x32_r4_s4_c2_L"),
        make_t2<64, 32, 4, 4, 2, 3>("t2_64x32_r4_s4_c2_X"),
        make_t2<64, 32, 4, 4, 2, 6>("t2_64x32_r4_s4_c2_Q"),      // "_P" with neighbour-to-neighbour barriers instead of the CTA barrier
        make_t2<64, 16, 2, 4, 3, 2>("t2_64x16_r2_s4_c3_P"),
        // 512 threads, one CTA per SM: smaller redundant level-1 ring (124x30 / 60x62 of 128x32 / 64x64 useful); measured
        // slower than two independent 256-thread CTAs per SM (one barrier domain of 16 warps), kept for experiments
        make_t2<128, 32, 4, 4, 1, 2>("t2_128x32_r4_s4_c1_P"),
        make_t2<64, 64, 4, 4, 1, 2>("t2_64x64_r4_s4_c1_P"),
    };
    return v;
}

struct Slab {
    int dev = 0;
    int gidx = 0;                 // global slab index
    uint32_t z0 = 0, z1 = 0;      // owned global planes [z0, z1)
    uint32_t G = 1;               // ghost depth
    uint32_t planes = 0;          // allocated planes = (z1-z0) + 2G
    long long zg0 = 0;            // global z of local plane 0 (= z0 - G)
    uint32_t pitch = 0;
    long long plane = 0;          // doubles per plane
    double* buf[2] = {nullptr, nullptr};
    CUtensorMap tmap[2];
    CUtensorMap tmap2[2];         // depth-2 kernel's box shape
    cudaStream_t compute = nullptr, comm = nullptr, copy = nullptr;
    bool own_compute = true;
    cudaEvent_t ev_boundary = nullptr, ev_put = nullptr, ev_start = nullptr, ev_stop = nullptr, ev_snap = nullptr,
                ev_tmp = nullptr;
    std::vector<cudaEvent_t> ev_wait0, ev_wait1;   // HEAT3D_FLAG_TIME_HALO: brackets of the ghost wait, ring of HALO_RING steps
    unsigned long long* d_res = nullptr;   // residual slot in use (bit pattern of a non-negative double); = d_res_base + 0 or 1
    unsigned long long* d_res_base = nullptr;
    cudaEvent_t ev_res[2] = {nullptr, nullptr};   // pipelined convergence loop: residual of slot i has reached the host
    double* d_partial = nullptr;           // error-norm scratch: partial sums
    double* d_scalar = nullptr;            // error-norm result / allreduce staging
    double* h_scalar = nullptr;            // pinned: 2 doubles
    int partial_blocks = 0;
    int num_sms = 0;
    ncclComm_t nccl = nullptr;
    // cross-process neighbours reached through CUDA IPC (side 0 = lower, 1 = upper): their two field buffers and
    // their arrival flags mapped into this process, plus this slab's own flags {written by lower, written by upper}
    bool ipc = false;
    bool peer_is_ipc[2] = {false, false};      // peer_buf/peer_flags of that side came from cudaIpcOpenMemHandle
    unsigned long long* d_side_done = nullptr;  // peer-store exchange: finished end-chunk CTAs per side {lower, upper}
    unsigned long long side_target[2] = {0, 0};
    double* peer_buf[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
    unsigned long long* peer_flags[2] = {nullptr, nullptr};
    unsigned long long* d_flags = nullptr;
    double* snap_dev = nullptr;              // device-side staging copy of the owned planes for async snapshots (lazily allocated)
    size_t snap_dev_bytes = 0;
    double* up_dev = nullptr;                // device-side landing buffer of an asynchronous upload (lazily allocated)
    size_t up_dev_bytes = 0;
    uint32_t up_a = 0, up_b = 0;             // owned global planes [up_a, up_b) staged there, empty when up_b <= up_a
    cudaStream_t h2d = nullptr;              // its own stream: uploads, snapshots (copy) and stepping all overlap
    cudaEvent_t ev_up = nullptr;
    unsigned long long* d_done = nullptr;    // end-chunk CTAs finished so far (device counter the exchange stream waits on)
    unsigned long long done_target = 0;      // value d_done reaches when every end chunk launched so far is in memory
    bool lower = false, upper = false;     // has a neighbouring slab below / above
    int za = 0, zb = 0;                    // local output plane range (owned ∩ global interior)
    std::map<long long, std::array<int, 4>> plan_cache;   // guided chunk plans by launch shape (plan_chunks)
};

}  // namespace

struct heat3d_ctx {
    heat3d_params p;
    std::vector<int32_t> devices;
    std::vector<Slab> slabs;
    int cur = 0;                       // buffer holding the current time level
    const TmaVariant* variant = nullptr;
    const T2Variant* variant2 = nullptr;   // set when temporal_k >= 2
    int t2_chunk = 64;
    bool t2_guided = true;             // long-then-short chunk plan (plan_guided); off when a chunk length is forced
    bool use_tma = true;
    bool multi_process = false;
    int snap_buf = -1;                 // buffer an async snapshot is still reading, or -1
    bool have_res = false;
    double last_res = 0.0;
    heat3d_timing tm;
    bool halo_timed = false;
    // CUDA graphs of GRAPH_BLOCKS consecutive passes of a single-slab context, one per (depth index, buffer parity):
    // small grids are launch-bound, and a graph launch costs about as much as one kernel launch
    cudaGraphExec_t step_graph[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
    bool graphs_ok = true;
    bool res_slot_prezeroed = false;   // pipelined convergence loop: residual slots are cleared off the compute stream
    bool store_mode = false;           // ghost planes are written by the stencil kernel itself into the neighbours' memory
    bool peers_reachable = true;       // every slab can address its in-process neighbours' memory (same device or peer access)
    bool end_signal = false;           // stream memory operations available: single-launch steps with end-chunk signalling
    unsigned long long seq = 0;        // number of ghost exchanges issued so far (same on every process)
    unsigned long long consumed_seq = 0;   // depth-4 blocks issued so far: the value neighbours publish once they have read their ghosts
    bool timing_pending = false;       // a call's ev_start/ev_stop have not been folded into tm yet
    uint32_t call_steps = 0;           // steps issued by the current/last call
    // ---- split along y (params.nslabs_y > 1): this context is then only a parent that owns one child context per
    // y-range; each child is an ordinary context (z-slabs and all) over rows [y_off, y_off + p.ny) of the global field
    int y_off = 0;                     // child: global row of local row 0
    uint32_t ny_global = 0;            // child: rows of the global field (0 = not a child)
    bool want_raw_error_sum = false;   // child: heat3d_mean_abs_error_vs_y returns the plain sum (the parent divides)
    std::vector<heat3d_ctx*> kids;     // parent: children in order of y
    struct YLink {                     // parent: one direction of one child-to-child face, per z-slab
        int from_kid, to_kid, slab;
        int src_row, dst_row;          // local row read in `from`, local ghost row written in `to`
        double* stage_src = nullptr;   // dense (plane, x) buffer on the sender's device ...
        double* stage_dst = nullptr;   // ... and on the receiver's (same pointer when both share a device)
        cudaEvent_t ev_packed = nullptr, ev_unpacked = nullptr;
    };
    std::vector<YLink> ylinks;
    cudaEvent_t g2_t0 = nullptr, g2_t1 = nullptr;   // parent: bracket of the last stepping call
    std::string g2_exchange;           // parent: what heat3d_exchange_name answers
};

namespace {

int set_dev(const Slab& s)
{
    CU_TRY(cudaSetDevice(s.dev));
    return HEAT3D_OK;
}

// Multi-process contexts wait on flags another process writes (cuStreamWaitValue64 in the streams): if that process has
// died the stream never drains and a plain cudaStreamSynchronize would hang with it.  These waits poll instead and
// give up after HEAT3D_WATCHDOG_S seconds (default 900, 0 = wait for ever) with the sequence numbers needed to see
// who is stuck.  Single-process contexts synchronise directly.
int watched_sync(heat3d_ctx* c, Slab& s, cudaStream_t st, const char* what)
{
    static const double limit = getenv("HEAT3D_WATCHDOG_S") ? atof(getenv("HEAT3D_WATCHDOG_S")) : 900.0;
    if (!c->multi_process || limit <= 0.0) {
        CU_TRY(cudaStreamSynchronize(st));
        return HEAT3D_OK;
    }
    const auto t0 = std::chrono::steady_clock::now();
    unsigned spins = 0;
    for (;;) {
        const cudaError_t e = cudaStreamQuery(st);
        if (e == cudaSuccess) return HEAT3D_OK;
        if (e != cudaErrorNotReady) return fail(HEAT3D_ERR_CUDA, "cudaStreamQuery(%s) failed: %s", what, cudaGetErrorString(e));
        if (++spins > 2000) std::this_thread::sleep_for(std::chrono::microseconds(50));      // busy-poll the first instants only
        if ((spins & 1023u) == 0 && std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > limit) {
            unsigned long long flags[4] = {0, 0, 0, 0};
            cudaMemcpyAsync(flags, s.d_flags, sizeof flags, cudaMemcpyDeviceToHost, s.copy);     // a stream that is not stuck
            cudaStreamSynchronize(s.copy);
            return fail(HEAT3D_ERR_STATE,
                        "slab %d: %s stream still busy after %.0f s - a neighbouring process has probably stopped.  Exchanges issued here: %llu; "
                        "arrival flags from below/above: %llu/%llu; consumed flags: %llu/%llu",
                        s.gidx, what, limit, c->seq, flags[0], flags[1], flags[2], flags[3]);
        }
    }
}

int make_tmap(Slab& s, CUtensorMap* map, int which, const heat3d_ctx* c, int box_x, int box_y)
{
    cuuint64_t dims[3] = {c->p.nx, c->p.ny, s.planes};
    cuuint64_t strides[2] = {(cuuint64_t)s.pitch * 8, (cuuint64_t)s.plane * 8};
    cuuint32_t box[3] = {(cuuint32_t)box_x, (cuuint32_t)box_y, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    // L2 promotion widens every L2 miss to 128/256 B; the staged rows start 16 B before a 128-B boundary, so
    // promotion would drag in whole extra lines per row.  Default: none.  (HEAT3D_TMA_L2PROMO=0..3 for experiments.)
    CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_NONE;
    if (const char* e = getenv("HEAT3D_TMA_L2PROMO")) promo = (CUtensorMapL2promotion)atoi(e);
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, s.buf[which], dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return HEAT3D_OK;
}

// Cuts output planes [za, zb) into an optional short end chunk per side whose neighbour is waiting (ends bit 0 =
// lower, bit 1 = upper) plus uniform middle chunks; returns the grid size and books the end-chunk CTAs on the
// slab's completion counter.
// Peer-store exchange: for side 0 (lower) / 1 (upper), where this slab's outer G planes go in the neighbour's memory.
struct MirrorInfo {
    double* base[2];                  // neighbour address of local plane z0[side] (row 0, column 0); nullptr = side off
    int z0[2], z1[2];
    unsigned long long* flag[2];
    unsigned long long flag_value;
};

// Long-then-short split of `nzp` middle planes for a launch of ntiles CTAs per chunk on `slots` resident CTA slots.
// Equal chunks leave a tail: the last, partly filled wave costs as much as a full one.  Model: CTAs are handed to free
// slots in index order, a CTA of L output planes costs L + ovh plane-times (ovh = the seam planes it loads and computes
// on top, plus its prologue); the makespan of every candidate (n_long chunks of L, then n_short equal short ones) is
// simulated and the shortest wins.  `first_ctas` CTAs of `first_cost` (the end chunks) come before everything else.
struct ChunkSplit { int len, n_long, len2, split; };   // split = planes covered by the long chunks

ChunkSplit plan_guided(int nzp, long long ntiles, long long slots, long long first_ctas, double first_cost, double ovh, int max_len)
{
    ChunkSplit best = {nzp, 1, nzp, nzp};
    double best_t = 1e300;
    std::vector<double> heap;
    auto makespan = [&](int L, int nl, int nzl, int L2, int ns) {
        heap.assign((size_t)slots, 0.0);           // min-heap of slot free times
        auto cmp = [](double a, double b) { return a > b; };
        auto give = [&](double cost) {
            std::pop_heap(heap.begin(), heap.end(), cmp);
            heap.back() += cost;
            std::push_heap(heap.begin(), heap.end(), cmp);
        };
        for (long long i = 0; i < first_ctas; ++i) give(first_cost);
        int left = nzl;
        for (int c = 0; c < nl; ++c) { const int n = std::min(L, left); left -= n; for (long long t = 0; t < ntiles; ++t) give(n + ovh); }
        left = nzp - nzl;
        for (int c = 0; c < ns; ++c) { const int n = std::min(L2, left); left -= n; for (long long t = 0; t < ntiles; ++t) give(n + ovh); }
        return *std::max_element(heap.begin(), heap.end());
    };
    auto consider = [&](int L, int nl, int ns) {
        const int nzl = std::min(nzp, L * nl);
        const int rest = nzp - nzl;
        if ((rest > 0) != (ns > 0)) return;
        const int L2 = ns ? (rest + ns - 1) / ns : L;
        if (ns && (L2 < 2 || L2 > L)) return;
        const int ns_eff = ns ? (rest + L2 - 1) / L2 : 0;
        const long long ctas = (long long)(nl + ns_eff) * ntiles;
        // no schedule beats total work / slots: most candidates with many short chunks end here
        const double bound = ((double)first_ctas * first_cost + (double)ntiles * (nzp + ovh * (nl + ns_eff))) / (double)slots;
        if (bound >= best_t || ctas > 60000) return;
        const double t = makespan(L, nl, nzl, L2, ns_eff);
        if (t < best_t - 1e-9) { best_t = t; best = {L, nl, L2, nzl}; }
    };
    const int lmax = std::min(max_len, nzp);
    // long candidates first: they set a good bound early
    for (int L = lmax; L >= 2; L -= (L > 64 ? 4 : L > 16 ? 2 : 1))
        for (int nl = 1; (nl - 1) * L < nzp; ++nl)
            for (int ns = 0; ns <= 6; ++ns) consider(L, nl, ns);
    return best;
}

long long plan_chunks(h3::TmaWork& w, int za, int zb, int ends, int end_len, int chunk_len, long long slots, Slab& s,
                      const MirrorInfo* mi = nullptr, bool guided = false)
{
    const long long ntiles = (long long)w.ntx * w.nty;
    w.n_end = 0;
    w.mid_za = za;
    w.mid_zb = zb;
    w.done = (ends && !mi) ? s.d_done : nullptr;
    w.flag_value = mi ? mi->flag_value : 0;
    for (int i = 0; i < 2; ++i) {
        w.end_za[i] = w.end_zb[i] = 0;
        w.mirror[i] = nullptr; w.mirror_z0[i] = w.mirror_z1[i] = 0;
        w.side_done[i] = nullptr; w.side_goal[i] = 0; w.peer_flag[i] = nullptr;
    }
    if (ends == 3) end_len = std::max((int)s.G, std::min(end_len, (zb - za) / 2));   // thin slabs: the two end chunks must not meet
    auto mirror_side = [&](int side) {
        if (!mi || !mi->base[side]) return;
        const int i = w.n_end;
        w.mirror[i] = mi->base[side]; w.mirror_z0[i] = mi->z0[side]; w.mirror_z1[i] = mi->z1[side];
        s.side_target[side] += (unsigned long long)ntiles;
        w.side_done[i] = s.d_side_done + side; w.side_goal[i] = s.side_target[side]; w.peer_flag[i] = mi->flag[side];
    };
    if ((ends & 1) && w.mid_zb > w.mid_za) {
        const int e = std::min(end_len, w.mid_zb - w.mid_za);
        w.end_za[w.n_end] = w.mid_za; w.end_zb[w.n_end] = w.mid_za + e;
        mirror_side(0);
        w.mid_za += e; w.n_end++;
    }
    if ((ends & 2) && w.mid_zb > w.mid_za) {
        const int e = std::min(end_len, w.mid_zb - w.mid_za);
        w.end_za[w.n_end] = w.mid_zb - e; w.end_zb[w.n_end] = w.mid_zb;
        mirror_side(1);
        w.mid_zb -= e; w.n_end++;
    }
    const int nzp = w.mid_zb - w.mid_za;
    long long nch = 0;
    w.chunk_len = w.chunk_len2 = chunk_len;
    w.mid_split = w.mid_zb;
    w.n_long = 0;
    if (nzp > 0 && guided) {
        // cached per (planes, end chunks): the search runs once per distinct launch shape of a slab
        const long long key = ((long long)nzp << 8) | (long long)(w.n_end << 4) | (end_len & 15);
        auto it = s.plan_cache.find(key);
        if (it == s.plan_cache.end()) {
            static const double ovh = getenv("HEAT3D_T2_OVH") ? atof(getenv("HEAT3D_T2_OVH")) : 3.0;
            static const int max_len = getenv("HEAT3D_T2_MAXLEN") ? atoi(getenv("HEAT3D_T2_MAXLEN")) : 160;
            const ChunkSplit cs = plan_guided(nzp, ntiles, slots, ntiles * w.n_end, end_len + ovh, ovh, max_len);
            it = s.plan_cache.emplace(key, std::array<int, 4>{cs.len, cs.n_long, cs.len2, cs.split}).first;
        }
        w.chunk_len = it->second[0];
        w.n_long = it->second[1];
        w.chunk_len2 = it->second[2];
        w.mid_split = w.mid_za + it->second[3];
        const int rest = w.mid_zb - w.mid_split;
        nch = w.n_long + (rest > 0 ? (rest + w.chunk_len2 - 1) / w.chunk_len2 : 0);
    } else if (nzp > 0) {
        nch = (nzp + chunk_len - 1) / chunk_len;
        if ((nch + w.n_end) * ntiles < slots) nch = (slots + ntiles - 1) / ntiles - w.n_end;
        if (nch < 1) nch = 1;
        if (nch > nzp) nch = nzp;
        w.chunk_len = w.chunk_len2 = (int)((nzp + nch - 1) / nch);
        nch = (nzp + w.chunk_len - 1) / w.chunk_len;
        w.n_long = (int)nch;
    }
    if (ends && !mi) s.done_target += (unsigned long long)ntiles * w.n_end;
    return ntiles * (w.n_end + nch);
}

// Where slab s's outer G planes of buffer `dstb` live in its neighbours' memory (peer-store exchange).
bool fill_mirror(const heat3d_ctx* c, const Slab& s, int dstb, int ends, MirrorInfo& mi)
{
    if (!c->store_mode || !ends) return false;
    const uint32_t nzl = s.z1 - s.z0;
    for (int side = 0; side < 2; ++side) { mi.base[side] = nullptr; mi.z0[side] = mi.z1[side] = 0; mi.flag[side] = nullptr; }
    mi.flag_value = c->seq;
    if ((ends & 1) && s.lower) {       // my first G owned planes -> the lower neighbour's upper ghost planes
        const uint32_t n_nzl = (uint32_t)((uint64_t)s.gidx * c->p.nz / c->p.nslabs) - (uint32_t)((uint64_t)(s.gidx - 1) * c->p.nz / c->p.nslabs);
        mi.base[0] = s.peer_buf[0][dstb] + (long long)(s.G + n_nzl) * s.plane;
        mi.z0[0] = (int)s.G; mi.z1[0] = (int)(2 * s.G);
        mi.flag[0] = s.peer_flags[0] + 1;
    }
    if ((ends & 2) && s.upper) {       // my last G owned planes -> the upper neighbour's lower ghost planes
        mi.base[1] = s.peer_buf[1][dstb];
        mi.z0[1] = (int)nzl; mi.z1[1] = (int)(nzl + s.G);
        mi.flag[1] = s.peer_flags[1] + 0;
    }
    return true;
}

// One stencil launch over local output planes [za, zb) of slab s, reading buffer `src`.
int launch_stencil(heat3d_ctx* c, Slab& s, int src, int za, int zb, bool res, int ends = 0)
{
    if (zb <= za) return HEAT3D_OK;
    h3::Geom g;
    g.nx = (int)c->p.nx; g.ny = (int)c->p.ny; g.pitch = (int)s.pitch; g.plane = s.plane;
    g.za = za; g.zb = zb; g.Dx = c->p.Dx; g.Dy = c->p.Dy; g.Dz = c->p.Dz;
    const int ar = c->p.arith == HEAT3D_ARITH_FMA ? 1 : 0;
    double* dst = s.buf[src ^ 1];
    if (c->use_tma) {
        const TmaVariant& v = *c->variant;
        h3::TmaWork w;
        w.ntx = ((int)c->p.nx + v.tx - 1) / v.tx;
        w.nty = ((int)c->p.ny + v.ty - 1) / v.ty;
        // z-chunks: short chunks (default 16 output planes, tune[5] overrides) make the grid sweep the domain
        // z-major, wave after wave, so the planes a wave touches - and the halos its tiles share - stay in L2;
        // but never fewer chunks than it takes to fill every resident CTA slot once.
        MirrorInfo mi;
        const bool mirrored = fill_mirror(c, s, src ^ 1, ends, mi);
        const long long grid = plan_chunks(w, za, zb, ends, /*end_len=*/8, c->p.tune[5] > 0 ? c->p.tune[5] : 16,
                                           (long long)s.num_sms * v.minb, s, mirrored ? &mi : nullptr);
        void* args[] = {(void*)&s.tmap[src], (void*)&dst, (void*)&g, (void*)&w, (void*)&s.d_res};
        CU_TRY(cudaLaunchKernel(v.fn[ar][res ? 1 : 0][mirrored ? 1 : 0], dim3((unsigned)grid), dim3(v.threads), args, v.smem, s.compute));
    } else {
        const dim3 block(64, 4);
        const int nzp = zb - za;
        const int xy_blocks = ((g.nx + 63) / 64) * ((g.ny + 3) / 4);
        int chunks = (4 * s.num_sms * 8 + xy_blocks - 1) / xy_blocks;      // aim for ~4 waves of 8 CTAs/SM
        if (chunks > nzp) chunks = nzp;
        if (chunks < 1) chunks = 1;
        const int zchunk = (nzp + chunks - 1) / chunks;
        const dim3 grid((g.nx + 63) / 64, (g.ny + 3) / 4, (nzp + zchunk - 1) / zchunk);
        const double* srcp = s.buf[src];
#define H3_LAUNCH_BASE(A, R) h3::stencil_baseline_kernel<A, R><<<grid, block, 0, s.compute>>>(srcp, dst, g, zchunk, s.d_res)
        if (ar == 0 && !res) H3_LAUNCH_BASE(h3::ARITH_STRICT, false);
        else if (ar == 0 && res) H3_LAUNCH_BASE(h3::ARITH_STRICT, true);
        else if (ar == 1 && !res) H3_LAUNCH_BASE(h3::ARITH_FMA, false);
        else H3_LAUNCH_BASE(h3::ARITH_FMA, true);
#undef H3_LAUNCH_BASE
        CU_TRY(cudaGetLastError());
    }
    c->tm.launches++;
    c->tm.last_call_launches++;
    c->tm.last_call_stencil_launches++;
    return HEAT3D_OK;
}

// One depth-2 launch: local output planes [za, zb) of slab s advance by TWO updates, reading buffer `src`.
int launch_stencil_t2(heat3d_ctx* c, Slab& s, int src, int za, int zb, bool res, int ends = 0)
{
    if (zb <= za) return HEAT3D_OK;
    const T2Variant& v = *c->variant2;
    h3::Geom g;
    g.nx = (int)c->p.nx; g.ny = (int)c->p.ny; g.pitch = (int)s.pitch; g.plane = s.plane;
    g.za = za; g.zb = zb; g.Dx = c->p.Dx; g.Dy = c->p.Dy; g.Dz = c->p.Dz;
    const int ar = c->p.arith == HEAT3D_ARITH_FMA ? 1 : 0;
    double* dst = s.buf[src ^ 1];
    h3::TmaWork w;
    // x tiles start at column 0 and must reach column nx-2; y tiles start at the first interior row
    w.ntx = ((int)c->p.nx - 1 + (v.cx - 4) - 1) / (v.cx - 4);
    w.nty = ((int)c->p.ny - 2 + (v.cy - 2) - 1) / (v.cy - 2);
    MirrorInfo mi;
    const bool mirrored = fill_mirror(c, s, src ^ 1, ends, mi);
    const long long grid = plan_chunks(w, za, zb, ends, /*end_len=*/16, c->t2_chunk, (long long)s.num_sms * v.minb, s, mirrored ? &mi : nullptr,
                                       c->t2_guided);
    int z1lo = (int)(1 - s.zg0), z1hi = (int)((long long)c->p.nz - 1 - s.zg0);   // level-1 planes that are global interior planes
    unsigned long long* first = res ? s.d_res_base + 2 : nullptr;     // residual of the first of the two updates
    void* args[] = {(void*)&s.tmap2[src], (void*)&dst, (void*)&g, (void*)&w, (void*)&z1lo, (void*)&z1hi, (void*)&s.d_res, (void*)&first};
    CU_TRY(cudaLaunchKernel(v.fn[ar][res ? 1 : 0][mirrored ? 1 : 0], dim3((unsigned)grid), dim3(v.threads), args, v.smem, s.compute));
    c->tm.launches++;
    c->tm.last_call_launches++;
    c->tm.last_call_stencil_launches++;
    return HEAT3D_OK;
}

Slab* local_slab(heat3d_ctx* c, int gidx)
{
    const int i = gidx - c->p.slab_first;
    return (i >= 0 && i < (int)c->slabs.size()) ? &c->slabs[i] : nullptr;
}

// Push the outermost G owned planes of buffer `b` into the neighbours' ghost planes of their buffer `b`
// (same-process neighbours: peer copy; other processes: NCCL send/recv), on each slab's comm stream,
// after that slab's ev_boundary.  Records ev_put per slab.
int exchange_ghosts(heat3d_ctx* c, int b, bool after_counted_launch = false, unsigned long long consumed = 0)
{
    bool any_nccl = false;
    c->seq++;
    for (Slab& s : c->slabs) {
        if (!s.lower && !s.upper) continue;
        int rc = set_dev(s);
        if (rc) return rc;
        if (after_counted_launch) {     // the end chunks of the launch just queued have counted themselves done
            CUresult r = g_stream_wait64((CUstream)s.comm, (CUdeviceptr)s.d_done, s.done_target, CU_STREAM_WAIT_VALUE_GEQ);
            if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWaitValue64 failed with CUresult %d", (int)r);
        } else {
            CU_TRY(cudaStreamWaitEvent(s.comm, s.ev_boundary, 0));
        }
        if (consumed) {
            // depth-4 blocks overwrite ghost planes of the buffer the neighbour's first pass is reading: wait until the
            // neighbour says it has read them (it writes `consumed` into my flags after that pass, in stream order)
            for (int side = 0; side < 2; ++side) {
                if (!(side == 0 ? s.lower : s.upper)) continue;
                if (!local_slab(c, s.gidx + (side == 0 ? -1 : 1)) && !s.ipc) continue;      // NCCL: the receiver posts the recv itself
                CUresult r = g_stream_wait64((CUstream)s.comm, (CUdeviceptr)(s.d_flags + 2 + side), consumed, CU_STREAM_WAIT_VALUE_GEQ);
                if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWaitValue64 failed with CUresult %d", (int)r);
            }
        }
        const size_t bytes = (size_t)s.G * s.plane * sizeof(double);
        const uint32_t nzl = s.z1 - s.z0;
        if (s.lower) {
            if (Slab* n = local_slab(c, s.gidx - 1)) {   // my first G owned planes -> n's upper ghost
                const double* from = s.buf[b] + (long long)s.G * s.plane;
                double* to = n->buf[b] + (long long)(n->G + (n->z1 - n->z0)) * n->plane;
                CU_TRY(cudaMemcpyPeerAsync(to, n->dev, from, s.dev, bytes, s.comm));
                if (c->store_mode) {   // keep the arrival-flag chain the kernels use consistent
                    CUresult r = g_stream_write64((CUstream)s.comm, (CUdeviceptr)(n->d_flags + 1), c->seq, 0);
                    if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWriteValue64 failed with CUresult %d", (int)r);
                }
            } else any_nccl = true;
        }
        if (s.upper) {
            if (Slab* n = local_slab(c, s.gidx + 1)) {   // my last G owned planes -> n's lower ghost
                const double* from = s.buf[b] + (long long)nzl * s.plane;
                double* to = n->buf[b];
                CU_TRY(cudaMemcpyPeerAsync(to, n->dev, from, s.dev, bytes, s.comm));
                if (c->store_mode) {
                    CUresult r = g_stream_write64((CUstream)s.comm, (CUdeviceptr)(n->d_flags + 0), c->seq, 0);
                    if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWriteValue64 failed with CUresult %d", (int)r);
                }
            } else any_nccl = true;
        }
    }
    if (any_nccl && c->slabs[0].ipc) {
        // one-sided puts into the neighbour process's ghost planes by the copy engines, then a flag write that is
        // ordered after the copy in the stream: no SM is needed while the interior kernel owns them all
        for (Slab& s : c->slabs) {
            int rc = set_dev(s);
            if (rc) return rc;
            const size_t bytes = (size_t)s.G * s.plane * sizeof(double);
            const uint32_t nzl = s.z1 - s.z0;
            if (s.lower && !local_slab(c, s.gidx - 1)) {      // my first G owned planes -> lower neighbour's upper ghost
                const uint32_t n_nzl = (uint32_t)((uint64_t)s.gidx * c->p.nz / c->p.nslabs) - (uint32_t)((uint64_t)(s.gidx - 1) * c->p.nz / c->p.nslabs);
                CU_TRY(cudaMemcpyAsync(s.peer_buf[0][b] + (long long)(s.G + n_nzl) * s.plane, s.buf[b] + (long long)s.G * s.plane, bytes,
                                       cudaMemcpyDeviceToDevice, s.comm));
                CUresult r = g_stream_write64((CUstream)s.comm, (CUdeviceptr)(s.peer_flags[0] + 1), c->seq, 0);
                if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWriteValue64 failed with CUresult %d", (int)r);
            }
            if (s.upper && !local_slab(c, s.gidx + 1)) {      // my last G owned planes -> upper neighbour's lower ghost
                CU_TRY(cudaMemcpyAsync(s.peer_buf[1][b], s.buf[b] + (long long)nzl * s.plane, bytes, cudaMemcpyDeviceToDevice, s.comm));
                CUresult r = g_stream_write64((CUstream)s.comm, (CUdeviceptr)(s.peer_flags[1] + 0), c->seq, 0);
                if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWriteValue64 failed with CUresult %d", (int)r);
            }
        }
    } else if (any_nccl) {
        NCCL_TRY(g_nccl.GroupStart());
        for (Slab& s : c->slabs) {
            const size_t cnt = (size_t)s.G * s.plane;
            const uint32_t nzl = s.z1 - s.z0;
            if (s.lower && !local_slab(c, s.gidx - 1)) {
                NCCL_TRY(g_nccl.Send(s.buf[b] + (long long)s.G * s.plane, cnt, ncclFloat64, s.gidx - 1, s.nccl, s.comm));
                NCCL_TRY(g_nccl.Recv(s.buf[b], cnt, ncclFloat64, s.gidx - 1, s.nccl, s.comm));
            }
            if (s.upper && !local_slab(c, s.gidx + 1)) {
                NCCL_TRY(g_nccl.Send(s.buf[b] + (long long)nzl * s.plane, cnt, ncclFloat64, s.gidx + 1, s.nccl, s.comm));
                NCCL_TRY(g_nccl.Recv(s.buf[b] + (long long)(s.G + nzl) * s.plane, cnt, ncclFloat64, s.gidx + 1, s.nccl, s.comm));
            }
        }
        NCCL_TRY(g_nccl.GroupEnd());
    }
    for (Slab& s : c->slabs) {
        if (!s.lower && !s.upper) continue;
        int rc = set_dev(s);
        if (rc) return rc;
        CU_TRY(cudaEventRecord(s.ev_put, s.comm));
    }
    return HEAT3D_OK;
}

// Make slab s's compute stream wait until its ghost planes (filled by the previous exchange) have landed.
int wait_ghosts(heat3d_ctx* c, Slab& s, unsigned long long seq)
{
    for (int side = 0; side < 2; ++side) {
        if (!(side == 0 ? s.lower : s.upper)) continue;
        Slab* n = local_slab(c, s.gidx + (side == 0 ? -1 : 1));
        if (!c->store_mode) {
            // own ev_put: my previous sends have left my planes; neighbour's ev_put: its planes have landed in my ghosts
            CU_TRY(cudaStreamWaitEvent(s.compute, n ? n->ev_put : s.ev_put, 0));
            if (n || !s.ipc) continue;
        }
        if (seq) {   // arrival flag written by the neighbour (its kernel in store mode, its exchange stream otherwise)
            CUresult r = g_stream_wait64((CUstream)s.compute, (CUdeviceptr)(s.d_flags + side), seq, CU_STREAM_WAIT_VALUE_GEQ);
            if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWaitValue64 failed with CUresult %d", (int)r);
        }
    }
    return HEAT3D_OK;
}

// One pass over HBM advancing every slab by k updates (k = 1, or 2 when the context was created with temporal_k >= 2).
int step_block(heat3d_ctx* c, int k, bool res)
{
    auto launch = [&](Slab& s, int src, int za, int zb, bool r, int ends) {
        return k == 2 ? launch_stencil_t2(c, s, src, za, zb, r, ends) : launch_stencil(c, s, src, za, zb, r, ends);
    };
    const int src = c->cur, dst = c->cur ^ 1;
    const bool overlap = !(c->p.flags & HEAT3D_FLAG_NO_OVERLAP);
    bool counted = false;
    const unsigned long long seq_before = c->seq;       // ghosts of `src` arrived with this sequence number
    if (c->store_mode) c->seq++;                        // the number this pass's kernels publish
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        if (c->snap_buf == dst) CU_TRY(cudaStreamWaitEvent(s.compute, s.ev_snap, 0));
        if (res && !c->res_slot_prezeroed) CU_TRY(cudaMemsetAsync(s.d_res, 0, sizeof(unsigned long long), s.compute));
        if (res && k == 2) CU_TRY(cudaMemsetAsync(s.d_res_base + 2, 0, sizeof(unsigned long long), s.compute));
        if (!s.lower && !s.upper) {
            if ((rc = launch(s, src, s.za, s.zb, res, 0))) return rc;
            continue;
        }
        if (c->halo_timed) CU_TRY(cudaEventRecord(s.ev_wait0[c->call_steps % HALO_RING], s.compute));
        if ((rc = wait_ghosts(c, s, seq_before))) return rc;
        if (c->halo_timed) CU_TRY(cudaEventRecord(s.ev_wait1[c->call_steps % HALO_RING], s.compute));
        if (c->store_mode) {
            // ONE launch and nothing else: the end chunks write their outer planes straight into the neighbours' ghost
            // planes and the last CTA of each publishes this pass's sequence number in the neighbour's arrival flag
            if ((rc = launch(s, src, s.za, s.zb, res, (s.lower ? 1 : 0) | (s.upper ? 2 : 0)))) return rc;
            continue;
        }
        if (overlap && c->use_tma && c->end_signal) {
            // ONE launch: the chunks next to a neighbour run first and count themselves done on s.d_done; the exchange
            // stream waits for that count, so the ghost planes leave while this kernel works on the rest of the slab
            if ((rc = launch(s, src, s.za, s.zb, res, (s.lower ? 1 : 0) | (s.upper ? 2 : 0)))) return rc;
            counted = true;
            continue;
        }
        // fallback (no stream memory operations, or the baseline kernel): outer planes in launches of their own,
        // an event after them, then the rest while the exchange is in flight
        const int G = (int)s.G;
        int lo_end = s.lower ? s.za + G : s.za;
        int hi_beg = s.upper ? s.zb - G : s.zb;
        if (lo_end > s.zb) lo_end = s.zb;
        if (hi_beg < lo_end) hi_beg = lo_end;
        if (!overlap) { lo_end = s.za; hi_beg = s.zb; }
        if ((rc = launch(s, src, s.za, lo_end, res, 0))) return rc;
        if ((rc = launch(s, src, hi_beg, s.zb, res, 0))) return rc;
        if (overlap) CU_TRY(cudaEventRecord(s.ev_boundary, s.compute));
        if ((rc = launch(s, src, lo_end, hi_beg, res, 0))) return rc;
        if (!overlap) CU_TRY(cudaEventRecord(s.ev_boundary, s.compute));
    }
    if (c->snap_buf == dst) c->snap_buf = -1;
    if (!c->store_mode && (c->slabs.size() > 1 || c->multi_process)) {
        int rc = exchange_ghosts(c, dst, counted);
        if (rc) return rc;
    }
    c->cur = dst;
    c->tm.steps += k;
    c->call_steps++;
    return HEAT3D_OK;
}

// Communication-avoiding block of FOUR updates on a context with ghost depth 4: two depth-2 passes and ONE exchange.
// Pass A advances the owned planes AND two ghost planes per neighbour side (their inputs, four ghost planes deep, are
// all there), pass B advances the owned planes again from that; then the outer four owned planes travel.  Both passes
// end in the buffer they started from, so `cur` does not change.  The exchange of this block overwrites, in the
// neighbour, ghost planes that ITS pass A of the same block reads, and nothing else orders the two: every slab
// publishes "consumed" to its neighbours after its pass A (a stream-ordered flag write) and puts wait for it.
int step_block4(heat3d_ctx* c, bool res)
{
    const int src = c->cur, mid = c->cur ^ 1;
    const unsigned long long seq_before = c->seq;
    const unsigned long long round = ++c->consumed_seq;
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        if (c->snap_buf >= 0) CU_TRY(cudaStreamWaitEvent(s.compute, s.ev_snap, 0));
        if (res && !c->res_slot_prezeroed) CU_TRY(cudaMemsetAsync(s.d_res, 0, sizeof(unsigned long long), s.compute));
        if (res) CU_TRY(cudaMemsetAsync(s.d_res_base + 2, 0, sizeof(unsigned long long), s.compute));
        if (!s.lower && !s.upper) {
            if ((rc = launch_stencil_t2(c, s, src, s.za, s.zb, false, 0))) return rc;
            if ((rc = launch_stencil_t2(c, s, mid, s.za, s.zb, res, 0))) return rc;
            continue;
        }
        if (c->halo_timed) CU_TRY(cudaEventRecord(s.ev_wait0[c->call_steps % HALO_RING], s.compute));
        if ((rc = wait_ghosts(c, s, seq_before))) return rc;
        if (c->halo_timed) CU_TRY(cudaEventRecord(s.ev_wait1[c->call_steps % HALO_RING], s.compute));
        if ((rc = launch_stencil_t2(c, s, src, s.lower ? s.za - 2 : s.za, s.upper ? s.zb + 2 : s.zb, false, 0))) return rc;
        for (int side = 0; side < 2; ++side) {
            if (!(side == 0 ? s.lower : s.upper) || !s.peer_flags[side]) continue;
            // I am the upper neighbour of the slab below me (its flag [3]) and the lower neighbour of the one above (its flag [2])
            CUresult r = g_stream_write64((CUstream)s.compute, (CUdeviceptr)(s.peer_flags[side] + 2 + (side == 0 ? 1 : 0)), round, 0);
            if (r != CUDA_SUCCESS) return fail(HEAT3D_ERR_CUDA, "cuStreamWriteValue64 failed with CUresult %d", (int)r);
        }
        if ((rc = launch_stencil_t2(c, s, mid, s.za, s.zb, res, (s.lower ? 1 : 0) | (s.upper ? 2 : 0)))) return rc;
    }
    c->snap_buf = -1;
    int rc = exchange_ghosts(c, src, /*after_counted_launch=*/true, round);
    if (rc) return rc;
    c->tm.steps += 4;
    c->call_steps++;
    return HEAT3D_OK;
}

// depth-4 blocks need slabs that talk (otherwise two depth-2 passes are the same thing) and stream memory operations
bool use_block4(const heat3d_ctx* c)
{
    return c->p.temporal_k == 4 && c->variant2 && (c->slabs.size() > 1 || c->multi_process) && c->end_signal && !c->store_mode &&
           c->peers_reachable && !(c->p.flags & HEAT3D_FLAG_NO_OVERLAP);
}

int finish_timing(heat3d_ctx* c);

int begin_call(heat3d_ctx* c)
{
    c->timing_pending = false;         // an unread previous call is dropped from total_ms, never waited for here
    c->call_steps = 0;
    c->tm.last_call_launches = 0;
    c->tm.last_call_stencil_launches = 0;
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        CU_TRY(cudaEventRecord(s.ev_start, s.compute));
    }
    return HEAT3D_OK;
}

int end_call(heat3d_ctx* c)
{
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        CU_TRY(cudaEventRecord(s.ev_stop, s.compute));
    }
    c->timing_pending = true;
    return HEAT3D_OK;
}

// Device-side residual of the last step -> host double (max over local slabs and, with NCCL, processes).
int fetch_residual(heat3d_ctx* c, double* out)
{
    double r = 0.0;
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        if (c->multi_process) {
            // bit patterns of non-negative doubles order like the doubles: reduce as fp64 max.
            CU_TRY(cudaStreamWaitEvent(s.comm, s.ev_stop, 0));
            NCCL_TRY(g_nccl.AllReduce(s.d_res, s.d_scalar, 1, ncclFloat64, ncclMax, s.nccl, s.comm));
            CU_TRY(cudaMemcpyAsync(s.h_scalar, s.d_scalar, sizeof(double), cudaMemcpyDeviceToHost, s.comm));
            if ((rc = watched_sync(c, s, s.comm, "residual all-reduce"))) return rc;
        } else {
            CU_TRY(cudaMemcpyAsync(s.h_scalar, s.d_res, sizeof(double), cudaMemcpyDeviceToHost, s.compute));
            CU_TRY(cudaStreamSynchronize(s.compute));
        }
        if (s.h_scalar[0] > r) r = s.h_scalar[0];
    }
    if (r < DBL_MIN) r = DBL_MIN;     // reference starts its max at numeric_limits<double>::min() (cu:1019)
    *out = r;
    c->have_res = true;
    c->last_res = r;
    return HEAT3D_OK;
}

int finish_timing(heat3d_ctx* c)
{
    if (!c->timing_pending) return HEAT3D_OK;
    double worst = 0.0, halo = 0.0;
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        if ((rc = watched_sync(c, s, s.compute, "compute"))) return rc;
        CU_TRY(cudaEventSynchronize(s.ev_stop));
        float ms = 0.f;
        CU_TRY(cudaEventElapsedTime(&ms, s.ev_start, s.ev_stop));
        if (ms > worst) worst = ms;
        if (c->halo_timed && (s.lower || s.upper) && c->call_steps) {
            const uint32_t n = c->call_steps < HALO_RING ? c->call_steps : HALO_RING;
            double sum = 0.0;
            for (uint32_t i = 0; i < n; ++i) {
                float w = 0.f;
                CU_TRY(cudaEventElapsedTime(&w, s.ev_wait0[i], s.ev_wait1[i]));
                sum += w;
            }
            const double est = sum / n * c->call_steps;    // ring average scaled to the whole call
            if (est > halo) halo = est;
        }
    }
    c->tm.last_call_ms = worst;
    c->tm.total_ms += worst;
    c->tm.exposed_halo_ms = halo;
    c->timing_pending = false;
    return HEAT3D_OK;
}

void destroy_slab(Slab& s)
{
    cudaSetDevice(s.dev);
    for (int side = 0; side < 2; ++side) {
        if (!s.peer_is_ipc[side]) continue;
        for (int b = 0; b < 2; ++b)
            if (s.peer_buf[side][b]) cudaIpcCloseMemHandle(s.peer_buf[side][b]);
        if (s.peer_flags[side]) cudaIpcCloseMemHandle(s.peer_flags[side]);
    }
    if (s.d_flags) cudaFree(s.d_flags);
    if (s.d_side_done) cudaFree(s.d_side_done);
    if (s.nccl && g_nccl.CommDestroy) g_nccl.CommDestroy(s.nccl);
    for (int b = 0; b < 2; ++b)
        if (s.buf[b]) cudaFree(s.buf[b]);
    if (s.d_res_base) cudaFree(s.d_res_base);
    for (cudaEvent_t e : s.ev_res)
        if (e) cudaEventDestroy(e);
    if (s.d_done) cudaFree(s.d_done);
    if (s.snap_dev) cudaFree(s.snap_dev);
    if (s.up_dev) cudaFree(s.up_dev);
    if (s.h2d) cudaStreamDestroy(s.h2d);
    if (s.ev_up) cudaEventDestroy(s.ev_up);
    if (s.d_partial) cudaFree(s.d_partial);
    if (s.d_scalar) cudaFree(s.d_scalar);
    if (s.h_scalar) cudaFreeHost(s.h_scalar);
    cudaEvent_t* evs[] = {&s.ev_boundary, &s.ev_put, &s.ev_start, &s.ev_stop, &s.ev_snap, &s.ev_tmp};
    for (cudaEvent_t* e : evs)
        if (*e) cudaEventDestroy(*e);
    for (cudaEvent_t e : s.ev_wait0) cudaEventDestroy(e);
    for (cudaEvent_t e : s.ev_wait1) cudaEventDestroy(e);
    if (s.compute && s.own_compute) cudaStreamDestroy(s.compute);
    if (s.comm) cudaStreamDestroy(s.comm);
    if (s.copy) cudaStreamDestroy(s.copy);
}

int setup_slab(heat3d_ctx* c, Slab& s)
{
    int rc = set_dev(s);
    if (rc) return rc;
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, s.dev));
    if (prop.major < 10)
        return fail(HEAT3D_ERR_CUDA, "device %d is sm_%d%d; libheat3d is built for sm_100a (B200) only", s.dev, prop.major, prop.minor);
    s.num_sms = prop.multiProcessorCount;
    const heat3d_params& p = c->p;
    s.pitch = (p.nx + 15u) / 16u * 16u;
    s.plane = (long long)s.pitch * p.ny;
    s.planes = (s.z1 - s.z0) + 2 * s.G;
    s.zg0 = (long long)s.z0 - s.G;
    const size_t bytes = (size_t)s.planes * s.plane * sizeof(double);
    for (int b = 0; b < 2; ++b) {
        cudaError_t e = cudaMalloc(&s.buf[b], bytes);
        if (e != cudaSuccess) return fail(HEAT3D_ERR_ALLOC, "cudaMalloc of %zu bytes on device %d failed: %s", bytes, s.dev, cudaGetErrorString(e));
        CU_TRY(cudaMemset(s.buf[b], 0, bytes));
    }
    if (p.stream && c->slabs.size() == 1) { s.compute = (cudaStream_t)p.stream; s.own_compute = false; }
    else CU_TRY(cudaStreamCreateWithFlags(&s.compute, cudaStreamNonBlocking));
    {   // the exchange must not queue behind the interior kernel's CTAs: give its stream the highest priority
        int lo = 0, hi = 0;
        CU_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU_TRY(cudaStreamCreateWithPriority(&s.comm, cudaStreamNonBlocking, hi));
    }
    CU_TRY(cudaStreamCreateWithFlags(&s.copy, cudaStreamNonBlocking));
    CU_TRY(cudaStreamCreateWithFlags(&s.h2d, cudaStreamNonBlocking));
    CU_TRY(cudaEventCreateWithFlags(&s.ev_up, cudaEventDisableTiming));
    cudaEvent_t* sync_evs[] = {&s.ev_boundary, &s.ev_put, &s.ev_snap, &s.ev_tmp};
    for (cudaEvent_t* e : sync_evs) CU_TRY(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    cudaEvent_t* time_evs[] = {&s.ev_start, &s.ev_stop};
    for (cudaEvent_t* e : time_evs) CU_TRY(cudaEventCreate(e));
    if (c->halo_timed) {
        s.ev_wait0.resize(HALO_RING);
        s.ev_wait1.resize(HALO_RING);
        for (uint32_t i = 0; i < HALO_RING; ++i) {
            CU_TRY(cudaEventCreate(&s.ev_wait0[i]));
            CU_TRY(cudaEventCreate(&s.ev_wait1[i]));
        }
    }
    CU_TRY(cudaMalloc(&s.d_res_base, 3 * sizeof(unsigned long long)));   // [0],[1]: slots of the depth-1 loop; [2]: first update of a depth-2 pass
    s.d_res = s.d_res_base;
    for (cudaEvent_t& e : s.ev_res) CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CU_TRY(cudaMalloc(&s.d_done, sizeof(unsigned long long)));
    CU_TRY(cudaMemset(s.d_done, 0, sizeof(unsigned long long)));
    CU_TRY(cudaMalloc(&s.d_side_done, 2 * sizeof(unsigned long long)));
    CU_TRY(cudaMemset(s.d_side_done, 0, 2 * sizeof(unsigned long long)));
    CU_TRY(cudaMalloc(&s.d_flags, 4 * sizeof(unsigned long long)));    // [0],[1] arrival from below / above; [2],[3] consumed by below / above
    CU_TRY(cudaMemset(s.d_flags, 0, 4 * sizeof(unsigned long long)));
    CU_TRY(cudaMemset(s.d_res_base, 0, 3 * sizeof(unsigned long long)));
    s.partial_blocks = s.num_sms * 8;
    CU_TRY(cudaMalloc(&s.d_partial, (size_t)s.partial_blocks * sizeof(double)));
    CU_TRY(cudaMalloc(&s.d_scalar, 8 * sizeof(double)));                // [0..1] reductions, [4..6] all-reduced residual slots
    CU_TRY(cudaHostAlloc(&s.h_scalar, 8 * sizeof(double), cudaHostAllocDefault));   // [0..1] queries, [2..3] pipelined residuals, [4..6] all-reduced slots
    // local output range: owned planes that are global interior planes
    long long a = (long long)s.G, b = (long long)s.G + (s.z1 - s.z0);
    const long long lo = 1 - s.zg0, hi = (long long)p.nz - 1 - s.zg0;
    if (a < lo) a = lo;
    if (b > hi) b = hi;
    if (b < a) b = a;
    s.za = (int)a;
    s.zb = (int)b;
    s.lower = s.gidx > 0;
    s.upper = s.gidx < p.nslabs - 1;
    if (c->use_tma) {
        const TmaVariant& v = *c->variant;
        for (int w = 0; w < 2; ++w)
            if ((rc = make_tmap(s, &s.tmap[w], w, c, v.tx + 4, v.ty + 2))) return rc;
        for (int ar = 0; ar < 2; ++ar)
            for (int r = 0; r < 2; ++r)
                for (int m = 0; m < 2; ++m)
                    CU_TRY(cudaFuncSetAttribute(v.fn[ar][r][m], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v.smem));
        if (c->variant2) {
            const T2Variant& v2 = *c->variant2;
            for (int w = 0; w < 2; ++w)
                if ((rc = make_tmap(s, &s.tmap2[w], w, c, v2.cx + 4, v2.cy + 2))) return rc;
            for (int ar = 0; ar < 2; ++ar)
                for (int r = 0; r < 2; ++r)
                    for (int m = 0; m < 2; ++m)
                        CU_TRY(cudaFuncSetAttribute(v2.fn[ar][r][m], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v2.smem));
        }
    }
    return HEAT3D_OK;
}

// Cross-process neighbours: map their field buffers and arrival flags into this process (CUDA IPC), so ghost planes
// move by copy-engine puts + stream-ordered flag writes instead of NCCL kernels that would have to queue for SMs
// behind the interior stencil.  Handles travel by one ncclAllGather.  Every process must reach the same verdict,
// so the per-process outcome is all-reduced (min) before the path is switched on; on any failure the NCCL
// send/recv path stays in use.
int setup_ipc(heat3d_ctx* c, bool want)
{
    Slab& s = c->slabs[0];
    int rc = set_dev(s);
    if (rc) return rc;
    struct Handles { cudaIpcMemHandle_t buf[2], flags; };
    const int P = c->p.nslabs;
    bool ok = want && load_stream_memops() == HEAT3D_OK;
    Handles mine;
    memset(&mine, 0, sizeof mine);
    if (ok) {
        ok = cudaIpcGetMemHandle(&mine.buf[0], s.buf[0]) == cudaSuccess && cudaIpcGetMemHandle(&mine.buf[1], s.buf[1]) == cudaSuccess &&
             cudaIpcGetMemHandle(&mine.flags, s.d_flags) == cudaSuccess;
        cudaGetLastError();
    }
    unsigned char* d_all = nullptr;
    CU_TRY(cudaMalloc(&d_all, (size_t)(P + 1) * sizeof(Handles)));
    CU_TRY(cudaMemcpy(d_all + (size_t)P * sizeof(Handles), &mine, sizeof mine, cudaMemcpyHostToDevice));
    NCCL_TRY(g_nccl.AllGather(d_all + (size_t)P * sizeof(Handles), d_all, sizeof(Handles), ncclInt8, s.nccl, s.comm));
    CU_TRY(cudaStreamSynchronize(s.comm));
    std::vector<Handles> all(P);
    CU_TRY(cudaMemcpy(all.data(), d_all, (size_t)P * sizeof(Handles), cudaMemcpyDeviceToHost));
    if (ok) {
        for (int side = 0; side < 2 && ok; ++side) {
            const int n = s.gidx + (side == 0 ? -1 : 1);
            if (n < 0 || n >= P) continue;
            for (int b = 0; b < 2 && ok; ++b)
                ok = cudaIpcOpenMemHandle((void**)&s.peer_buf[side][b], all[n].buf[b], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            if (ok) ok = cudaIpcOpenMemHandle((void**)&s.peer_flags[side], all[n].flags, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
            s.peer_is_ipc[side] = true;
        }
        cudaGetLastError();
    }
    // agree: IPC only if it worked everywhere
    double verdict = ok ? 1.0 : 0.0;
    CU_TRY(cudaMemcpy(s.d_scalar, &verdict, sizeof verdict, cudaMemcpyHostToDevice));
    NCCL_TRY(g_nccl.AllReduce(s.d_scalar, s.d_scalar + 1, 1, ncclFloat64, 3 /* ncclMin */, s.nccl, s.comm));
    CU_TRY(cudaMemcpyAsync(&verdict, s.d_scalar + 1, sizeof verdict, cudaMemcpyDeviceToHost, s.comm));
    CU_TRY(cudaStreamSynchronize(s.comm));
    CU_TRY(cudaFree(d_all));
    s.ipc = verdict > 0.5;
    return HEAT3D_OK;
}

// Multi-process contexts: every process has finished all work it queued so far (its puts into neighbours included).
int mp_barrier(heat3d_ctx* c)
{
    if (!c->multi_process) return HEAT3D_OK;
    Slab& s = c->slabs[0];
    int rc = set_dev(s);
    if (rc) return rc;
    if ((rc = watched_sync(c, s, s.compute, "compute"))) return rc;
    if ((rc = watched_sync(c, s, s.comm, "exchange"))) return rc;
    NCCL_TRY(g_nccl.AllReduce(s.d_scalar, s.d_scalar + 1, 1, ncclFloat64, ncclSum, s.nccl, s.comm));
    return watched_sync(c, s, s.comm, "barrier");
}

// compute + exchange streams only: snapshot (copy) and upload (h2d) transfers may keep running across this
int sync_compute(heat3d_ctx* c)
{
    for (Slab& s : c->slabs) {
        int rc = set_dev(s);
        if (rc) return rc;
        if ((rc = watched_sync(c, s, s.compute, "compute"))) return rc;
        if ((rc = watched_sync(c, s, s.comm, "exchange"))) return rc;
    }
    return HEAT3D_OK;
}

int check_ctx(const heat3d_ctx* c)
{
    if (!c) return fail(HEAT3D_ERR_INVALID, "null context");
    return HEAT3D_OK;
}

// rows of nx doubles: host dense <-> device pitched, planes [zl, zl+n) of slab s buffer b
int copy_planes(Slab& s, int b, uint32_t nx, uint32_t ny, uint32_t zl, uint32_t n, double* host, bool to_device, cudaStream_t st)
{
    double* dev = s.buf[b] + (long long)zl * s.plane;
    const size_t rows = (size_t)ny * n;
    if (s.pitch == nx) {   // rows are dense on the device too: one flat copy
        if (to_device) CU_TRY(cudaMemcpyAsync(dev, host, rows * nx * 8, cudaMemcpyHostToDevice, st));
        else CU_TRY(cudaMemcpyAsync(host, dev, rows * nx * 8, cudaMemcpyDeviceToHost, st));
        return HEAT3D_OK;
    }
    if (to_device)
        CU_TRY(cudaMemcpy2DAsync(dev, (size_t)s.pitch * 8, host, (size_t)nx * 8, (size_t)nx * 8, rows, cudaMemcpyHostToDevice, st));
    else
        CU_TRY(cudaMemcpy2DAsync(host, (size_t)nx * 8, dev, (size_t)s.pitch * 8, (size_t)nx * 8, rows, cudaMemcpyDeviceToHost, st));
    return HEAT3D_OK;
}

#include "grid2d.inl"

}  // namespace

int h3::set_last_error(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

// =========================================================================================================
extern "C" {

const char* heat3d_last_error(void) { return g_err; }
int heat3d_version(void) { return HEAT3D_VERSION; }

void heat3d_default_params(heat3d_params* p, uint32_t nx, uint32_t ny, uint32_t nz)
{
    memset(p, 0, sizeof *p);
    p->struct_size = (uint32_t)sizeof *p;
    p->nx = nx; p->ny = ny; p->nz = nz;
    p->arith = HEAT3D_ARITH_STRICT;
    p->temporal_k = 1;
    p->kernel = HEAT3D_KERNEL_AUTO;
    p->nslabs = 1; p->slab_first = 0; p->nlocal = 1;
}

int heat3d_nccl_unique_id(void* out128)
{
    if (!out128) return fail(HEAT3D_ERR_INVALID, "null output");
    int rc = load_nccl();
    if (rc) return rc;
    ncclUniqueId id;
    NCCL_TRY(g_nccl.GetUniqueId(&id));
    memcpy(out128, &id, sizeof id);
    return HEAT3D_OK;
}

int heat3d_create(const heat3d_params* pin, heat3d_ctx** out)
{
    if (!pin || !out) return fail(HEAT3D_ERR_INVALID, "null argument");
    *out = nullptr;
    if (pin->struct_size != sizeof(heat3d_params))
        return fail(HEAT3D_ERR_INVALID, "heat3d_params.struct_size %u != %zu (header/library mismatch)", pin->struct_size, sizeof(heat3d_params));
    heat3d_params p = *pin;
    if (p.nslabs_y > 1) {
        if (p.nx < 3 || p.ny < 3 || p.nz < 3) return fail(HEAT3D_ERR_INVALID, "grid %ux%ux%u: every axis needs >= 3 points", p.nx, p.ny, p.nz);
        return g2_create(p, out);
    }
    if (p.nx < 3 || p.ny < 3 || p.nz < 3) return fail(HEAT3D_ERR_INVALID, "grid %ux%ux%u: every axis needs >= 3 points", p.nx, p.ny, p.nz);
    if (p.nx > (1u << 30) || p.ny > (1u << 30) || p.nz > (1u << 30)) return fail(HEAT3D_ERR_INVALID, "grid axis too large");
    if (p.arith != HEAT3D_ARITH_STRICT && p.arith != HEAT3D_ARITH_FMA) return fail(HEAT3D_ERR_INVALID, "unknown arith %d", p.arith);
    if (p.temporal_k <= 0) p.temporal_k = 1;
    if (p.temporal_k != 1 && p.temporal_k != 2 && p.temporal_k != 4)
        return fail(HEAT3D_ERR_INVALID, "temporal_k=%d not available (1, 2 or 4)", p.temporal_k);
    if (p.temporal_k >= 2 && p.kernel == HEAT3D_KERNEL_BASELINE) return fail(HEAT3D_ERR_INVALID, "temporal_k=%d needs the TMA kernels", p.temporal_k);
    if (p.nslabs <= 0) p.nslabs = 1;
    if (p.nlocal <= 0) p.nlocal = p.nslabs;
    if (p.slab_first < 0 || p.slab_first + p.nlocal > p.nslabs) return fail(HEAT3D_ERR_INVALID, "slab range [%d,%d) outside [0,%d)", p.slab_first, p.slab_first + p.nlocal, p.nslabs);
    if ((uint32_t)p.nslabs > p.nz / (uint32_t)p.temporal_k) return fail(HEAT3D_ERR_INVALID, "%d slabs for %u planes: every slab needs >= %d planes", p.nslabs, p.nz, p.temporal_k);
    const bool multi_process = p.nlocal < p.nslabs;
    if (multi_process && !p.nccl_unique_id) return fail(HEAT3D_ERR_INVALID, "nccl_unique_id required when nlocal < nslabs");
    if (multi_process && p.nlocal != 1) return fail(HEAT3D_ERR_INVALID, "a multi-process run takes one slab per process (nlocal=1), got %d", p.nlocal);

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(HEAT3D_ERR_CUDA, "no usable CUDA device (%s); libheat3d has no CPU fallback", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));

    heat3d_ctx* c = new (std::nothrow) heat3d_ctx();
    if (!c) return fail(HEAT3D_ERR_ALLOC, "out of host memory");
    c->p = p;
    memset(&c->tm, 0, sizeof c->tm);
    c->multi_process = multi_process;
    c->halo_timed = (p.flags & HEAT3D_FLAG_TIME_HALO) != 0;
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    for (int i = 0; i < p.nlocal; ++i) c->devices.push_back(p.devices ? p.devices[i] : cur_dev);
    c->p.devices = nullptr;
    c->p.nccl_unique_id = nullptr;
    for (int d : c->devices)
        if (d < 0 || d >= ndev) { delete c; return fail(HEAT3D_ERR_INVALID, "device ordinal %d out of range [0,%d)", d, ndev); }

    // kernel selection
    c->use_tma = p.kernel != HEAT3D_KERNEL_BASELINE;
    int rc = HEAT3D_OK;
    if (c->use_tma) {
        if ((rc = load_encode())) { delete c; return rc; }
        const auto& vs = variants();
        c->variant = &vs[0];
        if (p.tune[0] || p.tune[1] || p.tune[2] || p.tune[3] || p.tune[4]) {
            c->variant = nullptr;
            for (const TmaVariant& v : vs)
                if ((!p.tune[0] || p.tune[0] == v.tx) && (!p.tune[1] || p.tune[1] == v.ty) && (!p.tune[2] || p.tune[2] == v.ry) &&
                    (!p.tune[3] || p.tune[3] == v.stages) && (!p.tune[4] || p.tune[4] == v.minb)) { c->variant = &v; break; }
            if (!c->variant) { delete c; return fail(HEAT3D_ERR_INVALID, "no TMA kernel variant matches tune={%d,%d,%d,%d,%d}", p.tune[0], p.tune[1], p.tune[2], p.tune[3], p.tune[4]); }
        }
    }

    if (c->use_tma && p.temporal_k >= 2) {
        const auto& vs = t2_variants();
        c->variant2 = &vs[0];
        if (const char* e = getenv("HEAT3D_T2_TUNE")) {
            int t[7] = {0, 0, 0, 0, 0, 0, vs[0].xmode};
            sscanf(e, "%d,%d,%d,%d,%d,%d,%d", &t[0], &t[1], &t[2], &t[3], &t[4], &t[5], &t[6]);
            c->variant2 = nullptr;
            for (const T2Variant& v : vs)
                if (t[0] == v.cx && t[1] == v.cy && t[2] == v.ry && t[3] == v.stages && t[4] == v.minb && t[6] == v.xmode) { c->variant2 = &v; break; }
            if (!c->variant2) { delete c; return fail(HEAT3D_ERR_INVALID, "no depth-2 kernel variant matches HEAT3D_T2_TUNE=%s", e); }
            if (t[5] > 0) { c->t2_chunk = t[5]; c->t2_guided = false; }      // a forced chunk length means uniform chunks
        }
        if (const char* e = getenv("HEAT3D_T2_GUIDED")) c->t2_guided = atoi(e) != 0;
    }

    if (p.nslabs > 1 && c->use_tma) {
        const char* e = getenv("HEAT3D_END_SIGNAL");
        c->end_signal = !(e && atoi(e) == 0) && load_stream_memops() == HEAT3D_OK;
    }

    c->slabs.resize(p.nlocal);
    for (int i = 0; i < p.nlocal; ++i) {
        Slab& s = c->slabs[i];
        s.dev = c->devices[i];
        s.gidx = p.slab_first + i;
        s.z0 = (uint32_t)((uint64_t)s.gidx * p.nz / p.nslabs);
        s.z1 = (uint32_t)((uint64_t)(s.gidx + 1) * p.nz / p.nslabs);
        s.G = (uint32_t)p.temporal_k;
    }
    for (Slab& s : c->slabs)
        if ((rc = setup_slab(c, s))) { heat3d_destroy(c); return rc; }

    // peer access between distinct devices of this context
    for (size_t i = 0; i + 1 < c->slabs.size(); ++i) {
        const int a = c->slabs[i].dev, b = c->slabs[i + 1].dev;
        if (a == b) continue;
        int can = 0;
        cudaDeviceCanAccessPeer(&can, a, b);
        if (can) {
            cudaSetDevice(a); cudaError_t e1 = cudaDeviceEnablePeerAccess(b, 0);
            cudaSetDevice(b); cudaError_t e2 = cudaDeviceEnablePeerAccess(a, 0);
            (void)e1; (void)e2;
            cudaGetLastError();   // "already enabled" is fine
        }
    }
    if (multi_process) {
        if ((rc = load_nccl())) { heat3d_destroy(c); return rc; }
        ncclUniqueId id;
        memcpy(&id, p.nccl_unique_id, sizeof id);
        int r = g_nccl.GroupStart();
        for (Slab& s : c->slabs) {
            cudaSetDevice(s.dev);
            if (r == ncclSuccess) r = g_nccl.CommInitRank(&s.nccl, p.nslabs, id, s.gidx);
        }
        if (r == ncclSuccess) r = g_nccl.GroupEnd();
        if (r != ncclSuccess) { fail(HEAT3D_ERR_NCCL, "ncclCommInitRank failed: %s", g_nccl.GetErrorString(r)); heat3d_destroy(c); return HEAT3D_ERR_NCCL; }
        const char* ex = getenv("HEAT3D_EXCHANGE");
        const bool want_ipc = !(ex && strcmp(ex, "nccl") == 0);
        if ((rc = setup_ipc(c, want_ipc))) { heat3d_destroy(c); return rc; }
    }
    // slabs of this process reach each other by plain pointers (same device, or peer access enabled above)
    bool reachable = true;
    for (Slab& s : c->slabs)
        for (int side = 0; side < 2; ++side) {
            Slab* n = local_slab(c, s.gidx + (side == 0 ? -1 : 1));
            if (!n) continue;
            int can = 1;
            if (n->dev != s.dev) cudaDeviceCanAccessPeer(&can, s.dev, n->dev);
            reachable = reachable && can;
            s.peer_buf[side][0] = n->buf[0];
            s.peer_buf[side][1] = n->buf[1];
            s.peer_flags[side] = n->d_flags;
        }
    c->peers_reachable = reachable;
    {
        // HEAT3D_EXCHANGE=store: the stencil kernel itself writes the outer planes into the neighbours' ghost planes
        // (peer stores over NVLink) and publishes their arrival - no copy, no exchange stream.  Needs every neighbour's
        // memory mapped here and slabs thick enough for two separate end chunks; otherwise the copy paths are used.
        const char* ex = getenv("HEAT3D_EXCHANGE");
        bool ok = ex && strcmp(ex, "store") == 0 && p.temporal_k != 4 && p.nslabs > 1 && c->end_signal && c->use_tma && !(p.flags & HEAT3D_FLAG_NO_OVERLAP) &&
                  reachable && (!multi_process || c->slabs[0].ipc);
        for (int g = 0; g < p.nslabs && ok; ++g) {      // same verdict on every process: look at all slabs, not just mine
            long long a = (long long)((uint64_t)g * p.nz / p.nslabs), b = (long long)((uint64_t)(g + 1) * p.nz / p.nslabs);
            if (a < 1) a = 1;
            if (b > (long long)p.nz - 1) b = (long long)p.nz - 1;
            ok = b - a >= 2 * p.temporal_k;
        }
        c->store_mode = ok;
    }
    cudaSetDevice(c->slabs[0].dev);
    *out = c;
    return HEAT3D_OK;
}

int heat3d_destroy(heat3d_ctx* c)
{
    if (!c) return HEAT3D_OK;
    if (!c->kids.empty()) {
        g2_sync(c);
        g2_destroy(c);
        delete c;
        return HEAT3D_OK;
    }
    for (Slab& s : c->slabs) {
        cudaSetDevice(s.dev);
        cudaDeviceSynchronize();
    }
    // neighbours may still be putting into this process's ghost planes: nobody frees until everybody has drained
    if (c->multi_process && !c->slabs.empty() && c->slabs[0].nccl && c->slabs[0].d_scalar) {
        Slab& s = c->slabs[0];
        cudaSetDevice(s.dev);
        if (g_nccl.AllReduce(s.d_scalar, s.d_scalar + 1, 1, ncclFloat64, ncclSum, s.nccl, s.comm) == ncclSuccess) cudaStreamSynchronize(s.comm);
    }
    for (auto& row : c->step_graph)
        for (cudaGraphExec_t& g : row)
            if (g) { cudaGraphExecDestroy(g); g = nullptr; }
    for (Slab& s : c->slabs) destroy_slab(s);
    delete c;
    return HEAT3D_OK;
}

int heat3d_local_range(const heat3d_ctx* c, uint32_t* z0, uint32_t* z1)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return heat3d_local_range(c->kids[0], z0, z1);
    if (z0) *z0 = c->slabs.front().z0;
    if (z1) *z1 = c->slabs.back().z1;
    return HEAT3D_OK;
}

int heat3d_layout(const heat3d_ctx* c, int32_t i, uint32_t* pitch, uint32_t* planes, uint32_t* ghost)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return heat3d_layout(c->kids[0], i, pitch, planes, ghost);
    if (i < 0 || i >= (int)c->slabs.size()) return fail(HEAT3D_ERR_INVALID, "local slab %d out of range", i);
    if (pitch) *pitch = c->slabs[i].pitch;
    if (planes) *planes = c->slabs[i].planes;
    if (ghost) *ghost = c->slabs[i].G;
    return HEAT3D_OK;
}

const char* heat3d_kernel_name(const heat3d_ctx* c)
{
    if (!c) return "";
    if (!c->kids.empty()) return heat3d_kernel_name(c->kids[0]);
    if (c->variant2) return c->variant2->name;
    return c->use_tma ? c->variant->name : "baseline_64x4";
}

const char* heat3d_exchange_name(const heat3d_ctx* c)
{
    if (!c) return "";
    if (!c->kids.empty()) return c->g2_exchange.c_str();
    if (c->store_mode) return "peer-store";
    if (c->multi_process) return c->slabs[0].ipc ? "ipc-put" : "nccl";
    return c->slabs.size() > 1 ? "peer-copy" : "none";
}

void* heat3d_compute_stream(heat3d_ctx* c, int32_t i)
{
    if (!c || i < 0 || i >= (int)c->slabs.size()) return nullptr;      // (a parent of a y split has no slabs of its own)
    return (void*)c->slabs[i].compute;
}

int heat3d_sync(heat3d_ctx* c)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_sync(c);
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        if ((rc = watched_sync(c, s, s.compute, "compute"))) return rc;
        if ((rc = watched_sync(c, s, s.comm, "exchange"))) return rc;
        CU_TRY(cudaStreamSynchronize(s.copy));
        CU_TRY(cudaStreamSynchronize(s.h2d));
    }
    return HEAT3D_OK;
}

int heat3d_init_reference_field(heat3d_ctx* c, double sp_y)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_init(c, sp_y);
    if ((rc = heat3d_sync(c))) return rc;
    if ((rc = mp_barrier(c))) return rc;      // no neighbour may still be writing into the ghosts this call rewrites
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        h3::init_reference_field_kernel<<<s.num_sms * 8, 256, 0, s.compute>>>(s.buf[0], s.buf[1], (int)c->p.nx, (int)c->p.ny, (int)s.pitch,
                                                                             (int)s.planes, s.zg0, (int)c->p.nz, sp_y, c->y_off, c->ny_global ? (int)c->ny_global : (int)c->p.ny);
        CU_TRY(cudaGetLastError());
        c->tm.launches++;
        // ghosts are filled by the formula itself, so the first step has nothing to wait for
        CU_TRY(cudaEventRecord(s.ev_put, s.compute));
    }
    c->cur = 0;
    c->have_res = false;
    if ((rc = heat3d_sync(c))) return rc;
    // One-sided exchanges (ipc-put, peer-store) write into the NEIGHBOUR's ghost planes, and the first step after an
    // init waits for no arrival flag: nobody may start stepping before every process has finished rewriting its ghosts.
    return mp_barrier(c);
}

int heat3d_upload_planes(heat3d_ctx* c, const double* host, uint32_t z0, uint32_t n)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return host ? g2_transfer(c, const_cast<double*>(host), z0, n, true) : fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (!host) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (z0 < c->slabs.front().z0 || z0 + n > c->slabs.back().z1)
        return fail(HEAT3D_ERR_INVALID, "planes [%u,%u) outside this context's range [%u,%u)", z0, z0 + n, c->slabs.front().z0, c->slabs.back().z1);
    if ((rc = heat3d_sync(c))) return rc;
    if ((rc = mp_barrier(c))) return rc;
    const size_t plane_host = (size_t)c->p.nx * c->p.ny;
    for (Slab& s : c->slabs) {
        const uint32_t a = z0 > s.z0 ? z0 : s.z0, b = (z0 + n) < s.z1 ? (z0 + n) : s.z1;
        if (b <= a) continue;
        if ((rc = set_dev(s))) return rc;
        const uint32_t zl = a - s.z0 + s.G;
        // one trip over PCIe into the current buffer, then an on-device copy: both buffers carry the Dirichlet shell
        if ((rc = copy_planes(s, c->cur, c->p.nx, c->p.ny, zl, b - a, const_cast<double*>(host) + (size_t)(a - z0) * plane_host, true, s.compute))) return rc;
        CU_TRY(cudaMemcpyAsync(s.buf[c->cur ^ 1] + (long long)zl * s.plane, s.buf[c->cur] + (long long)zl * s.plane,
                               (size_t)(b - a) * s.plane * sizeof(double), cudaMemcpyDeviceToDevice, s.compute));
    }
    // refresh ghost planes of both buffers from the neighbours
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        CU_TRY(cudaEventRecord(s.ev_boundary, s.compute));
    }
    if (c->slabs.size() > 1 || c->multi_process)
        for (int w = 0; w < 2; ++w)
            if ((rc = exchange_ghosts(c, w))) return rc;
    c->have_res = false;
    return heat3d_sync(c);
}

int heat3d_upload_planes_async(heat3d_ctx* c, const double* host, uint32_t z0, uint32_t n)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_fail_unsupported("heat3d_upload_planes_async");
    if (!host) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (z0 < c->slabs.front().z0 || z0 + n > c->slabs.back().z1)
        return fail(HEAT3D_ERR_INVALID, "planes [%u,%u) outside this context's range [%u,%u)", z0, z0 + n, c->slabs.front().z0, c->slabs.back().z1);
    const size_t plane_host = (size_t)c->p.nx * c->p.ny;
    for (Slab& s : c->slabs) {
        const uint32_t a = z0 > s.z0 ? z0 : s.z0, b = (z0 + n) < s.z1 ? (z0 + n) : s.z1;
        s.up_a = s.up_b = 0;
        if (b <= a) continue;
        if ((rc = set_dev(s))) return rc;
        const size_t bytes = (size_t)(s.z1 - s.z0) * s.plane * sizeof(double);
        if (s.up_dev_bytes < bytes) {
            if (s.up_dev) CU_TRY(cudaFree(s.up_dev));
            s.up_dev = nullptr; s.up_dev_bytes = 0;
            cudaError_t e = cudaMalloc(&s.up_dev, bytes);
            if (e != cudaSuccess) return fail(HEAT3D_ERR_ALLOC, "cudaMalloc of %zu bytes for the upload landing buffer failed: %s", bytes, cudaGetErrorString(e));
            s.up_dev_bytes = bytes;
        }
        // the landing buffer may still be the source of the previous commit's device-side copies
        CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
        CU_TRY(cudaStreamWaitEvent(s.h2d, s.ev_tmp, 0));
        double* dev = s.up_dev + (size_t)(a - s.z0) * s.plane;
        const double* src = host + (size_t)(a - z0) * plane_host;
        const size_t rows = (size_t)c->p.ny * (b - a), row_bytes = (size_t)c->p.nx * 8;
        if (s.pitch == c->p.nx) CU_TRY(cudaMemcpyAsync(dev, src, rows * row_bytes, cudaMemcpyHostToDevice, s.h2d));
        else CU_TRY(cudaMemcpy2DAsync(dev, (size_t)s.pitch * 8, src, row_bytes, row_bytes, rows, cudaMemcpyHostToDevice, s.h2d));
        CU_TRY(cudaEventRecord(s.ev_up, s.h2d));
        s.up_a = a; s.up_b = b;
    }
    return HEAT3D_OK;
}

int heat3d_upload_commit(heat3d_ctx* c)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_fail_unsupported("heat3d_upload_commit");
    bool any = false;
    for (Slab& s : c->slabs) any = any || s.up_b > s.up_a;
    if (!any) return fail(HEAT3D_ERR_STATE, "heat3d_upload_commit without a staged heat3d_upload_planes_async");
    if ((rc = sync_compute(c))) return rc;
    if ((rc = mp_barrier(c))) return rc;      // no neighbour may still be writing into the ghosts this call rewrites
    for (Slab& s : c->slabs) {
        if (s.up_b <= s.up_a) continue;
        if ((rc = set_dev(s))) return rc;
        CU_TRY(cudaStreamWaitEvent(s.compute, s.ev_up, 0));
        const size_t off_up = (size_t)(s.up_a - s.z0) * s.plane, off_f = (size_t)(s.up_a - s.z0 + s.G) * s.plane;
        const size_t bytes = (size_t)(s.up_b - s.up_a) * s.plane * sizeof(double);
        for (int b = 0; b < 2; ++b)     // both buffers carry the Dirichlet shell
            CU_TRY(cudaMemcpyAsync(s.buf[b] + off_f, s.up_dev + off_up, bytes, cudaMemcpyDeviceToDevice, s.compute));
        s.up_a = s.up_b = 0;
    }
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        CU_TRY(cudaEventRecord(s.ev_boundary, s.compute));
    }
    if (c->slabs.size() > 1 || c->multi_process)
        for (int w = 0; w < 2; ++w)
            if ((rc = exchange_ghosts(c, w))) return rc;
    c->have_res = false;
    if ((rc = sync_compute(c))) return rc;
    return mp_barrier(c);                     // ghosts of every process are in place before anybody steps
}

int heat3d_upload(heat3d_ctx* c, const double* host_global)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return host_global ? g2_transfer(c, const_cast<double*>(host_global), 0, c->p.nz, true) : fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (!host_global) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    const uint32_t z0 = c->slabs.front().z0, z1 = c->slabs.back().z1;
    return heat3d_upload_planes(c, host_global + (size_t)z0 * c->p.nx * c->p.ny, z0, z1 - z0);
}

int heat3d_download_planes(heat3d_ctx* c, double* host, uint32_t z0, uint32_t n)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return host ? g2_transfer(c, host, z0, n, false) : fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (!host) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (z0 < c->slabs.front().z0 || z0 + n > c->slabs.back().z1)
        return fail(HEAT3D_ERR_INVALID, "planes [%u,%u) outside this context's range [%u,%u)", z0, z0 + n, c->slabs.front().z0, c->slabs.back().z1);
    const size_t plane_host = (size_t)c->p.nx * c->p.ny;
    for (Slab& s : c->slabs) {
        const uint32_t a = z0 > s.z0 ? z0 : s.z0, b = (z0 + n) < s.z1 ? (z0 + n) : s.z1;
        if (b <= a) continue;
        if ((rc = set_dev(s))) return rc;
        if ((rc = copy_planes(s, c->cur, c->p.nx, c->p.ny, a - s.z0 + s.G, b - a, host + (size_t)(a - z0) * plane_host, false, s.compute))) return rc;
    }
    return heat3d_sync(c);
}

int heat3d_download(heat3d_ctx* c, double* host_global)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return host_global ? g2_transfer(c, host_global, 0, c->p.nz, false) : fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (!host_global) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    const uint32_t z0 = c->slabs.front().z0, z1 = c->slabs.back().z1;
    return heat3d_download_planes(c, host_global + (size_t)z0 * c->p.nx * c->p.ny, z0, z1 - z0);
}

// GRAPH_BLOCKS passes of depth k as one graph launch (single-slab contexts only; captured on first use).
static int step_graph_blocks(heat3d_ctx* c, int k)
{
    Slab& s = c->slabs[0];
    cudaGraphExec_t& exec = c->step_graph[k - 1][c->cur];
    if (!exec) {
        const uint64_t steps0 = c->tm.steps, launches0 = c->tm.launches, lc0 = c->tm.last_call_launches,
                       ls0 = c->tm.last_call_stencil_launches;
        const uint32_t cs0 = c->call_steps;
        const int cur0 = c->cur;
        const unsigned long long seq0 = c->seq;
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(s.compute, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return 1; }
        int rc = HEAT3D_OK;
        for (uint32_t i = 0; i < GRAPH_BLOCKS && !rc; ++i) rc = step_block(c, k, false);
        const cudaError_t e = cudaStreamEndCapture(s.compute, &graph);
        // capture only records work: undo the bookkeeping the recording did
        c->tm.steps = steps0; c->tm.launches = launches0; c->tm.last_call_launches = lc0; c->tm.last_call_stencil_launches = ls0;
        c->call_steps = cs0;
        c->cur = cur0;          // GRAPH_BLOCKS is even, so a complete capture ends where it began; a failed one may not
        c->seq = seq0;
        if (rc || e != cudaSuccess || !graph) { cudaGetLastError(); if (graph) cudaGraphDestroy(graph); return 1; }
        const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess) { cudaGetLastError(); exec = nullptr; return 1; }
    }
    CU_TRY(cudaGraphLaunch(exec, s.compute));
    c->tm.steps += (uint64_t)k * GRAPH_BLOCKS;
    c->tm.launches += GRAPH_BLOCKS;
    c->tm.last_call_launches += GRAPH_BLOCKS;
    c->tm.last_call_stencil_launches += GRAPH_BLOCKS;
    c->call_steps += GRAPH_BLOCKS;
    return HEAT3D_OK;
}

// `nsteps` plain updates (no residual): graph replays while whole graphs fit, then depth-2 passes, then single steps.
static int advance_plain(heat3d_ctx* c, uint32_t nsteps)
{
    static const bool force_res = getenv("HEAT3D_DEBUG_FORCE_RES") != nullptr;   // experiments: time the residual-fused variant
    static const bool no_graphs = getenv("HEAT3D_NO_GRAPHS") != nullptr;
    int rc = HEAT3D_OK;
    uint32_t left = nsteps;
    if (c->slabs.size() == 1 && !c->multi_process && c->use_tma && c->graphs_ok && !no_graphs && !force_res && c->snap_buf < 0) {
        const int k = c->variant2 ? 2 : 1;
        while (left >= (uint32_t)k * GRAPH_BLOCKS) {
            const int g = step_graph_blocks(c, k);
            if (g == 1) { c->graphs_ok = false; break; }     // capture not possible here: plain launches from now on
            if (g) return g;
            left -= (uint32_t)k * GRAPH_BLOCKS;
        }
    }
    for (; use_block4(c) && left >= 4; left -= 4)
        if ((rc = step_block4(c, force_res))) return rc;
    for (; c->variant2 && left >= 2; left -= 2)
        if ((rc = step_block(c, 2, force_res))) return rc;
    for (; left; --left)
        if ((rc = step_block(c, 1, force_res))) return rc;
    return HEAT3D_OK;
}

int heat3d_step(heat3d_ctx* c, uint32_t nsteps)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_step(c, nsteps, nullptr);
    if ((rc = begin_call(c))) return rc;
    if ((rc = advance_plain(c, nsteps))) return rc;
    if (nsteps) c->have_res = false;
    return end_call(c);
}

int heat3d_step_residual(heat3d_ctx* c, uint32_t nsteps, double* residual)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) {
        if (!residual || nsteps == 0) return fail(HEAT3D_ERR_INVALID, "heat3d_step_residual needs a residual pointer and nsteps >= 1");
        if ((rc = g2_step(c, nsteps, residual))) return rc;
        if (*residual < DBL_MIN) *residual = DBL_MIN;     // cu:1019
        c->last_res = *residual;
        return HEAT3D_OK;
    }
    if (!residual) return fail(HEAT3D_ERR_INVALID, "null residual pointer");
    if (nsteps == 0) return fail(HEAT3D_ERR_INVALID, "nsteps must be >= 1 to produce a residual");
    if ((rc = begin_call(c))) return rc;
    // all but the last block plain; the last block (2 updates if that leaves an even count, else 1) carries the residual
    const uint32_t last = (use_block4(c) && nsteps >= 4 && nsteps % 4 == 0) ? 4u : (c->variant2 && nsteps >= 2 && nsteps % 2 == 0) ? 2u : 1u;
    if ((rc = advance_plain(c, nsteps - last))) return rc;
    if ((rc = last == 4 ? step_block4(c, true) : step_block(c, (int)last, true))) return rc;
    if ((rc = end_call(c))) return rc;
    return fetch_residual(c, residual);
}

int heat3d_residual(heat3d_ctx* c, double* residual)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!residual) return fail(HEAT3D_ERR_INVALID, "null residual pointer");
    if (!c->have_res) return fail(HEAT3D_ERR_STATE, "no residual recorded: the last update was not made by heat3d_step_residual/heat3d_run");
    *residual = c->last_res;
    return HEAT3D_OK;
}

// The reference's loop with the residual examined EVERY iteration, without idling the GPU between iterations:
// iteration i+1 is queued before the host looks at the residual of iteration i (two residual slots, two events).
// If iteration i turns out to be the converging one, iteration i+1 is simply dropped: a depth-1 step never touches
// the buffer it read from, so flipping `cur` back restores the state after iteration i exactly.
static int run_every_iteration_loop(heat3d_ctx* c, uint32_t iter_max, double eps, uint32_t* fin, int32_t* conv, uint32_t* done)
{
    int rc = begin_call(c);
    if (rc) return rc;
    // Residual slot i&1: written by iteration i's kernels; then, on the COPY stream (so the compute stream goes
    // straight on to iteration i+1), copied to the host and cleared for iteration i+2, which waits for that event.
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        CU_TRY(cudaMemsetAsync(s.d_res_base, 0, 2 * sizeof(unsigned long long), s.compute));
    }
    c->res_slot_prezeroed = true;
    auto issue = [&](uint32_t i) -> int {
        const int slot = (int)(i & 1u);
        for (Slab& s : c->slabs) {
            int r0 = set_dev(s);
            if (r0) return r0;
            s.d_res = s.d_res_base + slot;
            if (i >= 2) CU_TRY(cudaStreamWaitEvent(s.compute, s.ev_res[slot], 0));    // slot cleared after iteration i-2
        }
        int r = step_block(c, 1, true);
        for (Slab& s : c->slabs) {
            if (r) break;
            if ((r = set_dev(s))) break;
            CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
            CU_TRY(cudaStreamWaitEvent(s.copy, s.ev_tmp, 0));
            CU_TRY(cudaMemcpyAsync(s.h_scalar + 2 + slot, s.d_res, sizeof(double), cudaMemcpyDeviceToHost, s.copy));
            CU_TRY(cudaMemsetAsync(s.d_res, 0, sizeof(unsigned long long), s.copy));
            CU_TRY(cudaEventRecord(s.ev_res[slot], s.copy));
        }
        for (Slab& s : c->slabs) s.d_res = s.d_res_base;
        return r;
    };
    double norm = 1.0;                                       // cu:325
    uint32_t issued = 0;
    if (iter_max && (rc = issue(issued++))) return rc;
    for (uint32_t time = 0; time < iter_max; ++time) {
        if (issued < iter_max && (rc = issue(issued++))) return rc;     // speculative next iteration
        double res = 0.0;
        for (Slab& s : c->slabs) {
            if ((rc = set_dev(s))) return rc;
            CU_TRY(cudaEventSynchronize(s.ev_res[time & 1u]));
            if (s.h_scalar[2 + (time & 1u)] > res) res = s.h_scalar[2 + (time & 1u)];
        }
        if (res < DBL_MIN) res = DBL_MIN;                    // cu:1019
        ++*done;
        c->have_res = true;
        c->last_res = res;
        if (time == 0 && res != 0.0) norm = res;             // cu:1030-1032
        if (res / norm < eps) {                              // cu:1055-1073
            *conv = 1;
            *fin = time;
            if (issued > time + 1) {                         // drop the speculative iteration
                if ((rc = heat3d_sync(c))) return rc;
                c->cur ^= 1;
                c->tm.steps -= 1;
            }
            break;
        }
    }
    return end_call(c);
}

static int run_every_iteration(heat3d_ctx* c, uint32_t iter_max, double eps, uint32_t* fin, int32_t* conv, uint32_t* done)
{
    const int rc = run_every_iteration_loop(c, iter_max, eps, fin, conv, done);
    // on EVERY exit path, errors included: later residual queries use slot 0 and clear it themselves on the compute stream
    c->res_slot_prezeroed = false;
    for (Slab& s : c->slabs) {
        s.d_res = s.d_res_base;
        if (cudaSetDevice(s.dev) == cudaSuccess) cudaStreamSynchronize(s.copy);
    }
    if (rc) cudaGetLastError();
    return rc;
}

// The same loop on a depth-2 context: one pass advances two iterations and its kernel reports BOTH residuals, so the
// reference's every-iteration check costs one host round trip per two iterations.  No speculation here (the next pass
// would overwrite the buffer a rollback needs): if the FIRST of the two iterations is the converging one, the pass is
// undone by flipping `cur` back and that single iteration is redone with the depth-1 kernel.
static int run_every_iteration_t2(heat3d_ctx* c, uint32_t iter_max, double eps, uint32_t* fin, int32_t* conv, uint32_t* done)
{
    int rc = begin_call(c);
    if (rc) return rc;
    double norm = 1.0;                                       // cu:325
    auto read_pair = [&](double& first, double& second) -> int {
        first = second = 0.0;
        for (Slab& s : c->slabs) {
            int r0 = set_dev(s);
            if (r0) return r0;
            if (c->multi_process) {
                // one all-reduce (max; bit patterns of non-negative doubles order like the doubles) brings every process the
                // residuals of BOTH updates: all of them then take the same decision, one host round trip per two iterations
                CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
                CU_TRY(cudaStreamWaitEvent(s.comm, s.ev_tmp, 0));
                NCCL_TRY(g_nccl.AllReduce(s.d_res_base, s.d_scalar + 4, 3, ncclFloat64, ncclMax, s.nccl, s.comm));
                CU_TRY(cudaMemcpyAsync(s.h_scalar + 4, s.d_scalar + 4, 3 * sizeof(double), cudaMemcpyDeviceToHost, s.comm));
                if ((r0 = watched_sync(c, s, s.comm, "residual all-reduce"))) return r0;
                second = s.h_scalar[4];
                first = s.h_scalar[6];
                break;
            }
            CU_TRY(cudaMemcpyAsync(s.h_scalar + 2, s.d_res_base, sizeof(double), cudaMemcpyDeviceToHost, s.compute));
            CU_TRY(cudaMemcpyAsync(s.h_scalar + 3, s.d_res_base + 2, sizeof(double), cudaMemcpyDeviceToHost, s.compute));
            CU_TRY(cudaStreamSynchronize(s.compute));
            if (s.h_scalar[2] > second) second = s.h_scalar[2];
            if (s.h_scalar[3] > first) first = s.h_scalar[3];
        }
        if (first < DBL_MIN) first = DBL_MIN;                // cu:1019
        if (second < DBL_MIN) second = DBL_MIN;
        return HEAT3D_OK;
    };
    uint32_t time = 0;
    while (time < iter_max) {
        double r1 = 0.0, r2 = 0.0;
        if (time + 1 < iter_max) {
            if ((rc = step_block(c, 2, true))) return rc;
            if ((rc = read_pair(r1, r2))) return rc;
            ++*done;
            if (time == 0 && r1 != 0.0) norm = r1;           // cu:1030-1032
            if (r1 / norm < eps) {                           // iteration `time` converged: keep only that one update
                if ((rc = heat3d_sync(c))) return rc;
                c->cur ^= 1;
                c->tm.steps -= 2;
                if ((rc = step_block(c, 1, true))) return rc;
                double single = 0.0;
                if ((rc = fetch_residual(c, &single))) return rc;
                *conv = 1;
                *fin = time;
                break;
            }
            ++*done;
            c->have_res = true;
            c->last_res = r2;
            if (r2 / norm < eps) { *conv = 1; *fin = time + 1; break; }     // cu:1055-1073
            time += 2;
        } else {
            if ((rc = step_block(c, 1, true))) return rc;
            if ((rc = fetch_residual(c, &r1))) return rc;
            ++*done;
            if (time == 0 && r1 != 0.0) norm = r1;
            if (r1 / norm < eps) { *conv = 1; *fin = time; }
            break;
        }
    }
    return end_call(c);
}

int heat3d_run(heat3d_ctx* c, uint32_t iter_max, double eps, uint32_t check_every, uint32_t* final_iters, int32_t* converged,
               uint32_t* applied)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (check_every == 0) check_every = 1;
    double norm = 1.0;                 // cu:325
    uint32_t fin = 0, done = 0;
    int32_t conv = 0;
    uint32_t time = 0;
    // (the speculative depth-1 loop is single-process; the dual-residual depth-2 loop also runs one process per slab)
    if (check_every == 1 && c->snap_buf < 0 && c->kids.empty() && (!c->multi_process || c->variant2)) {
        if ((rc = c->variant2 ? run_every_iteration_t2(c, iter_max, eps, &fin, &conv, &done)
                              : run_every_iteration(c, iter_max, eps, &fin, &conv, &done))) return rc;
        time = iter_max;
    }
    while (time < iter_max) {
        // next iteration index whose residual is examined: 0, then multiples of check_every (…-1), capped at the end
        uint32_t target = time == 0 ? 0 : ((time / check_every) + 1) * check_every - 1;
        if (target < time) target = time;
        if (target > iter_max - 1) target = iter_max - 1;
        double res = 0.0;
        if ((rc = heat3d_step_residual(c, target - time + 1, &res))) return rc;
        done += target - time + 1;
        if (target == 0 && res != 0.0) norm = res;            // cu:1030-1032
        if (res / norm < eps) { conv = 1; fin = target; break; }   // cu:1055-1073
        time = target + 1;
    }
    if (final_iters) *final_iters = fin;
    if (converged) *converged = conv;
    if (applied) *applied = done;
    return HEAT3D_OK;
}

int heat3d_mean_abs_error_vs_y(heat3d_ctx* c, double sp_y, double* err)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return err ? g2_error(c, sp_y, err) : fail(HEAT3D_ERR_INVALID, "null output pointer");
    if (!err) return fail(HEAT3D_ERR_INVALID, "null output pointer");
    double sum = 0.0;
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        if (s.zb > s.za) {
            h3::abs_error_partial_kernel<<<s.partial_blocks, 256, 0, s.compute>>>(s.buf[c->cur], (int)c->p.nx, (int)c->p.ny, (int)s.pitch,
                                                                                  s.plane, s.za, s.zb, sp_y, s.d_partial, c->y_off);
            CU_TRY(cudaGetLastError());
            h3::sum_partials_kernel<<<1, 256, 0, s.compute>>>(s.d_partial, s.partial_blocks, s.d_scalar);
            CU_TRY(cudaGetLastError());
            c->tm.launches += 2;
        } else {
            CU_TRY(cudaMemsetAsync(s.d_scalar, 0, sizeof(double), s.compute));
        }
        if (c->multi_process) {
            CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
            CU_TRY(cudaStreamWaitEvent(s.comm, s.ev_tmp, 0));
            NCCL_TRY(g_nccl.AllReduce(s.d_scalar, s.d_scalar + 1, 1, ncclFloat64, ncclSum, s.nccl, s.comm));
            CU_TRY(cudaMemcpyAsync(s.h_scalar, s.d_scalar + 1, sizeof(double), cudaMemcpyDeviceToHost, s.comm));
            CU_TRY(cudaStreamSynchronize(s.comm));
        } else {
            CU_TRY(cudaMemcpyAsync(s.h_scalar, s.d_scalar, sizeof(double), cudaMemcpyDeviceToHost, s.compute));
            CU_TRY(cudaStreamSynchronize(s.compute));
        }
        sum += s.h_scalar[0];
        if (c->multi_process) break;   // the all-reduced value is already global
    }
    // the reference divides by an `unsigned` product (cu:1102); keep its wrap-around for grids beyond 2^32 cells
    const unsigned denom = (c->p.nx - 2u) * (c->p.ny - 2u) * (c->p.nz - 2u);
    *err = c->want_raw_error_sum ? sum : sum / denom;
    return HEAT3D_OK;
}

int heat3d_snapshot_async(heat3d_ctx* c, double* pinned)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_fail_unsupported("heat3d_snapshot_async");
    if (!pinned) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    if ((rc = heat3d_snapshot_wait(c))) return rc;           // one snapshot in flight at a time
    const size_t plane_host = (size_t)c->p.nx * c->p.ny;
    bool staged = true;
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        const uint32_t nzl = s.z1 - s.z0;
        const size_t bytes = (size_t)nzl * s.plane * sizeof(double);
        if (s.snap_dev_bytes < bytes) {
            if (s.snap_dev) CU_TRY(cudaFree(s.snap_dev));
            s.snap_dev = nullptr; s.snap_dev_bytes = 0;
            if (cudaMalloc(&s.snap_dev, bytes) == cudaSuccess) s.snap_dev_bytes = bytes;
            else { cudaGetLastError(); s.snap_dev = nullptr; }
        }
        if (s.snap_dev) {
            // a device-to-device copy at HBM speed freezes the time level on the compute stream (milliseconds); the slow
            // trip over PCIe then reads the frozen copy on the copy stream and never holds a field buffer hostage
            CU_TRY(cudaMemcpyAsync(s.snap_dev, s.buf[c->cur] + (long long)s.G * s.plane, bytes, cudaMemcpyDeviceToDevice, s.compute));
            CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
            CU_TRY(cudaStreamWaitEvent(s.copy, s.ev_tmp, 0));
            double* host = pinned + (size_t)s.z0 * plane_host;
            const size_t rows = (size_t)c->p.ny * nzl, row_bytes = (size_t)c->p.nx * 8;
            if (s.pitch == c->p.nx) CU_TRY(cudaMemcpyAsync(host, s.snap_dev, rows * row_bytes, cudaMemcpyDeviceToHost, s.copy));
            else CU_TRY(cudaMemcpy2DAsync(host, row_bytes, s.snap_dev, (size_t)s.pitch * 8, row_bytes, rows, cudaMemcpyDeviceToHost, s.copy));
        } else {
            // no room for a staging copy: read the field buffer itself; the step that would overwrite it waits on the device
            staged = false;
            CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
            CU_TRY(cudaStreamWaitEvent(s.copy, s.ev_tmp, 0));
            if ((rc = copy_planes(s, c->cur, c->p.nx, c->p.ny, s.G, nzl, pinned + (size_t)s.z0 * plane_host, false, s.copy))) return rc;
        }
        CU_TRY(cudaEventRecord(s.ev_snap, s.copy));
    }
    c->snap_buf = staged ? -1 : c->cur;
    return HEAT3D_OK;
}

int heat3d_snapshot_strided_async(heat3d_ctx* c, double* pinned, uint32_t sx, uint32_t sy, uint32_t sz)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return g2_fail_unsupported("heat3d_snapshot_strided_async");
    if (!pinned) return fail(HEAT3D_ERR_INVALID, "null host pointer");
    if (sx == 0 || sy == 0 || sz == 0) return fail(HEAT3D_ERR_INVALID, "strides must be >= 1");
    if (sx == 1 && sy == 1 && sz == 1) return heat3d_snapshot_async(c, pinned);
    if ((rc = heat3d_snapshot_wait(c))) return rc;           // one snapshot in flight at a time
    const uint32_t nxs = (c->p.nx + sx - 1) / sx, nys = (c->p.ny + sy - 1) / sy;
    if (nys > 65535u * 32u) return fail(HEAT3D_ERR_INVALID, "too many output rows");
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        const uint32_t zg_first = (s.z0 + sz - 1) / sz * sz;       // first owned global plane that is a multiple of sz
        if (zg_first >= s.z1) continue;
        const uint32_t np = (s.z1 - 1 - zg_first) / sz + 1;
        const size_t bytes = (size_t)np * nys * nxs * sizeof(double);
        if (s.snap_dev_bytes < bytes) {
            if (s.snap_dev) CU_TRY(cudaFree(s.snap_dev));
            s.snap_dev = nullptr; s.snap_dev_bytes = 0;
            cudaError_t e = cudaMalloc(&s.snap_dev, bytes);
            if (e != cudaSuccess) return fail(HEAT3D_ERR_ALLOC, "cudaMalloc of %zu bytes for the snapshot staging buffer failed: %s", bytes, cudaGetErrorString(e));
            s.snap_dev_bytes = bytes;
        }
        // gather on the compute stream (freezes the time level, reads the field at HBM speed), PCIe copy on the copy stream
        for (uint32_t p0 = 0; p0 < np; p0 += 65535u) {
            const uint32_t n = std::min(65535u, np - p0);
            h3::gather_strided_kernel<<<dim3(nys, n), 256, 0, s.compute>>>(s.buf[c->cur], s.snap_dev + (size_t)p0 * nys * nxs, (int)nxs, (int)nys,
                                                                          (int)sx, (int)sy, (int)sz, (int)(zg_first - s.z0 + s.G + p0 * sz),
                                                                          (int)s.pitch, s.plane);
            CU_TRY(cudaGetLastError());
            c->tm.launches++;
        }
        CU_TRY(cudaEventRecord(s.ev_tmp, s.compute));
        CU_TRY(cudaStreamWaitEvent(s.copy, s.ev_tmp, 0));
        CU_TRY(cudaMemcpyAsync(pinned + (size_t)(zg_first / sz) * nys * nxs, s.snap_dev, bytes, cudaMemcpyDeviceToHost, s.copy));
        CU_TRY(cudaEventRecord(s.ev_snap, s.copy));
    }
    c->snap_buf = -1;       // the staging copy is the only thing the PCIe transfer reads
    return HEAT3D_OK;
}

int heat3d_device_count(int32_t* count)
{
    if (!count) return fail(HEAT3D_ERR_INVALID, "null output");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return fail(HEAT3D_ERR_CUDA, "cudaGetDeviceCount failed: %s (libheat3d has no CPU fallback)", cudaGetErrorString(e));
    *count = n;
    return HEAT3D_OK;
}

int heat3d_snapshot_wait(heat3d_ctx* c)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) return HEAT3D_OK;
    for (Slab& s : c->slabs) {
        if ((rc = set_dev(s))) return rc;
        CU_TRY(cudaStreamSynchronize(s.copy));
    }
    c->snap_buf = -1;
    return HEAT3D_OK;
}

int heat3d_host_alloc(void** out, size_t bytes)
{
    if (!out) return fail(HEAT3D_ERR_INVALID, "null output");
    cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocDefault);
    if (e != cudaSuccess) return fail(HEAT3D_ERR_ALLOC, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
    return HEAT3D_OK;
}

int heat3d_host_free(void* p)
{
    if (p) CU_TRY(cudaFreeHost(p));
    return HEAT3D_OK;
}

int heat3d_get_timing(heat3d_ctx* c, heat3d_timing* t)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    if (!c->kids.empty()) {
        if (!t) return fail(HEAT3D_ERR_INVALID, "null output");
        if ((rc = g2_finish_timing(c))) return rc;
        *t = c->tm;
        return HEAT3D_OK;
    }
    if (!t) return fail(HEAT3D_ERR_INVALID, "null output");
    if ((rc = finish_timing(c))) return rc;
    *t = c->tm;
    return HEAT3D_OK;
}

int heat3d_reset_timing(heat3d_ctx* c)
{
    int rc = check_ctx(c);
    if (rc) return rc;
    memset(&c->tm, 0, sizeof c->tm);
    return HEAT3D_OK;
}

}  // extern "C"
