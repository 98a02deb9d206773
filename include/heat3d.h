/* libheat3d — B200-native (sm_100a) heat3D hot path behind a C ABI.
 *
 * What this replaces.  The reference (fredrickhang/Cuda-aware-MPI-on-3D-heate-quation) has no
 * plugin / operator / FFI interface: its hot path is inline in main().  This header is the seam a
 * maintainer would cut there.  It stands in for everything between the reference's "GPU BEGIN" and
 * "GPU END" banners (HeatEquation3D/src/heat3D.cu:644-858) plus the buffer rotate (cu:545-548), halo
 * pack/exchange (cu:562-641, cu:724-755), residual/convergence (cu:1019-1073) and the error norm
 * (cu:1096-1106).  Every entry point below cites the reference lines it takes over ("cu" =
 * HeatEquation3D/src/heat3D.cu, "cpp" = HeatEquation3D/src/heat3D.cpp).  INTEGRATION.md shows the
 * reference-side call sites.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, no C++/torch types.  No exception crosses the boundary.
 *  - Every call returns 0 (HEAT3D_OK) or a negative heat3d_status; heat3d_last_error() gives the text
 *    for the calling thread's most recent failure.
 *  - There is NO CPU fallback: with no usable CUDA device heat3d_create fails with HEAT3D_ERR_CUDA.
 *  - A heat3d_ctx is not thread-safe (the reference is single-threaded per rank); contexts on
 *    disjoint devices are independent.
 *  - Host field layout everywhere: dense fp64, x fastest:  host[(k*ny + j)*nx + i]  (i=x, j=y, k=z) —
 *    the point order of the reference's dump (cu:1169-1171).
 *  - Grid sizes are POINT counts including both Dirichlet boundary points, as argv[1..3] of the
 *    reference (cu:332-335).
 *
 * Decomposition.  The global field is cut into `nslabs` z-slabs with non-overlapping ownership
 * (slab s owns global planes [s*nz/nslabs, (s+1)*nz/nslabs)), each carrying `temporal_k` ghost planes
 * per side.  A context owns `nlocal` consecutive slabs starting at `slab_first`, one CUDA device per
 * slab (devices may repeat).  Slabs in one context exchange ghost planes by peer copies; slabs in
 * different processes by copy-engine puts into CUDA-IPC-mapped ghost planes, or NCCL send/recv where IPC is not
 * available (either way `nccl_unique_id` is required: NCCL carries the bootstrap and the scalar reductions).
 * In a multi-process run heat3d_create, heat3d_init_reference_field, heat3d_upload*, heat3d_step_residual,
 * heat3d_run, heat3d_mean_abs_error_vs_y and heat3d_destroy are collective: every process calls them, in the
 * same order.  nslabs = nlocal = 1 is the plain single-GPU case.  `nslabs_y` > 1 adds a cut along y (row ranges x z-slabs, one
 * process): see heat3d_params.
 */
#ifndef HEAT3D_H
#define HEAT3D_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HEAT3D_VERSION 100 /* 0.1.0 */

typedef enum heat3d_status {
    HEAT3D_OK = 0,
    HEAT3D_ERR_INVALID = -1, /* bad argument / unsupported combination            */
    HEAT3D_ERR_CUDA = -2,    /* CUDA runtime/driver error (includes "no device")   */
    HEAT3D_ERR_NCCL = -3,    /* NCCL missing or failed                             */
    HEAT3D_ERR_STATE = -4,   /* call not valid in the context's current state      */
    HEAT3D_ERR_ALLOC = -5    /* host or device allocation failed                   */
} heat3d_status;

/* Arithmetic order of the update expression (cu:128-131).
 * STRICT: one rounding per operation in the host compiler's order — bit-exact against the reference's
 *         host program;  FMA: the contraction nvcc emits for the reference kernel (3 fused multiply-adds). */
enum { HEAT3D_ARITH_STRICT = 0, HEAT3D_ARITH_FMA = 1 };

/* Stencil kernel selection.  AUTO picks the TMA-staged z-streaming kernels (depth 1, and depth 2 when temporal_k = 2). */
enum { HEAT3D_KERNEL_AUTO = 0, HEAT3D_KERNEL_BASELINE = 1, HEAT3D_KERNEL_TMA = 2 };

typedef struct heat3d_ctx heat3d_ctx;

typedef struct heat3d_params {
    uint32_t struct_size;       /* = sizeof(heat3d_params); guards against ABI drift                      */
    uint32_t nx, ny, nz;        /* global point counts (cu:332-335), each >= 3                             */
    double Dx, Dy, Dz;          /* dt*alpha/sp^2 as the caller derives them (cu:359-367)                  */
    int32_t arith;              /* HEAT3D_ARITH_*                                                          */
    int32_t temporal_k;         /* ghost depth: 1 (0 means 1), 2 or 4; results are bit-identical in every case.
                                   2: two FTCS updates fused per HBM pass (temporal blocking), one ghost exchange per pass.
                                   4: the same kernel, communication-avoiding: blocks of four updates = two passes and ONE
                                      exchange of four planes per side (the first pass also advances two ghost planes)  */
    int32_t kernel;             /* HEAT3D_KERNEL_*                                                         */
    int32_t nslabs;             /* total z-slabs over all processes; 0/1 = single slab                     */
    int32_t slab_first;         /* first slab owned by this context                                        */
    int32_t nlocal;             /* slabs owned by this context; 0 = all (nslabs)                           */
    const int32_t* devices;     /* nlocal CUDA ordinals; NULL = the calling thread's current device        */
    const void* nccl_unique_id; /* 128 bytes from heat3d_nccl_unique_id(), same on every process; only
                                   needed when nlocal < nslabs                                             */
    void* stream;               /* optional cudaStream_t to use as compute stream (nlocal == 1 only)       */
    uint32_t flags;             /* HEAT3D_FLAG_*                                                           */
    int32_t tune[6];            /* kernel tuning overrides (0 = default): tile_x, tile_y, rows_per_thread,
                                   stages, ctas_per_sm, chunk_len                                         */
    int32_t nslabs_y;           /* 0/1 = z-slabs only.  > 1: 2-D decomposition, nslabs_y ranges of rows times nslabs
                                   z-slabs (the reference cuts all three axes, cu:243-263, 385-389); faces between row
                                   ranges are strided in memory and travel through pack / unpack kernels (the reference's
                                   host-side face packing, cu:562-605).  One process only, temporal_k = 1; `devices` then
                                   holds nslabs_y * nslabs ordinals, index iy * nslabs + iz.  Ownership is non-overlapping
                                   with one ghost row per side, so the field equals the single-domain result bit for bit -
                                   the reference's shared-interface scheme with extrapolated edges and averaged corners
                                   (cu:861-1011) is deliberately not reproduced (out of scope)            */
} heat3d_params;

#define HEAT3D_FLAG_NO_OVERLAP 0x1u /* serialise halo exchange after compute (for exposed-halo measurement) */
#define HEAT3D_FLAG_TIME_HALO 0x2u  /* bracket every ghost wait with CUDA events -> heat3d_timing.exposed_halo_ms */

typedef struct heat3d_timing {
    uint64_t steps;            /* time steps advanced since create/reset                                   */
    uint64_t launches;         /* kernels of this library launched since create/reset                      */
    double last_call_ms;       /* device time of the most recent heat3d_step/_residual call (CUDA events
                                  on the compute stream; max over local slabs)                             */
    double total_ms;           /* sum of last_call_ms over the calls that were followed by heat3d_get_timing */
    double exposed_halo_ms;    /* of last_call_ms: time the compute stream sat waiting for ghost planes
                                  (needs HEAT3D_FLAG_TIME_HALO; worst slab)                               */
    uint64_t last_call_launches;
    uint64_t last_call_stencil_launches; /* launches of the dominant (stencil) kernel in the last call    */
} heat3d_timing;

/* ---- lifecycle ---------------------------------------------------------------------------------------- */

/* Fill p with defaults for an nx*ny*nz single-slab strict-arithmetic run (D* left 0). */
void heat3d_default_params(heat3d_params* p, uint32_t nx, uint32_t ny, uint32_t nz);

/* Replaces cu:375-406 (partition + field allocation) and the per-step device setup cu:650-687:
 * allocates the pitched double-buffered fp64 slabs (+ghost planes), streams, events, tensor maps. */
int heat3d_create(const heat3d_params* p, heat3d_ctx** out);

/* Replaces cu:708-716.  Frees everything; NULL is accepted. */
int heat3d_destroy(heat3d_ctx* ctx);

/* 128-byte NCCL bootstrap id, good for ONE heat3d_create per process (process 0 calls this and ships the bytes
 * to the others by any means —
 * the reference used MPI_Init/MPI_Cart_create for the same purpose, cu:203-263). */
int heat3d_nccl_unique_id(void* out128);

/* ---- field ------------------------------------------------------------------------------------------- */

/* cu:409-453 on the device: zero; y-top plane = 1.0; then x/z faces = j*sp_y (in that order). Both buffers. */
int heat3d_init_reference_field(heat3d_ctx* ctx, double sp_y);

/* Whole global field in / out (host layout above).  A context that owns only some slabs reads / writes only
 * its own planes of the array.  Blocking.  Upload stands where the reference did the per-step
 * cudaMemcpy H2D (cu:686-687), download where it did the D2H (cu:697). */
int heat3d_upload(heat3d_ctx* ctx, const double* host_global);
int heat3d_download(heat3d_ctx* ctx, double* host_global);

/* Same for a run of global planes [z0, z0+nplanes) that must lie inside this context's ownership;
 * `host_planes` points at plane z0. */
int heat3d_upload_planes(heat3d_ctx* ctx, const double* host_planes, uint32_t z0, uint32_t nplanes);
int heat3d_download_planes(heat3d_ctx* ctx, double* host_planes, uint32_t z0, uint32_t nplanes);

/* Streaming variant of heat3d_upload_planes for callers that feed one field after another: _async starts the
 * host->device copy of PINNED host memory on a stream of its own into a landing buffer (allocated on first use, one
 * slab's worth per GPU) and returns at once - the copy overlaps heat3d_step calls already queued and a snapshot in
 * flight; heat3d_upload_commit waits for it and installs the planes as the current time level (device-side copies
 * into both buffers + ghost refresh; collective in a multi-process run).  One staged upload at a time; `host_planes`
 * must stay valid until the commit.  Together with heat3d_snapshot_async this keeps PCIe busy in both directions
 * while the GPU steps, where the reference copied the whole field both ways around every single update (cu:686-697). */
int heat3d_upload_planes_async(heat3d_ctx* ctx, const double* pinned_host_planes, uint32_t z0, uint32_t nplanes);
int heat3d_upload_commit(heat3d_ctx* ctx);

/* Global plane range [*z0, *z1) owned by this context. */
int heat3d_local_range(const heat3d_ctx* ctx, uint32_t* z0, uint32_t* z1);

/* ---- time stepping ----------------------------------------------------------------------------------- */

/* Advance exactly nsteps FTCS updates (cu:128-131 over 1..n-2 on every axis; cu:545-548 becomes a pointer
 * swap; cu:562-641 + cu:724-853 become the ghost-plane exchange).  Asynchronous: returns once the work is
 * queued; any query below or heat3d_sync waits. */
int heat3d_step(heat3d_ctx* ctx, uint32_t nsteps);

/* As heat3d_step, and returns max(DBL_MIN, max_interior |T_new - T_old|) of the LAST of the nsteps updates
 * (cu:1019-1024), reduced on the device inside the stencil kernel; max over all slabs/processes.  Blocking. */
int heat3d_step_residual(heat3d_ctx* ctx, uint32_t nsteps, double* residual);

/* The reference's time loop (cu:541-1074): iterate until res/norm < eps or iter_max; norm = residual of
 * iteration 0 (cu:1030-1032).  The residual is examined on iteration 0 and every `check_every`-th thereafter
 * (check_every = 1 reproduces the reference's iteration count exactly, and the field is then the state after
 * exactly that iteration).  With check_every = 1 the GPU is not left waiting for the host's check: on a depth-1
 * context the next iteration is queued ahead and dropped again if it turns out not to be needed; on a depth-2
 * context one pass yields the residuals of both its iterations (also one process per slab: a single 3-double all-reduce per pass
 * gives every process both values).  *final_iters receives the reference's
 * finalNumIterations (0-based index of the converging iteration, cu:1070-1073; 0 if not converged), *converged its
 * globalBreakCondition, *applied (optional) the number of updates performed. */
int heat3d_run(heat3d_ctx* ctx, uint32_t iter_max, double eps, uint32_t check_every,
               uint32_t* final_iters, int32_t* converged, uint32_t* applied);

/* Residual recorded by the most recent heat3d_step_residual / heat3d_run check. */
int heat3d_residual(heat3d_ctx* ctx, double* residual);

/* mean over interior cells of |T - j*sp_y| (cu:1096-1102; the reference prints 100x this as
 * "L2-norm error").  Deterministic two-pass device reduction; agrees with the reference's serial sum to
 * ~1e-15 relative, not bitwise.  In a multi-process run every process gets the global mean. */
int heat3d_mean_abs_error_vs_y(heat3d_ctx* ctx, double sp_y, double* err);

/* The same quantity for callers that may not allocate device memory (no counterpart in the reference; added for
 * hosts whose device memory is wholly owned by their own pool/allocator, or that cannot tolerate cudaMalloc's
 * implicit synchronisation).  Works on the CALLER's buffers on the current device: d_field = nz planes of ny rows
 * of pitch_doubles doubles (x fastest, plane 0 first), d_err = one double; both device pointers.  Launches on
 * `stream` (a cudaStream_t, NULL = default stream) and returns without waiting.  Scratch lives in a fixed array
 * inside the library: at most HEAT3D_NOALLOC_MAX_PLANES interior planes (nz - 2), larger fields are refused;
 * calls that overlap on one GPU race on that array, so serialise them; the first call on each device resolves the
 * array's address and should happen outside stream capture.  Never calls cudaMalloc/cudaFree. */
#define HEAT3D_NOALLOC_MAX_PLANES 8192
int heat3d_mean_abs_error_vs_y_noalloc(const double* d_field, uint32_t nx, uint32_t ny, uint32_t nz, uint32_t pitch_doubles,
                                       double sp_y, double* d_err, void* stream);

/* ---- snapshots, sync, diagnostics --------------------------------------------------------------------- */

/* Asynchronous copy of the current time level into PINNED host memory (global layout; own planes only), overlapping
 * further heat3d_step calls: the level is first frozen by a device-to-device copy into a staging buffer (allocated
 * on first use, one slab's worth of planes per GPU; milliseconds at HBM speed, ordered on the compute stream), then
 * travels over PCIe on a copy stream while stepping continues.  If the staging buffer cannot be allocated the field
 * buffer itself is read and the step that would overwrite it waits on the device, not the host.  One snapshot is
 * in flight at a time (a second call waits for the first).  Replaces the blocking gather cu:1114-1124.
 * heat3d_snapshot_wait blocks until the copy has landed. */
int heat3d_snapshot_async(heat3d_ctx* ctx, double* pinned_host_global);
int heat3d_snapshot_wait(heat3d_ctx* ctx);

/* Decimated variant for grids whose full dump is not practical (the reference's text format is 61 bytes per point,
 * cu:1165-1177: 65 GB at 1024^3): every sx-th / sy-th / sz-th point (indices 0, s, 2s, ...) of the current time level
 * is packed on the device and only the packed array crosses PCIe.  `pinned_host_strided` is the dense decimated GLOBAL
 * array [ceil(nz/sz)][ceil(ny/sy)][ceil(nx/sx)], x fastest; a context writes the planes it owns.  Same rules as
 * heat3d_snapshot_async (one in flight; heat3d_snapshot_wait).  Strides (1,1,1) are heat3d_snapshot_async. */
int heat3d_snapshot_strided_async(heat3d_ctx* ctx, double* pinned_host_strided, uint32_t sx, uint32_t sy, uint32_t sz);

/* Pinned host allocation helpers (cudaHostAlloc / cudaFreeHost). */
int heat3d_host_alloc(void** out, size_t bytes);
int heat3d_host_free(void* p);

int heat3d_sync(heat3d_ctx* ctx);
int heat3d_get_timing(heat3d_ctx* ctx, heat3d_timing* t);
int heat3d_reset_timing(heat3d_ctx* ctx);

/* cudaStream_t of local slab `i`'s compute stream, for callers that record their own events. */
void* heat3d_compute_stream(heat3d_ctx* ctx, int32_t local_slab);

/* Row pitch in doubles and planes-per-buffer of local slab i (diagnostics / tests). */
int heat3d_layout(const heat3d_ctx* ctx, int32_t local_slab, uint32_t* pitch_doubles, uint32_t* planes_alloc,
                  uint32_t* ghost_depth);

const char* heat3d_last_error(void);
int heat3d_version(void);

/* Number of CUDA devices this process can use (the reference asks cudaGetDeviceCount every step, cu:650-654); lets a
 * host-only caller such as the heat3D driver map ranks to devices without linking the CUDA runtime itself. */
int heat3d_device_count(int32_t* count);

/* Name of the stencil kernel variant the context dispatches to (static string). */
const char* heat3d_kernel_name(const heat3d_ctx* ctx);

/* How ghost planes travel in this context (static string): "none" (single slab), "peer-copy" (slabs of one process:
 * cudaMemcpyPeerAsync), "ipc-put" (other processes: copy-engine puts into CUDA-IPC-mapped ghost planes + stream-ordered
 * arrival flags; default), "nccl" (ncclSend/ncclRecv; used when IPC is unavailable or HEAT3D_EXCHANGE=nccl is set) or
 * "peer-store" (HEAT3D_EXCHANGE=store: the stencil kernel stores its outer planes directly into the neighbours' ghost
 * planes and raises their arrival flags itself). */
const char* heat3d_exchange_name(const heat3d_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* HEAT3D_H */
