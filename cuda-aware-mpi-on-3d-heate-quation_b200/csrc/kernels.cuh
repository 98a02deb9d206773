// Device code of libheat3d: sm_100a kernels for the heat3D hot path.
//
// Field layout in HBM (one slab): planes[z_local][y][x], x fastest, row pitch `pitch` doubles
// (multiple of 16 = 128 B), plane stride pitch*ny.  z_local 0 is global plane `zg0` (may be negative:
// ghost planes of the first slab lie outside the domain and are never touched).
//
// Arithmetic of the update (reference HeatEquation3D/src/heat3D.cu:128-131):
//   T' = T + Dx*(T[x+1] - 2T + T[x-1]) + Dy*(T[y+1] - 2T + T[y-1]) + Dz*(T[z+1] - 2T + T[z-1])
// STRICT reproduces the host compiler's evaluation, one rounding per operation:
//   s_d = (T[+1] - 2.0*T) + T[-1];   T' = ((T + Dx*s_x) + Dy*s_y) + Dz*s_z
// FMA reproduces nvcc's contraction of the same source (PTX embedded in the reference's object file):
//   c2 = T+T;  r = fma((T[x+1]-c2)+T[x-1], Dx, T);  r = fma(s_y, Dy, r);  r = fma(s_z, Dz, r)
// Both are split as  partial_xy (needs only plane z)  +  finish_z (needs planes z-1, z+1)  so the
// streaming kernels can touch each staged plane exactly once; the split does not change any rounding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace h3 {

enum { ARITH_STRICT = 0, ARITH_FMA = 1 };

// twice(c) is exact in both modes (2.0*c == c+c), so one value serves partial_xy and finish_z.
__device__ __forceinline__ double twice(double c) { return __dadd_rn(c, c); }

template <int ARITH>
__device__ __forceinline__ double partial_xy2(double c, double c2, double xm, double xp, double ym, double yp, double Dx, double Dy)
{
    const double sx = __dadd_rn(__dsub_rn(xp, c2), xm);
    const double sy = __dadd_rn(__dsub_rn(yp, c2), ym);
    if (ARITH == ARITH_STRICT) return __dadd_rn(__dadd_rn(c, __dmul_rn(Dx, sx)), __dmul_rn(Dy, sy));
    return __fma_rn(sy, Dy, __fma_rn(sx, Dx, c));
}

template <int ARITH>
__device__ __forceinline__ double finish_z2(double t, double c2, double zm, double zp, double Dz)
{
    const double sz = __dadd_rn(__dsub_rn(zp, c2), zm);
    if (ARITH == ARITH_STRICT) return __dadd_rn(t, __dmul_rn(Dz, sz));
    return __fma_rn(sz, Dz, t);
}

template <int ARITH>
__device__ __forceinline__ double partial_xy(double c, double xm, double xp, double ym, double yp, double Dx, double Dy)
{
    if (ARITH == ARITH_STRICT) {
        const double c2 = __dmul_rn(2.0, c);
        const double sx = __dadd_rn(__dsub_rn(xp, c2), xm);
        const double sy = __dadd_rn(__dsub_rn(yp, c2), ym);
        return __dadd_rn(__dadd_rn(c, __dmul_rn(Dx, sx)), __dmul_rn(Dy, sy));
    } else {
        const double c2 = __dadd_rn(c, c);
        const double sx = __dadd_rn(__dsub_rn(xp, c2), xm);
        const double sy = __dadd_rn(__dsub_rn(yp, c2), ym);
        return __fma_rn(sy, Dy, __fma_rn(sx, Dx, c));
    }
}

template <int ARITH>
__device__ __forceinline__ double finish_z(double t, double c, double zm, double zp, double Dz)
{
    if (ARITH == ARITH_STRICT) {
        const double sz = __dadd_rn(__dsub_rn(zp, __dmul_rn(2.0, c)), zm);
        return __dadd_rn(t, __dmul_rn(Dz, sz));
    } else {
        const double sz = __dadd_rn(__dsub_rn(zp, __dadd_rn(c, c)), zm);
        return __fma_rn(sz, Dz, t);
    }
}

// rmax = max(rmax, |o - c|) on the bit patterns: one FP64 subtract, then integer compare/select on the ALU pipe (for
// non-negative doubles the unsigned order of the bits is the order of the values), instead of a second FP64-pipe
// operation per cell in a kernel whose sustained speed is set by the power the FP64 pipe draws.
__device__ __forceinline__ void res_track(double& rmax, double o, double c)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(__dsub_rn(o, c)) & 0x7fffffffffffffffull;
    const unsigned long long m = (unsigned long long)__double_as_longlong(rmax);
    rmax = __longlong_as_double((long long)(b > m ? b : m));
}

// max over a warp of non-negative doubles, then one atomicMax on the bit pattern (order-preserving for >= 0).
__device__ __forceinline__ void warp_residual_commit(double r, unsigned long long* slot)
{
    unsigned long long b = (unsigned long long)__double_as_longlong(r);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, b, o);
        b = other > b ? other : b;
    }
    if ((threadIdx.x & 31) == 0 && b != 0ull) atomicMax(slot, b);
}

struct Geom {
    int nx, ny;            // points per row / rows per plane
    int pitch;             // doubles per row in memory
    long long plane;       // doubles per plane in memory (pitch*ny)
    int za, zb;            // local output plane range [za, zb); every plane in it is a global interior plane
    double Dx, Dy, Dz;
};

// Same, one atomic per CTA: with ~65 000 warps per launch all hitting one address, per-warp atomics showed up as
// 16 % of the kernel time at 512^3.  `scratch` is a shared-memory array of (blockDim/32) u64; every thread calls this.
__device__ __forceinline__ void cta_residual_commit(double r, unsigned long long* slot, unsigned long long* scratch)
{
    unsigned long long b = (unsigned long long)__double_as_longlong(r);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, b, o);
        b = other > b ? other : b;
    }
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = b;
    __syncthreads();
    if (threadIdx.x == 0) {
        const int nw = (blockDim.x * blockDim.y + 31) / 32;
        for (int i = 1; i < nw; ++i) b = scratch[i] > b ? scratch[i] : b;
        if (b != 0ull) atomicMax(slot, b);
    }
}

// ------------------------------------------------------------------------------------------------------
// IC + Dirichlet shell (reference heat3D.cu:409-453), written to both buffers, ghost planes included.
// Order of the reference's writes: zero; y-top plane = 1.0; then x=0 / x=nx-1 / z=0 / z=nz-1 faces =
// j*sp_y — so the four side faces win on the top edges.
// ------------------------------------------------------------------------------------------------------
// (y_off / ny_global: a sub-box of a y-split holds rows [y_off, y_off + ny) of a field with ny_global rows; 0 / ny otherwise)
__global__ void init_reference_field_kernel(double* __restrict__ a, double* __restrict__ b, int nx, int ny, int pitch,
                                            int planes, long long zg0, int nz_global, double sp_y, int y_off, int ny_global)
{
    const long long rows = (long long)ny * planes;
    for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
        const int zl = (int)(row / ny), jl = (int)(row % ny), j = jl + y_off;
        const long long zg = zg0 + zl;
        const bool in_domain = zg >= 0 && zg < nz_global;
        const bool zface = zg == 0 || zg == nz_global - 1;
        double* ra = a + row * pitch;
        double* rb = b + row * pitch;
        for (int i = threadIdx.x; i < pitch; i += blockDim.x) {
            double v = 0.0;
            if (in_domain && i < nx) {
                if (j == ny_global - 1) v = 1.0;
                if (i == 0 || i == nx - 1 || zface) v = (double)j * sp_y;
            }
            ra[i] = v;
            rb[i] = v;
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Baseline stencil: thread per (x,y) column of a z-chunk, three z values in registers, xy neighbours
// through L1.  First correct path; also the fallback for shapes the TMA kernel does not take.
// ------------------------------------------------------------------------------------------------------
template <int ARITH, bool RES>
__global__ void __launch_bounds__(256) stencil_baseline_kernel(const double* __restrict__ src, double* __restrict__ dst,
                                                               Geom g, int zchunk, unsigned long long* res_slot)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int z0 = g.za + blockIdx.z * zchunk;
    const int z1 = min(z0 + zchunk, g.zb);
    const bool active = x >= 1 && x <= g.nx - 2 && y >= 1 && y <= g.ny - 2;
    double rmax = 0.0;
    if (active) {
        const double* p = src + (long long)z0 * g.plane + (long long)y * g.pitch + x;
        double* q = dst + (long long)z0 * g.plane + (long long)y * g.pitch + x;
        double below = __ldg(p - g.plane), c = __ldg(p);
        for (int z = z0; z < z1; ++z) {
            const double above = __ldg(p + g.plane);
            const double t = partial_xy<ARITH>(c, __ldg(p - 1), __ldg(p + 1), __ldg(p - g.pitch), __ldg(p + g.pitch), g.Dx, g.Dy);
            const double r = finish_z<ARITH>(t, c, below, above, g.Dz);
            *q = r;
            if (RES) res_track(rmax, r, c);
            below = c;
            c = above;
            p += g.plane;
            q += g.plane;
        }
    }
    if (RES) warp_residual_commit(rmax, res_slot);
}

// ------------------------------------------------------------------------------------------------------
// TMA-staged z-streaming stencil.
//
// Work = xy tiles x z-chunks; CTA b takes tile b % ntiles of chunk b / ntiles, so CTAs that are resident
// together sweep neighbouring tiles through the same planes at the same time and each other's halo
// columns/rows are L2 hits rather than second trips to HBM.  For a chunk of n output planes the CTA
// TMA-loads n+2 input planes (box (TX+4) x (TY+2) x 1: the tile plus a 2-column / 1-row halo; 2 columns
// keep every thread's pair of cells 16-byte aligned in shared memory) through a STAGES-deep ring guarded by
// full/empty mbarriers.  There is no producer warp (a 17th warp would cost every warp registers): the
// refill duty rotates over the compute warps and runs one plane behind the readers, so it never spins on
// a stage that is still being read.  Each thread owns a 2(x) x RY(y) patch of the tile for the
// whole chunk: when plane p lands it reads its own cells and the four xy neighbours from shared memory
// once, releases the stage, forms partial_xy(p), and finishes plane p-1 with finish_z from the centre
// values of planes p-2, p-1, p it keeps in registers; the finished pair goes to HBM as one 128-bit store.
// Dirichlet cells are never written (both buffers carry the shell).
// ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
template <int TX, int TY, int STAGES>
struct TmaCfg {
    static constexpr int ROW = TX + 4;                                // doubles per staged row
    static constexpr int ROWS = TY + 2;                               // staged rows
    static constexpr int STAGE_BYTES = ROW * ROWS * 8;                // bytes TMA delivers per plane
    static constexpr int STAGE_STRIDE = (STAGE_BYTES + 127) / 128 * 128;
    static constexpr int SMEM_BYTES = STAGES * STAGE_STRIDE + 2 * STAGES * 8;
};

// How one launch covers its output planes.  "End" chunks are the ones whose results neighbouring slabs are waiting
// for: they take the lowest block indices (so they are dispatched first) and every end-chunk CTA bumps `done` when
// its planes are in memory; the exchange stream waits on that counter (a stream memory operation, no extra launch),
// so ghost planes leave while the same kernel is still working on the middle of the slab.
struct TmaWork {
    int ntx, nty;                 // tiles in x, y
    int n_end;                    // 0..2 end chunks, taken by CTAs [0, n_end * ntiles)
    int end_za[2], end_zb[2];     // their output plane ranges
    int mid_za, mid_zb;           // the rest: n_long chunks of chunk_len from mid_za up to mid_split, then chunks of
    int chunk_len;                //           chunk_len2 up to mid_zb (long chunks first, short ones last: the launch's
    int mid_split, n_long;        //           tail is made of short CTAs; mid_split == mid_zb means uniform chunks)
    int chunk_len2;
    unsigned long long* done;     // copy-engine exchange: bumped once per finished end-chunk CTA; nullptr when nobody listens
    // Peer-store exchange (optional): end chunk i also writes the planes [mirror_z0[i], mirror_z1[i]) it produces straight
    // into the neighbouring slab's ghost planes (mirror[i] = the address local plane mirror_z0[i], row 0, column 0 maps
    // to there: the neighbour's memory, reached over NVLink or on the same GPU), and the LAST CTA of that end chunk to
    // finish (side_done[i] reaching side_goal[i]) publishes flag_value in the neighbour's arrival flag.
    double* mirror[2];
    int mirror_z0[2], mirror_z1[2];
    unsigned long long* side_done[2];
    unsigned long long side_goal[2];
    unsigned long long* peer_flag[2];
    unsigned long long flag_value;
};

__device__ __forceinline__ bool chunk_of_cta(const TmaWork& w, int cidx, int& zs, int& cnt)
{
    const bool is_end = cidx < w.n_end;
    if (is_end) {      // (selects, not indexing: a dynamically indexed kernel parameter would be copied to local memory)
        zs = cidx == 0 ? w.end_za[0] : w.end_za[1];
        cnt = (cidx == 0 ? w.end_zb[0] : w.end_zb[1]) - zs;
    } else {
        const int m = cidx - w.n_end;
        if (m < w.n_long) {
            zs = w.mid_za + m * w.chunk_len;
            cnt = min(w.chunk_len, w.mid_split - zs);
        } else {
            zs = w.mid_split + (m - w.n_long) * w.chunk_len2;
            cnt = min(w.chunk_len2, w.mid_zb - zs);
        }
    }
    return is_end;
}

// End-chunk CTA epilogue: make this CTA's stores (local and peer) visible, then count it.
__device__ __forceinline__ void signal_end_chunk(const TmaWork& w, int cidx)
{
    __threadfence_system();       // this thread's stores are visible, also to the peer GPU ...
    __syncthreads();              // ... for every thread of the CTA ...
    if (threadIdx.x == 0) {       // ... before the CTA is counted
        if (w.done) atomicAdd(w.done, 1ULL);
        unsigned long long* sd = cidx == 0 ? w.side_done[0] : w.side_done[1];
        if (sd) {
            const unsigned long long goal = cidx == 0 ? w.side_goal[0] : w.side_goal[1];
            if (atomicAdd(sd, 1ULL) + 1ULL == goal) {      // last CTA of this end chunk: every mirrored plane has landed
                __threadfence_system();
                *(volatile unsigned long long*)(cidx == 0 ? w.peer_flag[0] : w.peer_flag[1]) = w.flag_value;
            }
        }
    }
}

// This thread's address in the neighbour's ghost planes for the chunk's first output plane zs, or nullptr.
__device__ __forceinline__ double* mirror_origin(const TmaWork& w, bool is_end, int cidx, int zs, long long plane, long long row_off)
{
    if (!is_end) return nullptr;
    double* m = cidx == 0 ? w.mirror[0] : w.mirror[1];
    if (!m) return nullptr;
    const int mz0 = cidx == 0 ? w.mirror_z0[0] : w.mirror_z0[1];
    return m + (long long)(zs - mz0) * plane + row_off;
}

// Per-thread store plan, fixed for the whole chunk: which of the thread's cells exist and are interior.
struct StorePlan {
    bool all;          // every cell of the 2 x RY patch is an interior cell: plain 128-bit stores
    bool x0, x1;       // first / second column is interior
    uint32_t rows;     // bit r: row r is interior
};

// Residual contribution of one finished pair (only cells that are actually stored count).
__device__ __forceinline__ void res_pair(double& rmax, const double2& o, const double2& c, const StorePlan& sp, int r)
{
    if (sp.all) {
        res_track(rmax, o.x, c.x);
        res_track(rmax, o.y, c.y);
    } else if ((sp.rows >> r) & 1u) {
        if (sp.x0) res_track(rmax, o.x, c.x);
        if (sp.x1) res_track(rmax, o.y, c.y);
    }
}

// Store one finished pair according to the thread's plan; `m` (nullptr = off) is the same cell in a neighbour's ghost plane.
__device__ __forceinline__ void store_pair(double* __restrict__ o_ptr, double* __restrict__ m, const double2& o, const StorePlan& sp, int r)
{
    if (sp.all) {
        *(double2*)o_ptr = o;
        if (m) *(double2*)m = o;
    } else if ((sp.rows >> r) & 1u) {
        if (sp.x0) {
            o_ptr[0] = o.x;
            if (m) m[0] = o.x;
        }
        if (sp.x1) {
            o_ptr[1] = o.y;
            if (m) m[1] = o.y;
        }
    }
}

template <int RY, int ARITH, bool RES, bool FINISH>
__device__ __forceinline__ void plane_update(const double* __restrict__ S, int row_pitch, double2 (&cn)[RY], const double2 (&cp1)[RY],
                                             const double2 (&cp2)[RY], double2 (&t)[RY], double2 (&tw)[RY],
                                             double* __restrict__ out, double* __restrict__ mir, int pitch, const StorePlan& sp,
                                             double Dx, double Dy, double Dz, double& rmax)
{
    // one visit to the staged plane: own cells, x neighbours (left of .x, right of .y), y neighbours above/below the patch
    double xl[RY], xr[RY];
    const double2 up = *(const double2*)(S - row_pitch);
    const double2 dn = *(const double2*)(S + RY * row_pitch);
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        cn[r] = *(const double2*)(S + r * row_pitch);
        xl[r] = S[r * row_pitch - 1];
        xr[r] = S[r * row_pitch + 2];
    }
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        if (FINISH) {      // plane p-1: t and tw (= 2 * centre) were formed when it was staged
            double2 o;
            o.x = finish_z2<ARITH>(t[r].x, tw[r].x, cp2[r].x, cn[r].x, Dz);
            o.y = finish_z2<ARITH>(t[r].y, tw[r].y, cp2[r].y, cn[r].y, Dz);
            store_pair(out + (long long)r * pitch, mir ? mir + (long long)r * pitch : nullptr, o, sp, r);
            if (RES) res_pair(rmax, o, cp1[r], sp, r);
        }
        const double2 ym = r == 0 ? up : cn[r - 1];
        const double2 yp = r == RY - 1 ? dn : cn[r + 1];
        tw[r].x = twice(cn[r].x);
        tw[r].y = twice(cn[r].y);
        t[r].x = partial_xy2<ARITH>(cn[r].x, tw[r].x, xl[r], cn[r].y, ym.x, yp.x, Dx, Dy);
        t[r].y = partial_xy2<ARITH>(cn[r].y, tw[r].y, cn[r].x, xr[r], ym.y, yp.y, Dx, Dy);
    }
}

// MIRROR = peer-store exchange compiled in (end chunks also write into the neighbours' ghost planes); the plain
// instantiation carries none of its registers or instructions.
template <int TX, int TY, int RY, int STAGES, int ARITH, bool RES, int MINB, bool MIRROR>
__global__ void __launch_bounds__((TX / 2) * (TY / RY), MINB)
    stencil_tma_kernel(const __grid_constant__ CUtensorMap tmap, double* __restrict__ dst, Geom g, TmaWork w,
                       unsigned long long* res_slot)
{
    using Cfg = TmaCfg<TX, TY, STAGES>;
    constexpr int NTHR = (TX / 2) * (TY / RY);
    constexpr int NWARP = NTHR / 32;
    constexpr int ROW = Cfg::ROW;
    static_assert(TX % 64 == 0 && TY % RY == 0 && NTHR % 32 == 0 && NTHR <= 1024, "tile shape");
    static_assert(STAGES >= 3, "the refill runs one plane behind the readers");
    static_assert((NWARP & (NWARP - 1)) == 0, "the refill duty rotates over the warps with a mask");

    extern __shared__ __align__(128) unsigned char smem[];      // stages | full barriers | empty barriers
    uint64_t* full = (uint64_t*)(smem + STAGES * Cfg::STAGE_STRIDE);
    uint64_t* empty = full + STAGES;

    const int tid = threadIdx.x;
    const int ntiles = w.ntx * w.nty;
    const int tile = blockIdx.x % ntiles;
    int zs, cnt;                                                // first output plane / number of output planes of this CTA
    const int cidx = blockIdx.x / ntiles;
    const bool is_end = chunk_of_cta(w, cidx, zs, cnt);
    if (cnt <= 0) return;
    const int niter = cnt + 2;                                  // input planes zs-1 .. zs+cnt
    const int bx = (tile % w.ntx) * TX - 2;                     // TMA box origin
    const int by = (tile / w.ntx) * TY - 1;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NWARP);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < STAGES; ++s)
            if (s < niter) {
                mbar_arrive_expect_tx(&full[s], Cfg::STAGE_BYTES);
                tma_load_3d(smem + s * Cfg::STAGE_STRIDE, &tmap, &full[s], bx, by, zs - 1 + s);
            }
    }
    __syncthreads();

    const int lx = tid % (TX / 2);
    const int ry = tid / (TX / 2);
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int gx = bx + 2 + 2 * lx;                             // global x of this thread's first cell
    const int gy = by + 1 + ry * RY;                            // global y of this thread's first row
    StorePlan sp;
    sp.x0 = gx >= 1 && gx <= g.nx - 2;
    sp.x1 = gx + 1 <= g.nx - 2;                                 // gx+1 >= 1 always
    sp.rows = 0;
#pragma unroll
    for (int r = 0; r < RY; ++r)
        if (gy + r >= 1 && gy + r <= g.ny - 2) sp.rows |= 1u << r;
    sp.all = sp.x0 && sp.x1 && sp.rows == (1u << RY) - 1u;
    double* out = dst + (long long)zs * g.plane + (long long)gy * g.pitch + gx;
    // peer-store exchange: where this thread's patch of output plane zs lives in the neighbour's ghost planes
    double* mir = MIRROR ? mirror_origin(w, is_end, cidx, zs, g.plane, (long long)gy * g.pitch + gx) : nullptr;
    const int mz0 = cidx == 0 ? w.mirror_z0[0] : w.mirror_z0[1], mz1 = cidx == 0 ? w.mirror_z1[0] : w.mirror_z1[1];
    int zout = zs;                                              // output plane the next finished plane goes to
    auto mir_now = [&]() -> double* { return (MIRROR && mir && zout >= mz0 && zout < mz1) ? mir : nullptr; };
    const int soff = (1 + ry * RY) * ROW + 2 + 2 * lx;          // own first cell inside a stage
    double rmax = 0.0;

    // centre values of the three most recent planes rotate through C[0..2] (the plane loop is unrolled by 3
    // so the rotation costs no register moves); t / tw hold partial_xy and 2*centre of the newest complete plane.
    double2 C[3][RY], t[RY], tw[RY];
#pragma unroll
    for (int r = 0; r < RY; ++r) t[r] = tw[r] = make_double2(0.0, 0.0);

    // Ring bookkeeping.  Stage s is read at iteration q (s = q % STAGES) and handed back through empty[s]
    // (one arrival per warp).  The refill runs ONE iteration behind and rotates over the warps: at iteration q,
    // warp (q % NWARP) waits - almost never actually - for empty[stage of q-1] and issues the TMA load of plane
    // q-1+STAGES into it.  Nobody spins on a stage that is still being read, and the bookkeeping is two
    // instructions per warp per plane for everyone else.
    uint32_t s = 0, ph = 0;            // stage / full-barrier parity of the current iteration
    uint32_t sp_prev = 0, ph_prev = 0; // stage / empty-barrier parity of the previous iteration
    auto stage_ptr = [&]() { return (const double*)(smem + s * Cfg::STAGE_STRIDE) + soff; };
    auto release_stage = [&](int q) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        if (q >= 1 && ((q - warp) & (NWARP - 1)) == 0 && q - 1 + STAGES < niter) {
            if (lane == 0) {
                mbar_wait(&empty[sp_prev], ph_prev);
                mbar_arrive_expect_tx(&full[sp_prev], Cfg::STAGE_BYTES);
                tma_load_3d(smem + sp_prev * Cfg::STAGE_STRIDE, &tmap, &full[sp_prev], bx, by, zs - 1 + q - 1 + STAGES);
            }
        }
        sp_prev = s;
        ph_prev = ph;
        if (++s == STAGES) { s = 0; ph ^= 1u; }
    };

    // q = 0: plane zs-1, only its centre values are needed
    {
        const double* S = stage_ptr();
        mbar_wait(&full[s], ph);
#pragma unroll
        for (int r = 0; r < RY; ++r) C[0][r] = *(const double2*)(S + r * ROW);
        release_stage(0);
    }
    // q = 1: plane zs, centre + partial
    {
        const double* S = stage_ptr();
        mbar_wait(&full[s], ph);
        plane_update<RY, ARITH, RES, false>(S, ROW, C[1], C[0], C[0], t, tw, out, mir_now(), g.pitch, sp, g.Dx, g.Dy, g.Dz, rmax);
        release_stage(1);
    }
    // q >= 2: plane p = zs-1+q arrives -> finish plane p-1 (index q-2 of the chunk) and form partial(p)
    int q = 2;
    for (; q + 2 < niter; q += 3) {
        {
            const double* S = stage_ptr();
            mbar_wait(&full[s], ph);
            plane_update<RY, ARITH, RES, true>(S, ROW, C[2], C[1], C[0], t, tw, out, mir_now(), g.pitch, sp, g.Dx, g.Dy, g.Dz, rmax);
            release_stage(q);
            out += g.plane;
            if (MIRROR) { if (mir) mir += g.plane; ++zout; }
        }
        {
            const double* S = stage_ptr();
            mbar_wait(&full[s], ph);
            plane_update<RY, ARITH, RES, true>(S, ROW, C[0], C[2], C[1], t, tw, out, mir_now(), g.pitch, sp, g.Dx, g.Dy, g.Dz, rmax);
            release_stage(q + 1);
            out += g.plane;
            if (MIRROR) { if (mir) mir += g.plane; ++zout; }
        }
        {
            const double* S = stage_ptr();
            mbar_wait(&full[s], ph);
            plane_update<RY, ARITH, RES, true>(S, ROW, C[1], C[0], C[2], t, tw, out, mir_now(), g.pitch, sp, g.Dx, g.Dy, g.Dz, rmax);
            release_stage(q + 2);
            out += g.plane;
            if (MIRROR) { if (mir) mir += g.plane; ++zout; }
        }
    }
    if (q < niter) {
        const double* S = stage_ptr();
        mbar_wait(&full[s], ph);
        plane_update<RY, ARITH, RES, true>(S, ROW, C[2], C[1], C[0], t, tw, out, mir_now(), g.pitch, sp, g.Dx, g.Dy, g.Dz, rmax);
        release_stage(q);
        out += g.plane;
        if (MIRROR) { if (mir) mir += g.plane; ++zout; }
        ++q;
        if (q < niter) {
            const double* S2 = stage_ptr();
            mbar_wait(&full[s], ph);
            plane_update<RY, ARITH, RES, true>(S2, ROW, C[0], C[2], C[1], t, tw, out, mir_now(), g.pitch, sp, g.Dx, g.Dy, g.Dz, rmax);
            release_stage(q);
        }
    }
    if (RES) {
        __syncthreads();                                      // every warp is done with the staged planes
        cta_residual_commit(rmax, res_slot, (unsigned long long*)smem);
    }
    if (is_end && (w.done || (MIRROR && (w.side_done[0] || w.side_done[1])))) signal_end_chunk(w, cidx);
}

// ------------------------------------------------------------------------------------------------------
// K3: pack / unpack of a strided face.  Under a split along y a face is `nrows` rows of every owned plane: nx doubles
// contiguous, then a jump of a whole plane (the reference gathers its Y faces the same way, element by element on the
// host, heat3D.cu:576-589).  pack gathers rows [row0, row0 + nrows) of planes [z0, z0 + nplanes) into a dense buffer
// ordered (plane, row, x); unpack scatters such a buffer into the ghost rows of the receiving sub-box.  One block per
// (row, plane); 128-bit accesses when nx is even (rows start 128-byte aligned: the pitch is a multiple of 16 doubles).
// ------------------------------------------------------------------------------------------------------
template <bool PACK>
__global__ void __launch_bounds__(256) face_rows_kernel(double* __restrict__ field, double* __restrict__ buf, int nx, int nrows, int row0,
                                                        int z0, int pitch, long long plane)
{
    const int r = blockIdx.x, p = blockIdx.y;
    double* frow = field + (long long)(z0 + p) * plane + (long long)(row0 + r) * pitch;
    double* brow = buf + ((long long)p * nrows + r) * nx;
    if ((nx & 1) == 0) {
        double2* f2 = (double2*)frow;
        double2* b2 = (double2*)brow;
        for (int i = threadIdx.x; i < nx / 2; i += 256) {
            if (PACK) b2[i] = f2[i];
            else f2[i] = b2[i];
        }
    } else {
        for (int i = threadIdx.x; i < nx; i += 256) {
            if (PACK) brow[i] = frow[i];
            else frow[i] = brow[i];
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Decimated snapshot: every sx-th / sy-th point of every listed plane, packed densely (x fastest) for the trip over
// PCIe.  Plane p of the output is local plane zl0 + p*sz of the field.  One block per (output row, output plane).
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gather_strided_kernel(const double* __restrict__ f, double* __restrict__ out, int nxs, int nys,
                                                             int sx, int sy, int sz, int zl0, int pitch, long long plane)
{
    const int j = blockIdx.x, p = blockIdx.y;
    const double* row = f + (long long)(zl0 + (long long)p * sz) * plane + (long long)j * sy * pitch;
    double* o = out + ((long long)p * nys + j) * nxs;
    for (int i = threadIdx.x; i < nxs; i += 256) o[i] = row[(long long)i * sx];
}

// ------------------------------------------------------------------------------------------------------
// mean |T - j*sp_y| over interior cells (reference heat3D.cu:1096-1102): pass 1 writes one partial sum per
// block into scratch (fixed block->cells mapping, fixed tree order => deterministic), pass 2 folds them.
// ------------------------------------------------------------------------------------------------------
template <int BLOCK>
__device__ __forceinline__ double block_sum(double v)
{
    __shared__ double sh[BLOCK / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x < 32) {
        t = threadIdx.x < BLOCK / 32 ? sh[threadIdx.x] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
    }
    __syncthreads();
    return t;   // valid in thread 0
}

// rows = (ny-2) * number of owned interior planes; row r -> plane za + r/(ny-2), y = 1 + r%(ny-2).
__global__ void __launch_bounds__(256) abs_error_partial_kernel(const double* __restrict__ f, int nx, int ny, int pitch,
                                                                long long plane, int za, int zb, double sp_y,
                                                                double* __restrict__ partial, int y_off)
{
    const long long rows = (long long)(ny - 2) * (zb - za);
    double acc = 0.0;
    for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
        const int z = za + (int)(row / (ny - 2)), j = 1 + (int)(row % (ny - 2));
        const double yj = (double)(j + y_off) * sp_y;
        const double* p = f + (long long)z * plane + (long long)j * pitch;
        for (int i = 1 + threadIdx.x; i <= nx - 2; i += 256) acc += fabs(__dsub_rn(p[i], yj));
    }
    const double s = block_sum<256>(acc);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

__global__ void __launch_bounds__(256) sum_partials_kernel(const double* __restrict__ partial, int n, double* __restrict__ out)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += partial[i];
    const double s = block_sum<256>(acc);
    if (threadIdx.x == 0) *out = s;
}

}  // namespace h3
