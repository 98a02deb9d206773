// Temporal blocking, depth 2: two FTCS updates per trip through HBM.
//
// Same z-streaming skeleton as stencil_tma_kernel (kernels.cuh), with a second stencil level stacked on the
// first inside the CTA:
//   level 0  planes of time n arrive by TMA, box (CX+4) x (CY+2), exactly as in the depth-1 kernel;
//   level 1  every thread forms its 2 x RY patch of the time n+1 plane over the whole CX x CY compute region
//            (Dirichlet cells and Dirichlet planes are carried over unchanged, as the reference never updates
//            them), keeps the centre values in registers and publishes the plane to a double-buffered shared
//            "mid" plane;
//   level 2  one iteration later each thread reads the xy neighbours of its level-1 cells from the mid plane
//            and forms the time n+2 plane over the inner (CX-4) x (CY-2) region, which is what goes to HBM.
// Nothing but centre values lives across iterations: a staged plane stays resident one iteration longer and the
// whole update of plane p-1 is formed when plane p has arrived (both levels), which is what lets a thread own 4 rows
// inside 128 registers.  Every cell sees exactly the operations, in exactly the order, of two separate depth-1
// sweeps (heat3D.cu:128-131 twice): results are bit-identical to stepping twice.
// HBM traffic is one read + one write per cell per TWO updates (8 B per cell-update); the price is the redundant
// level-1 ring (CX*CY / ((CX-4)*(CY-2)) of the level-1 work) and one CTA barrier per plane.  The barrier also tells
// the refill thread that the older stage is free, so this kernel needs no empty barriers.
//
// Shared-memory access modes (template XMODE).  A thread's x neighbours are 8-byte words at a 16-byte lane stride: read
// straight, a warp touches only half of the banks and every such load costs 4 wavefronts instead of 2 (ncu, round 1:
// LSU data pipe at 77 %, a third of its wavefronts bank conflicts).  Two remedies, selectable per level:
//   bit 0 (staged planes, level 1): "swapped" loads - lanes 8-15 and 24-31 fetch their RIGHT neighbour with the
//          instruction that gives lanes 0-7 and 16-23 their LEFT one (and the other way round with the second
//          instruction), so the 16 lanes that share a wavefront cover all 32 banks; two selects per value put the
//          words back in place (TMA writes the stage, so its layout cannot change);
//   bit 1 (mid plane, level 2): the mid plane is stored split by x parity (even-x cells in one array, odd-x cells in
//          another), so left / right neighbours are 8-byte words at consecutive lanes - no selects, two 64-bit stores
//          instead of one 128-bit store per row.
// Measured (profiles/r02_*): both bits set remove the conflicts (LSU pipe 77 % -> 56 %) but the selects cost as many issue
// slots as the wavefronts saved; bit 1 alone is the default ("_P": LSU 66 %, 216 M instead of 260 M instructions per
// 512^3 launch).  The kernel is then co-limited by FP64 issue (2 cycles per warp instruction, 208 per thread and plane)
// and instruction issue, see DESIGN.md 3.1b.
#pragma once
#include "kernels.cuh"

namespace h3 {

template <int CX, int CY, int STAGES>
struct Tma2Cfg {
    static constexpr int OX = CX - 4, OY = CY - 2;                    // output tile
    static constexpr int ROW = CX + 4;
    static constexpr int ROWS = CY + 2;
    static constexpr int STAGE_BYTES = ROW * ROWS * 8;
    static constexpr int STAGE_STRIDE = (STAGE_BYTES + 127) / 128 * 128;
    static constexpr int SMEM_BYTES = (STAGES + 2) * STAGE_STRIDE + 2 * STAGES * 8;   // stages | 2 mid planes | full barriers | empty barriers
};

// Per-thread constants of the two levels.
struct T2Plan {
    bool all1;            // every cell of the patch is a domain-interior cell (level 1 needs no carry-over select)
    bool i1x0, i1x1;      // column interior?
    uint32_t rows1;       // bit r: row interior?
    StorePlan st;         // level-2 stores: inside the CTA's output tile AND domain-interior
};

// whole update of one plane in one go (same operations, same order as partial_xy2 followed by finish_z2)
template <int RY, int ARITH>
__device__ __forceinline__ void level_full(const double2 (&c)[RY], const double (&xl)[RY], const double (&xr)[RY], const double2& up,
                                           const double2& dn, const double2 (&zm)[RY], const double2 (&zp)[RY], double2 (&o)[RY],
                                           double Dx, double Dy, double Dz)
{
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        const double2 ym = r == 0 ? up : c[r - 1];
        const double2 yp = r == RY - 1 ? dn : c[r + 1];
        const double twx = twice(c[r].x), twy = twice(c[r].y);
        const double tx = partial_xy2<ARITH>(c[r].x, twx, xl[r], c[r].y, ym.x, yp.x, Dx, Dy);
        const double ty = partial_xy2<ARITH>(c[r].y, twy, c[r].x, xr[r], ym.y, yp.y, Dx, Dy);
        o[r].x = finish_z2<ARITH>(tx, twx, zm[r].x, zp[r].x, Dz);
        o[r].y = finish_z2<ARITH>(ty, twy, zm[r].y, zp[r].y, Dz);
    }
}

// one lane of a converged warp (the PTX way to say "somebody do this once"; lets ptxas keep the TMA issue on the uniform path)
__device__ __forceinline__ bool elect_one_sync()
{
    uint32_t pred;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
    return pred != 0;
}

// named barrier 2 + pair between the two warps `pair` and `pair + 1`
__device__ __forceinline__ void pair_barrier(int pair)
{
    asm volatile("bar.sync %0, 64;" ::"r"(2 + pair) : "memory");
}

// xy neighbours of a thread's 2 x RY patch, read from a staged (or mid) plane; S points at the thread's first cell.
// XSW: `swp` lanes (8-15, 24-31) load right-then-left, the others left-then-right, see the header comment.
template <int RY, bool XSW>
__device__ __forceinline__ void load_neighbours(const double* __restrict__ S, int row_pitch, bool swp, double (&xl)[RY], double (&xr)[RY],
                                                double2& up, double2& dn)
{
    up = *(const double2*)(S - row_pitch);
    dn = *(const double2*)(S + RY * row_pitch);
    if (XSW) {
        const double* A = S + (swp ? 2 : -1);
        const double* B = S + (swp ? -1 : 2);
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            const double a = A[r * row_pitch], b = B[r * row_pitch];
            xl[r] = swp ? b : a;
            xr[r] = swp ? a : b;
        }
    } else {
#pragma unroll
        for (int r = 0; r < RY; ++r) {
            xl[r] = S[r * row_pitch - 1];
            xr[r] = S[r * row_pitch + 2];
        }
    }
}

// the same from a mid plane stored split by x parity: E / O point at the thread's even / odd cell of its first row,
// MP = words per row of one parity array
template <int RY>
__device__ __forceinline__ void load_neighbours_split(const double* __restrict__ E, const double* __restrict__ O, int MP, double (&xl)[RY],
                                                      double (&xr)[RY], double2& up, double2& dn)
{
    up = make_double2(E[-MP], O[-MP]);
    dn = make_double2(E[RY * MP], O[RY * MP]);
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        xl[r] = O[r * MP - 1];      // the odd cell of the pair to the left: consecutive lanes, consecutive words
        xr[r] = E[r * MP + 1];      // the even cell of the pair to the right
    }
}

template <int CX, int CY, int RY, int STAGES, int ARITH, bool RES, int MINB, int XMODE, bool MIRROR>
__global__ void __launch_bounds__((CX / 2) * (CY / RY), MINB)
    stencil_tma_t2_kernel(const __grid_constant__ CUtensorMap tmap, double* __restrict__ dst, Geom g, TmaWork w, int z1lo, int z1hi,
                          unsigned long long* res_slot, unsigned long long* res_slot_first)
{
    using Cfg = Tma2Cfg<CX, CY, STAGES>;
    constexpr int NTHR = (CX / 2) * (CY / RY);
    constexpr int NWARP = NTHR / 32;
    constexpr int ROW = Cfg::ROW;
    static_assert(CX % 64 == 0 && CY % RY == 0 && NTHR % 32 == 0 && NTHR <= 1024 && STAGES >= 3, "tile shape");
    static_assert((NWARP & (NWARP - 1)) == 0, "the refill duty rotates over the warps with a mask");

    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* mid = smem + STAGES * Cfg::STAGE_STRIDE;
    uint64_t* full = (uint64_t*)(smem + (STAGES + 2) * Cfg::STAGE_STRIDE);
    uint64_t* empty = full + STAGES;                            // PSYNC only
    constexpr bool PSYNC = (XMODE & 4) != 0;                    // neighbour-to-neighbour barriers instead of one CTA barrier per plane
    static_assert(!PSYNC || NWARP <= 14, "one named barrier per pair of neighbouring warps");

    const int tid = threadIdx.x;
    const int ntiles = w.ntx * w.nty;
    const int tile = blockIdx.x % ntiles;
    int zs, cnt;                                                // first output (time n+2) plane / number of output planes
    const int cidx = blockIdx.x / ntiles;
    const bool is_end = chunk_of_cta(w, cidx, zs, cnt);
    if (cnt <= 0) return;
    const int niter = cnt + 4;                                  // input planes zs-2 .. zs+cnt+1
    // output tile origin: x tiles start at column 0 (pairs stay 16-byte aligned), y tiles at the first interior row
    const int ox = (tile % w.ntx) * Cfg::OX, oy = 1 + (tile / w.ntx) * Cfg::OY;
    const int bx = ox - 4, by = oy - 2;                         // TMA box origin

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            if (PSYNC) mbar_init(&empty[s], NWARP);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int s = 0; s < STAGES; ++s)
            if (s < niter) {
                mbar_arrive_expect_tx(&full[s], Cfg::STAGE_BYTES);
                tma_load_3d(smem + s * Cfg::STAGE_STRIDE, &tmap, &full[s], bx, by, zs - 2 + s);
            }
    }
    __syncthreads();

    const int lx = tid % (CX / 2);
    const int ry = tid / (CX / 2);
    const int lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);     // known to be warp-uniform: branches on it stay uniform
    const bool swp = (lane >> 3) & 1;
    constexpr bool XSW1 = (XMODE & 1) != 0;                     // swapped x-neighbour loads from the staged planes
    constexpr bool MIDP = (XMODE & 2) != 0;                     // mid plane split by x parity
    const int gx = ox - 2 + 2 * lx;                             // global x of this thread's first cell
    const int gy = oy - 1 + ry * RY;                            // global y of this thread's first row
    // A warp is one 64-cell run of RY rows.  Warps that lie wholly outside the domain (partial tiles at the high
    // edges) only keep the barriers and the refill duty: their cells feed nothing that is stored.
    const bool active = gy <= g.ny - 1 && (ox - 2 + 2 * (lx & ~31)) <= g.nx - 1;
    T2Plan pl;
    pl.i1x0 = gx >= 1 && gx <= g.nx - 2;
    pl.i1x1 = gx + 1 >= 1 && gx + 1 <= g.nx - 2;
    pl.rows1 = 0;
    pl.st.rows = 0;
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        const bool in = gy + r >= 1 && gy + r <= g.ny - 2;
        const int cyi = ry * RY + r;
        if (in) pl.rows1 |= 1u << r;
        if (in && cyi >= 1 && cyi <= CY - 2) pl.st.rows |= 1u << r;
    }
    pl.all1 = pl.i1x0 && pl.i1x1 && pl.rows1 == (1u << RY) - 1u;
    const bool xin = lx >= 1 && lx <= CX / 2 - 2;
    pl.st.x0 = xin && pl.i1x0;
    pl.st.x1 = xin && pl.i1x1;
    pl.st.all = pl.st.x0 && pl.st.x1 && pl.st.rows == (1u << RY) - 1u;
    double* out = dst + (long long)zs * g.plane + (long long)gy * g.pitch + gx;
    double* mir = MIRROR ? mirror_origin(w, is_end, cidx, zs, g.plane, (long long)gy * g.pitch + gx) : nullptr;     // peer-store exchange
    const int mz0 = cidx == 0 ? w.mirror_z0[0] : w.mirror_z0[1], mz1 = cidx == 0 ? w.mirror_z1[0] : w.mirror_z1[1];
    int zout = zs;
    auto mir_now = [&]() -> double* { return (MIRROR && mir && zout >= mz0 && zout < mz1) ? mir : nullptr; };
    const int soff = (1 + ry * RY) * ROW + 2 + 2 * lx;
    constexpr int MP = CX / 2 + 2;                       // split mid plane: words per row of one parity array (a halo word each side)
    constexpr int MHALF = (CY + 2) * MP;                 // one parity array; both fit one mid buffer
    static_assert(2 * MHALF * 8 <= Cfg::STAGE_STRIDE, "split mid plane must fit the mid buffer");
    const int moff = (1 + ry * RY) * MP + 1 + lx;
    // Warp-uniform shortcuts.  Level 1: no lane of the warp holds a Dirichlet cell.  Level 2 stores: every lane stores
    // whole pairs or nothing (true for all warps but those that hold the domain's first / last interior column), so a
    // row is one predicated 128-bit store and the warp never splits over the store code.
    const bool warp_all1 = __all_sync(0xffffffffu, pl.all1);
    const bool warp_pairs = __all_sync(0xffffffffu, pl.st.x0 == pl.st.x1);
    const uint32_t pair_rows = pl.st.x0 ? pl.st.rows : 0u;
    double rmax = 0.0;      // residual of the second update (time n+1 -> n+2), over the cells this CTA stores
    double rmax1 = 0.0;     // residual of the first update (time n -> n+1), over every level-1 cell this CTA forms: cells
                            // shared with neighbouring CTAs are counted more than once, which a maximum does not mind

    double2 C0[3][RY];      // level 0: centres of the last three staged planes
    double2 C1[3][RY];      // level 1: centres of the last three time n+1 planes
#pragma unroll
    for (int r = 0; r < RY; ++r) C0[0][r] = C0[1][r] = C0[2][r] = C1[0][r] = C1[1][r] = C1[2][r] = make_double2(0.0, 0.0);

    // End of a plane.  Default: one CTA barrier (named, id 1, so that warps with nothing to compute can take part in it from
    // a loop of their own): afterwards plane q-1's stage has been read by everyone and this iteration's mid plane is complete.
    // PSYNC: nobody but a warp's two neighbours in y reads its rows of the mid plane (its x neighbours are its own lanes), so
    // each warp meets only them (named barriers 2.., one per pair, 64 threads each; lower neighbour first, which cannot
    // deadlock: every barrier is the first one for its upper warp), and the stage is handed back through an `empty` mbarrier
    // with one arrival per warp, which the refilling lane waits on.  Warps then drift apart by at most one plane per neighbour.
    auto end_of_plane = [&](int q, uint32_t sprev) {
        if (PSYNC) {
            __syncwarp();
            if (q >= 1 && lane == 0) mbar_arrive(&empty[sprev]);
            if (warp > 0) pair_barrier(warp - 1);
            if (warp < NWARP - 1) pair_barrier(warp);
        } else {
            asm volatile("bar.sync 1, %0;" ::"n"(NTHR) : "memory");
        }
    };
    // after that: plane q-1's stage is free, refill it with plane q-1+STAGES.  The duty rotates over the warps; `warp` is a
    // warp-uniform value and the issuing lane is elected, so the seven warps whose turn it is not pay one uniform
    // compare-and-branch.
    auto refill = [&](int q, uint32_t sprev) {
        if (warp == (q & (NWARP - 1)) && q >= 1 && q - 1 + STAGES < niter) {
            if (elect_one_sync()) {
                if (PSYNC) mbar_wait(&empty[sprev], (uint32_t)(((q - 1) / STAGES) & 1));
                mbar_arrive_expect_tx(&full[sprev], Cfg::STAGE_BYTES);
                tma_load_3d(smem + sprev * Cfg::STAGE_STRIDE, &tmap, &full[sprev], bx, by, zs - 2 + q - 1 + STAGES);
            }
        }
    };

    auto run = [&] {
        uint32_t s = 0, ph = 0;
        int par = 0;
            // One input plane.  q = index of the plane inside the chunk's input run: it sits in stage q % STAGES, plane q-1 is
            // still in the previous stage.  The level-1 plane formed here is local plane zs + q - 3, the level-2 plane zs + q - 4.
            auto body = [&](int q, double2 (&c0n)[RY], const double2 (&c0p1)[RY], const double2 (&c0p2)[RY], double2 (&c1n)[RY],
                            const double2 (&c1p1)[RY], const double2 (&c1p2)[RY]) {
                const uint32_t sprev = s == 0 ? STAGES - 1 : s - 1;
                const double* S = (const double*)(smem + s * Cfg::STAGE_STRIDE) + soff;
                const double* Sp = (const double*)(smem + sprev * Cfg::STAGE_STRIDE) + soff;
                mbar_wait(&full[s], ph);
    #pragma unroll
                for (int r = 0; r < RY; ++r) c0n[r] = *(const double2*)(S + r * ROW);
                double xl[RY], xr[RY];
                double2 up, dn;
                if (q >= 2) {   // level 1: plane q-1 of the run = local plane zs + q - 3
                    load_neighbours<RY, XSW1>(Sp, ROW, swp, xl, xr, up, dn);
                    level_full<RY, ARITH>(c0p1, xl, xr, up, dn, c0p2, c0n, c1n, g.Dx, g.Dy, g.Dz);
                    const int z1 = zs + q - 3;
                    const bool zint = z1 >= z1lo && z1 < z1hi;
                    if (!(zint && warp_all1)) {         // Dirichlet cells / planes keep their value
    #pragma unroll
                        for (int r = 0; r < RY; ++r) {
                            const bool row = zint && ((pl.rows1 >> r) & 1u);
                            if (!(row && pl.i1x0)) c1n[r].x = c0p1[r].x;
                            if (!(row && pl.i1x1)) c1n[r].y = c0p1[r].y;
                        }
                    }
                    if (RES) {
    #pragma unroll
                        for (int r = 0; r < RY; ++r) {
                            res_track(rmax1, c1n[r].x, c0p1[r].x);
                            res_track(rmax1, c1n[r].y, c0p1[r].y);
                        }
                    }
                    if (MIDP) {
                        double* ME = (double*)(mid + par * Cfg::STAGE_STRIDE) + moff;
                        double* MO = ME + MHALF;
    #pragma unroll
                        for (int r = 0; r < RY; ++r) {
                            ME[r * MP] = c1n[r].x;
                            MO[r * MP] = c1n[r].y;
                        }
                    } else {
                        double* M = (double*)(mid + par * Cfg::STAGE_STRIDE) + soff;
    #pragma unroll
                        for (int r = 0; r < RY; ++r) *(double2*)(M + r * ROW) = c1n[r];
                    }
                }
                if (q >= 4) {   // level 2: local plane zs + q - 4, from the mid plane published one iteration ago
                    if (MIDP) {
                        const double* ME = (const double*)(mid + (par ^ 1) * Cfg::STAGE_STRIDE) + moff;
                        load_neighbours_split<RY>(ME, ME + MHALF, MP, xl, xr, up, dn);
                    } else {
                        const double* Mp = (const double*)(mid + (par ^ 1) * Cfg::STAGE_STRIDE) + soff;
                        load_neighbours<RY, XSW1>(Mp, ROW, swp, xl, xr, up, dn);
                    }
                    double2 o[RY];
                    level_full<RY, ARITH>(c1p1, xl, xr, up, dn, c1p2, c1n, o, g.Dx, g.Dy, g.Dz);
                    double* const mnow = mir_now();
                    if (warp_pairs && !(MIRROR && mnow)) {      // the common case: one predicated 128-bit store per row
    #pragma unroll
                        for (int r = 0; r < RY; ++r)
                            if ((pair_rows >> r) & 1u) *(double2*)(out + (long long)r * g.pitch) = o[r];
                    } else {
    #pragma unroll
                        for (int r = 0; r < RY; ++r)
                            store_pair(out + (long long)r * g.pitch, mnow ? mnow + (long long)r * g.pitch : nullptr, o[r], pl.st, r);
                    }
                    if (RES) {      // residual of the second update over the cells this thread stores (same warp-uniform split)
                        if (warp_pairs) {
    #pragma unroll
                            for (int r = 0; r < RY; ++r)
                                if ((pair_rows >> r) & 1u) {
                                    res_track(rmax, o[r].x, c1p1[r].x);
                                    res_track(rmax, o[r].y, c1p1[r].y);
                                }
                        } else {
    #pragma unroll
                            for (int r = 0; r < RY; ++r) res_pair(rmax, o[r], c1p1[r], pl.st, r);
                        }
                    }
                    out += g.plane;
                    if (MIRROR) { if (mir) mir += g.plane; ++zout; }
                }
                end_of_plane(q, sprev);
                refill(q, sprev);
                if (++s == STAGES) { s = 0; ph ^= 1u; }
                par ^= 1;
            };


        // Register rotation: level 0 newest = C0[q % 3]; level 1 newest (plane q-1 of the run) = C1[(q + 2) % 3].
        body(0, C0[0], C0[2], C0[1], C1[2], C1[1], C1[0]);
        body(1, C0[1], C0[0], C0[2], C1[0], C1[2], C1[1]);
        body(2, C0[2], C0[1], C0[0], C1[1], C1[0], C1[2]);
        int q = 3;
        for (; q + 2 < niter; q += 3) {
            body(q, C0[0], C0[2], C0[1], C1[2], C1[1], C1[0]);
            body(q + 1, C0[1], C0[0], C0[2], C1[0], C1[2], C1[1]);
            body(q + 2, C0[2], C0[1], C0[0], C1[1], C1[0], C1[2]);
        }
        if (q < niter) {
            body(q, C0[0], C0[2], C0[1], C1[2], C1[1], C1[0]);
            if (q + 1 < niter) body(q + 1, C0[1], C0[0], C0[2], C1[0], C1[2], C1[1]);
        }
    };

    if (!active) {      // nothing to compute: keep the barriers and take the refill turns
        uint32_t s = 0;
        for (int q = 0; q < niter; ++q) {
            end_of_plane(q, s == 0 ? STAGES - 1 : s - 1);
            refill(q, s == 0 ? STAGES - 1 : s - 1);
            if (++s == STAGES) s = 0;
        }
    } else {
        run();
    }
    if (RES) {
        __syncthreads();
        cta_residual_commit(rmax, res_slot, (unsigned long long*)smem);
        if (res_slot_first) {
            __syncthreads();
            cta_residual_commit(rmax1, res_slot_first, (unsigned long long*)smem);
        }
    }
    if (is_end && (w.done || (MIRROR && (w.side_done[0] || w.side_done[1])))) signal_end_chunk(w, cidx);
}

}  // namespace h3
