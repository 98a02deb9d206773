// Temporal blocking, depth 2: two FTCS updates per trip through HBM.
//
// Same z-streaming skeleton as stencil_tma_kernel (kernels.cuh), with a second stencil level stacked on the
// first inside the CTA:
//   level 0  planes of time n arrive by TMA, box (CX+4) x (CY+2), exactly as in the depth-1 kernel;
//   level 1  every thread finishes its 2 x RY patch of the time n+1 plane over the whole CX x CY compute region
//            (Dirichlet cells and Dirichlet planes are carried over unchanged, as the reference never updates
//            them), keeps the centre values in registers and publishes the plane to a double-buffered shared
//            "mid" plane;
//   level 2  after one CTA barrier each thread reads the xy neighbours of its level-1 cells from the mid plane
//            and finishes the time n+2 plane over the inner (CX-4) x (CY-2) region, which is what goes to HBM.
// Each level uses the split partial_xy / finish_z evaluation, so every cell sees exactly the operations, in
// exactly the order, of two separate depth-1 sweeps: results are bit-identical to stepping twice.
// HBM traffic is one read + one write per cell per TWO updates (8 B per cell-update); the price is the
// redundant level-1 ring (CX*CY / ((CX-4)*(CY-2)) of the level-1 work) and one barrier per plane.  The barrier
// also tells the refill thread that the stage just read is free, so this kernel needs no empty barriers.
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
    static constexpr int SMEM_BYTES = (STAGES + 2) * STAGE_STRIDE + STAGES * 8;   // stages | 2 mid planes | full barriers
};

// Per-thread constants of the two levels.
struct T2Plan {
    bool all1;            // every cell of the patch is a domain-interior cell (level 1 needs no carry-over select)
    bool i1x0, i1x1;      // column interior?
    uint32_t rows1;       // bit r: row interior?
    StorePlan st;         // level-2 stores: inside the CTA's output tile AND domain-interior
};

// finish plane p-1 of one level (-> o) and start plane p (t, tw) from its centre values cn and xy neighbours
template <int RY, int ARITH, bool FINISH, bool PARTIAL>
__device__ __forceinline__ void level_step(const double2 (&cn)[RY], const double (&xl)[RY], const double (&xr)[RY], const double2& up,
                                           const double2& dn, const double2 (&cp2)[RY], double2 (&t)[RY], double2 (&tw)[RY],
                                           double2 (&o)[RY], double Dx, double Dy, double Dz)
{
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        if (FINISH) {
            o[r].x = finish_z2<ARITH>(t[r].x, tw[r].x, cp2[r].x, cn[r].x, Dz);
            o[r].y = finish_z2<ARITH>(t[r].y, tw[r].y, cp2[r].y, cn[r].y, Dz);
        }
        if (PARTIAL) {
            const double2 ym = r == 0 ? up : cn[r - 1];
            const double2 yp = r == RY - 1 ? dn : cn[r + 1];
            tw[r].x = twice(cn[r].x);
            tw[r].y = twice(cn[r].y);
            t[r].x = partial_xy2<ARITH>(cn[r].x, tw[r].x, xl[r], cn[r].y, ym.x, yp.x, Dx, Dy);
            t[r].y = partial_xy2<ARITH>(cn[r].y, tw[r].y, cn[r].x, xr[r], ym.y, yp.y, Dx, Dy);
        }
    }
}

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

// xy neighbours of a thread's 2 x RY patch, read from the staged plane.  SHFL = take the x neighbours from the
// adjacent lanes' registers by warp shuffle instead (edge lanes still read the halo column): the kernel is bound
// by shared-memory wavefronts (ncu: l1tex 80 %), and a strided 8-byte load costs 4 wavefronts per warp, but the four
// 32-bit shuffles per row it takes to replace them turned out slower on B200 (profiles/r01_sweep_k2_shfl.txt).
template <int RY, bool SHFL>
__device__ __forceinline__ void load_neighbours(const double* __restrict__ S, int row_pitch, const double2 (&cn)[RY], int lane,
                                                double (&xl)[RY], double (&xr)[RY], double2& up, double2& dn)
{
    up = *(const double2*)(S - row_pitch);
    dn = *(const double2*)(S + RY * row_pitch);
#pragma unroll
    for (int r = 0; r < RY; ++r) {
        if (SHFL) {
            xl[r] = __shfl_up_sync(0xffffffffu, cn[r].y, 1);
            xr[r] = __shfl_down_sync(0xffffffffu, cn[r].x, 1);
            if (lane == 0) xl[r] = S[r * row_pitch - 1];
            if (lane == 31) xr[r] = S[r * row_pitch + 2];
        } else {
            xl[r] = S[r * row_pitch - 1];
            xr[r] = S[r * row_pitch + 2];
        }
    }
}

// SCHED 0: each level keeps partial_xy of its newest plane in registers and finishes it when the next plane shows up.
// SCHED 1: nothing but centre values lives across iterations: a staged plane stays resident one iteration longer and
//          the whole update of plane p-1 is formed when plane p has arrived (level 1), likewise for level 2 from the
//          previous mid plane.  Same arithmetic, ~32 fewer registers per thread.
// SCHED 2: schedule 1 with the mid plane stored split by x parity (cheaper x-neighbour loads at level 2).
template <int CX, int CY, int RY, int STAGES, int ARITH, bool RES, int MINB, int SCHED, bool MIRROR>
__global__ void __launch_bounds__((CX / 2) * (CY / RY), MINB)
    stencil_tma_t2_kernel(const __grid_constant__ CUtensorMap tmap, double* __restrict__ dst, Geom g, TmaWork w, int z1lo, int z1hi,
                          unsigned long long* res_slot, unsigned long long* res_slot_first)
{
    using Cfg = Tma2Cfg<CX, CY, STAGES>;
    constexpr int NTHR = (CX / 2) * (CY / RY);
    constexpr int NWARP = NTHR / 32;
    constexpr int ROW = Cfg::ROW;
#ifdef HEAT3D_T2_SHFL_X
    constexpr bool SHFL_X = true;      // measured on B200: 433 GLUPS vs 536 with shared-memory loads - left off
#else
    constexpr bool SHFL_X = false;
#endif
    static_assert(CX % 64 == 0 && CY % RY == 0 && NTHR % 32 == 0 && NTHR <= 1024 && STAGES >= 2 + (SCHED ? 1 : 0), "tile shape");

    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* mid = smem + STAGES * Cfg::STAGE_STRIDE;
    uint64_t* full = (uint64_t*)(smem + (STAGES + 2) * Cfg::STAGE_STRIDE);

    const int tid = threadIdx.x;
    const int ntiles = w.ntx * w.nty;
    const int tile = blockIdx.x % ntiles;
    int zs, cnt;                                                // first output (time n+2) plane / number of output planes
    const int cidx = blockIdx.x / ntiles;
    const bool is_end = chunk_of_cta(w, cidx, zs, cnt);
    if (cnt <= 0) return;
    const int niter = cnt + 4;                                  // input planes zs-2 .. zs+cnt+1
    const int ox = (tile % w.ntx) * Cfg::OX, oy = (tile / w.ntx) * Cfg::OY;   // output tile origin
    const int bx = ox - 4, by = oy - 2;                         // TMA box origin

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
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
    const int warp = tid >> 5, lane = tid & 31;
    const int gx = ox - 2 + 2 * lx;                             // global x of this thread's first cell
    const int gy = oy - 1 + ry * RY;                            // global y of this thread's first row
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
    // SCHED 2 stores the mid plane split by x parity: a thread's left / right x neighbours are then 8-byte words at
    // consecutive lanes (2 wavefronts per warp load) instead of the second halves of 16-byte pairs (4 wavefronts).
    constexpr int MP = CX / 2 + 2;                       // words per row of one parity array (one halo pair each side)
    constexpr int MHALF = (CY + 2) * MP;                 // one parity array; two of them fit one STAGE_STRIDE
    static_assert(2 * MHALF * 8 <= Cfg::STAGE_STRIDE, "split mid plane must fit the mid buffer");
    const int moff = (1 + ry * RY) * MP + 1 + lx;
    double rmax = 0.0;      // residual of the second update (time n+1 -> n+2), over the cells this CTA stores
    double rmax1 = 0.0;     // residual of the first update (time n -> n+1), over every level-1 cell this CTA forms: cells
                            // shared with neighbouring CTAs are counted more than once, which a maximum does not mind

    double2 C0[3][RY], t0[RY], tw0[RY];      // level 0: centres of the last three planes, partial / 2*centre of the newest
    double2 C1[3][RY], t1[RY], tw1[RY];      // level 1: same
#pragma unroll
    for (int r = 0; r < RY; ++r) t0[r] = tw0[r] = t1[r] = tw1[r] = C1[0][r] = C1[1][r] = C1[2][r] = make_double2(0.0, 0.0);

    uint32_t s = 0, ph = 0;
    int par = 0;

    // One input plane.  q = index of the plane inside the chunk's input run; the level-1 plane finished here is
    // local plane zs + q - 3, the level-2 plane finished here is zs + q - 4.
    auto body = [&](int q, double2 (&c0n)[RY], const double2 (&c0p1)[RY], const double2 (&c0p2)[RY], double2 (&c1n)[RY],
                    const double2 (&c1p1)[RY], const double2 (&c1p2)[RY]) {
        const double* S = (const double*)(smem + s * Cfg::STAGE_STRIDE) + soff;
        mbar_wait(&full[s], ph);
        double xl[RY], xr[RY];
        double2 up, dn;
#pragma unroll
        for (int r = 0; r < RY; ++r) c0n[r] = *(const double2*)(S + r * ROW);
        load_neighbours<RY, SHFL_X>(S, ROW, c0n, lane, xl, xr, up, dn);

        // ---- level 1: finish plane q-1 of the input run, start plane q
        if (q >= 2) level_step<RY, ARITH, true, true>(c0n, xl, xr, up, dn, c0p2, t0, tw0, c1n, g.Dx, g.Dy, g.Dz);
        else if (q == 1) level_step<RY, ARITH, false, true>(c0n, xl, xr, up, dn, c0p2, t0, tw0, c1n, g.Dx, g.Dy, g.Dz);
        if (q >= 2) {
            const int z1 = zs + q - 3;
            const bool zint = z1 >= z1lo && z1 < z1hi;
            if (!(zint && pl.all1)) {           // Dirichlet cells / planes keep their value
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
            double* M = (double*)(mid + par * Cfg::STAGE_STRIDE) + soff;
#pragma unroll
            for (int r = 0; r < RY; ++r) *(double2*)(M + r * ROW) = c1n[r];
        }
        __syncthreads();        // the stage has been read by everyone; the mid plane is complete
        if (warp == (q & (NWARP - 1)) && lane == 0 && q + STAGES < niter) {
            mbar_arrive_expect_tx(&full[s], Cfg::STAGE_BYTES);
            tma_load_3d(smem + s * Cfg::STAGE_STRIDE, &tmap, &full[s], bx, by, zs - 2 + q + STAGES);
        }
        if (++s == STAGES) { s = 0; ph ^= 1u; }

        // ---- level 2: finish plane (zs+q-4), start plane (zs+q-3)
        if (q >= 3) {
            const double* M = (const double*)(mid + par * Cfg::STAGE_STRIDE) + soff;
            load_neighbours<RY, SHFL_X>(M, ROW, c1n, lane, xl, xr, up, dn);
            double2 o[RY];
            if (q >= 4) level_step<RY, ARITH, true, true>(c1n, xl, xr, up, dn, c1p2, t1, tw1, o, g.Dx, g.Dy, g.Dz);
            else level_step<RY, ARITH, false, true>(c1n, xl, xr, up, dn, c1p2, t1, tw1, o, g.Dx, g.Dy, g.Dz);
            if (q >= 4) {
                double* const mnow = mir_now();
#pragma unroll
                for (int r = 0; r < RY; ++r) {
                    store_pair(out + (long long)r * g.pitch, mnow ? mnow + (long long)r * g.pitch : nullptr, o[r], pl.st, r);
                    if (RES) res_pair(rmax, o[r], c1p1[r], pl.st, r);
                }
                out += g.plane;
                if (MIRROR) { if (mir) mir += g.plane; ++zout; }
            }
        }
        par ^= 1;
    };

    // SCHED 1 body.  Input plane q sits in stage q % STAGES, plane q-1 is still in the previous stage.
    auto body1 = [&](int q, double2 (&c0n)[RY], const double2 (&c0p1)[RY], const double2 (&c0p2)[RY], double2 (&c1n)[RY],
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
            load_neighbours<RY, false>(Sp, ROW, c0p1, lane, xl, xr, up, dn);
            level_full<RY, ARITH>(c0p1, xl, xr, up, dn, c0p2, c0n, c1n, g.Dx, g.Dy, g.Dz);
            const int z1 = zs + q - 3;
            const bool zint = z1 >= z1lo && z1 < z1hi;
            if (!(zint && pl.all1)) {           // Dirichlet cells / planes keep their value
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
            if (SCHED == 2) {   // even-x cells and odd-x cells in separate arrays (see MP below)
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
            if (SCHED == 2) {
                const double* ME = (const double*)(mid + (par ^ 1) * Cfg::STAGE_STRIDE) + moff;
                const double* MO = ME + MHALF;
                up = make_double2(ME[-MP], MO[-MP]);
                dn = make_double2(ME[RY * MP], MO[RY * MP]);
#pragma unroll
                for (int r = 0; r < RY; ++r) {
                    xl[r] = MO[r * MP - 1];      // the odd cell of the pair to the left: consecutive lanes, consecutive words
                    xr[r] = ME[r * MP + 1];      // the even cell of the pair to the right
                }
            } else {
                const double* Mp = (const double*)(mid + (par ^ 1) * Cfg::STAGE_STRIDE) + soff;
                load_neighbours<RY, false>(Mp, ROW, c1p1, lane, xl, xr, up, dn);
            }
            double2 o[RY];
            level_full<RY, ARITH>(c1p1, xl, xr, up, dn, c1p2, c1n, o, g.Dx, g.Dy, g.Dz);
            double* const mnow = mir_now();
#pragma unroll
            for (int r = 0; r < RY; ++r) {
                store_pair(out + (long long)r * g.pitch, mnow ? mnow + (long long)r * g.pitch : nullptr, o[r], pl.st, r);
                if (RES) res_pair(rmax, o[r], c1p1[r], pl.st, r);
            }
            out += g.plane;
            if (MIRROR) { if (mir) mir += g.plane; ++zout; }
        }
        __syncthreads();        // plane q-1's stage has been read by everyone; this iteration's mid plane is complete
        if (q >= 1 && warp == (q & (NWARP - 1)) && lane == 0 && q - 1 + STAGES < niter) {
            mbar_arrive_expect_tx(&full[sprev], Cfg::STAGE_BYTES);
            tma_load_3d(smem + sprev * Cfg::STAGE_STRIDE, &tmap, &full[sprev], bx, by, zs - 2 + q - 1 + STAGES);
        }
        if (++s == STAGES) { s = 0; ph ^= 1u; }
        par ^= 1;
    };

    // Register rotation: level 0 newest = C0[q % 3]; level 1 newest (plane q-1 of the run) = C1[(q + 2) % 3].
#define H3_T2_BODY(qq, a, b, c, d, e, f)                                  \
    do {                                                                  \
        if (SCHED >= 1) body1(qq, a, b, c, d, e, f);                      \
        else body(qq, a, b, c, d, e, f);                                  \
    } while (0)
    H3_T2_BODY(0, C0[0], C0[2], C0[1], C1[2], C1[1], C1[0]);
    H3_T2_BODY(1, C0[1], C0[0], C0[2], C1[0], C1[2], C1[1]);
    H3_T2_BODY(2, C0[2], C0[1], C0[0], C1[1], C1[0], C1[2]);
    int q = 3;
    for (; q + 2 < niter; q += 3) {
        H3_T2_BODY(q, C0[0], C0[2], C0[1], C1[2], C1[1], C1[0]);
        H3_T2_BODY(q + 1, C0[1], C0[0], C0[2], C1[0], C1[2], C1[1]);
        H3_T2_BODY(q + 2, C0[2], C0[1], C0[0], C1[1], C1[0], C1[2]);
    }
    if (q < niter) {
        H3_T2_BODY(q, C0[0], C0[2], C0[1], C1[2], C1[1], C1[0]);
        if (q + 1 < niter) H3_T2_BODY(q + 1, C0[1], C0[0], C0[2], C1[0], C1[2], C1[1]);
    }
#undef H3_T2_BODY
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
