// heat3D — drop-in driver for the reference executable of the same name
// (fredrickhang/Cuda-aware-MPI-on-3D-heate-quation, HeatEquation3D/src/heat3D.cu; "cu:" below).
//
// Same command line (cu:277-291):   heat3D NUM_CELLS_X NUM_CELLS_Y NUM_CELLS_Z ITER_MAX EPS
// Same stdout report (cu:294-302, cu:1083-1089, cu:1106) and the same Tecplot-ASCII dump in
// ./output/out.dat (cu:1163-1179, single-zone form: one process drives every GPU, so the output equals
// the single-domain result whatever the GPU count).  All field work goes through libheat3d's C ABI
// (include/heat3d.h); this file is host C++ only and has no compute path of its own.
//
// Additive knobs, all via environment so argv stays the reference's; defaults = reference behaviour:
//   HEAT3D_NDEV=P          z-slabs on GPUs 0..P-1 (default 1; the reference's `mpirun -n P`)
//   HEAT3D_ARITH=strict|fma   arithmetic order of the update (default strict = the host program's)
//   HEAT3D_CHECK_EVERY=N   examine the residual every N iterations (default 1 = reference semantics)
//   HEAT3D_OUT_EVERY=M     also write output/out_<iteration>.dat every M iterations (async snapshots)
//   HEAT3D_TEMPORAL_K=1|2  updates fused per HBM pass (bit-identical results either way).  Default: 2 for grids of
//                          2^25 points and more, where a pass is long against a host round trip, else 1
//   HEAT3D_KERNEL=auto|tma|baseline
//   HEAT3D_PERF=1          print GLUPS / algorithmic GB/s to stderr
#include "../../include/heat3d.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

namespace {

[[noreturn]] void die(const char* what)
{
    std::cerr << "heat3D: " << what << ": " << heat3d_last_error() << std::endl;
    std::exit(2);
}
#define H3(call) do { if ((call) != HEAT3D_OK) die(#call); } while (0)

unsigned env_unsigned(const char* name, unsigned dflt)
{
    const char* v = std::getenv(name);
    return (v && *v) ? (unsigned)std::strtoul(v, nullptr, 10) : dflt;
}

// Byte-for-byte what `out << std::scientific << std::setprecision(5) << std::setw(15) << v` produces.
inline void fmt15(char* dst, double v)
{
    char tmp[32];
    std::snprintf(tmp, sizeof tmp, "%15.5e", v);
    std::memcpy(dst, tmp, 15);
}

// Tecplot POINT dump (cu:1163-1179).  Coordinates index*spacing are formatted once per axis; planes are
// formatted in parallel and written in order.
bool write_dump(const std::string& path, const double* T, unsigned nx, unsigned ny, unsigned nz, const double sp[3])
{
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;    // the reference's ofstream fails silently when ./output is missing; so do we
    std::fprintf(f, "TITLE=\"out\"\nVARIABLES = \"X\", \"Y\", \"Z\", \"T\"\nZONE T = \"0\", I=%u, J=%u, K=%u, F=POINT\n", nx, ny, nz);
    std::vector<char> xs((size_t)nx * 15), ys((size_t)ny * 15), zs((size_t)nz * 15);
    for (unsigned i = 0; i < nx; ++i) fmt15(&xs[(size_t)i * 15], i * sp[0]);
    for (unsigned j = 0; j < ny; ++j) fmt15(&ys[(size_t)j * 15], j * sp[1]);
    for (unsigned k = 0; k < nz; ++k) fmt15(&zs[(size_t)k * 15], k * sp[2]);
    const size_t line = 61, plane_bytes = (size_t)nx * ny * line;
    const unsigned nthreads = std::max(1u, std::min(std::thread::hardware_concurrency(), 16u));
    const unsigned batch = std::max(1u, std::min(nz, (unsigned)std::min<size_t>(nthreads, std::max<size_t>(1, (256u << 20) / plane_bytes))));
    std::vector<char> buf((size_t)batch * plane_bytes);
    bool ok = true;
    for (unsigned k0 = 0; k0 < nz && ok; k0 += batch) {
        const unsigned nb = std::min(batch, nz - k0);
        std::vector<std::thread> pool;
        for (unsigned b = 0; b < nb; ++b)
            pool.emplace_back([&, b] {
                const unsigned k = k0 + b;
                char* o = &buf[(size_t)b * plane_bytes];
                const double* t = T + (size_t)k * nx * ny;
                for (unsigned j = 0; j < ny; ++j)
                    for (unsigned i = 0; i < nx; ++i, o += line) {
                        std::memcpy(o, &xs[(size_t)i * 15], 15);
                        std::memcpy(o + 15, &ys[(size_t)j * 15], 15);
                        std::memcpy(o + 30, &zs[(size_t)k * 15], 15);
                        fmt15(o + 45, t[(size_t)j * nx + i]);
                        o[60] = '\n';
                    }
            });
        for (auto& th : pool) th.join();
        ok = std::fwrite(buf.data(), 1, (size_t)nb * plane_bytes, f) == (size_t)nb * plane_bytes;
    }
    return std::fclose(f) == 0 && ok;
}

}  // namespace

int main(int argc, char** argv)
{
    // ---- argument check and echo, cu:284-305 --------------------------------------------------------
    if (argc != 6) {
        std::cout << "Incorrect number of command line arguments specified, use the following syntax:\n" << std::endl;
        std::cout << "bin/HeatEquation3D NUM_CELLS_X NUM_CELLS_Y NUM_CELLS_Z ITER_MAX EPS" << std::endl;
        std::cout << "\nor, using MPI, use the following syntax:\n" << std::endl;
        std::cout << "mpirun -n NUM_PROCS bin/HeatEquation3D NUM_CELLS_X NUM_CELLS_Y NUM_CELLS_Z ITER_MAX EPS" << std::endl;
        std::cout << "\nSee source code for additional informations!" << std::endl;
        std::abort();
    }
    std::cout << "Runnung HeatEquation3D with the following arguments: " << std::endl;
    std::cout << "executable:               " << argv[0] << std::endl;
    std::cout << "number of cells in x:     " << std::stoi(argv[1]) << std::endl;
    std::cout << "number of cells in y:     " << std::stoi(argv[2]) << std::endl;
    std::cout << "number of cells in z:     " << std::stoi(argv[3]) << std::endl;
    std::cout << "max number of iterations: " << std::stoi(argv[4]) << std::endl;
    std::cout << "convergence threshold:    " << std::stod(argv[5]) << "\n" << std::endl;

    const unsigned iterMax = std::stoi(argv[4]);        // cu:309
    const double eps = std::stod(argv[5]);              // cu:315
    unsigned numCells[3] = {(unsigned)std::stoi(argv[1]), (unsigned)std::stoi(argv[2]), (unsigned)std::stoi(argv[3])};

    // ---- derived constants, same expressions as cu:338-367 -------------------------------------------
    const double domainLength[3] = {1.0, 1.0, 1.0};
    const double alpha = 1.0;
    const double CFL = 0.4;
    double spacing[3];
    for (int d = 0; d < 3; ++d) spacing[d] = domainLength[d] / static_cast<double>(numCells[d] - 1.0);
    const double dt = CFL * 1.0 / (3 * 2) * std::pow(std::min({spacing[0], spacing[1], spacing[2]}), 2.0) / alpha;
    const double Dx = dt * alpha / (std::pow(spacing[0], 2.0));
    const double Dy = dt * alpha / (std::pow(spacing[1], 2.0));
    const double Dz = dt * alpha / (std::pow(spacing[2], 2.0));

    // ---- environment knobs ------------------------------------------------------------------------------
    const unsigned ndev = std::max(1u, env_unsigned("HEAT3D_NDEV", 1));
    const unsigned checkEvery = std::max(1u, env_unsigned("HEAT3D_CHECK_EVERY", 1));
    const unsigned outEvery = env_unsigned("HEAT3D_OUT_EVERY", 0);
    const bool perf = env_unsigned("HEAT3D_PERF", 0) != 0;
    const char* arithEnv = std::getenv("HEAT3D_ARITH");
    const char* kernEnv = std::getenv("HEAT3D_KERNEL");

    heat3d_params p;
    heat3d_default_params(&p, numCells[0], numCells[1], numCells[2]);
    p.Dx = Dx; p.Dy = Dy; p.Dz = Dz;
    p.arith = (arithEnv && std::string(arithEnv) == "fma") ? HEAT3D_ARITH_FMA : HEAT3D_ARITH_STRICT;
    if (kernEnv && std::string(kernEnv) == "baseline") p.kernel = HEAT3D_KERNEL_BASELINE;
    if (kernEnv && std::string(kernEnv) == "tma") p.kernel = HEAT3D_KERNEL_TMA;
    const size_t points = (size_t)numCells[0] * numCells[1] * numCells[2];
    p.temporal_k = (int32_t)std::max(1u, env_unsigned("HEAT3D_TEMPORAL_K", points >= ((size_t)1 << 25) ? 2u : 1u));
    if (p.kernel == HEAT3D_KERNEL_BASELINE) p.temporal_k = 1;
    std::vector<int32_t> devices(ndev);
    for (unsigned i = 0; i < ndev; ++i) devices[i] = (int32_t)i;
    p.nslabs = (int32_t)ndev; p.slab_first = 0; p.nlocal = (int32_t)ndev;
    p.devices = devices.data();

    heat3d_ctx* ctx = nullptr;
    H3(heat3d_create(&p, &ctx));
    H3(heat3d_init_reference_field(ctx, spacing[1]));     // cu:409-453

    const size_t cells = (size_t)numCells[0] * numCells[1] * numCells[2];
    double* snap[2] = {nullptr, nullptr};
    std::thread writer;
    if (outEvery)
        for (int b = 0; b < 2; ++b) H3(heat3d_host_alloc((void**)&snap[b], cells * sizeof(double)));

    // ---- time loop, cu:541-1074 --------------------------------------------------------------------------
    double norm = 1.0;                       // cu:325
    int globalBreakCondition = false;        // cu:329
    unsigned finalNumIterations = 0;         // cu:370
    unsigned applied = 0, nsnap = 0;
    const auto start = std::chrono::steady_clock::now();      // cu:527
    unsigned time = 0;
    if (!outEvery) {
        // the whole loop inside the library: same norm / break logic (cu:1030-1073), and with checkEvery == 1 the next
        // iteration is already queued while the host looks at a residual, so the GPU never waits for the host
        int32_t conv = 0;
        H3(heat3d_run(ctx, iterMax, eps, checkEvery, &finalNumIterations, &conv, &applied));
        globalBreakCondition = conv;
        time = iterMax;
    }
    while (time < iterMax) {
        // next iteration whose residual is examined: 0, then every checkEvery-th, and the last one
        unsigned target = time == 0 ? 0 : ((time / checkEvery) + 1) * checkEvery - 1;
        target = std::min(std::max(target, time), iterMax - 1);
        if (outEvery) target = std::min(target, ((time / outEvery) + 1) * outEvery - 1);
        double res = 0.0;
        H3(heat3d_step_residual(ctx, target - time + 1, &res));       // cu:545-1024
        applied += target - time + 1;
        if (target == 0 && res != 0.0) norm = res;                     // cu:1030-1032
        if (res / norm < eps) {                                        // cu:1055-1073
            globalBreakCondition = true;
            finalNumIterations = target;
            break;
        }
        time = target + 1;
        if (outEvery && time % outEvery == 0 && time < iterMax) {
            double* dst = snap[nsnap & 1];
            if (writer.joinable()) writer.join();                      // previous file finished before its buffer is reused next time
            H3(heat3d_snapshot_async(ctx, dst));
            H3(heat3d_snapshot_wait(ctx));
            const std::string path = "output/out_" + std::to_string(time) + ".dat";
            writer = std::thread([=] { write_dump(path, dst, numCells[0], numCells[1], numCells[2], spacing); });
            ++nsnap;
        }
    }
    H3(heat3d_sync(ctx));
    const auto end = std::chrono::steady_clock::now();                 // cu:1081
    const double seconds = std::chrono::duration<double>(end - start).count();

    // ---- report, cu:1082-1090 ----------------------------------------------------------------------------
    std::cout << "Computational time (parallel): " << std::fixed << seconds << "\n" << std::endl;
    if (globalBreakCondition) {
        std::cout << "Simulation has converged in " << finalNumIterations << " iterations";
        std::cout << " with a convergence threshold of " << std::scientific << eps << std::endl;
    } else
        std::cout << "Simulation did not converge within " << iterMax << " iterations." << std::endl;

    double error = 0.0;
    H3(heat3d_mean_abs_error_vs_y(ctx, spacing[1], &error));           // cu:1096-1104
    std::cout << "L2-norm error: " << std::fixed << std::setprecision(4) << 100 * error << " %" << std::endl;

    if (perf) {
        const double upd = (double)(numCells[0] - 2) * (numCells[1] - 2) * (numCells[2] - 2) * applied;
        std::fprintf(stderr, "[heat3D perf] %u GPU(s), kernel %s: %u updates in %.6f s = %.2f GLUPS (%.1f GB/s algorithmic at 16 B/cell)\n",
                     ndev, heat3d_kernel_name(ctx), applied, seconds, upd / seconds / 1e9, upd * 16 / seconds / 1e9);
    }

    // ---- dump, cu:1163-1179 --------------------------------------------------------------------------------
    if (writer.joinable()) writer.join();
    {
        std::vector<double> T(cells);
        H3(heat3d_download(ctx, T.data()));
        write_dump("output/out.dat", T.data(), numCells[0], numCells[1], numCells[2], spacing);
    }
    for (int b = 0; b < 2; ++b) heat3d_host_free(snap[b]);
    heat3d_destroy(ctx);
    return 0;
}
