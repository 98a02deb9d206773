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
//   HEAT3D_DEVICES=a,b,..  explicit CUDA ordinal per slab instead (ordinals may repeat: several slabs on one GPU)
//   HEAT3D_ARITH=strict|fma   arithmetic order of the update (default strict = the host program's)
//   HEAT3D_CHECK_EVERY=N   examine the residual every N iterations (default 1 = reference semantics)
//   HEAT3D_OUT_EVERY=M     also write output/out_<iteration>.dat every M iterations.  The snapshot leaves the GPU on a
//                          copy stream while stepping goes on; its file is formatted by a writer thread and appears one
//                          interval later (single-process runs)
//   HEAT3D_OUT_FORMAT=ascii|raw   ascii (default) = the reference's Tecplot text, 61 bytes per point (cu:1165-1177);
//                          raw = one text header line + the fp64 values, x fastest, in output/out.raw / out_<it>.raw
//   HEAT3D_OUT_STRIDE=s | sx,sy,sz   write every s-th point per axis (default 1 = every point, the reference's dump);
//                          for grids whose full dump is impractical (1024^3 in text is 65 GB).  Periodic snapshots are
//                          decimated on the GPU, so only the kept points cross PCIe
//   HEAT3D_RANK / HEAT3D_WORLD (or RANK / WORLD_SIZE, OMPI_COMM_WORLD_RANK / _SIZE, PMI_RANK / PMI_SIZE)
//                          one process per z-slab, the reference's `mpirun -n P heat3D ...` launch model (cu:203-263)
//                          without MPI: rank 0 writes the 128-byte NCCL bootstrap id to HEAT3D_ID_FILE (default
//                          /tmp/heat3d_id_<MASTER_PORT>), the others wait for it; rank r uses GPU HEAT3D_LOCAL_RANK /
//                          LOCAL_RANK / r mod device count; every rank formats its own planes and rank 0 joins the parts
//                          into ONE single-zone output/out.dat (the reference's rank-0 gather, cu:1114-1162, minus its
//                          per-rank zones: the file equals the single-domain dump whatever P is)
//   HEAT3D_TEMPORAL_K=1|2  updates fused per HBM pass (bit-identical results either way).  Default: 2 for grids of
//                          2^25 points and more, where a pass is long against a host round trip, else 1
//   HEAT3D_NSLABS_Y=Q      additionally cut the rows into Q ranges (2-D decomposition, Q x P sub-domains, HEAT3D_DEVICES then
//                          lists Q*P ordinals): strided y-faces travel through pack / unpack kernels; depth 1 only
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

#include <sys/stat.h>
#include <unistd.h>

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

// What to write: every (sx,sy,sz)-th point of an nx x ny x nz grid; T holds the ALREADY decimated values
// [nzs][nys][nxs] of planes whose decimated index runs over [k0, k1).
struct DumpSpec {
    unsigned nx, ny, nz;          // full grid
    unsigned sx, sy, sz;          // strides
    double sp[3];                 // spacing of the full grid
    unsigned nxs() const { return (nx + sx - 1) / sx; }
    unsigned nys() const { return (ny + sy - 1) / sy; }
    unsigned nzs() const { return (nz + sz - 1) / sz; }
};

// Tecplot POINT dump (cu:1163-1179): header (optional) + the lines of decimated planes [k0, k1).  Coordinates
// index*spacing are formatted once per axis; planes are formatted in parallel and written in order.
bool write_ascii(const std::string& path, const double* T, const DumpSpec& d, unsigned k0, unsigned k1, bool header, bool append = false)
{
    std::FILE* f = std::fopen(path.c_str(), append ? "ab" : "wb");
    if (!f) return false;    // the reference's ofstream fails silently when ./output is missing; so do we
    const unsigned nx = d.nxs(), ny = d.nys(), nz = d.nzs();
    if (header) std::fprintf(f, "TITLE=\"out\"\nVARIABLES = \"X\", \"Y\", \"Z\", \"T\"\nZONE T = \"0\", I=%u, J=%u, K=%u, F=POINT\n", nx, ny, nz);
    std::vector<char> xs((size_t)nx * 15), ys((size_t)ny * 15), zs((size_t)nz * 15);
    for (unsigned i = 0; i < nx; ++i) fmt15(&xs[(size_t)i * 15], (i * d.sx) * d.sp[0]);
    for (unsigned j = 0; j < ny; ++j) fmt15(&ys[(size_t)j * 15], (j * d.sy) * d.sp[1]);
    for (unsigned k = 0; k < nz; ++k) fmt15(&zs[(size_t)k * 15], (k * d.sz) * d.sp[2]);
    const size_t line = 61, plane_bytes = (size_t)nx * ny * line;
    const unsigned nthreads = std::max(1u, std::min(std::thread::hardware_concurrency(), 16u));
    const unsigned nplanes = k1 - k0;
    const unsigned batch = std::max(1u, std::min(nplanes, (unsigned)std::min<size_t>(nthreads, std::max<size_t>(1, (256u << 20) / plane_bytes))));
    std::vector<char> buf((size_t)batch * plane_bytes);
    bool ok = true;
    for (unsigned b0 = 0; b0 < nplanes && ok; b0 += batch) {
        const unsigned nb = std::min(batch, nplanes - b0);
        std::vector<std::thread> pool;
        for (unsigned b = 0; b < nb; ++b)
            pool.emplace_back([&, b] {
                const unsigned k = k0 + b0 + b;
                char* o = &buf[(size_t)b * plane_bytes];
                const double* t = T + (size_t)(b0 + b) * nx * ny;
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

// Raw side format (additive; nothing in the reference): one text line, then the fp64 values of decimated planes
// [k0, k1), x fastest, native byte order.
bool write_raw(const std::string& path, const double* T, const DumpSpec& d, unsigned k0, unsigned k1, bool header, unsigned iteration,
               bool append = false)
{
    std::FILE* f = std::fopen(path.c_str(), append ? "ab" : "wb");
    if (!f) return false;
    if (header)
        std::fprintf(f, "# heat3D raw fp64, x fastest: points %u %u %u stride %u %u %u iteration %u spacing %.17g %.17g %.17g\n", d.nxs(), d.nys(),
                     d.nzs(), d.sx, d.sy, d.sz, iteration, d.sp[0], d.sp[1], d.sp[2]);
    const size_t n = (size_t)(k1 - k0) * d.nxs() * d.nys();
    const bool ok = std::fwrite(T, sizeof(double), n, f) == n;
    return std::fclose(f) == 0 && ok;
}

// Keep every (sx,sy,sz)-th point of full planes [z0, z1) (host side; used for the final dump, which is downloaded whole).
// Returns the decimated plane index range through k0/k1.
std::vector<double> decimate(const double* T, const DumpSpec& d, unsigned z0, unsigned z1, unsigned& k0, unsigned& k1)
{
    k0 = (z0 + d.sz - 1) / d.sz;
    k1 = z1 == 0 ? 0 : (z1 - 1) / d.sz + 1;
    if (k1 < k0) k1 = k0;
    const unsigned nxs = d.nxs(), nys = d.nys();
    std::vector<double> out((size_t)(k1 - k0) * nxs * nys);
    for (unsigned k = k0; k < k1; ++k)
        for (unsigned j = 0; j < nys; ++j) {
            const double* row = T + ((size_t)(k * d.sz - z0) * d.ny + (size_t)j * d.sy) * d.nx;
            double* o = &out[((size_t)(k - k0) * nys + j) * nxs];
            for (unsigned i = 0; i < nxs; ++i) o[i] = row[(size_t)i * d.sx];
        }
    return out;
}

const char* first_env(std::initializer_list<const char*> names)
{
    for (const char* n : names) {
        const char* v = std::getenv(n);
        if (v && *v) return v;
    }
    return nullptr;
}

// The 128-byte NCCL bootstrap id travels through a file: rank 0 writes <file>.tmp and renames it, the others poll.
// A stamp (seconds since the epoch) in front lets readers ignore a file some crashed earlier run left behind.
struct IdFile { char magic[8]; long long stamp; unsigned char id[128]; };

bool publish_id(const std::string& path, const unsigned char* id)
{
    ::unlink(path.c_str());
    IdFile f;
    std::memcpy(f.magic, "H3DNCCL1", 8);
    f.stamp = (long long)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::system_clock::now().time_since_epoch()).count();
    std::memcpy(f.id, id, 128);
    const std::string tmp = path + ".tmp";
    std::FILE* fp = std::fopen(tmp.c_str(), "wb");
    if (!fp) return false;
    const bool ok = std::fwrite(&f, sizeof f, 1, fp) == 1;
    return std::fclose(fp) == 0 && ok && std::rename(tmp.c_str(), path.c_str()) == 0;
}

bool await_id(const std::string& path, unsigned char* id, double timeout_s)
{
    const auto t0 = std::chrono::steady_clock::now();
    const long long started = (long long)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::system_clock::now().time_since_epoch()).count();
    while (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() < timeout_s) {
        if (std::FILE* fp = std::fopen(path.c_str(), "rb")) {
            IdFile f;
            const bool got = std::fread(&f, sizeof f, 1, fp) == 1;
            std::fclose(fp);
            if (got && std::memcmp(f.magic, "H3DNCCL1", 8) == 0 && f.stamp >= started - 120) {
                std::memcpy(id, f.id, 128);
                return true;
            }
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
    return false;
}

bool file_exists(const std::string& p)
{
    struct stat st;
    return ::stat(p.c_str(), &st) == 0;
}

}  // namespace

int main(int argc, char** argv)
{
    // ---- process layout: one process (default) or one per z-slab, the reference's `mpirun -n P` model, cu:203-263 ----
    const char* rankEnv = first_env({"HEAT3D_RANK", "RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK"});
    const char* worldEnv = first_env({"HEAT3D_WORLD", "WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE"});
    const unsigned world = worldEnv ? std::max(1ul, std::strtoul(worldEnv, nullptr, 10)) : 1u;
    const unsigned rank = (world > 1 && rankEnv) ? (unsigned)std::strtoul(rankEnv, nullptr, 10) : 0u;
    const bool root = rank == 0;                            // rank 0 does all the talking (cu:284, cu:1082, cu:1105)

    // ---- argument check and echo, cu:284-305 --------------------------------------------------------
    if (argc != 6) {
        if (root) {
            std::cout << "Incorrect number of command line arguments specified, use the following syntax:\n" << std::endl;
            std::cout << "bin/HeatEquation3D NUM_CELLS_X NUM_CELLS_Y NUM_CELLS_Z ITER_MAX EPS" << std::endl;
            std::cout << "\nor, using MPI, use the following syntax:\n" << std::endl;
            std::cout << "mpirun -n NUM_PROCS bin/HeatEquation3D NUM_CELLS_X NUM_CELLS_Y NUM_CELLS_Z ITER_MAX EPS" << std::endl;
            std::cout << "\nSee source code for additional informations!" << std::endl;
        }
        std::abort();
    }
    if (root) {
        std::cout << "Runnung HeatEquation3D with the following arguments: " << std::endl;
        std::cout << "executable:               " << argv[0] << std::endl;
        std::cout << "number of cells in x:     " << std::stoi(argv[1]) << std::endl;
        std::cout << "number of cells in y:     " << std::stoi(argv[2]) << std::endl;
        std::cout << "number of cells in z:     " << std::stoi(argv[3]) << std::endl;
        std::cout << "max number of iterations: " << std::stoi(argv[4]) << std::endl;
        std::cout << "convergence threshold:    " << std::stod(argv[5]) << "\n" << std::endl;
    }

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
    unsigned outEvery = env_unsigned("HEAT3D_OUT_EVERY", 0);
    const bool perf = env_unsigned("HEAT3D_PERF", 0) != 0;
    const char* arithEnv = std::getenv("HEAT3D_ARITH");
    const char* kernEnv = std::getenv("HEAT3D_KERNEL");
    const char* fmtEnv = std::getenv("HEAT3D_OUT_FORMAT");
    const bool raw = fmtEnv && std::string(fmtEnv) == "raw";
    DumpSpec spec{numCells[0], numCells[1], numCells[2], 1, 1, 1, {spacing[0], spacing[1], spacing[2]}};
    if (const char* st = std::getenv("HEAT3D_OUT_STRIDE")) {
        unsigned a = 0, b = 0, c = 0;
        const int n = std::sscanf(st, "%u,%u,%u", &a, &b, &c);
        if (n == 1) b = c = a;
        if (n == 1 || n == 3) { spec.sx = std::max(1u, a); spec.sy = std::max(1u, b); spec.sz = std::max(1u, c); }
    }
    const std::string ext = raw ? ".raw" : ".dat";
    if (world > 1 && outEvery) {
        if (root) std::cerr << "heat3D: HEAT3D_OUT_EVERY is not available in one-process-per-slab runs; writing the final dump only" << std::endl;
        outEvery = 0;
    }

    heat3d_params p;
    heat3d_default_params(&p, numCells[0], numCells[1], numCells[2]);
    p.Dx = Dx; p.Dy = Dy; p.Dz = Dz;
    p.arith = (arithEnv && std::string(arithEnv) == "fma") ? HEAT3D_ARITH_FMA : HEAT3D_ARITH_STRICT;
    if (kernEnv && std::string(kernEnv) == "baseline") p.kernel = HEAT3D_KERNEL_BASELINE;
    if (kernEnv && std::string(kernEnv) == "tma") p.kernel = HEAT3D_KERNEL_TMA;
    const size_t points = (size_t)numCells[0] * numCells[1] * numCells[2];
    p.temporal_k = (int32_t)std::max(1u, env_unsigned("HEAT3D_TEMPORAL_K", points >= ((size_t)1 << 25) ? 2u : 1u));
    if (p.kernel == HEAT3D_KERNEL_BASELINE) p.temporal_k = 1;
    std::vector<int32_t> devices;
    unsigned char ncclId[128];
    std::string idFile;
    if (world > 1) {
        // one slab per process; the id that ties the processes together travels through a file
        if (rank >= world) { std::cerr << "heat3D: rank " << rank << " outside world " << world << std::endl; return 2; }
        int32_t have = 0;
        H3(heat3d_device_count(&have));
        const char* lr = first_env({"HEAT3D_LOCAL_RANK", "LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK"});
        devices.push_back((int32_t)((lr ? std::strtoul(lr, nullptr, 10) : rank) % (unsigned)std::max(1, have)));   // cu:650-654
        if (const char* f = std::getenv("HEAT3D_ID_FILE")) idFile = f;
        else if (const char* port = std::getenv("MASTER_PORT")) idFile = std::string("/tmp/heat3d_id_") + port;
        else { std::cerr << "heat3D: a multi-process run needs HEAT3D_ID_FILE (a path all ranks can reach)" << std::endl; return 2; }
        if (root) {
            H3(heat3d_nccl_unique_id(ncclId));
            if (!publish_id(idFile, ncclId)) { std::cerr << "heat3D: cannot write " << idFile << std::endl; return 2; }
        } else if (!await_id(idFile, ncclId, 300.0)) {
            std::cerr << "heat3D: rank " << rank << " gave up waiting for " << idFile << std::endl;
            return 2;
        }
        p.nslabs = (int32_t)world; p.slab_first = (int32_t)rank; p.nlocal = 1;
        p.nccl_unique_id = ncclId;
    } else {
        if (const char* list = std::getenv("HEAT3D_DEVICES")) {
            for (const char* q = list; *q;) {
                char* e = nullptr;
                const long v = std::strtol(q, &e, 10);
                if (e == q) break;
                devices.push_back((int32_t)v);
                q = *e == ',' ? e + 1 : e;
            }
        }
        if (devices.empty()) {
            devices.resize(ndev);
            for (unsigned i = 0; i < ndev; ++i) devices[i] = (int32_t)i;
        }
        const unsigned qy = std::max(1u, env_unsigned("HEAT3D_NSLABS_Y", 1));
        if (qy > 1) {
            p.nslabs_y = (int32_t)qy;
            p.temporal_k = 1;
            if (devices.size() % qy != 0 || std::getenv("HEAT3D_DEVICES") == nullptr) {      // default: every sub-domain on its own GPU, wrapping around
                int32_t have = 1;
                H3(heat3d_device_count(&have));
                devices.resize((size_t)qy * ndev);
                for (size_t i = 0; i < devices.size(); ++i) devices[i] = (int32_t)(i % (size_t)std::max(1, have));
            }
        }
        const unsigned pz = (unsigned)devices.size() / qy;
        p.nslabs = (int32_t)pz; p.slab_first = 0; p.nlocal = (int32_t)pz;
    }
    p.devices = devices.data();

    // stdout belongs to the reference's report: NCCL prints its version banner there while the communicator comes up
    // (under NCCL_DEBUG=VERSION and louder), so for that one call file descriptor 1 points at stderr
    heat3d_ctx* ctx = nullptr;
    {
        int saved = -1;
        if (world > 1) {
            std::cout.flush();
            std::fflush(stdout);
            saved = ::dup(1);
            if (saved >= 0) ::dup2(2, 1);
        }
        const int rc = heat3d_create(&p, &ctx);
        if (saved >= 0) {
            std::fflush(stdout);
            ::dup2(saved, 1);
            ::close(saved);
        }
        if (rc != HEAT3D_OK) die("heat3d_create(&p, &ctx)");
    }
    if (world > 1 && root) ::unlink(idFile.c_str());        // every rank has read it: heat3d_create is collective
    H3(heat3d_init_reference_field(ctx, spacing[1]));     // cu:409-453

    const size_t snapCells = (size_t)spec.nxs() * spec.nys() * spec.nzs();
    double* snap[2] = {nullptr, nullptr};
    std::thread writer;
    if (outEvery)
        for (int b = 0; b < 2; ++b) H3(heat3d_host_alloc((void**)&snap[b], snapCells * sizeof(double)));
    // a snapshot in flight: issued at iteration `pendingIter` into snap[pendingBuf]; its file is written once the copy has
    // landed, which is checked when the next one is due (or at the end) - stepping never waits for PCIe or for the formatter
    int pendingBuf = -1;
    unsigned pendingIter = 0;
    auto flush_pending = [&] {
        if (pendingBuf < 0) return;
        H3(heat3d_snapshot_wait(ctx));
        if (writer.joinable()) writer.join();              // previous file done before its buffer can be reused next time
        const double* src = snap[pendingBuf];
        const std::string path = "output/out_" + std::to_string(pendingIter) + ext;
        const unsigned it = pendingIter;
        writer = std::thread([=] {
            if (raw) write_raw(path, src, spec, 0, spec.nzs(), true, it);
            else write_ascii(path, src, spec, 0, spec.nzs(), true);
        });
        pendingBuf = -1;
    };

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
            flush_pending();                                           // the previous snapshot has had a whole interval to land
            pendingBuf = (int)(nsnap & 1);
            pendingIter = time;
            H3(heat3d_snapshot_strided_async(ctx, snap[pendingBuf], spec.sx, spec.sy, spec.sz));
            ++nsnap;
        }
    }
    H3(heat3d_sync(ctx));
    const auto end = std::chrono::steady_clock::now();                 // cu:1081
    const double seconds = std::chrono::duration<double>(end - start).count();
    flush_pending();

    // ---- report, cu:1082-1090 ----------------------------------------------------------------------------
    if (root) {
        std::cout << "Computational time (parallel): " << std::fixed << seconds << "\n" << std::endl;
        if (globalBreakCondition) {
            std::cout << "Simulation has converged in " << finalNumIterations << " iterations";
            std::cout << " with a convergence threshold of " << std::scientific << eps << std::endl;
        } else
            std::cout << "Simulation did not converge within " << iterMax << " iterations." << std::endl;
    }

    double error = 0.0;
    H3(heat3d_mean_abs_error_vs_y(ctx, spacing[1], &error));           // cu:1096-1104 (collective in a multi-process run)
    if (root) std::cout << "L2-norm error: " << std::fixed << std::setprecision(4) << 100 * error << " %" << std::endl;

    if (perf && root) {
        const double upd = (double)(numCells[0] - 2) * (numCells[1] - 2) * (numCells[2] - 2) * applied;
        std::fprintf(stderr, "[heat3D perf] %u GPU(s), kernel %s: %u updates in %.6f s = %.2f GLUPS (%.1f GB/s algorithmic at 16 B/cell)\n",
                     world > 1 ? world : (unsigned)devices.size(), heat3d_kernel_name(ctx), applied, seconds, upd / seconds / 1e9, upd * 16 / seconds / 1e9);
    }

    // ---- dump, cu:1112-1179 --------------------------------------------------------------------------------
    if (writer.joinable()) writer.join();
    {
        uint32_t z0 = 0, z1 = 0;
        H3(heat3d_local_range(ctx, &z0, &z1));
        std::vector<double> T((size_t)(z1 - z0) * numCells[0] * numCells[1]);
        H3(heat3d_download_planes(ctx, T.data(), z0, z1 - z0));
        const bool full = spec.sx == 1 && spec.sy == 1 && spec.sz == 1;
        unsigned k0 = z0, k1 = z1;
        std::vector<double> kept;
        if (!full) kept = decimate(T.data(), spec, z0, z1, k0, k1);
        const double* src = full ? T.data() : kept.data();
        const std::string final_path = "output/out" + ext;
        if (world == 1) {
            if (raw) write_raw(final_path, src, spec, k0, k1, true, applied);
            else write_ascii(final_path, src, spec, k0, k1, true);
        } else {
            // Every rank formats its own planes into a part file; rank 0 joins them in rank order behind one header.  The
            // dump is k-outer / x-fastest (cu:1169-1171), so z-slabs in rank order ARE the single-zone file.
            const std::string part = final_path + ".part" + std::to_string(rank);
            const bool ok = raw ? write_raw(part, src, spec, k0, k1, false, applied) : write_ascii(part, src, spec, k0, k1, false);
            if (ok) { if (std::FILE* m = std::fopen((part + ".done").c_str(), "wb")) std::fclose(m); }
            if (root) {
                bool all = ok;
                const auto t0 = std::chrono::steady_clock::now();
                for (unsigned r = 0; r < world && all; ++r)
                    while (!file_exists(final_path + ".part" + std::to_string(r) + ".done")) {
                        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 3600.0) { all = false; break; }
                        std::this_thread::sleep_for(std::chrono::milliseconds(20));
                    }
                if (all) {
                    if (raw) write_raw(final_path, nullptr, spec, 0, 0, true, applied);
                    else write_ascii(final_path, nullptr, spec, 0, 0, true);
                    if (std::FILE* out = std::fopen(final_path.c_str(), "ab")) {
                        std::vector<char> buf(8u << 20);
                        for (unsigned r = 0; r < world; ++r) {
                            const std::string pr = final_path + ".part" + std::to_string(r);
                            if (std::FILE* in = std::fopen(pr.c_str(), "rb")) {
                                size_t n;
                                while ((n = std::fread(buf.data(), 1, buf.size(), in)) > 0) std::fwrite(buf.data(), 1, n, out);
                                std::fclose(in);
                            }
                            ::unlink(pr.c_str());
                            ::unlink((pr + ".done").c_str());
                        }
                        std::fclose(out);
                    }
                } else
                    std::cerr << "heat3D: not every rank delivered its part of " << final_path << std::endl;
            }
        }
    }
    for (int b = 0; b < 2; ++b) heat3d_host_free(snap[b]);
    heat3d_destroy(ctx);
    return 0;
}
