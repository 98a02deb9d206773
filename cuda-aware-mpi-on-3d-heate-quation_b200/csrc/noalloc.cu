// heat3d_mean_abs_error_vs_y_noalloc — the error norm for callers that are not allowed to allocate.
//
// Who needs this.  heat3d_mean_abs_error_vs_y() works on a heat3d_ctx, and a context owns device memory
// that heat3d_create obtained with cudaMalloc (field buffers, partial-sum scratch).  Some hosts of this
// library cannot go through that door at all: a solver plugged into an application that handed ALL device
// memory to its own pool at start-up (another cudaMalloc would fail or fragment the pool); a framework
// whose tensors must all come from its caching allocator; a monitoring hook that runs inside a latency-
// critical loop where cudaMalloc's implicit device synchronisation is not acceptable.  Such a caller
// already holds the field in its own device buffer and only wants the number.  This entry point takes raw
// device pointers for input and output and keeps its scratch (one partial sum per interior z-plane) in a
// fixed array that is part of the library image, so it never calls cudaMalloc, cudaFree, cudaHostAlloc or
// any library routine that allocates.
//
// Consequences of the fixed array, all of which a caller should know:
//  - capacity: HEAT3D_NOALLOC_MAX_PLANES interior planes (64 KiB of device memory, held by every GPU that
//    loads the library until the process exits); larger fields are refused with HEAT3D_ERR_INVALID;
//  - it is shared by every call in the process: two calls that overlap on the same GPU (different streams,
//    or different host threads) race on it and give wrong sums - serialise them;
//  - it keeps whatever the previous call left there.  Nothing has to start at zero: pass 1 overwrites exactly
//    the slots pass 2 reads, and pass 2 reads no others;
//  - the first call on each device looks up the array's device address (cudaGetSymbolAddress; this may load
//    the module on that device), so make that first call outside any stream capture.
#include "../../include/heat3d.h"
#include "internal.h"

#include <cuda_runtime.h>

#include <atomic>
#include <climits>

namespace {

constexpr int MAX_DEVICES = 64;

// The scratch itself.  In host code this name is only a handle for the runtime, not a device address:
// the address is fetched per device below and handed to the kernels as an ordinary pointer.
static __device__ __align__(256) double g_plane_partials[HEAT3D_NOALLOC_MAX_PLANES];

// device address of g_plane_partials on each device, once known.  Per process and per primary context: after a
// cudaDeviceReset the cached value is stale.  Atomic because first calls may come from several host threads (they all
// resolve the same address, so a lost race only repeats the look-up).
std::atomic<double*> g_partials_addr[MAX_DEVICES];

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
    return t;   // valid in thread 0
}

// pass 1: block b sums |T - j*sp_y| over the interior cells of plane z = 1 + b and stores it in partials[b]
__global__ void __launch_bounds__(256) plane_abs_error_kernel(const double* __restrict__ f, int nx, int ny, long long pitch,
                                                              long long plane, double sp_y, double* __restrict__ partials)
{
    const double* p = f + (long long)(1 + blockIdx.x) * plane;
    double acc = 0.0;
    for (int j = 1 + (threadIdx.x >> 5); j <= ny - 2; j += 8) {      // one warp per row, lanes along x
        const double yj = (double)j * sp_y;
        const double* row = p + (long long)j * pitch;
        for (int i = 1 + (threadIdx.x & 31); i <= nx - 2; i += 32) acc += fabs(__dsub_rn(row[i], yj));
    }
    const double s = block_sum<256>(acc);
    if (threadIdx.x == 0) partials[blockIdx.x] = s;
}

// pass 2: fold the first n partials (and only those) and scale
__global__ void __launch_bounds__(256) fold_partials_kernel(const double* __restrict__ partials, int n, double inv_count,
                                                            double* __restrict__ out)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) acc += partials[i];
    const double s = block_sum<256>(acc);
    if (threadIdx.x == 0) *out = s * inv_count;
}

}  // namespace

extern "C" int heat3d_mean_abs_error_vs_y_noalloc(const double* d_field, uint32_t nx, uint32_t ny, uint32_t nz,
                                                  uint32_t pitch_doubles, double sp_y, double* d_err, void* stream)
{
    if (!d_field || !d_err) return h3::set_last_error(HEAT3D_ERR_INVALID, "null device pointer");
    if (nx < 3 || ny < 3 || nz < 3) return h3::set_last_error(HEAT3D_ERR_INVALID, "grid %ux%ux%u: every axis needs >= 3 points", nx, ny, nz);
    if (pitch_doubles < nx) return h3::set_last_error(HEAT3D_ERR_INVALID, "row pitch %u doubles is shorter than nx=%u", pitch_doubles, nx);
    if (nx > (uint32_t)INT_MAX || ny > (uint32_t)INT_MAX || pitch_doubles > (uint32_t)INT_MAX)
        return h3::set_last_error(HEAT3D_ERR_INVALID, "nx=%u, ny=%u, pitch=%u: each must fit a 32-bit signed int", nx, ny, pitch_doubles);
    const uint32_t planes = nz - 2;
    if (planes > HEAT3D_NOALLOC_MAX_PLANES)
        return h3::set_last_error(HEAT3D_ERR_INVALID,
                                  "nz=%u has %u interior planes; the no-allocation entry point holds scratch for at most %d "
                                  "(HEAT3D_NOALLOC_MAX_PLANES) - use heat3d_mean_abs_error_vs_y on a context instead",
                                  nz, planes, HEAT3D_NOALLOC_MAX_PLANES);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return h3::set_last_error(HEAT3D_ERR_CUDA, "cudaGetDevice failed: %s (libheat3d has no CPU fallback)", cudaGetErrorString(e));
    if (dev < 0 || dev >= MAX_DEVICES) return h3::set_last_error(HEAT3D_ERR_INVALID, "device ordinal %d not supported here", dev);
    double* partials = g_partials_addr[dev].load(std::memory_order_acquire);
    if (!partials) {
        // once per device, with that device current: the host-side symbol is not an address a kernel can use
        e = cudaGetSymbolAddress((void**)&partials, g_plane_partials);
        if (e != cudaSuccess || !partials)
            return h3::set_last_error(HEAT3D_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d failed: %s", dev, cudaGetErrorString(e));
        g_partials_addr[dev].store(partials, std::memory_order_release);
    }
    cudaStream_t st = (cudaStream_t)stream;
    // same divisor as the reference: an `unsigned` product (heat3D.cu:1102), kept with its wrap-around
    const unsigned count = (nx - 2u) * (ny - 2u) * (nz - 2u);
    if (count == 0u)
        return h3::set_last_error(HEAT3D_ERR_INVALID, "(nx-2)(ny-2)(nz-2) wraps to 0 in the reference's 32-bit divisor for %ux%ux%u", nx, ny, nz);
    plane_abs_error_kernel<<<planes, 256, 0, st>>>(d_field, (int)nx, (int)ny, (long long)pitch_doubles,
                                                   (long long)pitch_doubles * ny, sp_y, partials);
    fold_partials_kernel<<<1, 256, 0, st>>>(partials, (int)planes, 1.0 / (double)count, d_err);
    e = cudaGetLastError();
    if (e != cudaSuccess) return h3::set_last_error(HEAT3D_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return HEAT3D_OK;
}
