/* Single-rank MPI stand-in — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * The image has no MPI. The reference's host program (src/heat3D.cpp) only needs
 * the handful of MPI-3 names below, and with one rank every one of them is trivial:
 * rank 0 of 1, a 1x1x1 Cartesian grid, no neighbours (MPI_PROC_NULL on all six
 * sides), reductions are copies, point-to-point calls are never reached with a
 * real peer.  This header lets `g++` build the reference source where it lies
 * under /root/reference (oracle/Makefile) as the strict-arithmetic CPU baseline.
 * Names follow the MPI standard; nothing here is taken from an MPI implementation.
 */
#ifndef HEAT3D_ORACLE_MPI_SHIM_H
#define HEAT3D_ORACLE_MPI_SHIM_H
#include <chrono>
#include <cstring>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
struct MPI_Status { int MPI_SOURCE, MPI_TAG, MPI_ERROR; };

#define MPI_COMM_WORLD 0
#define MPI_PROC_NULL (-2)
#define MPI_SUCCESS 0
#define MPI_DOUBLE 8
#define MPI_INT 4
#define MPI_MIN 1
#define MPI_MAX 2
#define MPI_SUM 3
#define MPI_STATUS_IGNORE ((MPI_Status*)0)

static inline int MPI_Init(int*, char***) { return MPI_SUCCESS; }
static inline int MPI_Finalize() { return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return MPI_SUCCESS; }
static inline int MPI_Dims_create(int, int nd, int* dims) { for (int d = 0; d < nd; ++d) dims[d] = 1; return MPI_SUCCESS; }
static inline int MPI_Cart_create(MPI_Comm, int, const int*, const int*, int, MPI_Comm* c) { *c = 1; return MPI_SUCCESS; }
static inline int MPI_Cart_shift(MPI_Comm, int, int, int* lo, int* hi) { *lo = MPI_PROC_NULL; *hi = MPI_PROC_NULL; return MPI_SUCCESS; }
static inline int MPI_Cart_coords(MPI_Comm, int, int nd, int* c) { for (int d = 0; d < nd; ++d) c[d] = 0; return MPI_SUCCESS; }
static inline int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request* r) { *r = 0; return MPI_SUCCESS; }
static inline int MPI_Recv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*) { return MPI_SUCCESS; }
static inline int MPI_Send(const void*, int, MPI_Datatype, int, int, MPI_Comm) { return MPI_SUCCESS; }
static inline int MPI_Waitall(int, MPI_Request*, MPI_Status*) { return MPI_SUCCESS; }
static inline int MPI_Wait(MPI_Request*, MPI_Status*) { return MPI_SUCCESS; }
static inline int MPI_Iallreduce(const void* in, void* out, int n, MPI_Datatype t, MPI_Op, MPI_Comm, MPI_Request* r) {
    std::memcpy(out, in, (size_t)n * (size_t)t); *r = 0; return MPI_SUCCESS;
}
static inline double MPI_Wtime() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
#endif
