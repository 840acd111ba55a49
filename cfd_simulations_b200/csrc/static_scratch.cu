// static_scratch.cu — b200cfd_jacobi_static: the Jacobi pressure sweeps of solve_pressure_fast without any allocation.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  b200cfd_poisson(mode = JACOBI) needs a
// handle (b200cfd_create allocates) and a ping-pong buffer from cudaMalloc; this entry point needs neither: the ping-pong
// buffer is the file-scope device array below, B200CFD_STATIC_SCRATCH_BYTES big, which every GPU that loads the library
// holds until the process exits.
//
// Rules of the array:
//   * `g_jacobi_scratch` in HOST code is a host-side stand-in for the symbol, not a device address; the address a kernel
//     may use comes from cudaGetSymbolAddress, asked once per device with that device current (scratch_for_current_device).
//   * All calls on a device share the one array and it keeps the last call's contents.  Nothing in it has to start at
//     zero: the first sweep of a call writes every cell of the (ny, nx) region before any sweep reads it.
//   * Two calls whose kernels overlap on the GPU (different streams, no dependency between them) race on the array;
//     calls on one stream, or ordered by events, are safe.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <mutex>
#include "../../include/b200cfd.h"
#include "kernels.h"

namespace b200cfd {

static __device__ __align__(256) unsigned char g_jacobi_scratch[B200CFD_STATIC_SCRATCH_BYTES];

// One sweep, the arithmetic of k_jacobi (poisson.cu) operation for operation:  v5.py:338-346
//   phi_new = copy(phi); interior = 0.25 (E + W + N + S - dx^2 div / dt); masked -> 0.
template <typename T>
__global__ void __launch_bounds__(256)
k_jacobi_static(int ny, int nx, const T* __restrict__ in, T* __restrict__ out, const T* __restrict__ div,
                const uint8_t* __restrict__ mask, double dx2_over_dt) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y;
    if (i >= ny || j >= nx) return;
    const long long k = (long long)i * nx + j;
    double o;
    if (mask && mask[k]) o = 0.0;
    else if (i >= 1 && i < ny - 1 && j >= 1 && j < nx - 1)
        o = 0.25 * ((double)in[k + 1] + (double)in[k - 1] + (double)in[k + nx] + (double)in[k - nx] - dx2_over_dt * (double)div[k]);
    else o = (double)in[k];
    out[k] = (T)o;
}

// device address of the array on the CURRENT device (asked once per device); nullptr + message on failure
static void* scratch_for_current_device(int dev, char* msg, size_t msg_cap) {
    static std::mutex mu;
    static void* ptr[64] = {nullptr};
    if (dev < 0 || dev >= 64) { snprintf(msg, msg_cap, "device ordinal %d out of range", dev); return nullptr; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_jacobi_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            snprintf(msg, msg_cap, "cudaGetSymbolAddress(g_jacobi_scratch): %s", cudaGetErrorString(e));
            return nullptr;
        }
        ptr[dev] = p;
    }
    return ptr[dev];
}

}  // namespace b200cfd

using namespace b200cfd;

extern "C" int b200cfd_jacobi_static(int dtype, int ny, int nx, void* phi, const void* rhs_div, const uint8_t* mask,
                                     double dx, double dt, int sweeps, void* stream) {
    char msg[256];
    if (!phi || !rhs_div) { set_global_error("b200cfd_jacobi_static: null argument"); return B200CFD_EINVAL; }
    if ((dtype != B200CFD_F64 && dtype != B200CFD_F32) || ny < 3 || nx < 3 || sweeps < 0 || !(dx > 0) || !(dt != 0)) {
        set_global_error("b200cfd_jacobi_static: bad dtype/ny/nx/sweeps/dx/dt");
        return B200CFD_EINVAL;
    }
    const size_t es = dtype == B200CFD_F64 ? 8 : 4;
    const size_t bytes = (size_t)ny * (size_t)nx * es;
    if (bytes > (size_t)B200CFD_STATIC_SCRATCH_BYTES) {
        snprintf(msg, sizeof msg, "b200cfd_jacobi_static: %d x %d cells of %zu bytes need %zu bytes of scratch, the static array holds "
                 "%zu (B200CFD_STATIC_SCRATCH_BYTES); use b200cfd_poisson for larger grids", ny, nx, es, bytes,
                 (size_t)B200CFD_STATIC_SCRATCH_BYTES);
        set_global_error(msg);
        return B200CFD_EINVAL;
    }
    if (sweeps == 0) return B200CFD_OK;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        snprintf(msg, sizeof msg, "b200cfd_jacobi_static: cudaGetDevice: %s", cudaGetErrorString(e));
        set_global_error(msg);
        return B200CFD_ECUDA;
    }
    void* scratch = scratch_for_current_device(dev, msg, sizeof msg);
    if (!scratch) { set_global_error(msg); return B200CFD_ECUDA; }
    cudaStream_t s = (cudaStream_t)stream;
    const double k = dx * dx / dt;
    const dim3 block(32, 8), grid((unsigned)((nx + 31) / 32), (unsigned)((ny + 7) / 8));
    void* a = phi;
    void* b = scratch;
    for (int it = 0; it < sweeps; ++it) {
        if (dtype == B200CFD_F64)
            k_jacobi_static<double><<<grid, block, 0, s>>>(ny, nx, (const double*)a, (double*)b, (const double*)rhs_div, mask, k);
        else
            k_jacobi_static<float><<<grid, block, 0, s>>>(ny, nx, (const float*)a, (float*)b, (const float*)rhs_div, mask, k);
        void* t = a; a = b; b = t;
    }
    e = cudaPeekAtLastError();
    if (e == cudaSuccess && a != phi) e = cudaMemcpyAsync(phi, a, bytes, cudaMemcpyDeviceToDevice, s);   // odd count: result is in the array
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "b200cfd_jacobi_static: %s", cudaGetErrorString(e));
        set_global_error(msg);
        return B200CFD_ECUDA;
    }
    return B200CFD_OK;
}
