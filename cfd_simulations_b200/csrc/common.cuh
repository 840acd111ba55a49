// common.cuh — shared device helpers for libb200cfd (sm_100a).
// Layout: row-major (ny, nx) with row pitch `pitch` (elements), x fastest.  i = row (y), j = column (x).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include <cstdlib>

namespace b200cfd {

inline bool pdl_enabled() {
    // OFF by default — measured on B200 (4096x1024 fp64, CUDA-graph replay): attribute alone 0.3032 vs 0.3030 ms/step
    // (graph launches already hide the launch latency), with an early trigger 0.332 ms/step (waiting blocks of the next
    // kernel take the slots of the running one).  B200CFD_PDL=1 turns it on for experiments.
    static const bool on = getenv("B200CFD_PDL") != nullptr;
    return on;
}
// <<<grid, block, 0, s>>> with the programmatic-stream-serialization attribute (see pdl_wait below)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// the same with the attribute always on (B200CFD_PDL_CTL=0 turns it off): the predictor behind the control kernel
inline bool pdl_ctl_enabled() {
    static const bool on = [] { const char* e = getenv("B200CFD_PDL_CTL"); return !(e && *e == '0'); }();
    return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_ctl(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = (pdl_ctl_enabled() || pdl_enabled()) ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Geometry + precomputed coefficients (host-computed in double with the reference's expressions so
// the constants are bit-identical to the oracle's: v5.py:131-132,168-169,182-183,205-210).
struct Geom {
    int ny, nx;          // local array shape
    long long pitch;     // row stride in elements
    int jg0, nxg;        // global column of local column 0, global nx (x-slab decomposition)
    int own0, own1;      // local columns [own0, own1) are OWNED by this slab; the rest are halo columns that are
                         // recomputed redundantly (results there are never reduced, never sent)
    double dx, dy;
    double hdx, hdy;     // 0.5/dx, 0.5/dy            (divergence / gradient / SUPG dx_inv)
    double qdx, qdy;     // 0.5*hdx, 0.5*hdy          (SUPG first derivative, v5.py:137-138)
    double hdx2, hdy2;   // hdx*hdx, hdy*hdy          (SUPG second derivative, v5.py:141-142)
    double ax, ay;       // 1/(dx*dx), 1/(dy*dy)
    double dinv;         // 1/(2(ax+ay))
    double rdx, rdy;     // 1/dx, 1/dy                (upwind)
    double hmin;         // min(dx,dy)
    double two_dx, two_dy; // 2*dx, 2*dy              (vorticity, v5.py:369-370)
};

// Device-resident per-step control block: nothing on the host is needed between steps.
struct StepCtl {
    double dt;            // this step's dt (adaptive_time_step, v5.py:316-326)
    double fs;            // min(1, step/initial_steps)          (v5.py:406)
    double pert_scale;    // min(1, step/1000) * 0.01            (v5.py:351)
    double phase;         // 0.02 * step                         (v5.py:352)
    long long step;
    // order-preserving bit patterns of non-negative doubles (atomicMax on u64)
    unsigned long long max_div_pre, max_grad, max_div_post, max_vort, max_u, max_v;
    unsigned long long sum_nut_bits_unused;
    unsigned int nonfinite;
    unsigned int pad0;
    // adaptive-dt inputs = post-clip maxima of the state entering the step
    double in_max_u, in_max_v, in_mean_nut;
    // Poisson iteration control (tolerance mode runs without host round trips)
    unsigned long long pois_change;   // running max|dphi| of the current iteration
    double pois_last;                 // max|dphi| of the last finished iteration
    int pois_iters;                   // iterations finished
    int pois_done;                    // tolerance reached -> later launches exit immediately
    unsigned int pois_blocks_done;    // last-block election counter
    unsigned int pad1;
    double energy_sum, energy_cnt;    // global sum / count of the finite 0.5(u^2+v^2) values of this step
};

// Per-rank reduction vector exchanged between slabs at the end of a step (all_gather of RED_N doubles per rank).
enum { RED_MAX_DIV_PRE = 0, RED_MAX_GRAD, RED_MAX_DIV_POST, RED_MAX_VORT, RED_MAX_U, RED_MAX_V, RED_POIS_CHANGE,
       RED_NONFINITE, RED_ENERGY_SUM, RED_ENERGY_CNT, RED_NUT_SUM, RED_N = 16 };

struct StepScalars {   // must match b200cfd_step_scalars (12 doubles)
    double dt, max_div_pre, max_grad_phi, max_div_post, max_vorticity, mean_energy, max_abs_u, max_abs_v,
        poisson_iters, poisson_change, nonfinite, step;
};

template <typename T> struct Rnd;
template <> struct Rnd<double> { static __device__ __forceinline__ double r(double x) { return x; } };
// float32 storage mode: by default every operation the reference performs in float32 (NumPy / Numba on float32 arrays)
// is rounded to float32 here too, so the fp32 results track the reference's step by step.  -DB200CFD_F32_LAZY_ROUNDING
// keeps double precision between load and store instead (measured on B200: all parity tests still pass at the 1e-5
// criterion, the fp32 step gets 11 % faster: 6.88 -> 6.11 ms at 16384x8192, 104 -> 92 us at 600x180); not the default
// because fidelity to the reference's float32 arithmetic is the point of that mode.
#ifdef B200CFD_F32_LAZY_ROUNDING
template <> struct Rnd<float>  { static __device__ __forceinline__ double r(double x) { return x; } };
#else
template <> struct Rnd<float>  { static __device__ __forceinline__ double r(double x) { return (double)(float)x; } };
#endif

// a + s*(-c + l) evaluated in the field dtype (NumPy glue, v5.py:402-403)
template <typename T> __device__ __forceinline__ double axpy_pred(double a, double s, double c, double l);
template <> __device__ __forceinline__ double axpy_pred<double>(double a, double s, double c, double l) { return a + s * (-c + l); }
template <> __device__ __forceinline__ double axpy_pred<float>(double a, double s, double c, double l) {
    return (double)((float)a + (float)s * (-(float)c + (float)l));
}
// a - s*g evaluated in the field dtype (v5.py:416-417)
template <typename T> __device__ __forceinline__ double axmy(double a, double s, double g);
// Written with explicit intrinsics: two roundings, as NumPy evaluates `u_star - dt * grad`.  Left as `a - s * g` the compiler
// is free to contract it into one FMA or not, and it chose differently in the lean and in the general rows of K2+K3
// (tests/test_gpu_lean_paths.py found 3 cells of 786 432 one ulp apart after 5 steps).
template <> __device__ __forceinline__ double axmy<double>(double a, double s, double g) { return __dsub_rn(a, __dmul_rn(s, g)); }
template <> __device__ __forceinline__ double axmy<float>(double a, double s, double g) {
    return (double)__fsub_rn((float)a, __fmul_rn((float)s, (float)g));
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block max of a non-negative value -> one atomicMax per block (warp shuffle + smem tree).
// NaN inputs are dropped by fmax (nanmax semantics).
__device__ __forceinline__ void block_atomic_max(double v, unsigned long long* dst) {
    __shared__ double s_red[32];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int nw = (blockDim.x * blockDim.y + 31) >> 5;
    v = warp_max(v);
    __syncthreads();   // protect s_red reuse between consecutive calls
    if ((tid & 31) == 0) s_red[tid >> 5] = v;
    __syncthreads();
    if (tid < 32) {
        double w = tid < nw ? s_red[tid] : 0.0;
        w = warp_max(w);
        if (tid == 0 && w > 0.0) atomicMax(dst, (unsigned long long)__double_as_longlong(w));
    }
}

// Three block maxima of non-negative values with ONE barrier (1-D blocks of up to 32 warps): warp shuffles, one slot
// per warp, thread 0 folds the warps and issues the atomics.
__device__ __forceinline__ void block_atomic_max3(double a, double b, double c, unsigned long long* da,
                                                  unsigned long long* db, unsigned long long* dc) {
    __shared__ double s_m[3][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    a = warp_max(a); b = warp_max(b); c = warp_max(c);
    if (lane == 0) { s_m[0][w] = a; s_m[1][w] = b; s_m[2][w] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < nw; ++k) { a = fmax(a, s_m[0][k]); b = fmax(b, s_m[1][k]); c = fmax(c, s_m[2][k]); }
        if (a > 0.0) atomicMax(da, (unsigned long long)__double_as_longlong(a));
        if (b > 0.0) atomicMax(db, (unsigned long long)__double_as_longlong(b));
        if (c > 0.0) atomicMax(dc, (unsigned long long)__double_as_longlong(c));
    }
}

// Block sum with a fixed reduction tree (deterministic for a fixed launch shape); result valid in thread 0.
__device__ __forceinline__ double block_sum(double v) {
    __shared__ double s_sum[32];
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int nw = (blockDim.x * blockDim.y + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if ((tid & 31) == 0) s_sum[tid >> 5] = v;
    __syncthreads();
    double w = 0.0;
    if (tid < 32) {
        w = tid < nw ? s_sum[tid] : 0.0;
        w = warp_sum(w);
    }
    return w;
}

__device__ __forceinline__ double bits_to_double(unsigned long long b) { return __longlong_as_double((long long)b); }

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------------------
// The step is a chain of 7 kernels of 30-70 us (4096x1024) down to 6-40 us (600x180): launch latency and the drain /
// ramp between two kernels are a visible share.  A kernel launched with launch_pdl() may be SCHEDULED while its
// predecessor is still running; it must call pdl_wait() before it touches anything the predecessor reads or writes
// (first statement of the kernel), which returns once the predecessor has completed and its writes are visible.
// pdl_trigger() right after it lets the NEXT kernel of the chain be scheduled as soon as every block of this one is
// resident.  Without the launch attribute both are no-ops, so the same kernels work from any launcher.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#ifndef B200CFD_PDL_EARLY_TRIGGER
#define B200CFD_PDL_EARLY_TRIGGER 0   // measured on B200: triggering at kernel entry is SLOWER (4096x1024: 0.332 vs 0.302 ms/step)
#endif
// Unconditional release of the dependent launch (used by the one-block control kernel only: the predictor of the next step
// may then run its prologue - flag tiles, first rows of u, v, which the control kernel does not touch - while the control
// kernel computes the next dt; the predictor waits right before it reads the control block).
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() {
#if B200CFD_PDL_EARLY_TRIGGER
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}

}  // namespace b200cfd
