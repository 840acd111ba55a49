// step_fast.cu — row-marching fast versions of the step kernels for the common case (SUPG on, LES off, even nx,
// 16-byte aligned fields).  Same arithmetic as the general kernels in step_kernels.cu / poisson.cu (they remain the
// path for every other configuration), reorganised so that the kernels are HBM-bound instead of issue-bound:
//   * each thread owns 2 adjacent columns (one 128-bit load / store per row and field) and marches down a chunk of
//     rows keeping the rows i-1, i, i+1 in registers, so north/south neighbours cost nothing;
//   * the block reductions (max|div|, max|vorticity|, energy sum, ...) happen once per block, not once per 256 cells;
//   * the immersed-boundary mask (fp64, 8 B/cell) and the cylinder mask are only read where they are non-trivial:
//     a tile-flag map (32 columns x 8 rows per flag, built once when the fields are bound) says which tiles contain
//     ibm > 0 or masked cells (dilated by 2 cells); warps whose tile is flagged, and the few threads that hold a
//     border column, run the general per-cell code (cellmath.cuh), everything else runs the lean interior path.
// Reference: python/flow_over_cylinder (Fischer)/v5.py:127-176 (SUPG, convection, Laplacian), :178-226 (divergence,
// red-black iteration), :349-360 (BC), :228-237 (IBM), :421-438 (diagnostics, clip).
#include "kernels.h"
#include <cstdlib>
#include <type_traits>
#include "cellmath.cuh"

namespace b200cfd {

#define FT_W 32      // flag tile: columns
#define FT_H 8       // flag tile: rows
#define FBT 128      // threads per block of the marching kernels
#define FCOLS 256    // columns per block (2 per thread)

// optional unrolling of the row loops (would remove the register-rotation moves; see the measurement below)
#ifndef B200CFD_MARCH_UNROLL
#define B200CFD_MARCH_UNROLL 1   // measured on B200: 3 spills (40 B stack) and is slower; 1 = no unrolling
#endif
#define MARCH_PRAGMA(x) _Pragma(#x)
#define MARCH_UNROLL_N(n) MARCH_PRAGMA(unroll n)
#define MARCH_UNROLL MARCH_UNROLL_N(B200CFD_MARCH_UNROLL)
#define FPF 6        // L2 prefetch distance of the marching kernels, rows
#ifndef FIN_MINB
#define FIN_MINB 3   // resident blocks per SM the finalize kernel is compiled for
#endif
#ifndef FIN_DEEP
#define FIN_DEEP 0   // lean finalize rows: 1 = loads issued two rows ahead; measured: no gain (60.2 vs 59.8 us at 4096x1024, 2.23 vs 2.20 ms at 16384^2)
#endif
#ifndef FIN_LEAN
#define FIN_LEAN 1   // 0: k_finalize_fast without its lean interior rows (A/B builds)
#endif

__device__ __forceinline__ void pf_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

struct D2 { double x, y; };
template <typename T> __device__ __forceinline__ D2 ld2(const T* __restrict__ p, long long k);
template <> __device__ __forceinline__ D2 ld2<double>(const double* __restrict__ p, long long k) {
    const double2 t = *reinterpret_cast<const double2*>(p + k);
    return D2{t.x, t.y};
}
template <> __device__ __forceinline__ D2 ld2<float>(const float* __restrict__ p, long long k) {
    const float2 t = *reinterpret_cast<const float2*>(p + k);
    return D2{(double)t.x, (double)t.y};
}
template <typename T> __device__ __forceinline__ D2 ld2cg(const T* __restrict__ p, long long k);
template <> __device__ __forceinline__ D2 ld2cg<double>(const double* __restrict__ p, long long k) {
    const double2 t = __ldcg(reinterpret_cast<const double2*>(p + k));
    return D2{t.x, t.y};
}
template <> __device__ __forceinline__ D2 ld2cg<float>(const float* __restrict__ p, long long k) {
    const float2 t = __ldcg(reinterpret_cast<const float2*>(p + k));
    return D2{(double)t.x, (double)t.y};
}
template <typename T> __device__ __forceinline__ void st2(T* __restrict__ p, long long k, double a, double b);
template <> __device__ __forceinline__ void st2<double>(double* __restrict__ p, long long k, double a, double b) {
    *reinterpret_cast<double2*>(p + k) = make_double2(a, b);
}
template <> __device__ __forceinline__ void st2<float>(float* __restrict__ p, long long k, double a, double b) {
    *reinterpret_cast<float2*>(p + k) = make_float2((float)a, (float)b);
}

// ------------------------------------------------------------------------------------------------
// tile flags: bit 0 = some ibm > 0, bit 1 = some masked cell, in the flag tile dilated by 2 cells
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_tile_flags(int ny, int nx, long long pitch, const double* __restrict__ ibm, const uint8_t* __restrict__ mask,
             uint8_t* __restrict__ flags, int ntx) {
    const int tx = blockIdx.x, ty = blockIdx.y;
    const int c0 = max(tx * FT_W - 2, 0), c1 = min(tx * FT_W + FT_W + 2, nx);
    const int r0 = max(ty * FT_H - 2, 0), r1 = min(ty * FT_H + FT_H + 2, ny);
    const int w = c1 - c0, n = w * (r1 - r0);
    int f = 0;
    for (int t = threadIdx.x; t < n; t += 256) {
        const long long k = (long long)(r0 + t / w) * pitch + c0 + t % w;
        if (ibm && ibm[k] > 0.0) f |= 1;
        if (mask && mask[k]) f |= 2;
    }
    const int a = __syncthreads_or(f & 1), b = __syncthreads_or(f & 2);
    if (threadIdx.x == 0) flags[ty * ntx + tx] = (uint8_t)((a ? 1 : 0) | (b ? 2 : 0));
}

int tile_flags_dims(const Geom& g, int* ntx, int* nty) {
    *ntx = (g.nx + FT_W - 1) / FT_W;
    *nty = (g.ny + FT_H - 1) / FT_H;
    return *ntx * *nty;
}
void launch_tile_flags(const Geom& g, const double* ibm, const uint8_t* mask, uint8_t* flags, cudaStream_t s) {
    int ntx, nty;
    tile_flags_dims(g, &ntx, &nty);
    k_tile_flags<<<dim3(ntx, nty), 256, 0, s>>>(g.ny, g.nx, g.pitch, ibm, mask, flags, ntx);
}

// does any flag tile overlapping columns [c0, c1) x rows [r0, r1) carry one of `bits`?  Block-wide (all threads call).
__device__ __forceinline__ bool block_tile_flagged(const uint8_t* __restrict__ flags, int ntx, int c0, int c1, int r0, int r1,
                                                   int bits) {
    const int tx0 = c0 / FT_W, tx1 = (c1 - 1) / FT_W, ty0 = r0 / FT_H, ty1 = (r1 - 1) / FT_H;
    const int w = tx1 - tx0 + 1, n = w * (ty1 - ty0 + 1);
    int f = 0;
    for (int t = threadIdx.x; t < n; t += blockDim.x) f |= flags[(ty0 + t / w) * ntx + tx0 + t % w] & bits;
    return __syncthreads_or(f) != 0;
}

// ------------------------------------------------------------------------------------------------
// K1 fast: u*, v* = u + dt(-C + L) for plain interior cells (no BC, no IBM in the tile)      v5.py:388-407
// tau = h/(2|V|) min(1, Pe/2) with Pe = |V| h/(nu+1e-10) is evaluated as min(h/(2|V|), h^2/(4(nu+1e-10))) with one
// rsqrt (algebraically identical to v5.py:155-158; differs in the last ulp only).
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void predict_fast(const Geom& g, double nu_eff, double half_h, double tau_cap, double dt,
                                             double uc, double ue, double uw, double un, double us_, double vc, double ve,
                                             double vw, double vn, double vs_, double& ru, double& rv, double& tau_o) {
    const double s = uc * uc + vc * vc;
    // the minimum as a compare-and-select: fmin() is 7-8 instructions on sm_100a (NaN quieting), and no NaN gets here
    const double hv = half_h * rsqrt(s);
    double tau = (s > 1e-20) ? (hv < tau_cap ? hv : tau_cap) : dt / 2;
    tau = Rnd<T>::r(tau);
    tau_o = tau;
    const double cu = conv_supg_cell<T>(g, uc, vc, uc, ue, uw, un, us_, tau);
    const double cv = conv_supg_cell<T>(g, uc, vc, vc, ve, vw, vn, vs_, tau);
    const double lu = lap_cell<T>(g, nu_eff, uc, ue, uw, un, us_);
    const double lv = lap_cell<T>(g, nu_eff, vc, ve, vw, vn, vs_);
    ru = axpy_pred<T>(uc, dt, cu, lu);
    rv = axpy_pred<T>(vc, dt, cv, lv);
}


// The lean rows of k_predictor_fast: every column of the block is an interior column with both x-neighbours, no flag tile
// over the block carries ibm > 0, rows [lo - 1, hi] exist and hi - lo is a multiple of 4.  Then BC, IBM and the column
// classes drop out and what is left is written for the issue slots (the general loop: 437 instructions per lane-row, 151
// of them fp64 arithmetic): rows rotate through register arrays indexed by (row & 3) in a loop unrolled by 4, and eight
// lanes of a warp prefetch the eight 128-byte lines of the row FPF ahead instead of all 32 lanes prefetching twice.
// predict_fast() is the same function the general rows call, and the two paths agree bit for bit in tests/test_gpu_lean_paths.py.
template <typename T, bool TAU>
__device__ __forceinline__ void predictor_rows_lean(const Geom& g, const Phys& p, const StepCtl* __restrict__ ctl, const T* __restrict__ u,
                                                    const T* __restrict__ v, T* __restrict__ us, T* __restrict__ vs,
                                                    T* __restrict__ tau_out, const int lo, const int hi, const int j0) {
    const double half_h = 0.5 * g.hmin, tau_cap = 0.25 * g.hmin * g.hmin / (p.nu_eff + 1e-10);
    const long long pitch = g.pitch;
    long long k = (long long)lo * pitch + j0;
    D2 U[4], V[4];                                    // rows i-1, i, i+1 and the row in flight: slot (row - lo + 1) & 3
    double UW[2], UE[2], VW[2], VE[2];                // x-neighbours of the lane's column pair: slot (row - lo) & 1
    // The first rows are requested BEFORE the wait for the control kernel (which this kernel may overlap, see pdl_release):
    // u, v were finished by the finalize kernel before the control kernel started.  ld.global.cg: past L1, whose state under
    // an overlapped launch is not worth reasoning about for twelve loads.
    U[0] = ld2cg<T>(u, k - pitch); V[0] = ld2cg<T>(v, k - pitch);
    U[1] = ld2cg<T>(u, k); V[1] = ld2cg<T>(v, k);
    U[2] = ld2cg<T>(u, k + pitch); V[2] = ld2cg<T>(v, k + pitch);
    UW[0] = (double)__ldcg(u + k - 1); UE[0] = (double)__ldcg(u + k + 2); VW[0] = (double)__ldcg(v + k - 1); VE[0] = (double)__ldcg(v + k + 2);
    pdl_wait();
    const double dt = ctl->dt;
    const int lane = threadIdx.x & 31;
    const T* pf = (lane < 4 ? u : v) + (long long)(lo + FPF) * pitch + (j0 - 2 * lane) + (lane & 3) * (int)(128 / sizeof(T));
    for (int i4 = lo; i4 < hi; i4 += 4) {
        const bool more = i4 + 4 < hi;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (r < 3 || more) {                      // row i+2, and the x-neighbours of row i+1
                U[(r + 3) & 3] = ld2<T>(u, k + 2 * pitch); V[(r + 3) & 3] = ld2<T>(v, k + 2 * pitch);
                UW[(r + 1) & 1] = ld(u, k + pitch - 1); UE[(r + 1) & 1] = ld(u, k + pitch + 2);
                VW[(r + 1) & 1] = ld(v, k + pitch - 1); VE[(r + 1) & 1] = ld(v, k + pitch + 2);
            }
            if (lane < 8 && i4 + r + FPF < g.ny) pf_l2(pf);
            pf += pitch;
            const D2 um = U[r], uc = U[(r + 1) & 3], un = U[(r + 2) & 3], vm = V[r], vc = V[(r + 1) & 3], vn = V[(r + 2) & 3];
            double ru0, rv0, t0, ru1, rv1, t1;
            predict_fast<T>(g, p.nu_eff, half_h, tau_cap, dt, uc.x, uc.y, UW[r & 1], un.x, um.x, vc.x, vc.y, VW[r & 1], vn.x, vm.x, ru0, rv0, t0);
            predict_fast<T>(g, p.nu_eff, half_h, tau_cap, dt, uc.y, UE[r & 1], uc.x, un.y, um.y, vc.y, VE[r & 1], vc.x, vn.y, vm.y, ru1, rv1, t1);
            st2<T>(us, k, ru0, ru1);
            st2<T>(vs, k, rv0, rv1);
            if (TAU) st2<T>(tau_out, k, t0, t1);
            k += pitch;
        }
    }
}

template <typename T, bool TAU>
__global__ void __launch_bounds__(FBT, 5)
k_predictor_fast(Geom g, Phys p, const StepCtl* __restrict__ ctl, const T* __restrict__ u, const T* __restrict__ v,
                 T* __restrict__ us, T* __restrict__ vs, T* __restrict__ tau_out, const double* __restrict__ ibm,
                 const double* __restrict__ y, const uint8_t* __restrict__ flags, int ntx, int rpb, int lean_on) {
    // no wait at the top: what precedes this kernel is the one-block control kernel, which releases it at once (pdl_release);
    // the flag tiles are static and u, v are final, only the control block must be waited for
    const int r0 = blockIdx.y * rpb, r1 = min(r0 + rpb, g.ny);
    const int cb = blockIdx.x * FCOLS;
    const int j0 = cb + 2 * threadIdx.x;
    const bool has_ibm = block_tile_flagged(flags, ntx, cb, min(cb + FCOLS, g.nx), r0, r1, 1);
    // lean rows (see predictor_rows_lean): a block-uniform choice
    if (lean_on && !has_ibm && r0 >= 1 && r1 <= g.ny - 1 && ((r1 - r0) & 3) == 0 &&
        __syncthreads_and(j0 >= 1 && j0 + 2 < g.nx && col_class(g, j0) == COL_INTERIOR && col_class(g, j0 + 1) == COL_INTERIOR)) {
        predictor_rows_lean<T, TAU>(g, p, ctl, u, v, us, vs, tau_out, r0, r1, j0);
        return;
    }
    pdl_wait();
    if (j0 >= g.nx) return;
    const int cx = col_class(g, j0), cy = col_class(g, j0 + 1);
    const bool hw = j0 >= 1, he = j0 + 2 < g.nx;             // west / east neighbour columns exist
    const int lo = max(r0, 1), hi = min(r1, g.ny - 1);
    if (r0 == 0) {
        st2<T>(us, j0, 0.0, 0.0); st2<T>(vs, j0, 0.0, 0.0);
        if (TAU) st2<T>(tau_out, j0, 0.0, 0.0);
    }
    if (r1 == g.ny) {
        const long long k = (long long)(g.ny - 1) * g.pitch + j0;
        st2<T>(us, k, 0.0, 0.0); st2<T>(vs, k, 0.0, 0.0);
        if (TAU) st2<T>(tau_out, k, 0.0, 0.0);
    }
    if (lo >= hi) return;
    const double dt = ctl->dt, fs = ctl->fs, pert = ctl->pert_scale, phase = ctl->phase;
    const double half_h = 0.5 * g.hmin, tau_cap = 0.25 * g.hmin * g.hmin / (p.nu_eff + 1e-10);
    long long k = (long long)lo * g.pitch + j0;
    D2 um = ld2<T>(u, k - g.pitch), uc = ld2<T>(u, k), vm = ld2<T>(v, k - g.pitch), vc = ld2<T>(v, k);
    D2 un = ld2<T>(u, k + g.pitch), vn = ld2<T>(v, k + g.pitch);
    double uw = hw ? ld(u, k - 1) : 0.0, ue = he ? ld(u, k + 2) : 0.0, vw = hw ? ld(v, k - 1) : 0.0, ve = he ? ld(v, k + 2) : 0.0;
    MARCH_UNROLL
    for (int i = lo; i < hi; ++i, k += g.pitch) {
        // software pipeline: the loads of the NEXT iteration (row i+2, and the x-neighbours of row i+1) are issued now
        D2 un2 = {0, 0}, vn2 = {0, 0};
        double uw1 = 0.0, ue1 = 0.0, vw1 = 0.0, ve1 = 0.0;
        if (i + 1 < hi) {
            un2 = ld2<T>(u, k + 2 * g.pitch); vn2 = ld2<T>(v, k + 2 * g.pitch);
            if (hw) { uw1 = ld(u, k + g.pitch - 1); vw1 = ld(v, k + g.pitch - 1); }
            if (he) { ue1 = ld(u, k + g.pitch + 2); ve1 = ld(v, k + g.pitch + 2); }
        }
        if (i + FPF < g.ny) { pf_l2(u + k + FPF * g.pitch); pf_l2(v + k + FPF * g.pitch); }
        double ru0, rv0, t0 = 0.0, ru1, rv1, t1 = 0.0;
        if (cx == COL_INTERIOR)
            predict_fast<T>(g, p.nu_eff, half_h, tau_cap, dt, uc.x, uc.y, uw, un.x, um.x, vc.x, vc.y, vw, vn.x, vm.x, ru0, rv0, t0);
        else if (cx == COL_INFLOW) { ru0 = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[i], pert, phase)); rv0 = 0.0; }
        else { ru0 = uc.x; rv0 = vc.x; }
        if (cy == COL_INTERIOR)
            predict_fast<T>(g, p.nu_eff, half_h, tau_cap, dt, uc.y, ue, uc.x, un.y, um.y, vc.y, ve, vc.x, vn.y, vm.y, ru1, rv1, t1);
        else if (cy == COL_OUTFLOW) { ru1 = ru0; rv1 = rv0; }        // BC copies the already predicted column nx-2 (v5.py:355-356)
        else { ru1 = uc.y; rv1 = vc.y; }
        if (has_ibm) {
            const double2 m = *reinterpret_cast<const double2*>(ibm + k);
            if (m.x > 0) { ru0 = Rnd<T>::r(ru0 * (1.0 - m.x * fs)); rv0 = Rnd<T>::r(rv0 * (1.0 - m.x * fs)); }
            if (m.y > 0) { ru1 = Rnd<T>::r(ru1 * (1.0 - m.y * fs)); rv1 = Rnd<T>::r(rv1 * (1.0 - m.y * fs)); }
        }
        st2<T>(us, k, ru0, ru1);
        st2<T>(vs, k, rv0, rv1);
        if (TAU) st2<T>(tau_out, k, t0, t1);
        um = uc; uc = un; vm = vc; vc = vn; un = un2; vn = vn2;
        uw = uw1; ue = ue1; vw = vw1; ve = ve1;
    }
}

// Rows per block of a row-marching kernel from a cost model instead of "about 8 blocks per SM": a launch takes
// waves x (rows marched per block), waves = ceil(blocks / resident slots), rows marched = rpb + the halo rows a block
// re-reads + a fixed prologue worth ~6 rows.  At 4096x1024 the old rule gave 2.6-2.8 waves of 8-16-row blocks (the last wave
// 60-80 % full, 12-25 % halo rows); one full wave of ~40-row blocks does the same work in fewer row-times.  Large grids
// (many waves either way) get long chunks.
// OFF by default - measured on B200 (profiles/r02k_phases.txt): the model is wrong where it matters.  4096x1024: predictor
// 44.4 -> 46.8 us, K2+K3 54.4 -> 57.5, finalize 77 -> 94 us (one wave of long blocks hides latency worse than 2.6 waves of
// short ones: every block starts with a dependent prologue of loads, and block turnover is what overlaps them); 16384^2:
// finalize 3.09 -> 2.99 ms, the other two 1 % slower.  B200CFD_RPB_MODEL=1 turns it on.
static inline int pick_rpb(int bx, int ny, int slots, int halo_rows, int fallback) {
    static const int on = getenv("B200CFD_RPB_MODEL") ? atoi(getenv("B200CFD_RPB_MODEL")) : 0;
    if (!on || (long long)bx * ((ny + 7) / 8) < 2ll * slots) return fallback;       // tiny grids: keep the latency rule
    int best = fallback;
    double best_cost = 1e300;
    for (int rpb = 8; rpb <= 128; ++rpb) {
        const long long nblk = (long long)bx * ((ny + rpb - 1) / rpb);
        const long long waves = (nblk + slots - 1) / slots;
        const double cost = (double)waves * (rpb + halo_rows + 6);
        if (cost < best_cost) { best_cost = cost; best = rpb; }
    }
    return best;
}

// The halo-lane kernels stop halving at 16 rows per block.  Where that still leaves fewer than 8 blocks per SM (4096x1024:
// 1152 blocks, 2.6 waves of the finalize kernel) 12-row blocks measured faster (tools/r02v.sh: finalize 59.8 -> 58.8 us,
// K2+K3 51.5 -> 49.3 us; 8, 20, 24 and longer are slower): more, shorter blocks fill the last wave better than the
// extra halo rows cost.  Not for grids so small that the latency rule below takes over.
static inline int rpb_12(int bx, int ny, int n_sm, int rpb) {
    if (rpb == 16 && (long long)bx * ((ny + 15) / 16) < 8ll * n_sm && (long long)bx * ((ny + 15) / 16) >= 2ll * n_sm) return 12;
    return rpb;
}

static inline int rows_per_block(const Geom& g, int n_sm) {
    // enough blocks for ~8 blocks per SM, chunks of 8..64 rows (multiples of the flag tile height)
    const int bx = (g.nx + FCOLS - 1) / FCOLS;
    int rpb = 64;
    while (rpb > 8 && (long long)bx * ((g.ny + rpb - 1) / rpb) < 8ll * n_sm) rpb >>= 1;
    // small grids (the shipped 600x180 case): the kernel is pure latency, so trade redundant halo rows for parallelism
    while (rpb > 4 && (long long)bx * ((g.ny + rpb - 1) / rpb) < 2ll * n_sm) rpb >>= 1;
    static const int force = getenv("B200CFD_PRED_RPB") ? atoi(getenv("B200CFD_PRED_RPB")) : 0;   // developer knob
    if (force > 0) return force;
    return pick_rpb(bx, g.ny, 5 * n_sm, 2, rpb);             // k_predictor_fast: 5 blocks per SM (launch bounds)
}

bool fast_path_ok(const Geom& g, const void* const* ptrs, int n) {
    if (g.nx % 2 || g.pitch % 2 || g.nx < 8 || g.ny < 4) return false;
    for (int k = 0; k < n; ++k)
        if (ptrs[k] && ((uintptr_t)ptrs[k] & 15)) return false;
    return true;
}

void launch_predictor_fast(int dtype, const Geom& g, const Phys& p, const StepCtl* ctl, const void* u, const void* v,
                           void* us, void* vs, void* tau, const double* ibm, const double* y, const uint8_t* flags,
                           int n_sm, cudaStream_t s, bool after_ctl) {
    const int rpb = rows_per_block(g, n_sm);
    const dim3 grid((g.nx + FCOLS - 1) / FCOLS, (g.ny + rpb - 1) / rpb);
    const int ntx = (g.nx + FT_W - 1) / FT_W;
#define PF_ARGS(T) g, p, ctl, (const T*)u, (const T*)v, (T*)us, (T*)vs, (T*)tau, ibm, y, flags, ntx, rpb, lean_on
#define PF(T, TAU) do { if (after_ctl) launch_pdl_ctl(k_predictor_fast<T, TAU>, grid, dim3(FBT), s, PF_ARGS(T)); \
                        else launch_pdl(k_predictor_fast<T, TAU>, grid, dim3(FBT), s, PF_ARGS(T)); } while (0)
    static const int lean_on = getenv("B200CFD_PRED_LEAN") ? atoi(getenv("B200CFD_PRED_LEAN")) : 1;   // developer knob (A/B)
    if (dtype == 0) { if (tau) PF(double, true); else PF(double, false); }
    else { if (tau) PF(float, true); else PF(float, false); }
#undef PF
#undef PF_ARGS
}

// ------------------------------------------------------------------------------------------------
// K2 fast: D = div(u*, v*) and the ONE red-black iteration from phi = 0 ("as executed")      v5.py:178-187,202-226
// Marching thread, 2 columns: keeps D of rows i-1, i, i+1 (computed once each).  An even cell needs phi of its four
// odd neighbours = D there * dinv / dt_c; three of them are in registers, the fourth (west of .x or east of .y,
// alternating with the row parity) belongs to the neighbouring thread and is re-derived from 3 extra loads.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ double div_raw(const Geom& g, double ue, double uw, double vn, double vs_) {
    return Rnd<T>::r((ue - uw) * g.hdx + (vn - vs_) * g.hdy);
}
// phi of an odd cell after the first colour: (a*0 + b*0 + D/dt_c) d, written exactly as the general kernel does
template <typename T>
__device__ __forceinline__ double phi_odd(const Geom& g, double D, double inv_dt) {
    const double rhs = -D * inv_dt;
    return Rnd<T>::r((g.ax * 0.0 + g.ay * 0.0 - rhs) * g.dinv);
}

template <typename T, bool DIV>
__global__ void __launch_bounds__(FBT, 6)
k_div_rb1_fast(Geom g, const T* __restrict__ us, const T* __restrict__ vs, const uint8_t* __restrict__ mask,
               T* __restrict__ phi, T* __restrict__ div_out, double inv_dt, StepCtl* __restrict__ ctl,
               const uint8_t* __restrict__ flags, int ntx, int rpb) {
    const int r0 = blockIdx.y * rpb, r1 = min(r0 + rpb, g.ny);
    const int cb = blockIdx.x * FCOLS;
    const int j0 = cb + 2 * threadIdx.x;
    const bool has_mask = mask && block_tile_flagged(flags, ntx, cb, min(cb + FCOLS, g.nx), r0, r1, 2);
    double dabs = 0.0, ch = 0.0;
    const int lo = max(r0, 1), hi = min(r1, g.ny - 1);
    if (j0 < g.nx) {
        if (r0 == 0) { st2<T>(phi, j0, 0.0, 0.0); if (DIV) st2<T>(div_out, j0, 0.0, 0.0); }
        if (r1 == g.ny) {
            const long long kk = (long long)(g.ny - 1) * g.pitch + j0;
            st2<T>(phi, kk, 0.0, 0.0); if (DIV) st2<T>(div_out, kk, 0.0, 0.0);
        }
    }
    if (j0 < g.nx && lo < hi) {
        // interior columns (local and global); border columns keep D = 0, phi = 0
        const bool xin = col_class(g, j0) == COL_INTERIOR, yin = col_class(g, j0 + 1) == COL_INTERIOR;
        // the foreign odd neighbour exists and is interior: west of .x is column j0-1, east of .y is column j0+2
        const bool fw = xin && j0 - 1 >= 1 && g.jg0 + j0 - 1 >= 1;
        const bool fe = yin && j0 + 2 < g.nx - 1 && g.jg0 + j0 + 2 < g.nxg - 1;
        const bool ownx = j0 >= g.own0 && j0 < g.own1, owny = j0 + 1 >= g.own0 && j0 + 1 < g.own1;
        long long k = (long long)lo * g.pitch + j0;
        // D of one row of my two columns from u* of that row (pair + the two x-neighbours) and v* above / below
        auto div_row = [&](long long kr, const D2& vN, const D2& vS) {
            const D2 uc = ld2<T>(us, kr);
            D2 d = {0.0, 0.0};
            if (xin) d.x = div_raw<T>(g, uc.y, ld(us, kr - 1), vN.x, vS.x);
            if (yin) d.y = div_raw<T>(g, ld(us, kr + 2), uc.x, vN.y, vS.y);
            return d;
        };
        D2 vm = ld2<T>(vs, k - g.pitch), vc = ld2<T>(vs, k), vn = ld2<T>(vs, k + g.pitch);
        D2 Dm = {0.0, 0.0};
        if (lo - 1 >= 1) Dm = div_row(k - g.pitch, vc, ld2<T>(vs, k - 2 * g.pitch));
        D2 Dc = div_row(k, vn, vm);
        uchar2 mm = make_uchar2(0, 0), mc = mm, mn = mm;
        if (has_mask) {
            mm = *reinterpret_cast<const uchar2*>(mask + k - g.pitch);
            mc = *reinterpret_cast<const uchar2*>(mask + k);
        }
        for (int i = lo; i < hi; ++i, k += g.pitch) {
            if (i + FPF < g.ny) { pf_l2(us + k + FPF * g.pitch); pf_l2(vs + k + FPF * g.pitch); }
            D2 Dn = {0.0, 0.0}, vnn = {0.0, 0.0};
            if (i + 1 < g.ny - 1) {                        // D of row i+1 (0 if that is the border row)
                vnn = ld2<T>(vs, k + 2 * g.pitch);
                Dn = div_row(k + g.pitch, vnn, vc);
            }
            if (has_mask) mn = *reinterpret_cast<const uchar2*>(mask + k + g.pitch);
            // which of my two cells is even ((i + jg) even)?  .x if (i + jg0 + j0) is even, else .y
            const bool x_even = ((i + g.jg0 + j0) & 1) == 0;
            double pf = 0.0;                               // phi of the foreign odd neighbour (0: border or masked)
            if (x_even ? fw : fe) {
                const long long kf = x_even ? k - 1 : k + 2;
                const double Df = div_raw<T>(g, ld(us, kf + 1), ld(us, kf - 1), ld(vs, kf + g.pitch), ld(vs, kf - g.pitch));
                if (!(has_mask && mask[kf])) pf = phi_odd<T>(g, Df, inv_dt);
            }
            double px = 0.0, py = 0.0;
            if (x_even) {
                if (yin && !mc.y) py = phi_odd<T>(g, Dc.y, inv_dt);
                if (xin && !mc.x) {
                    const double pn = mn.x ? 0.0 : phi_odd<T>(g, Dn.x, inv_dt), ps = mm.x ? 0.0 : phi_odd<T>(g, Dm.x, inv_dt);
                    px = (g.ax * (py + pf) + g.ay * (pn + ps) - (-Dc.x * inv_dt)) * g.dinv;
                }
            } else {
                if (xin && !mc.x) px = phi_odd<T>(g, Dc.x, inv_dt);
                if (yin && !mc.y) {
                    const double pn = mn.y ? 0.0 : phi_odd<T>(g, Dn.y, inv_dt), ps = mm.y ? 0.0 : phi_odd<T>(g, Dm.y, inv_dt);
                    py = (g.ax * (pf + px) + g.ay * (pn + ps) - (-Dc.y * inv_dt)) * g.dinv;
                }
            }
            st2<T>(phi, k, px, py);
            if (DIV) st2<T>(div_out, k, Dc.x, Dc.y);
            if (ownx) { dabs = fmax(dabs, fabs(Dc.x)); ch = fmax(ch, fabs(px)); }
            if (owny) { dabs = fmax(dabs, fabs(Dc.y)); ch = fmax(ch, fabs(py)); }
            Dm = Dc; Dc = Dn; vm = vc; vc = vn; vn = vnn; mm = mc; mc = mn;
        }
    }
    if (ctl) {
        block_atomic_max(dabs, &ctl->max_div_pre);
        block_atomic_max(ch, &ctl->pois_change);
    }
}

// ------------------------------------------------------------------------------------------------
// K5 fast (+ optionally the last clean-divergence correction fused in front of it):
//   [u,v -= grad(psi)  v5.py:254-256]  ->  max|div| (:421-422) -> BC (:424) -> IBM (:425) -> nanmax|vorticity| (:427-428)
//   -> nanmean energy (:431-432) -> clip (:437-438) -> post-clip max|u|,|v|.
// One warp owns a 64-column window (2 columns per lane) and marches down its rows; lanes 0 and 31 are halo lanes (their
// velocities after the pending correction are only needed as x-neighbours, exchanged with shuffles), lanes 1..30 own 60
// columns.  Rows i-1, i, i+1 of the corrected velocity and of psi live in registers.  Warps whose tile is flagged
// (ibm > 0 or masked cells nearby) and the lanes that hold border columns run the general per-cell code.
// ------------------------------------------------------------------------------------------------
#define FW_OWN 60                 // columns owned by one warp
#define FB_OWN (4 * FW_OWN)       // columns owned by one block (4 warps)

template <typename T> __device__ __forceinline__ D2 ld2g(const T* __restrict__ p, long long row_off, int c, bool inb) {
    return inb ? ld2<T>(p, row_off + c) : D2{0.0, 0.0};
}
__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double shfl_dn1(double v) { return __shfl_down_sync(0xffffffffu, v, 1); }

// one row of the clean-divergence potentials a lane holds: p = the LAST iteration's psi, q = the first iteration's psi_1
// (NPSI == 2: both corrections are pending, the sweep of the second iteration applied the first one only on its load side)
struct PsiRow { D2 p, q; };

// running maximum of |x|; a NaN never wins (the nanmax of the reference, like fmax) — one compare and a select instead
// of the NaN-quieting sequence fmax() expands to
__device__ __forceinline__ void max_abs(double& acc, double x) { const double a = fabs(x); acc = a > acc ? a : acc; }

// The lean rows of k_finalize_fast: the whole 64-column window of the warp lies in interior columns, no flag tile over it
// carries ibm / mask bits, rows [lo, hi) and the three rows either side of them are interior rows, hi - lo is a multiple of
// 4 and each lane's two columns are owned together or not at all.  Then BC, IBM and mask do nothing, every guard of the
// general rows is constant, and what is left is written for the issue slots (the ncu source page of the general loop: 614
// instructions per lane-row, 150 of them register-rotation moves, 83 fp64 operations):
//   * rows rotate through register arrays indexed by (row & 3) in a loop unrolled by 4, so rotation is renaming;
//   * reductions run unweighted per lane and the lane's ownership is applied once after the loop;
//   * clip / NaN bookkeeping (v5.py:431-438) is one warp vote per row: only a row that holds a value outside
//     [-vmax, vmax] or a NaN takes the exact per-cell code;
//   * one prefetch instruction per row covers the four input arrays (lane -> array, 128-byte line).
// Same operations in the same order per cell as the general rows (shared helpers); the two paths agree bit for bit in
// tests/test_gpu_lean_paths.py.
template <typename T, int NPSI>
__device__ __forceinline__ void finalize_rows_lean(const Geom& g, const double vmax, const T* __restrict__ ua, const T* __restrict__ va,
                                                   const T* __restrict__ psi, const T* __restrict__ psi1, T* __restrict__ u,
                                                   T* __restrict__ v, const int lo, const int hi, const int c0, const int lane,
                                                   const bool own, FinAcc& acc) {
    constexpr bool PSI = NPSI > 0;
    const bool owner = (unsigned)(lane - 1) < 30u;                   // every column of the window is in bounds here
    const double i2dx = 1.0 / Rnd<T>::r(g.two_dx), i2dy = 1.0 / Rnd<T>::r(g.two_dy);
    const double hdx = g.hdx, hdy = g.hdy;
    const long long pitch = g.pitch;
    long long ro = (long long)lo * pitch + c0;                       // element offset of (row i, this lane's first column)
    PsiRow P[4];                                                      // psi rows, slot = (row - lo) & 3
    D2 SU[2], SV[2], FU[4], FV[4];                                    // stored rows (slot (row - lo) & 1), corrected rows (& 3)
    auto ldpsi = [&](PsiRow& r, long long off) {
        if (PSI) r.p = ld2<T>(psi, off);
        if (NPSI == 2) r.q = ld2<T>(psi1, off);
    };
    auto apply = [&](const D2& ur, const D2& vr, const PsiRow& pS, const PsiRow& pC, const PsiRow& pN, D2& fu, D2& fv) {
        fu = ur; fv = vr;
        if (PSI) {
            if (NPSI == 2) {
                const double qW = shfl_up1(pC.q.y), qE = shfl_dn1(pC.q.x);
                fu.x = Rnd<T>::r(fu.x - Rnd<T>::r((pC.q.y - qW) * hdx));
                fv.x = Rnd<T>::r(fv.x - Rnd<T>::r((pN.q.x - pS.q.x) * hdy));
                fu.y = Rnd<T>::r(fu.y - Rnd<T>::r((qE - pC.q.x) * hdx));
                fv.y = Rnd<T>::r(fv.y - Rnd<T>::r((pN.q.y - pS.q.y) * hdy));
            }
            const double pW = shfl_up1(pC.p.y), pE = shfl_dn1(pC.p.x);
            fu.x = Rnd<T>::r(fu.x - Rnd<T>::r((pC.p.y - pW) * hdx));
            fv.x = Rnd<T>::r(fv.x - Rnd<T>::r((pN.p.x - pS.p.x) * hdy));
            fu.y = Rnd<T>::r(fu.y - Rnd<T>::r((pE - pC.p.x) * hdx));
            fv.y = Rnd<T>::r(fv.y - Rnd<T>::r((pN.p.y - pS.p.y) * hdy));
        }
    };
    {   // rows lo-1 and lo
        PsiRow pmm, pm;
        ldpsi(pmm, ro - 2 * pitch); ldpsi(pm, ro - pitch); ldpsi(P[0], ro); ldpsi(P[1], ro + pitch);
        const D2 u1 = ld2<T>(ua, ro - pitch), v1 = ld2<T>(va, ro - pitch), u0 = ld2<T>(ua, ro), v0 = ld2<T>(va, ro);
        apply(u1, v1, pmm, pm, P[0], FU[3], FV[3]);
        apply(u0, v0, pm, P[0], P[1], FU[0], FV[0]);
    }
    SU[1] = ld2<T>(ua, ro + pitch); SV[1] = ld2<T>(va, ro + pitch);
    ldpsi(P[2], ro + 2 * pitch);
    if (FIN_DEEP) {      // two rows of loads in flight per warp: an L2 hit takes longer than one row of this loop
        SU[0] = ld2<T>(ua, ro + 2 * pitch); SV[0] = ld2<T>(va, ro + 2 * pitch);
        ldpsi(P[3], ro + 3 * pitch);
    }
    // prefetch: lane -> (array, 128-byte line of the window's row); 6 lines cover the 64 columns at any alignment
    const T* pf;
    {
        const int a = min(lane / 6, NPSI + 1), line = lane % 6;
        const T* base = a == 0 ? ua : a == 1 ? va : a == 2 ? psi : psi1;
        const int col = min(c0 - 2 * lane + line * (int)(128 / sizeof(T)), g.nx - 1);
        pf = base + (long long)(lo + FPF) * pitch + col;
    }
    double dmax = 0.0, wmax = 0.0, e = 0.0, mu = 0.0, mv = 0.0;
    unsigned cnt = 0, bad = 0;
    for (int i4 = lo; i4 < hi; i4 += 4) {
        const bool more = i4 + 4 < hi;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (!FIN_DEEP && (k < 3 || more)) {                       // stored row i+2, psi row i+3
                SU[k & 1] = ld2<T>(ua, ro + 2 * pitch); SV[k & 1] = ld2<T>(va, ro + 2 * pitch);
                ldpsi(P[(k + 3) & 3], ro + 3 * pitch);
            }
            if (i4 + k + FPF < g.ny) pf_l2(pf);
            pf += pitch;
            apply(SU[(k + 1) & 1], SV[(k + 1) & 1], P[k], P[(k + 1) & 3], P[(k + 2) & 3], FU[(k + 1) & 3], FV[(k + 1) & 3]);
            if (FIN_DEEP && (k < 2 || more)) {                        // stored row i+3, psi row i+4 into the slots apply() just freed
                SU[(k + 1) & 1] = ld2<T>(ua, ro + 3 * pitch); SV[(k + 1) & 1] = ld2<T>(va, ro + 3 * pitch);
                ldpsi(P[k], ro + 4 * pitch);
            }
            const D2 fuc = FU[k], fvc = FV[k], fum = FU[(k + 3) & 3], fvm = FV[(k + 3) & 3], fun = FU[(k + 1) & 3], fvn = FV[(k + 1) & 3];
            const double fuW = shfl_up1(fuc.y), fuE = shfl_dn1(fuc.x), fvW = shfl_up1(fvc.y), fvE = shfl_dn1(fvc.x);
            // The three diagnostics only feed max / mean reductions that are logged and compared at 1e-5 in float32 mode, never
            // the fields: in float32 storage their intermediate float32 roundings (26 of the 42 emulated roundings of a
            // lane-row, two XU conversions each) are skipped here (RD = identity); the general rows keep them.  fp64: no change.
            auto RD = [](double x) { return x; };
            const double Dx = RD((fuc.y - fuW) * hdx + (fvn.x - fvm.x) * hdy);
            const double Dy = RD((fuE - fuc.x) * hdx + (fvn.y - fvm.y) * hdy);
            max_abs(dmax, Dx); max_abs(dmax, Dy);
            const double Wx = RD(RD(RD(fvc.y - fvW) * i2dx) - RD(RD(fun.x - fum.x) * i2dy));
            const double Wy = RD(RD(RD(fvE - fvc.x) * i2dx) - RD(RD(fun.y - fum.y) * i2dy));
            max_abs(wmax, Wx); max_abs(wmax, Wy);
            const double ex = RD(0.5 * RD(RD(fuc.x * fuc.x) + RD(fvc.x * fvc.x)));
            const double ey = RD(0.5 * RD(RD(fuc.y * fuc.y) + RD(fvc.y * fvc.y)));
            const bool odd = !(fabs(fuc.x) <= vmax) || !(fabs(fuc.y) <= vmax) || !(fabs(fvc.x) <= vmax) || !(fabs(fvc.y) <= vmax);
            if (!__any_sync(0xffffffffu, odd)) {
                if (owner) { st2<T>(u, ro, fuc.x, fuc.y); st2<T>(v, ro, fvc.x, fvc.y); }
                e += ex; e += ey; cnt += 2;
                max_abs(mu, fuc.x); max_abs(mu, fuc.y); max_abs(mv, fvc.x); max_abs(mv, fvc.y);
            } else {
                const double cux = fmin(fmax(fuc.x, -vmax), vmax), cuy = fmin(fmax(fuc.y, -vmax), vmax);
                const double cvx = fmin(fmax(fvc.x, -vmax), vmax), cvy = fmin(fmax(fvc.y, -vmax), vmax);
                if (owner) {     // np.clip propagates NaN; fmin/fmax would drop it
                    st2<T>(u, ro, fuc.x == fuc.x ? cux : fuc.x, fuc.y == fuc.y ? cuy : fuc.y);
                    st2<T>(v, ro, fvc.x == fvc.x ? cvx : fvc.x, fvc.y == fvc.y ? cvy : fvc.y);
                }
                if (ex == ex) { e += ex; cnt += 1; }
                if (ey == ey) { e += ey; cnt += 1; }
                bad |= (!isfinite(fuc.x) || !isfinite(fvc.x) || !isfinite(fuc.y) || !isfinite(fvc.y)) ? 1u : 0u;
                max_abs(mu, cux); max_abs(mu, cuy); max_abs(mv, cvx); max_abs(mv, cvy);
            }
            ro += pitch;
        }
    }
    if (own) { acc.dmax = dmax; acc.wmax = wmax; acc.e = e; acc.cnt = cnt; acc.bad = bad; acc.mu = mu; acc.mv = mv; }
}

template <typename T, int NPSI>
__global__ void __launch_bounds__(FBT, FIN_MINB)
k_finalize_fast(Geom g, Phys p, StepCtl* __restrict__ ctl, const T* __restrict__ ua, const T* __restrict__ va,
                const T* __restrict__ psi, const T* __restrict__ psi1, T* __restrict__ u, T* __restrict__ v,
                const double* __restrict__ ibm, const uint8_t* __restrict__ mask, const double* __restrict__ y,
                double* __restrict__ ke_part, unsigned long long* __restrict__ cnt_part, const uint8_t* __restrict__ flags,
                int ntx, int rpb, int bx_cut, int bx_off_a, int bx_off_b, int nbx_total, int lean_on) {
    constexpr bool PSI = NPSI > 0;
    pdl_wait(); pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r0 = blockIdx.y * rpb, r1 = min(r0 + rpb, g.ny);
    // column block of this CUDA block: a launch may cover a subset of the column blocks (x-slab edge / interior split)
    const int bx = (int)blockIdx.x < bx_cut ? (int)blockIdx.x + bx_off_a : (int)blockIdx.x + bx_off_b;
    const int own_c0 = bx * FB_OWN + warp * FW_OWN;                // first column owned by this warp
    const int c0 = own_c0 - 2 + 2 * lane;                          // this lane's two columns: c0, c0+1
    const bool inb = c0 >= 0 && c0 + 1 < g.nx;
    const bool owner = inb && lane >= 1 && lane <= 30;
    FinAcc acc;
    if (own_c0 < g.nx) {
        // warp-uniform: do the flag tiles over the owned columns (they are dilated by 2 cells) carry ibm / mask bits?
        int f = 0;
        {
            const int cB = min(own_c0 + FW_OWN, g.nx);
            const int tx0 = own_c0 / FT_W, tx1 = (cB - 1) / FT_W, ty0 = r0 / FT_H, ty1 = (r1 - 1) / FT_H;
            const int w = tx1 - tx0 + 1, n = w * (ty1 - ty0 + 1);
            for (int t = lane; t < n; t += 32) f |= flags[(ty0 + t / w) * ntx + tx0 + t % w];
        }
        const bool has_ibm = __any_sync(0xffffffffu, f & 1), has_mask = mask && __any_sync(0xffffffffu, f & 2);
        const int cx = inb ? col_class(g, c0) : COL_COPY, cy = inb ? col_class(g, c0 + 1) : COL_COPY;
        const bool inx = inb && cx == COL_INTERIOR, iny = inb && cy == COL_INTERIOR;
        const bool ownx = owner && c0 >= g.own0 && c0 < g.own1, owny = owner && c0 + 1 >= g.own0 && c0 + 1 < g.own1;
        const double i2dx = 1.0 / Rnd<T>::r(g.two_dx), i2dy = 1.0 / Rnd<T>::r(g.two_dy);
        const double vmax = p.max_velocity, fs = ctl->fs, pert = ctl->pert_scale, phase = ctl->phase;
        // reduction weights: interior cells of the owned columns (divergence, vorticity) / owned columns (energy, maxima)
        const double wdx = (inx && ownx) ? 1.0 : 0.0, wdy = (iny && owny) ? 1.0 : 0.0, wox = ownx ? 1.0 : 0.0, woy = owny ? 1.0 : 0.0;
        const int lo = max(r0, 1), hi = min(r1, g.ny - 1);
        // the lean rows (see finalize_rows_lean): a warp-uniform choice
        const bool lean = FIN_LEAN && lean_on && !has_ibm && !has_mask && r0 >= 2 && r1 + 3 <= g.ny && ((r1 - r0) & 3) == 0 &&
                          __all_sync(0xffffffffu, inx && iny && ownx == owny);
        if (lean) {
            finalize_rows_lean<T, NPSI>(g, p.max_velocity, ua, va, psi, psi1, u, v, lo, hi, c0, lane, ownx, acc);
        } else {
        // wall rows: u = v = 0 (v5.py:357-360); they count in the energy mean
        if (owner && r0 == 0) { st2<T>(u, c0, 0.0, 0.0); st2<T>(v, c0, 0.0, 0.0); acc.cnt += (ownx ? 1 : 0) + (owny ? 1 : 0); }
        if (owner && r1 == g.ny) {
            const long long kk = (long long)(g.ny - 1) * g.pitch + c0;
            st2<T>(u, kk, 0.0, 0.0); st2<T>(v, kk, 0.0, 0.0); acc.cnt += (ownx ? 1 : 0) + (owny ? 1 : 0);
        }
        const PsiRow PZ = {{0, 0}, {0, 0}};
        auto ldpsi = [&](long long row_off) {
            PsiRow r = PZ;
            if (PSI) r.p = ld2g<T>(psi, row_off, c0, inb);
            if (NPSI == 2) r.q = ld2g<T>(psi1, row_off, c0, inb);
            return r;
        };
        // velocity of one row after the pending correction(s), from its stored values and three psi rows (all lanes:
        // the x-neighbours of psi come from shuffles).  NPSI == 2: first u,v -= grad(psi_1), then -= grad(psi): the two
        // roundings of the reference's two in-place updates (v5.py:254-256, executed twice).
        auto apply_row = [&](int row, const D2& ur, const D2& vr, const PsiRow& pS, const PsiRow& pC, const PsiRow& pN, D2& fu, D2& fv) {
            fu = ur; fv = vr;
            if (PSI) {
                const double pW = shfl_up1(pC.p.y), pE = shfl_dn1(pC.p.x);
                double qW = 0.0, qE = 0.0;
                if (NPSI == 2) { qW = shfl_up1(pC.q.y); qE = shfl_dn1(pC.q.x); }
                if (row >= 1 && row < g.ny - 1) {
                    if (inx) {
                        if (NPSI == 2) {
                            fu.x = Rnd<T>::r(fu.x - Rnd<T>::r((pC.q.y - qW) * g.hdx));
                            fv.x = Rnd<T>::r(fv.x - Rnd<T>::r((pN.q.x - pS.q.x) * g.hdy));
                        }
                        fu.x = Rnd<T>::r(fu.x - Rnd<T>::r((pC.p.y - pW) * g.hdx));
                        fv.x = Rnd<T>::r(fv.x - Rnd<T>::r((pN.p.x - pS.p.x) * g.hdy));
                    }
                    if (iny) {
                        if (NPSI == 2) {
                            fu.y = Rnd<T>::r(fu.y - Rnd<T>::r((qE - pC.q.x) * g.hdx));
                            fv.y = Rnd<T>::r(fv.y - Rnd<T>::r((pN.q.y - pS.q.y) * g.hdy));
                        }
                        fu.y = Rnd<T>::r(fu.y - Rnd<T>::r((pE - pC.p.x) * g.hdx));
                        fv.y = Rnd<T>::r(fv.y - Rnd<T>::r((pN.p.y - pS.p.y) * g.hdy));
                    }
                }
            }
        };
        // IBM multiplier 1 - ibm*fs of one row (1 where ibm <= 0; exact: x*1.0 == x)
        auto mult_row = [&](long long ro_) {
            D2 m = {1.0, 1.0};
            if (has_ibm && inb) {
                const double2 t = *reinterpret_cast<const double2*>(ibm + ro_ + c0);
                if (t.x > 0) m.x = 1.0 - t.x * fs;
                if (t.y > 0) m.y = 1.0 - t.y * fs;
            }
            return m;
        };
        if (lo < hi) {
            long long ro = (long long)lo * g.pitch;
            PsiRow pm = PZ, pc = PZ, pn = PZ;             // psi rows (i), (i+1), (i+2) while row i is processed
            D2 gum = {0, 0}, fvm, fuc, fvc, mc;
            {   // rows lo-1 and lo
                PsiRow pmm = PZ;
                D2 fum;
                if (PSI) {
                    if (lo - 2 >= 0) pmm = ldpsi(ro - 2 * g.pitch);
                    pm = ldpsi(ro - g.pitch);
                    pc = ldpsi(ro);
                    pn = ldpsi(ro + g.pitch);
                }
                apply_row(lo - 1, ld2g<T>(ua, ro - g.pitch, c0, inb), ld2g<T>(va, ro - g.pitch, c0, inb), pmm, pm, pc, fum, fvm);
                apply_row(lo, ld2g<T>(ua, ro, c0, inb), ld2g<T>(va, ro, c0, inb), pm, pc, pn, fuc, fvc);
                pm = pc; pc = pn;      // now pm = psi(lo), pc = psi(lo+1)
                if (lo - 1 >= 1) {     // processed u of row lo-1 (0 on the wall row); only interior columns use it
                    const D2 mm = mult_row(ro - g.pitch);
                    gum.x = Rnd<T>::r(fum.x * mm.x); gum.y = Rnd<T>::r(fum.y * mm.y);
                }
                mc = mult_row(ro);
            }
            // software pipeline: stored values of row i+1 (and psi row i+2) are loaded one iteration ahead
            D2 urn = ld2g<T>(ua, ro + g.pitch, c0, inb), vrn = ld2g<T>(va, ro + g.pitch, c0, inb);
            if (PSI) pn = (lo + 2 < g.ny) ? ldpsi(ro + 2 * g.pitch) : PZ;
            D2 mn = (lo + 1 < g.ny - 1) ? mult_row(ro + g.pitch) : D2{1.0, 1.0};
            MARCH_UNROLL
            for (int i = lo; i < hi; ++i, ro += g.pitch) {
                // issue the loads of the NEXT iteration: stored row i+2, psi row i+3, ibm row i+2
                D2 urn2 = {0, 0}, vrn2 = {0, 0}, mn2 = {1.0, 1.0};
                PsiRow pn2 = PZ;
                if (i + 1 < hi) {
                    urn2 = ld2g<T>(ua, ro + 2 * g.pitch, c0, inb); vrn2 = ld2g<T>(va, ro + 2 * g.pitch, c0, inb);
                    if (PSI && i + 3 < g.ny) pn2 = ldpsi(ro + 3 * g.pitch);
                    if (i + 2 < g.ny - 1) mn2 = mult_row(ro + 2 * g.pitch);
                }
                if (inb && i + FPF < g.ny) {
                    const long long kp = ro + FPF * g.pitch + c0;
                    pf_l2(ua + kp); pf_l2(va + kp);
                    if (PSI) pf_l2(psi + kp);
                    if (NPSI == 2) pf_l2(psi1 + kp);
                }
                D2 fun, fvn;
                apply_row(i + 1, urn, vrn, pm, pc, pn, fun, fvn);
                // BC (v5.py:349-360) + IBM (v5.py:228-237) of row i, pointwise
                D2 guc = fuc, gvc = fvc;
                if (cx == COL_INFLOW) { guc.x = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[i], pert, phase)); gvc.x = 0.0; }
                if (cy == COL_OUTFLOW) { guc.y = fuc.x; gvc.y = fvc.x; }     // copies column nx-2 (before its IBM)
                if (has_ibm) {
                    guc.x = Rnd<T>::r(guc.x * mc.x); gvc.x = Rnd<T>::r(gvc.x * mc.x);
                    guc.y = Rnd<T>::r(guc.y * mc.y); gvc.y = Rnd<T>::r(gvc.y * mc.y);
                }
                // processed u of row i+1 (north neighbour of the vorticity stencil): 0 on the wall row
                D2 gun = {0, 0};
                if (i + 1 < g.ny - 1) { gun.x = Rnd<T>::r(fun.x * mn.x); gun.y = Rnd<T>::r(fun.y * mn.y); }
                const double fuW = shfl_up1(fuc.y), fuE = shfl_dn1(fuc.x), gvW = shfl_up1(gvc.y), gvE = shfl_dn1(gvc.x);
                // Branch-free from here on (the ncu source page of the branchy form: 15 % of the issue slots on BSSY / BSYNC /
                // BRA, 6 % on ISETP): every lane does the arithmetic, ownership enters the reductions as 0 / 1 weights (a NaN or
                // Inf times 0 is NaN, which fmax drops exactly like the nanmax it replaces), only the stores are predicated.
                uchar2 mk = make_uchar2(0, 0);
                if (has_mask && owner) mk = *reinterpret_cast<const uchar2*>(mask + ro + c0);
                const double Dx = Rnd<T>::r((fuc.y - fuW) * g.hdx + (fvn.x - fvm.x) * g.hdy);
                const double Dy = Rnd<T>::r((fuE - fuc.x) * g.hdx + (fvn.y - fvm.y) * g.hdy);
                acc.dmax = fmax(acc.dmax, fmax(fabs(Dx) * wdx, fabs(Dy) * wdy));
                const double Wx = Rnd<T>::r(Rnd<T>::r(Rnd<T>::r(gvc.y - gvW) * i2dx) - Rnd<T>::r(Rnd<T>::r(gun.x - gum.x) * i2dy));
                const double Wy = Rnd<T>::r(Rnd<T>::r(Rnd<T>::r(gvE - gvc.x) * i2dx) - Rnd<T>::r(Rnd<T>::r(gun.y - gum.y) * i2dy));
                acc.wmax = fmax(acc.wmax, fmax(fabs(Wx) * (mk.x ? 0.0 : wdx), fabs(Wy) * (mk.y ? 0.0 : wdy)));
                const double ex = Rnd<T>::r(0.5 * Rnd<T>::r(Rnd<T>::r(guc.x * guc.x) + Rnd<T>::r(gvc.x * gvc.x)));
                const double ey = Rnd<T>::r(0.5 * Rnd<T>::r(Rnd<T>::r(guc.y * guc.y) + Rnd<T>::r(gvc.y * gvc.y)));
                const double cux = fmin(fmax(guc.x, -vmax), vmax), cuy = fmin(fmax(guc.y, -vmax), vmax);
                const double cvx = fmin(fmax(gvc.x, -vmax), vmax), cvy = fmin(fmax(gvc.y, -vmax), vmax);
                if (owner) {
                    // np.clip propagates NaN; fmin/fmax would drop it
                    st2<T>(u, ro + c0, guc.x == guc.x ? cux : guc.x, guc.y == guc.y ? cuy : guc.y);
                    st2<T>(v, ro + c0, gvc.x == gvc.x ? cvx : gvc.x, gvc.y == gvc.y ? cvy : gvc.y);
                }
                const bool okx = ownx && ex == ex, oky = owny && ey == ey;           // nanmean: NaN cells are not counted
                acc.e += okx ? ex : 0.0; acc.cnt += okx ? 1u : 0u;
                acc.e += oky ? ey : 0.0; acc.cnt += oky ? 1u : 0u;
                acc.bad |= (ownx && (!isfinite(guc.x) || !isfinite(gvc.x))) || (owny && (!isfinite(guc.y) || !isfinite(gvc.y))) ? 1u : 0u;
                acc.mu = fmax(acc.mu, fmax(fabs(cux) * wox, fabs(cuy) * woy));
                acc.mv = fmax(acc.mv, fmax(fabs(cvx) * wox, fabs(cvy) * woy));
                gum = guc; fvm = fvc; fuc = fun; fvc = fvn; mc = mn; mn = mn2;
                pm = pc; pc = pn; pn = pn2; urn = urn2; vrn = vrn2;
            }
        }
        }   // general rows
    }
    finalize_block_reduce1(acc, ctl, ke_part, cnt_part, blockIdx.y * nbx_total + bx);
}

void launch_div_rb1_fast(int dtype, const Geom& g, const void* us, const void* vs, const uint8_t* mask, void* phi,
                         void* div_out, double inv_dt, StepCtl* ctl, const uint8_t* flags, int n_sm, cudaStream_t s) {
    const int rpb = rows_per_block(g, n_sm);
    const dim3 grid((g.nx + FCOLS - 1) / FCOLS, (g.ny + rpb - 1) / rpb);
    const int ntx = (g.nx + FT_W - 1) / FT_W;
#define DF(T, DIV) k_div_rb1_fast<T, DIV><<<grid, FBT, 0, s>>>(g, (const T*)us, (const T*)vs, mask, (T*)phi, (T*)div_out, inv_dt, ctl, flags, ntx, rpb)
    if (dtype == 0) { if (div_out) DF(double, true); else DF(double, false); }
    else { if (div_out) DF(float, true); else DF(float, false); }
#undef DF
}

static inline dim3 finalize_fast_grid(const Geom& g, int n_sm, int* rpb_out) {
    const int bx = (g.nx + FB_OWN - 1) / FB_OWN;
    static const int rpb_min = getenv("B200CFD_FIN_RPB_MIN") ? atoi(getenv("B200CFD_FIN_RPB_MIN")) : 16;   // developer knob
    int rpb = 64;
    while (rpb > rpb_min && (long long)bx * ((g.ny + rpb - 1) / rpb) < 8ll * n_sm) rpb >>= 1;
    rpb = rpb_12(bx, g.ny, n_sm, rpb);
    // small grids (the shipped 600x180 case): the kernel is pure latency, so trade redundant halo rows for parallelism
    while (rpb > 4 && (long long)bx * ((g.ny + rpb - 1) / rpb) < 2ll * n_sm) rpb >>= 1;
    rpb = pick_rpb(bx, g.ny, 3 * n_sm, 3, rpb);              // k_finalize_fast: 3 blocks per SM
    static const int force = getenv("B200CFD_FIN_RPB") ? atoi(getenv("B200CFD_FIN_RPB")) : 0;      // developer knob
    if (force > 0) rpb = force;
    *rpb_out = rpb;
    return dim3(bx, (g.ny + rpb - 1) / rpb);
}
int finalize_fast_num_partials(const Geom& g, int n_sm) {
    int rpb;
    const dim3 gr = finalize_fast_grid(g, n_sm, &rpb);
    return (int)(gr.x * gr.y);
}
// psi != NULL: the pending correction u,v -= grad(psi) of the last clean-divergence iteration is applied on the fly;
// psi1 != NULL as well: the first iteration's correction u,v -= grad(psi1) is pending too (applied first).
void launch_finalize_fast(int dtype, const Geom& g, const Phys& p, StepCtl* ctl, const void* ua, const void* va,
                          const void* psi, const void* psi1, void* u, void* v, const double* ibm, const uint8_t* mask,
                          const double* y, double* ke_partials, unsigned long long* cnt_partials, const uint8_t* flags,
                          int n_sm, cudaStream_t s, int part, int c_left, int c_right) {
    int rpb;
    dim3 grid = finalize_fast_grid(g, n_sm, &rpb);
    const int ntx = (g.nx + FT_W - 1) / FT_W;
    const int np = psi ? (psi1 ? 2 : 1) : 0;
    const int nbx = (int)grid.x;
    // column blocks [0, bl) overlap [0, c_left); column blocks [br, nbx) overlap [c_right, nx)
    int bl = part ? (c_left + FB_OWN - 1) / FB_OWN : 0, br = part ? c_right / FB_OWN : nbx;
    if (bl > nbx) bl = nbx;
    if (br < bl) br = bl;
    int cut = nbx, off_a = 0, off_b = 0;
    if (part == 1) { grid.x = bl + (nbx - br); cut = bl; off_a = 0; off_b = br - bl; }        // edges: [0, bl) and [br, nbx)
    else if (part == 2) { grid.x = br - bl; cut = nbx; off_a = bl; }                            // interior: [bl, br)
    if (grid.x == 0) return;
#define FF(T, NP) launch_pdl(k_finalize_fast<T, NP>, grid, dim3(FBT), s, g, p, ctl, (const T*)ua, (const T*)va, (const T*)psi, (const T*)psi1, (T*)u, (T*)v, ibm, mask, y, ke_partials, cnt_partials, flags, ntx, rpb, cut, off_a, off_b, nbx, lean_on)
    static const int lean_on = getenv("B200CFD_FIN_LEAN") ? atoi(getenv("B200CFD_FIN_LEAN")) : 1;   // developer knob (A/B)
    if (dtype == 0) { if (np == 2) FF(double, 2); else if (np == 1) FF(double, 1); else FF(double, 0); }
    else { if (np == 2) FF(float, 2); else if (np == 1) FF(float, 1); else FF(float, 0); }
#undef FF
}

// ------------------------------------------------------------------------------------------------
// K2 + K3 fused: D = div(u*, v*), the ONE red-black iteration from phi = 0, and the velocity correction
//   ua = u* - dt (phi_E - phi_W)/(2dx),  va = v* - dt (phi_N - phi_S)/(2dy),  max|grad phi|      v5.py:409-417
// in one pass: u*, v* are read once (16 B/cell) instead of twice, phi is never re-read: 41 B/cell instead of 25 + 40.
// Layout as the finalize kernel: one warp owns 60 columns of a 64-column window (2 per lane, lanes 0 / 31 are halo
// lanes whose phi is only the x-neighbour of lanes 1 / 30), and marches down its rows with D of rows i-1, i, i+1 and
// phi of rows i-2, i-1, i in registers; the correction of row i-1 is applied as soon as phi of row i exists.  A block
// of rows [r0, r1) computes phi of rows r0-1 and r1 redundantly (they belong to the neighbouring blocks).
// ------------------------------------------------------------------------------------------------
// The lean rows of k_div_rb1_correct_fast (same idea as finalize_rows_lean): the warp's whole 64-column window lies in
// interior columns, no flag tile over it carries a mask bit, rows [lo-3, hi+3) are interior rows of the array, hi - lo is a
// multiple of 4, each lane's two columns are owned together or not at all.  What the general rows spend on guards, clamps,
// per-cell selects and register rotation (365 instructions per lane-row for 48 fp64 operations) goes away:
//   * rows rotate through register arrays indexed by (row & 3) in a loop unrolled by 4, the colour of each row is a
//     compile-time constant of the unrolled body (two instances of the loop, one per colour of row lo);
//   * phi of an odd cell, (D/dt_c) d, is computed ONCE per cell: it is that cell's final value and the north / south
//     neighbour of the even cells above and below (the general rows evaluate it three times);
//   * the two halo lanes fetch their outer u* neighbour with one predicated load per row;
//   * reductions run unweighted per lane, ownership is applied once after the loop.
// Per cell the operations and their order are those of the general rows.  The bits agree too, except that the redundant phi
// row above a block is computed by the prologue's copy of the formula here and by the loop's in the general rows, and the
// compiler contracts a*x + b*y - c into FMAs differently in the two: a handful of cells of v end up one ulp apart
// (tests/test_gpu_lean_paths.py: 3 of 786 432 after 5 steps).
__device__ __forceinline__ void max_pos(double& acc, double x) { acc = x > acc ? x : acc; }

template <typename T, bool DIV>
__device__ __forceinline__ void div_rb1_correct_rows_lean(const Geom& g, const T* __restrict__ us, const T* __restrict__ vs,
                                                          T* __restrict__ phi, T* __restrict__ div_out, T* __restrict__ ua,
                                                          T* __restrict__ va, const double inv_dt, const double dt, const int lo,
                                                          const int hi, const int c0, const int lane, const bool own,
                                                          double& dabs_o, double& ch_o, double& gmax_o) {
    const bool owner = (unsigned)(lane - 1) < 30u;
    const bool halo_lane = lane == 0 || lane == 31;
    const long long pitch = g.pitch;
    const double hdx = g.hdx, hdy = g.hdy;
    long long ro = (long long)lo * pitch + c0;                       // element offset of (row j, this lane's first column)
    const T* __restrict__ hp = us + (lane == 0 ? -1 : 2);            // + ro: the outer x-neighbour of a halo lane
    auto halo = [&](long long off) { return halo_lane ? ld(hp, off) : 0.0; };
    // D of one row from its u* pair (x-neighbours by shuffle, the halo lanes' outer one from `h`) and v* of the rows N, S
    auto drow = [&](const D2& uc, const double h, const D2& vN, const D2& vS) {
        double uW = shfl_up1(uc.y), uE = shfl_dn1(uc.x);
        uW = lane == 0 ? h : uW;
        uE = lane == 31 ? h : uE;
        D2 d;
        d.x = div_raw<T>(g, uc.y, uW, vN.x, vS.x);
        d.y = div_raw<T>(g, uE, uc.x, vN.y, vS.y);
        return d;
    };
    // final phi of a row: po = phi of its odd cell, pn / ps = phi of the (odd) cells north / south of its even cell
    auto final_row = [&](const bool xe, const double po, const double pn, const double ps, const D2& Dc) {
        D2 q;
        if (xe) {
            const double pW = shfl_up1(po);
            q.y = po;
            q.x = Rnd<T>::r((g.ax * (po + pW) + g.ay * (pn + ps) - (-Dc.x * inv_dt)) * g.dinv);
        } else {
            const double pE = shfl_dn1(po);
            q.x = po;
            q.y = Rnd<T>::r((g.ax * (pE + po) + g.ay * (pn + ps) - (-Dc.y * inv_dt)) * g.dinv);
        }
        return q;
    };
    double dabs = 0.0, ch = 0.0, gmax = 0.0;
    auto run = [&](auto XE_LO) {
        constexpr bool xe_lo = decltype(XE_LO)::value;               // the .x cells of row lo are the even (black) ones
        D2 U[4], V[4], PH[4], DS[2];                                 // slots: u*, v* (row - lo) & 3; phi (row - lo + 1) & 3; D (row - lo) & 1
        double HL[2], PO[2];                                         // halo-lane neighbour, phi of the odd cell: (row - lo) & 1
        {   // rows lo-2 .. lo+1: D and the odd cells' phi; phi of rows lo-1 and lo
            const D2 vm3 = ld2<T>(vs, ro - 3 * pitch), vm2 = ld2<T>(vs, ro - 2 * pitch), vm1 = ld2<T>(vs, ro - pitch);
            V[0] = ld2<T>(vs, ro); V[1] = ld2<T>(vs, ro + pitch); V[2] = ld2<T>(vs, ro + 2 * pitch);
            const D2 um2 = ld2<T>(us, ro - 2 * pitch), um1 = ld2<T>(us, ro - pitch);
            U[0] = ld2<T>(us, ro); U[1] = ld2<T>(us, ro + pitch);
            const double hm2 = halo(ro - 2 * pitch), hm1 = halo(ro - pitch), h0 = halo(ro), h1 = halo(ro + pitch);
            const D2 Dm2 = drow(um2, hm2, vm1, vm3), Dm1 = drow(um1, hm1, V[0], vm2), D0 = drow(U[0], h0, V[1], vm1);
            DS[1] = drow(U[1], h1, V[2], V[0]);
            // row parity: lo-2 and lo as xe_lo, lo-1 and lo+1 the other colour
            const double pom2 = phi_odd<T>(g, xe_lo ? Dm2.y : Dm2.x, inv_dt), pom1 = phi_odd<T>(g, xe_lo ? Dm1.x : Dm1.y, inv_dt);
            const double po0 = phi_odd<T>(g, xe_lo ? D0.y : D0.x, inv_dt);
            PO[1] = phi_odd<T>(g, xe_lo ? DS[1].x : DS[1].y, inv_dt);
            PH[0] = final_row(!xe_lo, pom1, po0, pom2, Dm1);
            PH[1] = final_row(xe_lo, po0, PO[1], xe_lo ? PH[0].x : PH[0].y, D0);
            if (owner) {
                st2<T>(phi, ro, PH[1].x, PH[1].y);
                if (DIV) st2<T>(div_out, ro, D0.x, D0.y);
            }
            max_abs(dabs, D0.x); max_abs(dabs, D0.y); max_abs(ch, PH[1].x); max_abs(ch, PH[1].y);
        }
        U[2] = ld2<T>(us, ro + 2 * pitch); HL[0] = halo(ro + 2 * pitch); V[3] = ld2<T>(vs, ro + 3 * pitch);
        const T* pf = (lane < 6 ? us : vs) + (long long)(lo + FPF) * pitch + min(c0 - 2 * lane + (lane % 6) * (int)(128 / sizeof(T)), g.nx - 1);
        for (int j4 = lo; j4 < hi; j4 += 4) {
            const bool more = j4 + 4 < hi;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const bool xe_j1 = (k & 1) ? xe_lo : !xe_lo;         // colour of row j+1 (row j+2: the other one)
                const D2 vk = V[k];                                  // v*(j): its slot takes v*(j+4)
                if (k < 3 || more) {
                    V[k] = ld2<T>(vs, ro + 4 * pitch);
                    U[(k + 3) & 3] = ld2<T>(us, ro + 3 * pitch);
                    HL[(k + 1) & 1] = halo(ro + 3 * pitch);
                }
                if (lane < 12 && j4 + k + FPF < g.ny) pf_l2(pf);
                pf += pitch;
                // D(j+2), phi of its odd cell
                const D2 D2n = drow(U[(k + 2) & 3], HL[k & 1], V[(k + 3) & 3], V[(k + 1) & 3]);
                const double po2 = phi_odd<T>(g, xe_j1 ? D2n.x : D2n.y, inv_dt);
                // final phi(j+1)
                const D2 D1 = DS[(k + 1) & 1];
                const D2 pj = PH[(k + 1) & 3];
                const D2 pc = final_row(xe_j1, PO[(k + 1) & 1], po2, xe_j1 ? pj.x : pj.y, D1);
                PH[(k + 2) & 3] = pc; DS[k & 1] = D2n; PO[k & 1] = po2;
                if (k < 3 || more) {
                    if (owner) {
                        st2<T>(phi, ro + pitch, pc.x, pc.y);
                        if (DIV) st2<T>(div_out, ro + pitch, D1.x, D1.y);
                    }
                    max_abs(dabs, D1.x); max_abs(dabs, D1.y); max_abs(ch, pc.x); max_abs(ch, pc.y);
                }
                // correction of row j from phi rows j-1, j, j+1
                const D2 pmm = PH[k];
                const double pW = shfl_up1(pj.y), pE = shfl_dn1(pj.x);
                const double gxx = Rnd<T>::r((pj.y - pW) * hdx), gyx = Rnd<T>::r((pc.x - pmm.x) * hdy);
                const double gxy = Rnd<T>::r((pE - pj.x) * hdx), gyy = Rnd<T>::r((pc.y - pmm.y) * hdy);
                if (owner) {
                    st2<T>(ua, ro, axmy<T>(U[k].x, dt, gxx), axmy<T>(U[k].y, dt, gxy));
                    st2<T>(va, ro, axmy<T>(vk.x, dt, gyx), axmy<T>(vk.y, dt, gyy));
                }
                max_pos(gmax, Rnd<T>::r(Rnd<T>::r(gxx * gxx) + Rnd<T>::r(gyx * gyx)));
                max_pos(gmax, Rnd<T>::r(Rnd<T>::r(gxy * gxy) + Rnd<T>::r(gyy * gyy)));
                ro += pitch;
            }
        }
    };
    if (((lo + g.jg0 + c0) & 1) == 0) run(std::true_type{}); else run(std::false_type{});
    if (own) { dabs_o = dabs; ch_o = ch; gmax_o = gmax; }
}

template <typename T, bool DIV>
__global__ void __launch_bounds__(FBT, 4)
k_div_rb1_correct_fast(Geom g, const T* __restrict__ us, const T* __restrict__ vs, const uint8_t* __restrict__ mask,
                       T* __restrict__ phi, T* __restrict__ div_out, T* __restrict__ ua, T* __restrict__ va,
                       double inv_dt, StepCtl* __restrict__ ctl, const uint8_t* __restrict__ flags, int ntx, int rpb,
                       int lean_on) {
    pdl_wait(); pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r0 = blockIdx.y * rpb, r1 = min(r0 + rpb, g.ny);
    const int own_c0 = blockIdx.x * FB_OWN + warp * FW_OWN;        // first column owned by this warp
    const int c0 = own_c0 - 2 + 2 * lane;                          // this lane's two columns: c0, c0+1
    const bool inb = c0 >= 0 && c0 + 1 < g.nx;
    const bool owner = inb && lane >= 1 && lane <= 30;
    double dabs = 0.0, ch = 0.0, gmax = 0.0;
    if (own_c0 < g.nx) {
        int f = 0;
        {
            const int cB = min(own_c0 + FW_OWN, g.nx);
            const int tx0 = own_c0 / FT_W, tx1 = (cB - 1) / FT_W, ty0 = r0 / FT_H, ty1 = (r1 - 1) / FT_H;
            const int w = tx1 - tx0 + 1, n = w * (ty1 - ty0 + 1);
            for (int t = lane; t < n; t += 32) f |= flags[(ty0 + t / w) * ntx + tx0 + t % w];
        }
        const bool has_mask = mask && __any_sync(0xffffffffu, f & 2);
        const bool xin = inb && col_class(g, c0) == COL_INTERIOR, yin = inb && col_class(g, c0 + 1) == COL_INTERIOR;
        const bool ownx = owner && c0 >= g.own0 && c0 < g.own1, owny = owner && c0 + 1 >= g.own0 && c0 + 1 < g.own1;
        const double dt = ctl->dt;
        const int lo = max(r0, 1), hi = min(r1, g.ny - 1);
        const bool lean = lean_on && !has_mask && r0 >= 3 && r1 + 3 <= g.ny && ((r1 - r0) & 3) == 0 &&
                          __all_sync(0xffffffffu, xin && yin && ownx == owny);
        if (lean) {
            div_rb1_correct_rows_lean<T, DIV>(g, us, vs, phi, div_out, ua, va, inv_dt, dt, lo, hi, c0, lane, ownx, dabs, ch, gmax);
        } else {
        // wall rows: D = phi = 0, no gradient: ua = u*, va = v*
        if (owner && r0 == 0) {
            st2<T>(phi, c0, 0.0, 0.0); if (DIV) st2<T>(div_out, c0, 0.0, 0.0);
            const D2 a = ld2<T>(us, c0), b = ld2<T>(vs, c0);
            st2<T>(ua, c0, a.x, a.y); st2<T>(va, c0, b.x, b.y);
        }
        if (owner && r1 == g.ny) {
            const long long kk = (long long)(g.ny - 1) * g.pitch + c0;
            st2<T>(phi, kk, 0.0, 0.0); if (DIV) st2<T>(div_out, kk, 0.0, 0.0);
            const D2 a = ld2<T>(us, kk), b = ld2<T>(vs, kk);
            st2<T>(ua, kk, a.x, a.y); st2<T>(va, kk, b.x, b.y);
        }
        if (lo < hi) {                                             // warp-uniform
            // Guard-free accessors: rows are clamped into [0, ny-1] and a lane outside the array reads a clamped column.
            // Nothing read that way is ever used: D is forced to 0 on border rows / non-interior columns, and stores are
            // owner-only.  (Selects and zero-fills per load were a quarter of this kernel's instructions.)
            const int cs = min(max(c0, 0), g.nx - 2);
            const T* __restrict__ usp = us + cs;
            const T* __restrict__ vsp = vs + cs;
            const int nym1 = g.ny - 1;
            auto rowoff = [&](int r) { return (long long)min(max(r, 0), nym1) * g.pitch; };
            auto vrow = [&](int r) { return ld2<T>(vsp, rowoff(r)); };
            auto urow = [&](int r) { return ld2<T>(usp, rowoff(r)); };
            auto mrow = [&](int r) {
                return has_mask ? *reinterpret_cast<const uchar2*>(mask + rowoff(r) + cs) : make_uchar2(0, 0);
            };
            // D of row r from its u* pair (x-neighbours by shuffle; the two halo lanes load theirs) and v* of rows r+1, r-1
            auto drow = [&](int r, const D2& uc, const D2& vN, const D2& vS) {
                const bool rin = r >= 1 && r < nym1;                // warp-uniform
                double uW = shfl_up1(uc.y), uE = shfl_dn1(uc.x);
                const long long kr = (long long)r * g.pitch + c0;
                if (lane == 0 && xin && rin) uW = ld(us, kr - 1);
                if (lane == 31 && yin && rin) uE = ld(us, kr + 2);
                D2 d;
                d.x = (xin && rin) ? div_raw<T>(g, uc.y, uW, vN.x, vS.x) : 0.0;
                d.y = (yin && rin) ? div_raw<T>(g, uE, uc.x, vN.y, vS.y) : 0.0;
                return d;
            };
            // phi of row i after the red (odd) then black (even) colour of the first iteration from phi = 0
            auto prow = [&](int i, const D2& Dm, const D2& Dc, const D2& Dn, uchar2 mm, uchar2 mc, uchar2 mn) {
                D2 q = {0.0, 0.0};
                if (i >= 1 && i < g.ny - 1) {                       // warp-uniform
                    if (((i + g.jg0 + c0) & 1) == 0) {              // warp-uniform: .x cells are even, .y cells odd
                        if (yin && !mc.y) q.y = phi_odd<T>(g, Dc.y, inv_dt);
                        const double pW = shfl_up1(q.y);
                        if (xin && !mc.x) {
                            const double pn = mn.x ? 0.0 : phi_odd<T>(g, Dn.x, inv_dt), ps = mm.x ? 0.0 : phi_odd<T>(g, Dm.x, inv_dt);
                            q.x = Rnd<T>::r((g.ax * (q.y + pW) + g.ay * (pn + ps) - (-Dc.x * inv_dt)) * g.dinv);
                        }
                    } else {
                        if (xin && !mc.x) q.x = phi_odd<T>(g, Dc.x, inv_dt);
                        const double pE = shfl_dn1(q.x);
                        if (yin && !mc.y) {
                            const double pn = mn.y ? 0.0 : phi_odd<T>(g, Dn.y, inv_dt), ps = mm.y ? 0.0 : phi_odd<T>(g, Dm.y, inv_dt);
                            q.y = Rnd<T>::r((g.ax * (pE + q.x) + g.ay * (pn + ps) - (-Dc.y * inv_dt)) * g.dinv);
                        }
                    }
                }
                return q;
            };
            int i = lo - 1;
            D2 v_m = vrow(i - 1), v_c = vrow(i), v_n = vrow(i + 1);
            D2 u_m = {0.0, 0.0}, u_c = urow(i);
            D2 Dm = drow(i - 1, urow(i - 1), v_c, vrow(i - 2));
            D2 Dc = drow(i, u_c, v_n, v_m);
            uchar2 mm = mrow(i - 1), mc = mrow(i);
            D2 p_mm = {0.0, 0.0}, p_m = {0.0, 0.0};
            // software pipeline: the rows the NEXT iteration consumes are requested one iteration ahead
            D2 v_nn = vrow(i + 2), u_n = urow(i + 1);
            uchar2 mn = mrow(i + 1);
            MARCH_UNROLL
            for (; i <= hi; ++i) {
                const long long k = (long long)i * g.pitch + c0;
                if (inb && i + FPF < g.ny) { pf_l2(us + k + FPF * g.pitch); pf_l2(vs + k + FPF * g.pitch); }
                const D2 v_n3 = vrow(i + 3), u_n2 = urow(i + 2);     // (clamped rows: harmless on the last iterations)
                const uchar2 mn2 = mrow(i + 2);
                const D2 Dn = drow(i + 1, u_n, v_nn, v_c);
                const D2 p_c = prow(i, Dm, Dc, Dn, mm, mc, mn);
                if (owner && i >= lo && i < hi) {
                    st2<T>(phi, k, p_c.x, p_c.y);
                    if (DIV) st2<T>(div_out, k, Dc.x, Dc.y);
                    if (ownx) { dabs = fmax(dabs, fabs(Dc.x)); ch = fmax(ch, fabs(p_c.x)); }
                    if (owny) { dabs = fmax(dabs, fabs(Dc.y)); ch = fmax(ch, fabs(p_c.y)); }
                }
                if (i - 1 >= lo) {                                  // correct row i-1: phi rows i-2 (p_mm), i-1 (p_m), i (p_c)
                    const double pW = shfl_up1(p_m.y), pE = shfl_dn1(p_m.x);
                    double gxx = 0.0, gxy = 0.0, gyx = 0.0, gyy = 0.0;
                    if (xin) { gxx = Rnd<T>::r((p_m.y - pW) * g.hdx); gyx = Rnd<T>::r((p_c.x - p_mm.x) * g.hdy); }
                    if (yin) { gxy = Rnd<T>::r((pE - p_m.x) * g.hdx); gyy = Rnd<T>::r((p_c.y - p_mm.y) * g.hdy); }
                    if (owner) {
                        st2<T>(ua, k - g.pitch, axmy<T>(u_m.x, dt, gxx), axmy<T>(u_m.y, dt, gxy));
                        st2<T>(va, k - g.pitch, axmy<T>(v_m.x, dt, gyx), axmy<T>(v_m.y, dt, gyy));
                    }
                    // max |grad phi| = sqrt(max |grad phi|^2): the square root is taken once per thread after the loop
                    if (ownx) gmax = fmax(gmax, Rnd<T>::r(Rnd<T>::r(gxx * gxx) + Rnd<T>::r(gyx * gyx)));
                    if (owny) gmax = fmax(gmax, Rnd<T>::r(Rnd<T>::r(gxy * gxy) + Rnd<T>::r(gyy * gyy)));
                }
                Dm = Dc; Dc = Dn; v_m = v_c; v_c = v_n; v_n = v_nn; v_nn = v_n3; u_m = u_c; u_c = u_n; u_n = u_n2;
                p_mm = p_m; p_m = p_c; mm = mc; mc = mn; mn = mn2;
            }
        }
        }   // general rows
    }
    if (ctl) {
        block_atomic_max3(dabs, ch, sqrt(gmax), &ctl->max_div_pre, &ctl->pois_change, &ctl->max_grad);
    }
}

void launch_div_rb1_correct_fast(int dtype, const Geom& g, const void* us, const void* vs, const uint8_t* mask, void* phi,
                                 void* div_out, void* ua, void* va, double inv_dt, StepCtl* ctl, const uint8_t* flags,
                                 int n_sm, cudaStream_t s) {
    // chunks of 16..64 rows: two of a chunk's phi rows are recomputed by its neighbours
    const int bx = (g.nx + FB_OWN - 1) / FB_OWN;
    int rpb = 64;
    while (rpb > 16 && (long long)bx * ((g.ny + rpb - 1) / rpb) < 8ll * n_sm) rpb >>= 1;
    rpb = rpb_12(bx, g.ny, n_sm, rpb);
    // small grids (the shipped 600x180 case): the kernel is pure latency, so trade redundant halo rows for parallelism
    while (rpb > 4 && (long long)bx * ((g.ny + rpb - 1) / rpb) < 2ll * n_sm) rpb >>= 1;
    rpb = pick_rpb(bx, g.ny, 4 * n_sm, 4, rpb);              // k_div_rb1_correct_fast: 4 blocks per SM
    static const int force = getenv("B200CFD_K23_RPB") ? atoi(getenv("B200CFD_K23_RPB")) : 0;      // developer knob
    if (force > 0) rpb = force;
    const dim3 grid(bx, (g.ny + rpb - 1) / rpb);
    const int ntx = (g.nx + FT_W - 1) / FT_W;
#define DCF(T, DIV) launch_pdl(k_div_rb1_correct_fast<T, DIV>, grid, dim3(FBT), s, g, (const T*)us, (const T*)vs, mask, (T*)phi, (T*)div_out, (T*)ua, (T*)va, inv_dt, ctl, flags, ntx, rpb, lean_on)
    static const int lean_on = getenv("B200CFD_K23_LEAN") ? atoi(getenv("B200CFD_K23_LEAN")) : 1;   // developer knob (A/B)
    if (dtype == 0) { if (div_out) DCF(double, true); else DCF(double, false); }
    else { if (div_out) DCF(float, true); else DCF(float, false); }
#undef DCF
}

}  // namespace b200cfd
