// cleandiv_tma.cu — the lexicographic Gauss-Seidel sweep of clean_divergence_fast (v5.py:239-257) with its rows staged in
// shared memory by TMA.  Same tiling and the same decay-apron argument as cleandiv.cu (one warp = one 128-column window
// marching down its rows, tiles independent); what changes is what bounded that kernel on B200 (ncu r02a: 2 warps per
// scheduler, 6.4 stall cycles per issued instruction, a third of them waiting on global loads issued one row ahead):
//   * rows of u, v (and psi_old) arrive through a per-warp ring of TMA boxes (3 rows x 128 columns per array and slot,
//     cp.async.bulk.tensor.2d + mbarrier, zero fill outside the array), issued NSLOT chunks ahead by one lane: no
//     register staging, no address arithmetic, no clamping, and memory latency is off the row-to-row dependency chain;
//   * the chain itself is shorter: r = cn*x_north + s with s (everything that does not depend on the previous row)
//     computed off the chain, the four in-lane columns combined as a depth-2 tree, the Kogge-Stone scan of the lane
//     carries cut to the levels whose weight alpha^(4*2^k) is still above one ulp (3 for fp64 instead of 5), and the
//     `lane >= d` selects folded into per-lane coefficient registers: 56 + 35*LEVELS cycles per row instead of ~320
//     (tools/ubench_chain.cu: DFMA 8.2, SHFL64+DFMA 35 cycles dependent issue on B200);
//   * MODE 2: the sweep of the SECOND clean-divergence iteration applies the first iteration's correction
//     u,v -= grad(psi_1) (v5.py:254-256) on its load side (psi_1 is the psi_old it streams anyway), so the separate
//     k_clean_apply pass (40 B/cell) disappears; the finalize kernel applies both corrections.
#include "kernels.h"
#include "tma.cuh"
#include <cstdlib>
#include <type_traits>

namespace b200cfd {

#define TW 128        // window width (columns) of one warp = TMA box width
#define TPL 4         // columns per lane
#define TWARPS 4      // warps (tiles) per block
#define TKR 3         // rows per ring slot = unroll of the row loop = rotation period of the register rows

template <typename T> struct Sm4;
template <> struct Sm4<double> {
    static __device__ __forceinline__ void ld(const double* p, double (&o)[TPL]) {
        const double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
        o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
    }
    static __device__ __forceinline__ void st(double* p, const double (&x)[TPL]) {
        *reinterpret_cast<double2*>(p) = make_double2(x[0], x[1]);
        *reinterpret_cast<double2*>(p + 2) = make_double2(x[2], x[3]);
    }
};
template <> struct Sm4<float> {
    static __device__ __forceinline__ void ld(const float* p, double (&o)[TPL]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
    }
    static __device__ __forceinline__ void st(float* p, const double (&x)[TPL]) {
        *reinterpret_cast<float4*>(p) = make_float4((float)x[0], (float)x[1], (float)x[2], (float)x[3]);
    }
};

#define FULLM 0xffffffffu

// MODE 0: psi_old = 0 (first iteration).  MODE 1: psi_old streamed.  MODE 2: psi_old = psi_1 streamed AND the correction
// u,v -= grad(psi_1) applied to u, v as they are loaded (second iteration of the fused step).
template <typename T, int MODE, int LEVELS, int NSLOT, int MINB>
__global__ void __launch_bounds__(32 * TWARPS, MINB)
k_clean_sweep_tma(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
                  const __grid_constant__ CUtensorMap tm_p, Geom g, CleanPlan cp, T* __restrict__ pnew) {
    constexpr int NARR = MODE == 0 ? 2 : 3;
    constexpr int RA = MODE == 2 ? 2 : 1;          // the row computed at ingest step q is i = q - RA
    constexpr int WUC = MODE == 2 ? 2 : 1;         // warm-up chunks (ingest only)
    constexpr int DU = MODE == 2 ? 2 : 1;          // at ingest step q the slot row holds u[q - DU], v[q - DV], psi_old[q]
    constexpr int DV = MODE == 2 ? 1 : 0;
    constexpr int ARR = TKR * TW;                  // elements per array per slot
    constexpr int SLOT = NARR * ARR;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    pdl_wait(); pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.x * TWARPS + warp;
    if (tile >= cp.n_strips * cp.n_tiles) return;              // whole warp exits together (warps are independent)
    T* ring = reinterpret_cast<T*>(smem_raw) + (size_t)warp * NSLOT * SLOT;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)TWARPS * NSLOT * SLOT * sizeof(T)) + warp * NSLOT;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < NSLOT; ++k) mbar_init(&bars[k], 1);
        mbar_fence_init();
    }
    __syncwarp();
    const int s = tile % cp.n_strips, t = tile / cp.n_strips;
    const int js = cp.j_first + s * cp.strip_w;
    const int je = min(js + cp.strip_w, cp.j_end);
    const int is = 1 + t * cp.tile_h;
    const int ie = min(is + cp.tile_h, g.ny - 1);
    const int iw0 = max(1, is - cp.apron_v);
    const int R = ie - iw0;                                    // rows marched (apron + owned)
    const int NCH = WUC + (R + TKR - 1) / TKR;
    const int q0 = iw0 + RA - WUC * TKR;                       // first ingest row (may be negative: zero fill)
    const int jw0 = js - cp.apron_h;                           // window origin, a multiple of 4 (may be negative)
    const int c0 = jw0 + TPL * lane;

    auto issue = [&](int k) {                                  // lane 0: TMA boxes of chunk k into its slot
        const int slot = k % NSLOT;
        T* dst = ring + slot * SLOT;
        const int qk = q0 + TKR * k;
        mbar_expect_tx(&bars[slot], (uint32_t)(SLOT * sizeof(T)));
        tma_load_2d(dst, &tm_u, jw0, qk - DU, &bars[slot]);
        tma_load_2d(dst + ARR, &tm_v, jw0, qk - DV, &bars[slot]);
        if (NARR == 3) tma_load_2d(dst + 2 * ARR, &tm_p, jw0, qk, &bars[slot]);
    };
    if (lane == 0) {
        const int n0 = NCH < NSLOT ? NCH : NSLOT;
        for (int k = 0; k < n0; ++k) issue(k);
    }

    // ---- per-lane constants
    const double alpha = g.ax * g.dinv;
    const double ce = alpha, cn = g.ay * g.dinv, ndinv = -g.dinv, hdx = g.hdx, hdy = g.hdy;
    const double a2 = alpha * alpha;
    double ac[TPL], cl[LEVELS];
    {
        const double a3 = a2 * alpha, a4 = a2 * a2;
        const double z = lane > 0 ? 1.0 : 0.0;
        ac[0] = z * alpha; ac[1] = z * a2; ac[2] = z * a3; ac[3] = z * a4;
        double w = a4;
#pragma unroll
        for (int k = 0; k < LEVELS; ++k) { cl[k] = lane >= (1 << k) ? w : 0.0; w *= w; }
    }
    // Only global-interior columns are ever updated (border psi stays 0).  The predicate enters as a multiplicative mask
    // cm = 1.0 / 0.0 on the chain-independent part s (everything it multiplies is finite: inputs are fields or TMA zero
    // fill): one DMUL per column off the chain instead of re-evaluated compare/select chains.  Non-updated columns WEST of
    // the updated range then stay exactly 0 by themselves (x_north = 0, s = 0, carry-in = 0); the ones EAST of it collect
    // finite values that nothing reads (the recurrence runs west to east, they are not stored, and psi_old east of an
    // updated column comes from memory).
    double cm[TPL];
    bool stm[TPL];
#pragma unroll
    for (int m = 0; m < TPL; ++m) {
        const int c = c0 + m, cg = g.jg0 + c;
        cm[m] = (c >= 1 && c < g.nx - 1 && cg >= 1 && cg < g.nxg - 1) ? 1.0 : 0.0;
        stm[m] = (c >= js && c < je);
    }
    const bool st_all = (c0 >= js && c0 + TPL <= je);
    T* qn = pnew + (long long)iw0 * g.pitch + c0;               // row pointer of the row being computed (owned cells only)
    const long long pitch = g.pitch;

    // ---- the row-to-row chain: x = scan_alpha(cn * x_north + s)
    auto scan = [&](const double (&sv)[TPL], double (&xp)[TPL]) {
        const double r0 = fma(cn, xp[0], sv[0]), r1 = fma(cn, xp[1], sv[1]), r2 = fma(cn, xp[2], sv[2]), r3 = fma(cn, xp[3], sv[3]);
        const double P1 = fma(alpha, r0, r1), Q = fma(alpha, r2, r3);
        const double P2 = fma(alpha, P1, r2), P3 = fma(a2, P1, Q);
        double B = P3;
#pragma unroll
        for (int k = 0; k < LEVELS; ++k) { const double tmp = __shfl_up_sync(FULLM, B, 1 << k); B = fma(cl[k], tmp, B); }
        const double carry = __shfl_up_sync(FULLM, B, 1);
        xp[0] = Rnd<T>::r(fma(ac[0], carry, r0));
        xp[1] = Rnd<T>::r(fma(ac[1], carry, P1));
        xp[2] = Rnd<T>::r(fma(ac[2], carry, P2));
        xp[3] = Rnd<T>::r(fma(ac[3], carry, P3));
    };
    auto store_row = [&](int i, const double (&x)[TPL]) {
        if (i >= is && i < ie) {
            if (st_all) Sm4<T>::st(qn, x);
            else {
#pragma unroll
                for (int m = 0; m < TPL; ++m) if (stm[m]) qn[m] = (T)x[m];
            }
        }
        qn += pitch;
    };

    double xprev[TPL] = {0.0, 0.0, 0.0, 0.0};
    double vA[TPL], vB[TPL], vC[TPL];          // MODE 0/1: v rows i-1, i, i+1.  MODE 2: the CORRECTED v rows
    double pA[TPL], pB[TPL], pC[TPL];          // psi_old rows (MODE 1: two live, MODE 2: three live)
#pragma unroll
    for (int m = 0; m < TPL; ++m) { vA[m] = vB[m] = vC[m] = 0.0; pA[m] = pB[m] = pC[m] = 0.0; }

    // One ingest step.  sp = this lane's 4 elements of the slot row (u array; v at +ARR, psi_old at +2*ARR).
    // MODE 0/1: V0 = v[i-1], V1 = v[i] in, V2 = v[i+1] out;  P0 = psi_old[i] in, P1 = psi_old[i+1] out (MODE 1).
    // MODE 2:   V0 = v'[i-1], V1 = v'[i] in, V2 = v'[i+1] out; P0 = psi_1[i], P1 = psi_1[i+1] in, P2 = psi_1[i+2] out.
    auto step = [&](auto compute_tag, const T* sp, int q, double (&V0)[TPL], double (&V1)[TPL], double (&V2)[TPL],
                    double (&P0)[TPL], double (&P1)[TPL], double (&P2)[TPL]) {
        constexpr bool COMPUTE = decltype(compute_tag)::value;
        (void)V1;
        double uq[TPL], sv[TPL];
        Sm4<T>::ld(sp, uq);
        if (MODE == 0) {
            Sm4<T>::ld(sp + ARR, V2);
            if (COMPUTE) {
                const double uW = __shfl_up_sync(FULLM, uq[3], 1), uE = __shfl_down_sync(FULLM, uq[0], 1);
#pragma unroll
                for (int m = 0; m < TPL; ++m) {
                    const double e = m < TPL - 1 ? uq[(m + 1) % TPL] : uE, w = m > 0 ? uq[(m + TPL - 1) % TPL] : uW;
                    const double D = Rnd<T>::r((e - w) * hdx + (V2[m] - V0[m]) * hdy);
                    sv[m] = (D * ndinv) * cm[m];
                }
            }
        } else if (MODE == 1) {
            Sm4<T>::ld(sp + ARR, V2);
            Sm4<T>::ld(sp + 2 * ARR, P1);
            if (COMPUTE) {
                const double uW = __shfl_up_sync(FULLM, uq[3], 1), uE = __shfl_down_sync(FULLM, uq[0], 1);
                const double pE3 = __shfl_down_sync(FULLM, P0[0], 1);
#pragma unroll
                for (int m = 0; m < TPL; ++m) {
                    const double e = m < TPL - 1 ? uq[(m + 1) % TPL] : uE, w = m > 0 ? uq[(m + TPL - 1) % TPL] : uW;
                    const double D = Rnd<T>::r((e - w) * hdx + (V2[m] - V0[m]) * hdy);
                    const double pE = m < TPL - 1 ? P0[(m + 1) % TPL] : pE3;
                    sv[m] = fma(ce, pE, fma(cn, P1[m], D * ndinv)) * cm[m];
                }
            }
        } else {
            double vq[TPL];
            Sm4<T>::ld(sp + ARR, vq);
            Sm4<T>::ld(sp + 2 * ARR, P2);
            // corrected v of row q-1 = i+1 (the correction touches interior cells only, v5.py:254-256)
            const double vin = ((q - 1 >= 1) && (q - 1 < g.ny - 1)) ? 1.0 : 0.0;        // warp-uniform
#pragma unroll
            for (int m = 0; m < TPL; ++m)
                V2[m] = Rnd<T>::r(vq[m] - Rnd<T>::r((P2[m] - P0[m]) * hdy) * (vin * cm[m]));
            if (COMPUTE) {
                // corrected u of row i from psi_1 row i (x-neighbours by shuffle), then its own x-neighbours by shuffle
                const double pW = __shfl_up_sync(FULLM, P0[3], 1), pE3 = __shfl_down_sync(FULLM, P0[0], 1);
                double uc[TPL];
#pragma unroll
                for (int m = 0; m < TPL; ++m) {
                    const double e = m < TPL - 1 ? P0[(m + 1) % TPL] : pE3, w = m > 0 ? P0[(m + TPL - 1) % TPL] : pW;
                    uc[m] = Rnd<T>::r(uq[m] - Rnd<T>::r((e - w) * hdx) * cm[m]);
                }
                const double uW = __shfl_up_sync(FULLM, uc[3], 1), uE = __shfl_down_sync(FULLM, uc[0], 1);
#pragma unroll
                for (int m = 0; m < TPL; ++m) {
                    const double e = m < TPL - 1 ? uc[(m + 1) % TPL] : uE, w = m > 0 ? uc[(m + TPL - 1) % TPL] : uW;
                    const double D = Rnd<T>::r((e - w) * hdx + (V2[m] - V0[m]) * hdy);
                    const double pE = m < TPL - 1 ? P0[(m + 1) % TPL] : pE3;
                    sv[m] = fma(ce, pE, fma(cn, P1[m], D * ndinv)) * cm[m];
                }
            }
        }
        if (COMPUTE) {
            scan(sv, xprev);
            store_row(q - RA, xprev);
        }
    };
    // three steps of one chunk with the register rows rotated by renaming (period 3 = TKR)
    auto chunk = [&](auto compute_tag, int k) {
        const int slot = k % NSLOT;
        mbar_wait(&bars[slot], (uint32_t)((k / NSLOT) & 1));
        const T* sp = ring + slot * SLOT + TPL * lane;
        const int q = q0 + TKR * k;
        step(compute_tag, sp, q, vA, vB, vC, pA, pB, pC);
        step(compute_tag, sp + TW, q + 1, vB, vC, vA, pB, pC, pA);
        step(compute_tag, sp + 2 * TW, q + 2, vC, vA, vB, pC, pA, pB);
        __syncwarp();                                          // every lane has read the slot
        if (lane == 0 && k + NSLOT < NCH) issue(k + NSLOT);
    };
    for (int k = 0; k < WUC; ++k) chunk(std::false_type{}, k);
    // inflow guess for the row above the window = psi_old there (0 on the border row, in non-updated columns and outside
    // the grid: those cells of psi are never written).  After the warm-up chunks that row, psi_old[iw0 - 1], sits in pC.
    if (MODE != 0) {
#pragma unroll
        for (int m = 0; m < TPL; ++m) xprev[m] = pC[m];
    }
    for (int k = WUC; k < NCH; ++k) chunk(std::true_type{}, k);
}

// smallest number of Kogge-Stone levels L for which what the truncated scan drops, alpha^(4 * 2^L) / (1 - alpha) relative,
// is below 2^-(bits + 4)
static int scan_levels(double alpha, int bits) {
    for (int L = 3; L < 4; ++L)
        if ((double)(4 << L) * -log2(alpha) >= bits + 4 - log2(1.0 - alpha)) return L;
    return 4;
}

// ring slots per warp (3 rows each).  Measured on B200 (profiles/r02d_knobs.txt): the two-array sweep of the first
// iteration likes the deepest ring (4 slots, 2 blocks per SM: 1.11 ms against 1.25 ms with 2 slots at 16384^2, equal at
// 4096x1024); the three-array sweeps are better off with more resident warps than with more rows in flight (2 slots,
// 3 blocks per SM: 1.45 / 0.046 ms against 1.76 / 0.068 ms with 4 slots at 16384^2 / 4096x1024).
int clean_tma_default_slots(int dtype, int mode) {
    (void)dtype;
    return mode == 0 ? 4 : 2;
}

static int tune_env(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

template <typename T, int MODE, int LEVELS, int NSLOT>
static bool launch_one(const CUtensorMap& mu, const CUtensorMap& mv, const CUtensorMap& mp, const Geom& g, const CleanPlan& cp,
                       void* pnew, int nblk, cudaStream_t s) {
    constexpr int NARR = MODE == 0 ? 2 : 3;
    // resident blocks per SM the ring allows (227 KB of shared memory per SM, 1 KB reserved per block); the register cap
    // of __launch_bounds__ follows it
    constexpr size_t SMEM = (size_t)TWARPS * NSLOT * NARR * TKR * TW * sizeof(T) + TWARPS * NSLOT * sizeof(uint64_t);
    constexpr int FIT = (int)((227 * 1024) / (SMEM + 1024));
    constexpr int MINB = FIT > 4 ? 4 : (FIT < 1 ? 1 : FIT);
    auto kern = k_clean_sweep_tma<T, MODE, LEVELS, NSLOT, MINB>;
    static bool configured = false;
    if (!configured) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        configured = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nblk); cfg.blockDim = dim3(32 * TWARPS); cfg.dynamicSmemBytes = SMEM; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, mu, mv, mp, g, cp, (T*)pnew) == cudaSuccess;
}

template <typename T, int MODE, int NSLOT>
static bool launch_levels(int L, const CUtensorMap& mu, const CUtensorMap& mv, const CUtensorMap& mp, const Geom& g,
                          const CleanPlan& cp, void* pnew, int nblk, cudaStream_t s) {
    if (L <= 3) return launch_one<T, MODE, 3, NSLOT>(mu, mv, mp, g, cp, pnew, nblk, s);
    return launch_one<T, MODE, 4, NSLOT>(mu, mv, mp, g, cp, pnew, nblk, s);
}
template <typename T, int MODE>
static bool launch_slots(int nslot, int L, const CUtensorMap& mu, const CUtensorMap& mv, const CUtensorMap& mp, const Geom& g,
                         const CleanPlan& cp, void* pnew, int nblk, cudaStream_t s) {
    if (nslot <= 2) return launch_levels<T, MODE, 2>(L, mu, mv, mp, g, cp, pnew, nblk, s);
    if (nslot == 3) return launch_levels<T, MODE, 3>(L, mu, mv, mp, g, cp, pnew, nblk, s);
    return launch_levels<T, MODE, 4>(L, mu, mv, mp, g, cp, pnew, nblk, s);
}

// Is the TMA sweep applicable?  One-pass plan, rows of a multiple of 4 cells, 16-byte aligned bases and pitch.
bool clean_sweep_tma_ok(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v, const void* psi_old,
                        const void* psi_new) {
    static const int on = tune_env("B200CFD_CLEAN_TMA", 1);
    const int es = dtype == 0 ? 8 : 4;
    if (!on || cp.two_pass || g.nx % 4 || g.nx < 8 || (g.pitch * es) % 16 || (cp.j_first - cp.apron_h) % 4 || cp.strip_w % 4) return false;
    if (cp.strip_w + cp.apron_h > TW) return false;
    if (((uintptr_t)u | (uintptr_t)v | (uintptr_t)psi_old | (uintptr_t)psi_new) & 15) return false;
    return tma_encoder() != nullptr;
}

// mode 0: psi_old = 0;  1: psi_old streamed;  2: psi_old = psi_1 streamed and u,v -= grad(psi_1) applied on the load side
bool launch_clean_sweep_tma(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v, const void* psi_old,
                            void* psi_new, int mode, cudaStream_t s) {
    if (!clean_sweep_tma_ok(dtype, g, cp, u, v, psi_old, psi_new)) return false;
    CUtensorMap mu, mv, mp;
    if (!tma_make_2d(&mu, dtype, u, g.nx, g.ny, g.pitch, TW, TKR) || !tma_make_2d(&mv, dtype, v, g.nx, g.ny, g.pitch, TW, TKR) ||
        !tma_make_2d(&mp, dtype, mode ? psi_old : u, g.nx, g.ny, g.pitch, TW, TKR))
        return false;
    const int ntile = cp.n_strips * cp.n_tiles;
    const int nblk = (ntile + TWARPS - 1) / TWARPS;
    if (nblk <= 0) return true;
    const double alpha = g.ax * g.dinv;
    // alpha < 1/2, so 4 levels (64 columns) always suffice; 3 when alpha^32 is already below one ulp (dx >= ~0.8 dy)
    const int L = tune_env("B200CFD_CLEAN_LEVELS", scan_levels(alpha, dtype == 0 ? 53 : 30));
    // ring slots per warp: more slots = more rows in flight per warp but fewer resident warps (shared memory)
    const int nslot = tune_env("B200CFD_CLEAN_NSLOT", clean_tma_default_slots(dtype, mode));
    if (dtype == 0) {
        if (mode == 0) return launch_slots<double, 0>(nslot, L, mu, mv, mp, g, cp, psi_new, nblk, s);
        if (mode == 1) return launch_slots<double, 1>(nslot, L, mu, mv, mp, g, cp, psi_new, nblk, s);
        return launch_slots<double, 2>(nslot, L, mu, mv, mp, g, cp, psi_new, nblk, s);
    }
    if (mode == 0) return launch_slots<float, 0>(nslot, L, mu, mv, mp, g, cp, psi_new, nblk, s);
    if (mode == 1) return launch_slots<float, 1>(nslot, L, mu, mv, mp, g, cp, psi_new, nblk, s);
    return launch_slots<float, 2>(nslot, L, mu, mv, mp, g, cp, psi_new, nblk, s);
}

}  // namespace b200cfd
