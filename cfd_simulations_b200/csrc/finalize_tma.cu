// finalize_tma.cu — K5 of the time step with its rows staged in shared memory by TMA:
//   [u,v -= grad(psi_1)] [u,v -= grad(psi)]  (the pending clean-divergence corrections, v5.py:254-256)
//   -> max|div| (:421-422) -> BC (:424, :349-360) -> IBM (:425, :228-237) -> nanmax|vorticity| (:427-428)
//   -> nanmean energy (:431-432) -> clip (:437-438) -> post-clip max|u|, |v|.
// Same arithmetic, in the same order, as k_finalize_fast (step_fast.cu); what changes is the shape.  One warp owns 120
// columns of a 128-column window (4 columns per lane, lanes 0 and 31 are halo lanes) and streams down a chunk of rows:
// rows of ua, va, psi (and psi_1) arrive through a per-warp ring of TMA boxes (3 rows per slot, zero fill outside the
// array), so there are no software-pipelined prefetch registers, no address arithmetic and no bounds tests in the row
// loop; column predicates are multiplicative masks.  A row runs through two stages held in registers: stage 1 forms the
// corrected velocity f of row q-1 and its boundary / immersed-boundary processed value g, stage 2 finishes row q-2
// (divergence from f rows q-3, q-1; vorticity from g rows q-3, q-1 and the x-neighbours of g row q-2; energy, clip, store).
// 128 columns per warp-row instead of 64 and ~2.6x fewer instructions per cell than the register-marching kernel.
#include <cstdlib>
#include <type_traits>

#include "kernels.h"
#include "tma.cuh"
#include "cellmath.cuh"

namespace b200cfd {

#define FW 128        // window width = TMA box width
#define FPL 4         // columns per lane
#define FOWN 120      // owned columns per warp (lanes 1..30)
#define FWARPS 4
#define FKR 3         // rows per ring slot = unroll of the row loop
#define FT_W 32       // flag tile (step_fast.cu): columns
#define FT_H 8        // rows
#define FULLM 0xffffffffu

struct FinPlan { int n_strips, n_rchunks, rpb; };

template <typename T> struct FRow;
template <> struct FRow<double> {
    static __device__ __forceinline__ void ld(const double* p, double (&o)[FPL]) {
        const double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
        o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
    }
    static __device__ __forceinline__ void st(double* p, const double (&x)[FPL]) {
        *reinterpret_cast<double2*>(p) = make_double2(x[0], x[1]);
        *reinterpret_cast<double2*>(p + 2) = make_double2(x[2], x[3]);
    }
};
template <> struct FRow<float> {
    static __device__ __forceinline__ void ld(const float* p, double (&o)[FPL]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
    }
    static __device__ __forceinline__ void st(float* p, const double (&x)[FPL]) {
        *reinterpret_cast<float4*>(p) = make_float4((float)x[0], (float)x[1], (float)x[2], (float)x[3]);
    }
};

template <typename T, int NPSI, int NSLOT>
__global__ void __launch_bounds__(32 * FWARPS, 2)
k_finalize_tma(const __grid_constant__ CUtensorMap tm_u, const __grid_constant__ CUtensorMap tm_v,
               const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_q, Geom g, Phys p,
               StepCtl* __restrict__ ctl, T* __restrict__ u, T* __restrict__ v, const double* __restrict__ ibm,
               const uint8_t* __restrict__ mask, const double* __restrict__ y, double* __restrict__ ke_part,
               unsigned long long* __restrict__ cnt_part, const uint8_t* __restrict__ flags, int ntx, FinPlan fp) {
    constexpr int NARR = 2 + NPSI;
    constexpr int ARR = FKR * FW;
    constexpr int SLOT = NARR * ARR;                  // ua rows, va rows, psi rows, (psi_1 rows)
    extern __shared__ __align__(128) unsigned char smem_raw[];
    pdl_wait(); pdl_trigger();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.x * FWARPS + warp;
    FinAcc acc;
    if (tile < fp.n_strips * fp.n_rchunks) {
        T* ring = reinterpret_cast<T*>(smem_raw) + (size_t)warp * NSLOT * SLOT;
        uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)FWARPS * NSLOT * SLOT * sizeof(T)) + warp * NSLOT;
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < NSLOT; ++k) mbar_init(&bars[k], 1);
            mbar_fence_init();
        }
        __syncwarp();
        const int s = tile % fp.n_strips, rc = tile / fp.n_strips;
        const int x0 = s * FOWN, jw0 = x0 - FPL, c0 = jw0 + FPL * lane;
        const int r0 = rc * fp.rpb, r1 = min(r0 + fp.rpb, g.ny);
        // ingest step q: slot row = psi[q] (and psi_1[q]), ua[q-1], va[q-1]; stage 1 makes f, g of row q-1, stage 2 finishes q-2
        const int q0 = r0 - 2;
        const int NCH = (r1 - r0 + 4 + FKR - 1) / FKR;
        auto issue = [&](int k) {
            const int slot = k % NSLOT;
            T* dst = ring + slot * SLOT;
            const int qk = q0 + FKR * k;
            mbar_expect_tx(&bars[slot], (uint32_t)(SLOT * sizeof(T)));
            tma_load_2d(dst, &tm_u, jw0, qk - 1, &bars[slot]);
            tma_load_2d(dst + ARR, &tm_v, jw0, qk - 1, &bars[slot]);
            tma_load_2d(dst + 2 * ARR, &tm_p, jw0, qk, &bars[slot]);
            if (NPSI == 2) tma_load_2d(dst + 3 * ARR, &tm_q, jw0, qk, &bars[slot]);
        };
        if (lane == 0) {
            const int n0 = NCH < NSLOT ? NCH : NSLOT;
            for (int k = 0; k < n0; ++k) issue(k);
        }
        // ---- warp-uniform: do the flag tiles over the owned columns (dilated by 2 cells) carry ibm / mask bits?
        int f = 0;
        {
            const int cB = min(x0 + FOWN, g.nx);
            const int tx0 = x0 / FT_W, tx1 = (cB - 1) / FT_W, ty0 = r0 / FT_H, ty1 = (r1 - 1) / FT_H;
            const int w = tx1 - tx0 + 1, n = w * (ty1 - ty0 + 1);
            for (int t = lane; t < n; t += 32) f |= flags[(ty0 + t / w) * ntx + tx0 + t % w];
        }
        const bool has_ibm = __any_sync(FULLM, f & 1), has_mask = mask && __any_sync(FULLM, f & 2);
        // ---- per-lane column data.  c0 and nx are multiples of 4: a lane's four columns are all inside the array or all outside
        const bool inb = c0 >= 0 && c0 < g.nx;
        const bool st_lane = inb && lane >= 1 && lane <= 30;       // stores: the owned lanes
        double cm[FPL], wo[FPL], wi[FPL];                           // interior column / slab-owned / both (as 1.0 or 0.0)
        int cls[FPL];
#pragma unroll
        for (int m = 0; m < FPL; ++m) {
            const int c = c0 + m;
            cls[m] = inb ? col_class(g, c) : COL_COPY;
            cm[m] = (inb && cls[m] == COL_INTERIOR) ? 1.0 : 0.0;
            wo[m] = (st_lane && c >= g.own0 && c < g.own1) ? 1.0 : 0.0;
            wi[m] = cm[m] * wo[m];
        }
        const bool edge_lane = inb && (cls[0] == COL_INFLOW || cls[3] == COL_OUTFLOW);
        const double hdx = g.hdx, hdy = g.hdy;
        const double i2dx = 1.0 / Rnd<T>::r(g.two_dx), i2dy = 1.0 / Rnd<T>::r(g.two_dy);
        const double vmax = p.max_velocity, fs = ctl->fs, pert = ctl->pert_scale, phase = ctl->phase;
        const long long pitch = g.pitch;

        // register rows (rotated by renaming, period 3)
        double P[3][FPL], Q[3][FPL];          // psi / psi_1 rows q-2, q-1, q
        double FV[3][FPL], GU[3][FPL];        // corrected v / processed u of rows q-3, q-2, q-1
        double FU[3][FPL], GV[3][FPL];        // corrected u / processed v of rows q-2, q-1 (third name: renaming slack)
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int m = 0; m < FPL; ++m) { P[a][m] = Q[a][m] = FV[a][m] = GU[a][m] = FU[a][m] = GV[a][m] = 0.0; }

        auto step = [&](const T* sp, int q, double (&P0)[FPL], double (&P1)[FPL], double (&P2)[FPL], double (&Q0)[FPL],
                        double (&Q1)[FPL], double (&Q2)[FPL], double (&fvA)[FPL], double (&fvC)[FPL], double (&guA)[FPL],
                        double (&guB)[FPL], double (&guC)[FPL], double (&fuB)[FPL], double (&fuC)[FPL], double (&gvB)[FPL],
                        double (&gvC)[FPL]) {
            double uq[FPL], vq[FPL];
            FRow<T>::ld(sp, uq);
            FRow<T>::ld(sp + ARR, vq);
            FRow<T>::ld(sp + 2 * ARR, P2);
            if (NPSI == 2) FRow<T>::ld(sp + 3 * ARR, Q2);
            // ---- stage 1: corrected velocity f of row q-1 (the corrections touch interior cells only)
            const int rf = q - 1;
            const double rin = (rf >= 1 && rf < g.ny - 1) ? 1.0 : 0.0;             // warp-uniform
            const double pW = __shfl_up_sync(FULLM, P1[3], 1), pE = __shfl_down_sync(FULLM, P1[0], 1);
            double qW = 0.0, qE = 0.0;
            if (NPSI == 2) { qW = __shfl_up_sync(FULLM, Q1[3], 1); qE = __shfl_down_sync(FULLM, Q1[0], 1); }
#pragma unroll
            for (int m = 0; m < FPL; ++m) {
                const double mk = cm[m] * rin;
                double fu = uq[m], fv = vq[m];
                if (NPSI == 2) {
                    const double e = m < FPL - 1 ? Q1[(m + 1) % FPL] : qE, w = m > 0 ? Q1[(m + FPL - 1) % FPL] : qW;
                    fu = Rnd<T>::r(fu - Rnd<T>::r((e - w) * hdx) * mk);
                    fv = Rnd<T>::r(fv - Rnd<T>::r((Q2[m] - Q0[m]) * hdy) * mk);
                }
                {
                    const double e = m < FPL - 1 ? P1[(m + 1) % FPL] : pE, w = m > 0 ? P1[(m + FPL - 1) % FPL] : pW;
                    fu = Rnd<T>::r(fu - Rnd<T>::r((e - w) * hdx) * mk);
                    fv = Rnd<T>::r(fv - Rnd<T>::r((P2[m] - P0[m]) * hdy) * mk);
                }
                fuC[m] = fu; fvC[m] = fv;
            }
            // ---- BC (v5.py:349-360) + IBM (v5.py:228-237) of row q-1, pointwise: g
            const bool wall = rf <= 0 || rf >= g.ny - 1;                            // warp-uniform: u = v = 0 (v5.py:357-360)
#pragma unroll
            for (int m = 0; m < FPL; ++m) { guC[m] = wall ? 0.0 : fuC[m]; gvC[m] = wall ? 0.0 : fvC[m]; }
            if (edge_lane && !wall) {
                if (cls[0] == COL_INFLOW) { guC[0] = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[rf], pert, phase)); gvC[0] = 0.0; }
                if (cls[3] == COL_OUTFLOW) { guC[3] = fuC[2]; gvC[3] = fvC[2]; }    // copies column nx-2 (before its IBM)
            }
            if (has_ibm && inb && !wall) {
                const double* ib = ibm + (long long)rf * pitch + c0;
                const double2 a = *reinterpret_cast<const double2*>(ib), b = *reinterpret_cast<const double2*>(ib + 2);
                const double mm[FPL] = {a.x, a.y, b.x, b.y};
#pragma unroll
                for (int m = 0; m < FPL; ++m)
                    if (mm[m] > 0) { const double w = 1.0 - mm[m] * fs; guC[m] = Rnd<T>::r(guC[m] * w); gvC[m] = Rnd<T>::r(gvC[m] * w); }
            }
            // ---- stage 2: finish row i = q-2 (the shuffles stay outside the branch: no convergence protocol)
            const int i = q - 2;
            const double fuW = __shfl_up_sync(FULLM, fuB[3], 1), fuE = __shfl_down_sync(FULLM, fuB[0], 1);
            const double gvW = __shfl_up_sync(FULLM, gvB[3], 1), gvE = __shfl_down_sync(FULLM, gvB[0], 1);
            if (i >= r0 && i < r1) {                                                 // warp-uniform
                const double iin = (i >= 1 && i < g.ny - 1) ? 1.0 : 0.0;            // interior row
                unsigned mk4 = 0u;
                if (has_mask && inb) mk4 = *reinterpret_cast<const unsigned*>(mask + (long long)i * pitch + c0);
                double cu[FPL], cv[FPL];
#pragma unroll
                for (int m = 0; m < FPL; ++m) {
                    const double e = m < FPL - 1 ? fuB[(m + 1) % FPL] : fuE, w = m > 0 ? fuB[(m + FPL - 1) % FPL] : fuW;
                    const double D = Rnd<T>::r((e - w) * hdx + (fvC[m] - fvA[m]) * hdy);
                    acc.dmax = fmax(acc.dmax, fabs(D) * (wi[m] * iin));
                    const double ge = m < FPL - 1 ? gvB[(m + 1) % FPL] : gvE, gw = m > 0 ? gvB[(m + FPL - 1) % FPL] : gvW;
                    const double om = Rnd<T>::r(Rnd<T>::r(Rnd<T>::r(ge - gw) * i2dx) - Rnd<T>::r(Rnd<T>::r(guC[m] - guA[m]) * i2dy));
                    const double wm = ((mk4 >> (8 * m)) & 0xffu) ? 0.0 : wi[m] * iin;
                    acc.wmax = fmax(acc.wmax, fabs(om) * wm);                        // NaN is dropped by fmax (nanmax)
                    const double gu = guB[m], gv = gvB[m];
                    const double ex = Rnd<T>::r(0.5 * Rnd<T>::r(Rnd<T>::r(gu * gu) + Rnd<T>::r(gv * gv)));
                    const double a = fmin(fmax(gu, -vmax), vmax), b = fmin(fmax(gv, -vmax), vmax);
                    cu[m] = gu == gu ? a : gu;                                        // np.clip propagates NaN
                    cv[m] = gv == gv ? b : gv;
                    if (wo[m] != 0.0) {
                        if (ex == ex) { acc.e += ex; acc.cnt += 1; }
                        if (!isfinite(gu) || !isfinite(gv)) acc.bad = 1;
                        acc.mu = fmax(acc.mu, fabs(a)); acc.mv = fmax(acc.mv, fabs(b));
                    }
                }
                if (st_lane) {
                    const long long k = (long long)i * pitch + c0;
                    FRow<T>::st(u + k, cu);
                    FRow<T>::st(v + k, cv);
                }
            }
        };
        for (int k = 0; k < NCH; ++k) {
            const int slot = k % NSLOT;
            mbar_wait(&bars[slot], (uint32_t)((k / NSLOT) & 1));
            const T* sp = ring + slot * SLOT + FPL * lane;
            const int q = q0 + FKR * k;
            //            P0    P1    P2    Q0    Q1    Q2    fvA    fvC    guA    guB    guC    fuB    fuC    gvB    gvC
            step(sp,          q,     P[0], P[1], P[2], Q[0], Q[1], Q[2], FV[0], FV[2], GU[0], GU[1], GU[2], FU[1], FU[2], GV[1], GV[2]);
            step(sp + FW,     q + 1, P[1], P[2], P[0], Q[1], Q[2], Q[0], FV[1], FV[0], GU[1], GU[2], GU[0], FU[2], FU[0], GV[2], GV[0]);
            step(sp + 2 * FW, q + 2, P[2], P[0], P[1], Q[2], Q[0], Q[1], FV[2], FV[1], GU[2], GU[0], GU[1], FU[0], FU[1], GV[0], GV[1]);
            __syncwarp();
            if (lane == 0 && k + NSLOT < NCH) issue(k + NSLOT);
        }
    }
    finalize_block_reduce(acc, ctl, ke_part, cnt_part, blockIdx.x);
}

static int fin_env(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

static FinPlan fin_plan(const Geom& g, int n_sm) {
    FinPlan fp;
    fp.n_strips = (g.nx + FOWN - 1) / FOWN;
    // every tile resident at once: 2 blocks of 4 warps per SM
    const int want = fin_env("B200CFD_FIN_WARPS_PER_SM", 8) * n_sm;
    int nrc = want / fp.n_strips;
    if (nrc < 1) nrc = 1;
    int rpb = (g.ny + nrc - 1) / nrc;
    if (rpb < 8) rpb = 8;
    fp.rpb = rpb;
    fp.n_rchunks = (g.ny + rpb - 1) / rpb;
    return fp;
}

bool finalize_tma_ok(int dtype, const Geom& g, const void* ua, const void* va, const void* psi, const void* psi1) {
    // OFF by default - measured on B200 (profiles/r02g_phases.txt): parity-green, but SLOWER than the register-marching
    // k_finalize_fast (4096x1024: 86 us against 78 us; 16384^2: 4.6 ms against 3.2 ms).  With four staged arrays a warp
    // needs 24 KB of ring and 220-248 registers, so only 8 warps fit an SM, and a single in-order instruction stream per
    // warp (1000 instructions per 3 rows, a third of them register moves of the rotated row arrays) cannot hide its own
    // latencies; the marching kernel runs 12 warps per SM.  B200CFD_FIN_TMA=1 selects this kernel for experiments.
    static const int on = fin_env("B200CFD_FIN_TMA", 0);
    const int es = dtype == 0 ? 8 : 4;
    if (!on || !psi || g.nx % 4 || g.nx < 8 || (g.pitch * es) % 16) return false;
    if (((uintptr_t)ua | (uintptr_t)va | (uintptr_t)psi | (uintptr_t)psi1) & 15) return false;
    return tma_encoder() != nullptr;
}
int finalize_tma_num_partials(const Geom& g, int n_sm) {
    const FinPlan fp = fin_plan(g, n_sm);
    return (fp.n_strips * fp.n_rchunks + FWARPS - 1) / FWARPS;
}

template <typename T, int NPSI, int NSLOT>
static bool fin_launch(const CUtensorMap& mu, const CUtensorMap& mv, const CUtensorMap& mp, const CUtensorMap& mq, const Geom& g,
                       const Phys& p, StepCtl* ctl, void* u, void* v, const double* ibm, const uint8_t* mask, const double* y,
                       double* ke, unsigned long long* cnt, const uint8_t* flags, const FinPlan& fp, cudaStream_t s) {
    constexpr size_t SMEM = (size_t)FWARPS * NSLOT * (2 + NPSI) * FKR * FW * sizeof(T) + FWARPS * NSLOT * sizeof(uint64_t);
    auto kern = k_finalize_tma<T, NPSI, NSLOT>;
    static bool configured = false;
    if (!configured) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM) != cudaSuccess) { cudaGetLastError(); return false; }
        configured = true;
    }
    const int ntile = fp.n_strips * fp.n_rchunks, ntx = (g.nx + FT_W - 1) / FT_W;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((ntile + FWARPS - 1) / FWARPS); cfg.blockDim = dim3(32 * FWARPS); cfg.dynamicSmemBytes = SMEM; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, mu, mv, mp, mq, g, p, ctl, (T*)u, (T*)v, ibm, mask, y, ke, cnt, flags, ntx, fp) == cudaSuccess;
}

// psi = the last clean-divergence potential (pending correction), psi1 = the first one when it is pending too (or NULL)
bool launch_finalize_tma(int dtype, const Geom& g, const Phys& p, StepCtl* ctl, const void* ua, const void* va, const void* psi,
                         const void* psi1, void* u, void* v, const double* ibm, const uint8_t* mask, const double* y,
                         double* ke_partials, unsigned long long* cnt_partials, const uint8_t* flags, int n_sm, cudaStream_t s) {
    if (!finalize_tma_ok(dtype, g, ua, va, psi, psi1)) return false;
    CUtensorMap mu, mv, mp, mq;
    if (!tma_make_2d(&mu, dtype, ua, g.nx, g.ny, g.pitch, FW, FKR) || !tma_make_2d(&mv, dtype, va, g.nx, g.ny, g.pitch, FW, FKR) ||
        !tma_make_2d(&mp, dtype, psi, g.nx, g.ny, g.pitch, FW, FKR) ||
        !tma_make_2d(&mq, dtype, psi1 ? psi1 : psi, g.nx, g.ny, g.pitch, FW, FKR))
        return false;
    const FinPlan fp = fin_plan(g, n_sm);
    const int nslot = fin_env("B200CFD_FIN_NSLOT", 2);
#define FIN(T, NP)                                                                                                              \
    (nslot >= 3 ? fin_launch<T, NP, 3>(mu, mv, mp, mq, g, p, ctl, u, v, ibm, mask, y, ke_partials, cnt_partials, flags, fp, s) \
                : fin_launch<T, NP, 2>(mu, mv, mp, mq, g, p, ctl, u, v, ibm, mask, y, ke_partials, cnt_partials, flags, fp, s))
    if (dtype == 0) return psi1 ? FIN(double, 2) : FIN(double, 1);
    return psi1 ? FIN(float, 2) : FIN(float, 1);
#undef FIN
}

}  // namespace b200cfd
