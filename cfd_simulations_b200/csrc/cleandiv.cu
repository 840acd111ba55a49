// cleandiv.cu — clean_divergence_fast (v5.py:239-257): lexicographic Gauss-Seidel sweep of
//   psi[i,j] = (a(psi[i,j+1] + psi[i,j-1]) + b(psi[i+1,j] + psi[i-1,j]) - div[i,j]) d       (in place, row-major order)
// followed by u,v -= grad(psi), twice, psi carried across the two iterations.
//
// Lexicographic order means cell (i,j) sees NEW west/south and OLD east/north values: a recurrence
//   x[i,j] = alpha x[i,j-1] + beta x[i-1,j] + r[i,j],   alpha = a d, beta = b d, alpha + beta = 1/2 exactly.
// The influence of a value n cells upstream (Manhattan distance) is bounded by 2^-n, so a tile that starts
// L cells above/left of the cells it owns with *guessed* inflow reproduces the global sweep to
// (alpha/(1-beta))^L resp. (beta/(1-alpha))^L relative error.  L is chosen so this is below
// one ulp of the storage type: the tiles are independent, no wavefront ordering, no grid sync.
//
// One warp owns one tile.  Each lane holds CPL = 4 adjacent columns of a 128-column window and marches down
// the rows; within a row the west-recurrence is a first-order linear scan: 4 serial FMAs per lane, a 5-level
// Kogge-Stone scan of the lane carries with warp shuffles, then 4 FMAs to apply the carry.  div(u,v) is
// formed on the fly from rotating rows of v and the current row of u, so psi_new costs
// R(u, v, psi_old) + W(psi_new) = 32 B/cell in fp64 (24 B/cell in the first iteration where psi_old = 0).
#include "kernels.h"

namespace b200cfd {

#define CW 128          // window width (columns) of one warp
#define CPL 4           // columns per lane
#define CWARPS 4        // warps (tiles) per block
#define CPFD 6          // L2 prefetch distance in rows

template <typename T> __device__ __forceinline__ double ldg0(const T* __restrict__ p, long long row_off, int c, int nx) {
    return (c >= 0 && c < nx) ? (double)p[row_off + c] : 0.0;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

template <typename T, bool HAS_OLD>
__global__ void __launch_bounds__(32 * CWARPS)
k_clean_sweep(Geom g, CleanPlan cp, const T* __restrict__ u, const T* __restrict__ v, const T* __restrict__ pold,
              T* __restrict__ pnew) {
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x * CWARPS + (threadIdx.x >> 5);
    if (tile >= cp.n_strips * cp.n_tiles) return;          // whole warp exits together
    const int s = tile % cp.n_strips, t = tile / cp.n_strips;
    const int js = cp.j_first + s * cp.strip_w;             // first owned column
    const int je = min(js + cp.strip_w, cp.j_end);          // one past the last owned column
    const int jw0 = js - cp.apron_h;                        // window origin (may be < 1: forced-zero cells)
    const int is = 1 + t * cp.tile_h;
    const int ie = min(is + cp.tile_h, g.ny - 1);
    const int iw0 = max(1, is - cp.apron_v);
    const int c0 = jw0 + CPL * lane;
    const int nx = g.nx;
    const double alpha = g.ax * g.dinv;
    const double a2 = alpha * alpha, a3 = a2 * alpha, a4 = a2 * a2;
    const double a8 = a4 * a4, a16 = a8 * a8, a32 = a16 * a16, a64 = a32 * a32;
    const double ce = g.ax * g.dinv, cn = g.ay * g.dinv;

    // column predicates: only global-interior columns are ever updated (border psi stays 0)
    bool colin[CPL];
#pragma unroll
    for (int m = 0; m < CPL; ++m) {
        const int c = c0 + m, cg = g.jg0 + c;
        colin[m] = (c >= 1 && c < nx - 1 && cg >= 1 && cg < g.nxg - 1);
    }

    double xprev[CPL], v_im1[CPL], v_i[CPL], pcur[CPL];
    double ucur[CPL], vnext[CPL], pnext[CPL];
    {
        const long long rm = (long long)(iw0 - 1) * g.pitch, r0 = (long long)iw0 * g.pitch;
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            v_im1[m] = ldg0(v, rm, c0 + m, nx);
            v_i[m] = ldg0(v, r0, c0 + m, nx);
            // inflow guess for the rows above the window: psi_old there (exact 0 when iw0 == 1: border row)
            xprev[m] = (HAS_OLD && iw0 > 1 && colin[m]) ? ldg0(pold, rm, c0 + m, nx) : 0.0;
            pcur[m] = HAS_OLD ? ldg0(pold, r0, c0 + m, nx) : 0.0;
            ucur[m] = ldg0(u, r0, c0 + m, nx);
            vnext[m] = ldg0(v, r0 + g.pitch, c0 + m, nx);
            pnext[m] = HAS_OLD ? ldg0(pold, r0 + g.pitch, c0 + m, nx) : 0.0;
        }
    }

    for (int i = iw0; i < ie; ++i) {
        // ---- issue next row's loads (row i+1: u row i+1, v row i+2, psi_old row i+2) and an L2 prefetch ahead
        double un_[CPL], vn2[CPL], pn2[CPL];
        {
            const bool more = (i + 1 < ie);
            const long long r1 = (long long)(i + 1) * g.pitch, r2 = r1 + g.pitch;
#pragma unroll
            for (int m = 0; m < CPL; ++m) {
                un_[m] = more ? ldg0(u, r1, c0 + m, nx) : 0.0;
                vn2[m] = more ? ldg0(v, r2, c0 + m, nx) : 0.0;     // row i+2 <= ny-1 because i+1 < ie <= ny-1
                pn2[m] = (HAS_OLD && more) ? ldg0(pold, r2, c0 + m, nx) : 0.0;
            }
            const int ip = i + CPFD;
            if (ip < ie && c0 >= 0 && c0 < nx) {
                const long long rp = (long long)ip * g.pitch + c0;
                prefetch_l2(u + rp);
                prefetch_l2(v + rp + g.pitch);
                if (HAS_OLD) prefetch_l2(pold + rp + g.pitch);
            }
        }
        // ---- right-hand side of row i
        const double uE3 = __shfl_down_sync(0xffffffffu, ucur[0], 1);
        const double uW0 = __shfl_up_sync(0xffffffffu, ucur[CPL - 1], 1);
        const double pE3 = HAS_OLD ? __shfl_down_sync(0xffffffffu, pcur[0], 1) : 0.0;
        double r[CPL];
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            const double uE = (m < CPL - 1) ? ucur[(m + 1) % CPL] : (lane < 31 ? uE3 : 0.0);
            const double uW = (m > 0) ? ucur[(m + CPL - 1) % CPL] : (lane > 0 ? uW0 : 0.0);
            const double D = Rnd<T>::r((uE - uW) * g.hdx + (vnext[m] - v_im1[m]) * g.hdy);
            const double pE = (m < CPL - 1) ? pcur[(m + 1) % CPL] : (lane < 31 ? pE3 : 0.0);
            const double rr = ce * pE + cn * (pnext[m] + xprev[m]) - D * g.dinv;
            r[m] = colin[m] ? rr : 0.0;
        }
        // ---- west recurrence x[c] = alpha x[c-1] + r[c]: serial in the lane, Kogge-Stone across lanes
        double t0 = r[0];
        double t1 = fma(alpha, t0, r[1]);
        double t2 = fma(alpha, t1, r[2]);
        double t3 = fma(alpha, t2, r[3]);
        double B = t3, tmp;
        tmp = __shfl_up_sync(0xffffffffu, B, 1);  if (lane >= 1)  B = fma(a4, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 2);  if (lane >= 2)  B = fma(a8, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 4);  if (lane >= 4)  B = fma(a16, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 8);  if (lane >= 8)  B = fma(a32, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 16); if (lane >= 16) B = fma(a64, tmp, B);
        double carry = __shfl_up_sync(0xffffffffu, B, 1);
        if (lane == 0) carry = 0.0;
        double x[CPL];
        x[0] = Rnd<T>::r(fma(alpha, carry, t0));
        x[1] = Rnd<T>::r(fma(a2, carry, t1));
        x[2] = Rnd<T>::r(fma(a3, carry, t2));
        x[3] = Rnd<T>::r(fma(a4, carry, t3));
        // a non-updated column keeps psi = 0 but the scan leaks alpha^k x into it: force it
#pragma unroll
        for (int m = 0; m < CPL; ++m) if (!colin[m]) x[m] = 0.0;
        // ---- store owned cells
        if (i >= is) {
            const long long ro = (long long)i * g.pitch;
#pragma unroll
            for (int m = 0; m < CPL; ++m) {
                const int c = c0 + m;
                if (c >= js && c < je) pnew[ro + c] = (T)x[m];
            }
        }
        // ---- rotate
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            xprev[m] = x[m];
            v_im1[m] = v_i[m]; v_i[m] = vnext[m]; vnext[m] = vn2[m];
            pcur[m] = pnext[m]; pnext[m] = pn2[m];
            ucur[m] = un_[m];
        }
    }
}

// u,v[1:-1,1:-1] -= grad(psi)[1:-1,1:-1]   (v5.py:254-256), arithmetic in the field dtype
template <typename T>
__global__ void __launch_bounds__(256)
k_clean_apply(Geom g, T* __restrict__ u, T* __restrict__ v, const T* __restrict__ psi) {
    const int j = blockIdx.x * 64 + threadIdx.x, i = blockIdx.y * 4 + threadIdx.y;
    if (i < 1 || i >= g.ny - 1 || j < 1 || j >= g.nx - 1) return;
    const int jg = g.jg0 + j;
    if (jg < 1 || jg >= g.nxg - 1) return;
    const long long k = i * g.pitch + j;
    const double gx = Rnd<T>::r(((double)psi[k + 1] - (double)psi[k - 1]) * g.hdx);
    const double gy = Rnd<T>::r(((double)psi[k + g.pitch] - (double)psi[k - g.pitch]) * g.hdy);
    u[k] = (T)((double)u[k] - gx);
    v[k] = (T)((double)v[k] - gy);
}

static int apron_for(double ratio, int bits) {
    // smallest L with ratio^L <= 2^-bits
    if (!(ratio > 0.0)) return 1;
    if (ratio >= 0.5) ratio = 0.5;
    return (int)ceil((double)bits / -log2(ratio)) + 1;
}

CleanPlan clean_plan(const Geom& g, int dtype, int n_sm) {
    CleanPlan cp;
    const double alpha = g.ax * g.dinv, beta = g.ay * g.dinv;
    const int bits = dtype == 0 ? 53 : 30;
    int Lh = apron_for(alpha / (1.0 - beta), bits);
    int Lv = apron_for(beta / (1.0 - alpha), bits);
    // window origin must stay 4-aligned relative to column 1 + s*strip_w: make apron_h = 4k + 1
    Lh = ((Lh + 2) / 4) * 4 + 1;
    if (Lh < 5) Lh = 5;
    cp.apron_h = Lh;
    cp.apron_v = Lv;
    cp.strip_w = CW - Lh - 3;            // multiple of 4; the last 3 window columns are never owned
    cp.j_first = 1;
    cp.j_end = g.nx - 1;
    const int ncols = g.nx - 2, nrows = g.ny - 2;
    cp.n_strips = (ncols + cp.strip_w - 1) / cp.strip_w;
    // enough tiles for ~8 warps per SM, tile height in [32, 512]
    const int want = 8 * n_sm;
    int nt = (want + cp.n_strips - 1) / cp.n_strips;
    int th = (nrows + nt - 1) / (nt > 0 ? nt : 1);
    if (th < 32) th = 32;
    if (th > 512) th = 512;
    if (th > nrows) th = nrows > 0 ? nrows : 1;
    cp.tile_h = th;
    cp.n_tiles = (nrows + th - 1) / th;
    return cp;
}

void launch_clean_sweep(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v,
                        const void* psi_old, void* psi_new, cudaStream_t s) {
    const int ntile = cp.n_strips * cp.n_tiles;
    const int nblk = (ntile + CWARPS - 1) / CWARPS;
    if (nblk <= 0) return;
    if (dtype == 0) {
        if (psi_old) k_clean_sweep<double, true><<<nblk, 32 * CWARPS, 0, s>>>(g, cp, (const double*)u, (const double*)v, (const double*)psi_old, (double*)psi_new);
        else k_clean_sweep<double, false><<<nblk, 32 * CWARPS, 0, s>>>(g, cp, (const double*)u, (const double*)v, nullptr, (double*)psi_new);
    } else {
        if (psi_old) k_clean_sweep<float, true><<<nblk, 32 * CWARPS, 0, s>>>(g, cp, (const float*)u, (const float*)v, (const float*)psi_old, (float*)psi_new);
        else k_clean_sweep<float, false><<<nblk, 32 * CWARPS, 0, s>>>(g, cp, (const float*)u, (const float*)v, nullptr, (float*)psi_new);
    }
}

void launch_clean_apply(int dtype, const Geom& g, void* u, void* v, const void* psi, cudaStream_t s) {
    dim3 grid((g.nx + 63) / 64, (g.ny + 3) / 4), block(64, 4);
    if (dtype == 0) k_clean_apply<double><<<grid, block, 0, s>>>(g, (double*)u, (double*)v, (const double*)psi);
    else k_clean_apply<float><<<grid, block, 0, s>>>(g, (float*)u, (float*)v, (const float*)psi);
}

}  // namespace b200cfd
