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
#include <cstdlib>

namespace b200cfd {

#define CW 128          // window width (columns) of one warp
#define CPL 4           // columns per lane
#define CWARPS 4        // warps (tiles) per block
#define CPFD 6          // L2 prefetch distance in rows

template <typename T> __device__ __forceinline__ double ldg0(const T* __restrict__ p, long long row_off, int c, int nx) {
    return (c >= 0 && c < nx) ? (double)p[row_off + c] : 0.0;
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// One lane's CPL = 4 adjacent cells of a row.  The window origin is a multiple of 4 columns, so with a pitch that is a
// multiple of 4 and 32-byte aligned fields (VEC) the four cells are ONE 256-bit (fp64: LDG.E.256, sm_100) or 128-bit
// (fp32) access: a warp then touches each 32-byte sector once per row instead of four times (the scalar form is
// L1-wavefront bound).  Lanes whose four columns are not all inside [0, nx) take the scalar, zero-filled path.
template <typename T, int VEC> struct Row4;   // VEC: 0 scalar, 1 one 256-bit access, 2 two 128-bit accesses
template <int VEC> struct Row4<double, VEC> {
    static __device__ __forceinline__ void ld(const double* __restrict__ p, long long row_off, int c, int nx, double (&o)[CPL]) {
        if (VEC == 1 && c >= 0 && c + CPL <= nx) {
            asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                         : "=d"(o[0]), "=d"(o[1]), "=d"(o[2]), "=d"(o[3]) : "l"(p + row_off + c));
        } else if (VEC == 2 && c >= 0 && c + CPL <= nx) {
            const double2 t0 = __ldg(reinterpret_cast<const double2*>(p + row_off + c));
            const double2 t1 = __ldg(reinterpret_cast<const double2*>(p + row_off + c + 2));
            o[0] = t0.x; o[1] = t0.y; o[2] = t1.x; o[3] = t1.y;
        } else {
#pragma unroll
            for (int m = 0; m < CPL; ++m) o[m] = ldg0(p, row_off, c + m, nx);
        }
    }
    static __device__ __forceinline__ void st(double* __restrict__ p, long long k, const double (&x)[CPL]) {
        if (VEC == 1) {
            asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p + k), "d"(x[0]), "d"(x[1]), "d"(x[2]), "d"(x[3]) : "memory");
        } else {
            *reinterpret_cast<double2*>(p + k) = make_double2(x[0], x[1]);
            *reinterpret_cast<double2*>(p + k + 2) = make_double2(x[2], x[3]);
        }
    }
};
template <int VEC> struct Row4<float, VEC> {
    static __device__ __forceinline__ void ld(const float* __restrict__ p, long long row_off, int c, int nx, double (&o)[CPL]) {
        if (VEC && c >= 0 && c + CPL <= nx) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(p + row_off + c));
            o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
        } else {
#pragma unroll
            for (int m = 0; m < CPL; ++m) o[m] = ldg0(p, row_off, c + m, nx);
        }
    }
    static __device__ __forceinline__ void st(float* __restrict__ p, long long k, const double (&x)[CPL]) {
        *reinterpret_cast<float4*>(p + k) = make_float4((float)x[0], (float)x[1], (float)x[2], (float)x[3]);
    }
};

// APPLY: the correction u,v -= grad(psi_new) (v5.py:254-256) of the SAME iteration is fused in as a lagged epilogue: once
// row i of psi_new exists, row i-1 of the corrected velocity (psi rows i-2, i-1, i; x-neighbours by shuffle) is written
// to uo, vo (separate buffers: neighbouring tiles still read u, v in their aprons).  The tile then marches one row past
// its owned rows; the first / last strip also carry the local border columns and the first / last tile the wall rows
// (copied unchanged), so uo, vo are complete fields.  Saves the 40 B/cell pass of k_clean_apply minus 16 B/cell.
template <typename T, bool HAS_OLD, int VEC, bool APPLY>
__global__ void __launch_bounds__(32 * CWARPS)
k_clean_sweep(Geom g, CleanPlan cp, const T* __restrict__ u, const T* __restrict__ v, const T* __restrict__ pold,
              T* __restrict__ pnew, T* __restrict__ uo, T* __restrict__ vo) {
    pdl_wait(); pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x * CWARPS + (threadIdx.x >> 5);
    if (tile >= cp.n_strips * cp.n_tiles) return;          // whole warp exits together
    const int s = tile % cp.n_strips, t = tile / cp.n_strips;
    const int js = cp.j_first + s * cp.strip_w;             // first owned column
    const int je = min(js + cp.strip_w, cp.j_end);          // one past the last owned column
    const int jw0 = js - cp.apron_h;                        // window origin (may be < 1: forced-zero cells)
    const int is = 1 + t * cp.tile_h;
    const int ie = min(is + cp.tile_h, g.ny - 1);
    const int iw0 = cp.two_pass ? is : max(1, is - cp.apron_v);   // two-pass: own rows only, zero inflow
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
    double xprev2[CPL] = {0.0, 0.0, 0.0, 0.0}, uprev[CPL] = {0.0, 0.0, 0.0, 0.0};   // APPLY: psi row i-2, u row i-1
    // APPLY: columns this warp writes uo, vo for = owned columns + the local border column of the first / last strip
    const int jsA = (s == 0) ? 0 : js, jeA = (s == cp.n_strips - 1) ? nx : je;
    if (APPLY) {
        if (t == 0 || t == cp.n_tiles - 1) {                  // wall rows: copied unchanged
            const long long rw = (t == 0) ? 0 : (long long)(g.ny - 1) * g.pitch;
#pragma unroll
            for (int m = 0; m < CPL; ++m) {
                const int c = c0 + m;
                if (c >= jsA && c < jeA) {
                    uo[rw + c] = u[rw + c]; vo[rw + c] = v[rw + c];
                    if (cp.n_tiles == 1) { const long long r2 = (long long)(g.ny - 1) * g.pitch; uo[r2 + c] = u[r2 + c]; vo[r2 + c] = v[r2 + c]; }
                }
            }
        }
    }
    const int iend = APPLY ? ie : ie - 1;                      // last row marched (APPLY: one past the owned rows)
    {
        const long long rm = (long long)(iw0 - 1) * g.pitch, r0 = (long long)iw0 * g.pitch;
        typedef Row4<T, VEC> R4;
        R4::ld(v, rm, c0, nx, v_im1);
        R4::ld(v, r0, c0, nx, v_i);
        R4::ld(u, r0, c0, nx, ucur);
        R4::ld(v, r0 + g.pitch, c0, nx, vnext);
        if (HAS_OLD) {
            // inflow guess for the rows above the window: psi_old there (exact 0 when iw0 == 1: border row)
            R4::ld(pold, rm, c0, nx, xprev);
            R4::ld(pold, r0, c0, nx, pcur);
            R4::ld(pold, r0 + g.pitch, c0, nx, pnext);
        }
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            if (!(HAS_OLD && iw0 > 1 && colin[m]) || cp.two_pass) xprev[m] = 0.0;
            if (!HAS_OLD) { pcur[m] = 0.0; pnext[m] = 0.0; }
        }
    }

    for (int i = iw0; i <= iend; ++i) {
        // ---- issue next row's loads (row i+1: u row i+1, v row i+2, psi_old row i+2) and an L2 prefetch ahead
        double un_[CPL], vn2[CPL], pn2[CPL];
        {
            const bool more = (i + 1 <= iend) && (i + 1 < g.ny - 1);
            const long long r1 = (long long)(i + 1) * g.pitch, r2 = r1 + g.pitch;
            if (more) {                                            // row i+2 <= ny-1 because i+1 <= ny-2
                Row4<T, VEC>::ld(u, r1, c0, nx, un_);
                Row4<T, VEC>::ld(v, r2, c0, nx, vn2);
                if (HAS_OLD) Row4<T, VEC>::ld(pold, r2, c0, nx, pn2);
            }
#pragma unroll
            for (int m = 0; m < CPL; ++m) {
                if (!more) { un_[m] = 0.0; vn2[m] = 0.0; }
                if (!(HAS_OLD && more)) pn2[m] = 0.0;
            }
            const int ip = i + CPFD;
            if (ip < ie && c0 >= 0 && c0 < nx) {   // (APPLY marches to ie: that row is prefetched by the tile that owns it)
                const long long rp = (long long)ip * g.pitch + c0;
                prefetch_l2(u + rp);
                prefetch_l2(v + rp + g.pitch);
                if (HAS_OLD) prefetch_l2(pold + rp + g.pitch);
            }
        }
        double x[CPL] = {0.0, 0.0, 0.0, 0.0};
        if (!APPLY || i < g.ny - 1) {                              // (APPLY: row ny-1 is the wall, psi = 0 there)
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
        x[0] = Rnd<T>::r(fma(alpha, carry, t0));
        x[1] = Rnd<T>::r(fma(a2, carry, t1));
        x[2] = Rnd<T>::r(fma(a3, carry, t2));
        x[3] = Rnd<T>::r(fma(a4, carry, t3));
        // a non-updated column keeps psi = 0 but the scan leaks alpha^k x into it: force it
#pragma unroll
        for (int m = 0; m < CPL; ++m) if (!colin[m]) x[m] = 0.0;
        }
        // ---- store owned cells
        if (i >= is && i < ie) {
            const long long ro = (long long)i * g.pitch;
            if (VEC && c0 >= js && c0 + CPL <= je) {
                Row4<T, VEC>::st(pnew, ro + c0, x);
            } else {
#pragma unroll
                for (int m = 0; m < CPL; ++m) {
                    const int c = c0 + m;
                    if (c >= js && c < je) pnew[ro + c] = (T)x[m];
                }
            }
        }
        // ---- APPLY: corrected velocity of row i-1 from psi rows i-2 (xprev2), i-1 (xprev), i (x)
        if (APPLY && i - 1 >= is) {
            const double xE3 = __shfl_down_sync(0xffffffffu, xprev[0], 1);
            const double xW0 = __shfl_up_sync(0xffffffffu, xprev[CPL - 1], 1);
            const long long ro = (long long)(i - 1) * g.pitch;
#pragma unroll
            for (int m = 0; m < CPL; ++m) {
                const int c = c0 + m;
                double gx = 0.0, gy = 0.0;
                if (colin[m]) {
                    const double xE = (m < CPL - 1) ? xprev[(m + 1) % CPL] : xE3;
                    const double xW = (m > 0) ? xprev[(m + CPL - 1) % CPL] : xW0;
                    gx = Rnd<T>::r((xE - xW) * g.hdx);
                    gy = Rnd<T>::r((x[m] - xprev2[m]) * g.hdy);
                }
                if (c >= jsA && c < jeA) {
                    uo[ro + c] = (T)(uprev[m] - gx);
                    vo[ro + c] = (T)(v_im1[m] - gy);
                }
            }
        }
        // ---- rotate
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            xprev2[m] = xprev[m]; uprev[m] = ucur[m];
            xprev[m] = x[m];
            v_im1[m] = v_i[m]; v_i[m] = vnext[m]; vnext[m] = vn2[m];
            pcur[m] = pnext[m]; pnext[m] = pn2[m];
            ucur[m] = un_[m];
        }
    }
}

// Lean form of the one-pass sweep for grids whose row length is a multiple of 4 (every shipped configuration).  Same
// arithmetic, same order, bit-identical results; what changes is the instruction count per row (ncu, r01f: 369 SASS
// instructions per row of which 64 were fp64 math or shuffles — the rest were per-element bounds tests, 64-bit address
// arithmetic, zero fills and register rotation):
//   * the window origin and nx are multiples of 4, so a lane's four columns are all inside [0, nx) or all outside: the
//     lane's base column is CLAMPED into the row once, every load is an unguarded 128-bit access off a row pointer that
//     advances by the pitch.  A clamped lane reads finite garbage, but all of its columns are non-updated (colin false:
//     r and x are selected to 0, never multiplied), and the only in-range columns that see it are c = 0 and c = nx-1,
//     which are non-updated themselves;
//   * on the last row the look-ahead pointers stay where they are (a re-read that hits L1) instead of zero-filling;
//   * the row loop is unrolled by 6 = lcm(3 rows of v, 2 rows of psi_old in flight): the rotation is register renaming.
template <typename T> struct Ld4;
template <> struct Ld4<double> {
    static __device__ __forceinline__ void ld(const double* __restrict__ p, double (&o)[CPL]) {
        const double2 a = __ldg(reinterpret_cast<const double2*>(p)), b = __ldg(reinterpret_cast<const double2*>(p + 2));
        o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
    }
};
template <> struct Ld4<float> {
    static __device__ __forceinline__ void ld(const float* __restrict__ p, double (&o)[CPL]) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(p));
        o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
    }
};

template <typename T, bool HAS_OLD, int UNR, int MINB>
__global__ void __launch_bounds__(32 * CWARPS, MINB)
k_clean_sweep_lean(Geom g, CleanPlan cp, const T* __restrict__ u, const T* __restrict__ v, const T* __restrict__ pold,
                   T* __restrict__ pnew) {
    pdl_wait(); pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x * CWARPS + (threadIdx.x >> 5);
    if (tile >= cp.n_strips * cp.n_tiles) return;          // whole warp exits together
    const int s = tile % cp.n_strips, t = tile / cp.n_strips;
    const int js = cp.j_first + s * cp.strip_w;
    const int je = min(js + cp.strip_w, cp.j_end);
    const int is = 1 + t * cp.tile_h;
    const int ie = min(is + cp.tile_h, g.ny - 1);
    const int iw0 = max(1, is - cp.apron_v);
    const int iend = ie - 1;
    const int nx = g.nx;
    const int c0 = js - cp.apron_h + CPL * lane;            // true first column of the lane (multiple of 4)
    const int c0c = min(max(c0, 0), nx - CPL);              // clamped: where the lane loads from
    const long long pitch = g.pitch;
    const double alpha = g.ax * g.dinv;
    const double a2 = alpha * alpha, a3 = a2 * alpha, a4 = a2 * a2;
    const double a8 = a4 * a4, a16 = a8 * a8, a32 = a16 * a16, a64 = a32 * a32;
    const double ce = g.ax * g.dinv, cn = g.ay * g.dinv;
    const double hdx = g.hdx, hdy = g.hdy, dinv = g.dinv;
    const int pfd = cp.pfd;
    const long long pfo = (long long)(pfd - 1) * pitch;

    bool colin[CPL], stm[CPL];
#pragma unroll
    for (int m = 0; m < CPL; ++m) {
        const int c = c0 + m, cg = g.jg0 + c;
        colin[m] = (c >= 1 && c < nx - 1 && cg >= 1 && cg < g.nxg - 1);
        stm[m] = (c >= js && c < je);
    }
    const bool st_all = (c0 >= js && c0 + CPL <= je);

    double xprev[CPL], v_im1[CPL], v_i[CPL], pcur[CPL], ucur[CPL], vnext[CPL], pnext[CPL];
    const T* qv = v + (long long)(iw0 - 1) * pitch + c0c;
    const T* qu = u + (long long)iw0 * pitch + c0c;
    const T* qp = HAS_OLD ? pold + (long long)(iw0 - 1) * pitch + c0c : nullptr;
    Ld4<T>::ld(qv, v_im1);
    Ld4<T>::ld(qv + pitch, v_i);
    Ld4<T>::ld(qv + 2 * pitch, vnext);
    Ld4<T>::ld(qu, ucur);
    if (HAS_OLD) {
        Ld4<T>::ld(qp, xprev);
        Ld4<T>::ld(qp + pitch, pcur);
        Ld4<T>::ld(qp + 2 * pitch, pnext);
    }
#pragma unroll
    for (int m = 0; m < CPL; ++m) {
        if (!(HAS_OLD && iw0 > 1 && colin[m])) xprev[m] = 0.0;
        if (!HAS_OLD) { pcur[m] = 0.0; pnext[m] = 0.0; }
    }
    // look-ahead pointers: u row min(i+1, iend), v and psi_old row min(i+2, iend+1)  (iend + 1 <= ny - 1)
    {
        const long long du = (iw0 + 1 <= iend) ? 1 : 0;
        qu += du * pitch;
        qv += (2 + du) * pitch;
        if (HAS_OLD) qp += (2 + du) * pitch;
    }
    T* qn = pnew + (long long)iw0 * pitch + c0;             // dereferenced only where stm[] holds

#pragma unroll UNR
    for (int i = iw0; i <= iend; ++i) {
        double un_[CPL], vn2[CPL], pn2[CPL];
        Ld4<T>::ld(qu, un_);
        Ld4<T>::ld(qv, vn2);
        if (HAS_OLD) Ld4<T>::ld(qp, pn2);
        if (i + pfd < ie) {
            prefetch_l2(qu + pfo);
            prefetch_l2(qv + pfo);
            if (HAS_OLD) prefetch_l2(qp + pfo);
        }
        {
            const long long adv = (i + 2 <= iend) ? pitch : 0;
            qu += adv; qv += adv;
            if (HAS_OLD) qp += adv;
        }
        // ---- right-hand side of row i
        const double uE3 = __shfl_down_sync(0xffffffffu, ucur[0], 1);
        const double uW0 = __shfl_up_sync(0xffffffffu, ucur[CPL - 1], 1);
        const double pE3 = HAS_OLD ? __shfl_down_sync(0xffffffffu, pcur[0], 1) : 0.0;
        double r[CPL], x[CPL];
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            const double uE = (m < CPL - 1) ? ucur[(m + 1) % CPL] : (lane < 31 ? uE3 : 0.0);
            const double uW = (m > 0) ? ucur[(m + CPL - 1) % CPL] : (lane > 0 ? uW0 : 0.0);
            const double D = Rnd<T>::r((uE - uW) * hdx + (vnext[m] - v_im1[m]) * hdy);
            const double pE = (m < CPL - 1) ? pcur[(m + 1) % CPL] : (lane < 31 ? pE3 : 0.0);
            const double rr = ce * pE + cn * (pnext[m] + xprev[m]) - D * dinv;
            r[m] = colin[m] ? rr : 0.0;
        }
        // ---- west recurrence x[c] = alpha x[c-1] + r[c]: serial in the lane, Kogge-Stone across lanes
        const double t0 = r[0];
        const double t1 = fma(alpha, t0, r[1]);
        const double t2 = fma(alpha, t1, r[2]);
        const double t3 = fma(alpha, t2, r[3]);
        double B = t3, tmp;
        tmp = __shfl_up_sync(0xffffffffu, B, 1);  if (lane >= 1)  B = fma(a4, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 2);  if (lane >= 2)  B = fma(a8, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 4);  if (lane >= 4)  B = fma(a16, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 8);  if (lane >= 8)  B = fma(a32, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 16); if (lane >= 16) B = fma(a64, tmp, B);
        double carry = __shfl_up_sync(0xffffffffu, B, 1);
        if (lane == 0) carry = 0.0;
        x[0] = Rnd<T>::r(fma(alpha, carry, t0));
        x[1] = Rnd<T>::r(fma(a2, carry, t1));
        x[2] = Rnd<T>::r(fma(a3, carry, t2));
        x[3] = Rnd<T>::r(fma(a4, carry, t3));
#pragma unroll
        for (int m = 0; m < CPL; ++m) if (!colin[m]) x[m] = 0.0;
        // ---- store owned cells
        if (i >= is) {
            if (st_all) {
                Row4<T, 2>::st(qn, 0, x);
            } else {
#pragma unroll
                for (int m = 0; m < CPL; ++m) if (stm[m]) qn[m] = (T)x[m];
            }
        }
        qn += pitch;
        // ---- rotate
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            xprev[m] = x[m];
            v_im1[m] = v_i[m]; v_i[m] = vnext[m]; vnext[m] = vn2[m];
            if (HAS_OLD) { pcur[m] = pnext[m]; pnext[m] = pn2[m]; }     // (no psi_old: both stay 0)
            ucur[m] = un_[m];
        }
    }
}

// Second pass of the two-pass sweep.  The recurrence is linear, so the sweep of a tile from its TRUE inflow row equals
// its zero-inflow sweep y (first pass, k_clean_sweep with cp.two_pass) plus the homogeneous response c to that row:
//   c[i,j] = alpha c[i,j-1] + beta c[i-1,j],  c[is-1,:] = psi[is-1,:],  psi[i,j] = y[i,j] + c[i,j].
// c decays by beta/(1-alpha) per row: after apron_v rows it is below one ulp, so only the first apron_v rows of a tile
// are touched, and because tile_h > apron_v the inflow row (the LAST row of the tile above) is never one of them — it
// already holds its final value when this kernel starts.  No u, v, psi_old traffic, no right-hand side: the scan only.
// Same 128-column windows and horizontal apron as the first pass (c is guessed 0 west of the window).
template <typename T>
__global__ void __launch_bounds__(32 * CWARPS)
k_clean_corr(Geom g, CleanPlan cp, T* __restrict__ psi) {
    const int lane = threadIdx.x & 31;
    const int tile = blockIdx.x * CWARPS + (threadIdx.x >> 5);
    if (tile >= cp.n_strips * (cp.n_tiles - 1)) return;
    const int s = tile % cp.n_strips, t = 1 + tile / cp.n_strips;
    const int js = cp.j_first + s * cp.strip_w;
    const int je = min(js + cp.strip_w, cp.j_end);
    const int c0 = js - cp.apron_h + CPL * lane;
    const int is = 1 + t * cp.tile_h;
    const int ie = min(min(is + cp.tile_h, g.ny - 1), is + cp.apron_v);
    const int nx = g.nx;
    const double alpha = g.ax * g.dinv, cn = g.ay * g.dinv;
    const double a2 = alpha * alpha, a3 = a2 * alpha, a4 = a2 * a2;
    const double a8 = a4 * a4, a16 = a8 * a8, a32 = a16 * a16, a64 = a32 * a32;
    bool colin[CPL], own[CPL];
    double cprev[CPL];
    const long long rm = (long long)(is - 1) * g.pitch;
#pragma unroll
    for (int m = 0; m < CPL; ++m) {
        const int c = c0 + m, cg = g.jg0 + c;
        colin[m] = (c >= 1 && c < nx - 1 && cg >= 1 && cg < g.nxg - 1);
        own[m] = (c >= js && c < je);
        cprev[m] = colin[m] ? ldg0(psi, rm, c, nx) : 0.0;
    }
    double ycur[CPL];
#pragma unroll
    for (int m = 0; m < CPL; ++m) ycur[m] = own[m] ? (double)psi[(long long)is * g.pitch + c0 + m] : 0.0;
    for (int i = is; i < ie; ++i) {
        const long long ro = (long long)i * g.pitch;
        double ynext[CPL];
#pragma unroll
        for (int m = 0; m < CPL; ++m) ynext[m] = (own[m] && i + 1 < ie) ? (double)psi[ro + g.pitch + c0 + m] : 0.0;
        double r[CPL];
#pragma unroll
        for (int m = 0; m < CPL; ++m) r[m] = colin[m] ? cn * cprev[m] : 0.0;
        const double t0 = r[0];
        const double t1 = fma(alpha, t0, r[1]);
        const double t2 = fma(alpha, t1, r[2]);
        const double t3 = fma(alpha, t2, r[3]);
        double B = t3, tmp;
        tmp = __shfl_up_sync(0xffffffffu, B, 1);  if (lane >= 1)  B = fma(a4, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 2);  if (lane >= 2)  B = fma(a8, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 4);  if (lane >= 4)  B = fma(a16, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 8);  if (lane >= 8)  B = fma(a32, tmp, B);
        tmp = __shfl_up_sync(0xffffffffu, B, 16); if (lane >= 16) B = fma(a64, tmp, B);
        double carry = __shfl_up_sync(0xffffffffu, B, 1);
        if (lane == 0) carry = 0.0;
        double x[CPL];
        x[0] = fma(alpha, carry, t0);
        x[1] = fma(a2, carry, t1);
        x[2] = fma(a3, carry, t2);
        x[3] = fma(a4, carry, t3);
#pragma unroll
        for (int m = 0; m < CPL; ++m) {
            if (!colin[m]) x[m] = 0.0;
            if (own[m]) psi[ro + c0 + m] = (T)(ycur[m] + x[m]);
            cprev[m] = x[m];
            ycur[m] = ynext[m];
        }
    }
}

// u,v[1:-1,1:-1] -= grad(psi)[1:-1,1:-1]   (v5.py:254-256), arithmetic in the field dtype
template <typename T>
__global__ void __launch_bounds__(256)
k_clean_apply(Geom g, T* __restrict__ u, T* __restrict__ v, const T* __restrict__ psi) {
    pdl_wait(); pdl_trigger();
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

// developer tuning knob (environment), read once per process
static int tune_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
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
    // accuracy of the guessed-inflow decay, in bits of psi: 53 / 30 = one ulp of the storage type (default, exact
    // reproduction of the sequential sweep); B200CFD_CLEAN_BITS is a developer knob for measuring what fewer bits would buy
    const int bits = tune_int("B200CFD_CLEAN_BITS", dtype == 0 ? 53 : 30);
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
    // enough tiles for ~8 warps per SM, tile height in [32 (8 on small grids), 512]
    // (rounded DOWN: every tile must be resident at once - one warp per tile, all tiles the same length - or the few
    // blocks of a second wave double the kernel's time: measured r02c, 304 blocks on 296 slots)
    const int want = tune_int("B200CFD_CLEAN_WARPS_PER_SM", 8) * n_sm;
    int nt = want / cp.n_strips;
    if (nt < 1) nt = 1;
    int th = (nrows + nt - 1) / (nt > 0 ? nt : 1);
    // at least 32 rows per tile when there are enough tiles anyway; on small grids (not even 2 warps per SM) latency
    // is all that matters: go down to 8 owned rows per tile
    const int th_min = ((long long)cp.n_strips * ((nrows + 31) / 32) < 2ll * n_sm) ? 8 : 32;
    if (th < th_min) th = th_min;
    if (th > 512) th = 512;
    if (th > nrows) th = nrows > 0 ? nrows : 1;
    cp.tile_h = th;
    cp.n_tiles = (nrows + th - 1) / th;
    // two-pass sweep (k_clean_corr): needs tiles taller than the decay length so that a tile's last row is final after the
    // first pass; worth it when such tiles still fill the machine (otherwise short tiles + re-swept aprons have less latency)
    cp.two_pass = 0;
    cp.pfd = tune_int("B200CFD_CLEAN_PFD", CPFD);
    if (cp.pfd < 1) cp.pfd = 1;
    // OFF by default — measured on B200: parity-green, but no faster (4096x1024: 60.7 + 68.8 us against 56.6 + 71.5 us for
    // the two sweeps; 16384^2: 1.59 + 1.91 ms against 1.40 + 2.06 ms).  The sweep is bound by the latency of the scan chain
    // per row at 2 warps per scheduler, and a correction row has the same chain as a full row.
    if (tune_int("B200CFD_CLEAN_TWO_PASS", 0)) {
        int th2 = th > Lv + 1 ? th : Lv + 1;
        if (th2 > nrows) th2 = nrows > 0 ? nrows : 1;
        const int nt2 = (nrows + th2 - 1) / th2;
        if (nt2 > 1 && (long long)cp.n_strips * nt2 >= 4ll * n_sm) { cp.two_pass = 1; cp.tile_h = th2; cp.n_tiles = nt2; }
    }
    return cp;
}

template <typename T>
static void launch_clean_sweep_t(const Geom& g, const CleanPlan& cp, const void* u, const void* v, const void* psi_old,
                                 void* psi_new, void* uo, void* vo, int nblk, cudaStream_t s) {
    // vector path: the window origin 1 + s*strip_w - apron_h is a multiple of 4 by construction (clean_plan)
    const size_t al = 4 * sizeof(T);
    const bool vec = g.pitch % 4 == 0 && (cp.j_first - cp.apron_h) % 4 == 0 && cp.strip_w % 4 == 0 &&
                     ((size_t)u % al) == 0 && ((size_t)v % al) == 0 && ((size_t)psi_old % al) == 0 && ((size_t)psi_new % al) == 0;
    const T *up = (const T*)u, *vp = (const T*)v, *po = (const T*)psi_old;
    T *pn = (T*)psi_new, *uop = (T*)uo, *vop = (T*)vo;
    // measured on B200: the scalar form is the fastest (the kernel is latency-, not L1-wavefront-bound); 1 / 2 kept as knobs
    const int mode = vec ? tune_int("B200CFD_CLEAN_VEC", 0) : 0;
#define SWEEP_LAUNCH(V, A)                                                                                          \
    do {                                                                                                            \
        if (po) launch_pdl(k_clean_sweep<T, true, V, A>, dim3(nblk), dim3(32 * CWARPS), s, g, cp, up, vp, po, pn, uop, vop);         \
        else launch_pdl(k_clean_sweep<T, false, V, A>, dim3(nblk), dim3(32 * CWARPS), s, g, cp, up, vp, (const T*)nullptr, pn, uop, vop);      \
    } while (0)
    // lean kernel (default): one-pass plan, 4-aligned rows of a length that is a multiple of 4; B200CFD_CLEAN_LEAN=0
    // falls back to the general kernel (which also serves every other shape)
    static const int lean_on = tune_int("B200CFD_CLEAN_LEAN", 1);
    if (uo && vo) {
        SWEEP_LAUNCH(0, true);                 // (callers only fuse the correction with a one-pass plan)
    } else if (lean_on && vec && !cp.two_pass && g.nx % 4 == 0 && g.nx >= 4) {
        static const int unr = tune_int("B200CFD_CLEAN_UNROLL", 6);    // developer knobs: 2 or 6 rows per loop trip,
        static const int minb = tune_int("B200CFD_CLEAN_MINB", 3);     // 3 or 4 resident blocks per SM (register cap)
#define LEAN_LAUNCH(U, B)                                                                                           \
    do {                                                                                                            \
        if (po) launch_pdl(k_clean_sweep_lean<T, true, U, B>, dim3(nblk), dim3(32 * CWARPS), s, g, cp, up, vp, po, pn); \
        else launch_pdl(k_clean_sweep_lean<T, false, U, B>, dim3(nblk), dim3(32 * CWARPS), s, g, cp, up, vp, (const T*)nullptr, pn); \
    } while (0)
        if (unr == 2) { if (minb == 4) LEAN_LAUNCH(2, 4); else LEAN_LAUNCH(2, 3); }
        else { if (minb == 4) LEAN_LAUNCH(6, 4); else LEAN_LAUNCH(6, 3); }
#undef LEAN_LAUNCH
    } else {
        if (mode == 1) SWEEP_LAUNCH(1, false);
        else if (mode == 2) SWEEP_LAUNCH(2, false);
        else SWEEP_LAUNCH(0, false);
    }
#undef SWEEP_LAUNCH
}

// uo, vo != NULL: also write the corrected velocity u,v - grad(psi_new) there (complete fields, borders copied)
void launch_clean_sweep(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v,
                        const void* psi_old, void* psi_new, void* uo, void* vo, cudaStream_t s) {
    const int ntile = cp.n_strips * cp.n_tiles;
    const int nblk = (ntile + CWARPS - 1) / CWARPS;
    if (nblk <= 0) return;
    // default: rows staged by TMA (cleandiv_tma.cu); falls through to the register-marching kernels when not applicable
    if (!uo && !vo && launch_clean_sweep_tma(dtype, g, cp, u, v, psi_old, psi_new, psi_old ? 1 : 0, s)) return;
    if (dtype == 0) launch_clean_sweep_t<double>(g, cp, u, v, psi_old, psi_new, uo, vo, nblk, s);
    else launch_clean_sweep_t<float>(g, cp, u, v, psi_old, psi_new, uo, vo, nblk, s);
    if (cp.two_pass && cp.n_tiles > 1) {
        const int nb2 = (cp.n_strips * (cp.n_tiles - 1) + CWARPS - 1) / CWARPS;
        if (dtype == 0) k_clean_corr<double><<<nb2, 32 * CWARPS, 0, s>>>(g, cp, (double*)psi_new);
        else k_clean_corr<float><<<nb2, 32 * CWARPS, 0, s>>>(g, cp, (float*)psi_new);
    }
}
int clean_sweep_launches(const CleanPlan& cp) { return (cp.two_pass && cp.n_tiles > 1) ? 2 : 1; }

void launch_clean_apply(int dtype, const Geom& g, void* u, void* v, const void* psi, cudaStream_t s) {
    dim3 grid((g.nx + 63) / 64, (g.ny + 3) / 4), block(64, 4);
    if (dtype == 0) launch_pdl(k_clean_apply<double>, grid, block, s, g, (double*)u, (double*)v, (const double*)psi);
    else launch_pdl(k_clean_apply<float>, grid, block, s, g, (float*)u, (float*)v, (const float*)psi);
}

}  // namespace b200cfd
