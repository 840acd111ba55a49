// cellmath.cuh — per-cell arithmetic of the projection step shared by the general kernels (step_kernels.cu) and the
// row-marching fast kernels (step_fast.cu).  Reference: python/flow_over_cylinder (Fischer)/v5.py:112-176,228-237,349-360.
// All arithmetic is done in double after the load; values the reference materialises as arrays of the field dtype are
// rounded to T at the same points (Rnd<T>), so fp32 storage follows the reference too.
#pragma once
#include "kernels.h"

namespace b200cfd {

template <typename T> __device__ __forceinline__ double ld(const T* p, long long k) { return (double)p[k]; }

// interior in the GLOBAL grid and with all four neighbours present locally
__device__ __forceinline__ bool is_interior(const Geom& g, int i, int j) {
    const int jg = g.jg0 + j;
    return i >= 1 && i < g.ny - 1 && j >= 1 && j < g.nx - 1 && jg >= 1 && jg < g.nxg - 1;
}

// ------------------------------------------------------------------------------------------------
// shared per-cell math
// ------------------------------------------------------------------------------------------------
// v5.py:149-162
template <typename T>
__device__ __forceinline__ double tau_cell(const Geom& g, double uc, double vc, double dt, double nu_eff) {
    const double vel = sqrt(uc * uc + vc * vc);
    double t;
    if (vel > 1e-10) {
        const double Pe = vel * g.hmin / (nu_eff + 1e-10);
        t = g.hmin / (2 * vel) * fmin(1.0, Pe / 2.0);
    } else {
        t = dt / 2;
    }
    return Rnd<T>::r(t);
}
// v5.py:127-147
template <typename T>
__device__ __forceinline__ double conv_supg_cell(const Geom& g, double uc, double vc, double pc, double pe, double pw,
                                                 double pn, double ps, double tau) {
    double c = uc * ((pe - pw) * g.qdx) + vc * ((pn - ps) * g.qdy);
    if (tau > 0) c -= tau * (uc * ((pe - 2 * pc + pw) * g.hdx2) + vc * ((pn - 2 * pc + ps) * g.hdy2));
    return Rnd<T>::r(c);
}
// v5.py:112-125
template <typename T>
__device__ __forceinline__ double conv_upwind_cell(const Geom& g, double uc, double vc, double pc, double pe, double pw,
                                                   double pn, double ps) {
    const double dpx = uc > 0 ? (pc - pw) * g.rdx : (pe - pc) * g.rdx;
    const double dpy = vc > 0 ? (pc - ps) * g.rdy : (pn - pc) * g.rdy;
    return Rnd<T>::r(uc * dpx + vc * dpy);
}
// v5.py:164-176
template <typename T>
__device__ __forceinline__ double lap_cell(const Geom& g, double nu_eff, double pc, double pe, double pw, double pn, double ps) {
    return Rnd<T>::r(nu_eff * ((pe - 2 * pc + pw) * g.ax + (pn - 2 * pc + ps) * g.ay));
}

// inflow profile value, v5.py:351-353
__device__ __forceinline__ double inflow_u(double V_inf, double y_max, double yi, double pert_scale, double phase) {
    return V_inf * (1 + pert_scale * sin(2 * M_PI * yi / y_max + phase));
}

template <typename T>
__device__ __forceinline__ void predict_cell(const Geom& g, const Phys& p, const T* __restrict__ u,
                                             const T* __restrict__ v, const T* __restrict__ nu_t, long long k,
                                             double dt, double& ru, double& rv, double& tau_o) {
    const double uc = ld(u, k), vc = ld(v, k);
    const double ue = ld(u, k + 1), uw = ld(u, k - 1), un = ld(u, k + g.pitch), us = ld(u, k - g.pitch);
    const double ve = ld(v, k + 1), vw = ld(v, k - 1), vn = ld(v, k + g.pitch), vs = ld(v, k - g.pitch);
    const double ne = nu_t ? Rnd<T>::r(p.nu_eff + ld(nu_t, k)) : p.nu_eff;
    double cu, cv;
    if (p.use_supg) {
        const double tau = tau_cell<T>(g, uc, vc, dt, ne);
        tau_o = tau;
        cu = conv_supg_cell<T>(g, uc, vc, uc, ue, uw, un, us, tau);
        cv = conv_supg_cell<T>(g, uc, vc, vc, ve, vw, vn, vs, tau);
    } else {
        tau_o = 0.0;
        cu = conv_upwind_cell<T>(g, uc, vc, uc, ue, uw, un, us);
        cv = conv_upwind_cell<T>(g, uc, vc, vc, ve, vw, vn, vs);
    }
    const double lu = lap_cell<T>(g, ne, uc, ue, uw, un, us);
    const double lv = lap_cell<T>(g, ne, vc, ve, vw, vn, vs);
    ru = axpy_pred<T>(uc, dt, cu, lu);
    rv = axpy_pred<T>(vc, dt, cv, lv);
}

// Field accessors for K5: the velocity entering the finalize stage is either stored (PlainVel) or still has the last
// clean-divergence correction u,v -= grad(psi) pending (ApplyVel: v5.py:254-256 evaluated on the fly, same rounding as
// k_clean_apply), which saves one full read + write of u, v per step.
template <typename T>
struct PlainVel {
    const T* __restrict__ ua; const T* __restrict__ va;
    __device__ __forceinline__ double u(const Geom& g, int i, int j) const { return ld(ua, i * g.pitch + j); }
    __device__ __forceinline__ double v(const Geom& g, int i, int j) const { return ld(va, i * g.pitch + j); }
};
template <typename T>
struct ApplyVel {
    const T* __restrict__ ua; const T* __restrict__ va; const T* __restrict__ psi;
    __device__ __forceinline__ double u(const Geom& g, int i, int j) const {
        const long long k = i * g.pitch + j;
        const double a = ld(ua, k);
        if (!is_interior(g, i, j)) return a;
        return Rnd<T>::r(a - Rnd<T>::r((ld(psi, k + 1) - ld(psi, k - 1)) * g.hdx));
    }
    __device__ __forceinline__ double v(const Geom& g, int i, int j) const {
        const long long k = i * g.pitch + j;
        const double a = ld(va, k);
        if (!is_interior(g, i, j)) return a;
        return Rnd<T>::r(a - Rnd<T>::r((ld(psi, k + g.pitch) - ld(psi, k - g.pitch)) * g.hdy));
    }
};

// BC (v5.py:349-360) + IBM (v5.py:228-237) of the final velocity at one cell, pointwise in the entering velocity
template <typename T, typename A>
__device__ __forceinline__ void bcibm_cell(const Geom& g, const Phys& p, const StepCtl* ctl, const A& acc,
                                           const double* __restrict__ ibm, const double* __restrict__ y, int i, int j,
                                           double& fu, double& fv) {
    const int jg = g.jg0 + j;
    const long long k = i * g.pitch + j;
    if (i == 0 || i == g.ny - 1) { fu = 0.0; fv = 0.0; }
    else if (jg == 0) { fu = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[i], ctl->pert_scale, ctl->phase)); fv = 0.0; }
    else if (jg == g.nxg - 1) {
        if (jg - 1 == 0) { fu = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[i], ctl->pert_scale, ctl->phase)); fv = 0.0; }
        else { fu = acc.u(g, i, j - 1); fv = acc.v(g, i, j - 1); }
    } else { fu = acc.u(g, i, j); fv = acc.v(g, i, j); }
    const double m = ibm[k];
    if (m > 0) {
        fu = Rnd<T>::r(fu * (1.0 - m * ctl->fs));
        fv = Rnd<T>::r(fv * (1.0 - m * ctl->fs));
    }
}

// ------------------------------------------------------------------------------------------------
// K1 at one cell, any position: u*, v* = IBM(BC(u + dt(-C + L)))            v5.py:388-407
// The thread of the global outflow column evaluates the predictor at column nx-2 (BC copies the already predicted
// neighbour, v5.py:355-356); wall rows are 0; the global inflow column takes the perturbed profile.  The outermost
// local column of a slab that is not a global edge has no stencil: it keeps its input (it is a halo cell).
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void predictor_point(const Geom& g, const Phys& p, const StepCtl* __restrict__ ctl,
                                                const T* __restrict__ u, const T* __restrict__ v, T* __restrict__ us,
                                                T* __restrict__ vs, T* __restrict__ tau_out, const T* __restrict__ nu_t,
                                                const double* __restrict__ ibm, const double* __restrict__ y, int i, int j) {
    const int jg = g.jg0 + j;
    const long long k = i * g.pitch + j;
    const bool wall = (i == 0 || i == g.ny - 1);
    const bool inflow = (jg == 0), outflow = (jg == g.nxg - 1);
    const bool ledge = !inflow && !outflow && (j == 0 || j == g.nx - 1);
    const double dt = ctl->dt, fs = ctl->fs;
    double ru, rv, tau_o = 0.0;
    if (wall) {
        ru = 0.0; rv = 0.0;
    } else if (ledge) {
        ru = ld(u, k); rv = ld(v, k);
    } else if (inflow) {
        ru = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[i], ctl->pert_scale, ctl->phase));
        rv = 0.0;
    } else if (outflow) {
        double t;
        if (jg - 1 == 0) { ru = Rnd<T>::r(inflow_u(p.V_inf, p.y_max, y[i], ctl->pert_scale, ctl->phase)); rv = 0.0; }
        else if (j - 1 == 0) { ru = ld(u, k - 1); rv = ld(v, k - 1); }     // 2-column slab remnant: no stencil either
        else predict_cell<T>(g, p, u, v, nu_t, k - 1, dt, ru, rv, t);
    } else {
        predict_cell<T>(g, p, u, v, nu_t, k, dt, ru, rv, tau_o);
    }
    const double m = ibm[k];
    if (m > 0) {
        ru = Rnd<T>::r(ru * (1.0 - m * fs));
        rv = Rnd<T>::r(rv * (1.0 - m * fs));
    }
    us[k] = (T)ru;
    vs[k] = (T)rv;
    if (tau_out) tau_out[k] = (T)tau_o;
}

// column classes of a column (fast kernels)
enum { COL_INTERIOR = 0, COL_INFLOW = 1, COL_OUTFLOW = 2, COL_COPY = 3 };
__device__ __forceinline__ int col_class(const Geom& g, int c) {
    const int cg = g.jg0 + c;
    if (cg == 0) return COL_INFLOW;
    if (cg == g.nxg - 1) return COL_OUTFLOW;
    if (c == 0 || c == g.nx - 1) return COL_COPY;        // outermost halo column of a slab: no stencil
    return COL_INTERIOR;
}

// ------------------------------------------------------------------------------------------------
// K5 at one cell.  Reads ua, va (after clean_divergence), writes u, v (different buffers):
//   max|div(ua,va)| (v5.py:421-422) -> BC (v5.py:424) -> IBM (:425) -> nanmax|vorticity| (:427-428)
//   -> nanmean energy (:431-432) -> clip (:437-438) -> post-clip max|u|,|v| for the next adaptive dt.
// BC+IBM are pointwise in (ua, va), so the vorticity stencil re-evaluates them at the 4 neighbours.
// Every local cell is computed (halo columns redundantly); only OWNED columns enter the reductions.
// ------------------------------------------------------------------------------------------------
struct FinAcc {
    double dmax = 0.0, wmax = 0.0, e = 0.0, mu = 0.0, mv = 0.0;
    unsigned cnt = 0, bad = 0;
};

template <typename T, typename A>
__device__ __forceinline__ void finalize_point(const Geom& g, const Phys& p, const StepCtl* __restrict__ ctl,
                                               const A& acc, T* __restrict__ u,
                                               T* __restrict__ v, const double* __restrict__ ibm,
                                               const uint8_t* __restrict__ mask, const double* __restrict__ y, int i, int j,
                                               FinAcc& a) {
    const bool own = j >= g.own0 && j < g.own1;
    const long long k = i * g.pitch + j;
    double fu, fv;
    bcibm_cell<T>(g, p, ctl, acc, ibm, y, i, j, fu, fv);
    if (own && is_interior(g, i, j)) {
        a.dmax = fmax(a.dmax, fabs(Rnd<T>::r((acc.u(g, i, j + 1) - acc.u(g, i, j - 1)) * g.hdx + (acc.v(g, i + 1, j) - acc.v(g, i - 1, j)) * g.hdy)));
        if (!(mask && mask[k])) {
            double eu, ev, wu, wv, nu_, nv, su, sv;
            bcibm_cell<T>(g, p, ctl, acc, ibm, y, i, j + 1, eu, ev);
            bcibm_cell<T>(g, p, ctl, acc, ibm, y, i, j - 1, wu, wv);
            bcibm_cell<T>(g, p, ctl, acc, ibm, y, i + 1, j, nu_, nv);
            bcibm_cell<T>(g, p, ctl, acc, ibm, y, i - 1, j, su, sv);
            const double w = Rnd<T>::r(Rnd<T>::r(Rnd<T>::r(ev - wv) / Rnd<T>::r(g.two_dx)) -
                                       Rnd<T>::r(Rnd<T>::r(nu_ - su) / Rnd<T>::r(g.two_dy)));
            a.wmax = fmax(a.wmax, fabs(w));   // NaN is dropped by fmax (nanmax)
        }
    }
    const double en = Rnd<T>::r(0.5 * Rnd<T>::r(Rnd<T>::r(fu * fu) + Rnd<T>::r(fv * fv)));
    if (own && en == en) { a.e += en; a.cnt += 1; }
    if (own && (!isfinite(fu) || !isfinite(fv))) a.bad = 1;
    const double cu = fmin(fmax(fu, -p.max_velocity), p.max_velocity);
    const double cv = fmin(fmax(fv, -p.max_velocity), p.max_velocity);
    // np.clip propagates NaN; fmin/fmax would drop it
    u[k] = (T)(fu == fu ? cu : fu);
    v[k] = (T)(fv == fv ? cv : fv);
    if (own) { a.mu = fmax(a.mu, fabs(cu)); a.mv = fmax(a.mv, fabs(cv)); }
}

// one block's accumulators -> the step's running maxima (one atomicMax each) and the block's energy partial
__device__ __forceinline__ void finalize_block_reduce(const FinAcc& a, StepCtl* __restrict__ ctl, double* __restrict__ ke_part,
                                                      unsigned long long* __restrict__ cnt_part, int block_id) {
    block_atomic_max(a.dmax, &ctl->max_div_post);
    block_atomic_max(a.wmax, &ctl->max_vort);
    block_atomic_max(a.mu, &ctl->max_u);
    block_atomic_max(a.mv, &ctl->max_v);
    const double bs = block_sum(a.e);
    const double bc = block_sum((double)a.cnt);
    const int nbad = __syncthreads_or((int)a.bad);
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        ke_part[block_id] = bs;
        cnt_part[block_id] = (unsigned long long)(bc + 0.5);
        if (nbad) atomicOr(&ctl->nonfinite, 1u);
    }
}

// The same for 1-D blocks with ONE barrier instead of thirteen: every warp reduces its seven values with shuffles and
// leaves them in shared memory, warp 0 folds the warps.  The energy and count sums use the tree of block_sum (warp_sum of
// the lanes, then warp_sum of the per-warp values padded with zeros), so the partial a block writes has the same bits.
__device__ __forceinline__ void finalize_block_reduce1(const FinAcc& a, StepCtl* __restrict__ ctl, double* __restrict__ ke_part,
                                                       unsigned long long* __restrict__ cnt_part, int block_id) {
    __shared__ double s_r[6][32];
    __shared__ unsigned s_bad[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    const double m0 = warp_max(a.dmax), m1 = warp_max(a.wmax), m2 = warp_max(a.mu), m3 = warp_max(a.mv);
    const double e = warp_sum(a.e), c = warp_sum((double)a.cnt);
    const unsigned bad = __any_sync(0xffffffffu, a.bad != 0) ? 1u : 0u;
    if (lane == 0) { s_r[0][w] = m0; s_r[1][w] = m1; s_r[2][w] = m2; s_r[3][w] = m3; s_r[4][w] = e; s_r[5][w] = c; s_bad[w] = bad; }
    __syncthreads();
    if (w != 0) return;
    const bool in = lane < nw;
    const double M0 = warp_max(in ? s_r[0][lane] : 0.0), M1 = warp_max(in ? s_r[1][lane] : 0.0);
    const double M2 = warp_max(in ? s_r[2][lane] : 0.0), M3 = warp_max(in ? s_r[3][lane] : 0.0);
    const double E = warp_sum(in ? s_r[4][lane] : 0.0), Cn = warp_sum(in ? s_r[5][lane] : 0.0);
    const bool nbad = __any_sync(0xffffffffu, in && s_bad[lane] != 0);
    if (lane == 0) {
        if (M0 > 0.0) atomicMax(&ctl->max_div_post, (unsigned long long)__double_as_longlong(M0));
        if (M1 > 0.0) atomicMax(&ctl->max_vort, (unsigned long long)__double_as_longlong(M1));
        if (M2 > 0.0) atomicMax(&ctl->max_u, (unsigned long long)__double_as_longlong(M2));
        if (M3 > 0.0) atomicMax(&ctl->max_v, (unsigned long long)__double_as_longlong(M3));
        ke_part[block_id] = E;
        cnt_part[block_id] = (unsigned long long)(Cn + 0.5);
        if (nbad) atomicOr(&ctl->nonfinite, 1u);
    }
}

// ------------------------------------------------------------------------------------------------
// K2 at one cell, any position: D = div(u*, v*) and exactly ONE red-black iteration from phi = 0 ("as executed",
// SURVEY finding 3).  Odd cells: phi = (D/dt_c) d.  Even cells: phi = (a(pE+pW) + b(pN+pS) + D/dt_c) d with the four
// (odd) neighbours evaluated in place from D at those cells.  Masked / border cells stay 0.   v5.py:178-187,202-226
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ double div_at(const Geom& g, const T* __restrict__ us, const T* __restrict__ vs, int i, int j) {
    // 0 outside the global interior (compute_divergence_fast leaves the border 0)
    if (i < 1 || i >= g.ny - 1 || g.jg0 + j < 1 || g.jg0 + j >= g.nxg - 1 || j < 1 || j >= g.nx - 1) return 0.0;
    const long long k = i * g.pitch + j;
    return Rnd<T>::r((ld(us, k + 1) - ld(us, k - 1)) * g.hdx + (ld(vs, k + g.pitch) - ld(vs, k - g.pitch)) * g.hdy);
}
template <typename T>
__device__ __forceinline__ double odd_phi_at(const Geom& g, const T* __restrict__ us, const T* __restrict__ vs,
                                             const uint8_t* __restrict__ mask, double inv_dt, int i, int j) {
    // value after the first colour of an odd cell, 0 if it is border / masked
    if (i < 1 || i >= g.ny - 1 || g.jg0 + j < 1 || g.jg0 + j >= g.nxg - 1) return 0.0;
    if (mask && mask[i * g.pitch + j]) return 0.0;
    const double rhs = -div_at<T>(g, us, vs, i, j) * inv_dt;
    return Rnd<T>::r((g.ax * 0.0 + g.ay * 0.0 - rhs) * g.dinv);
}
template <typename T>
__device__ __forceinline__ void divrb1_point(const Geom& g, const T* __restrict__ us, const T* __restrict__ vs,
                                             const uint8_t* __restrict__ mask, T* __restrict__ phi, T* __restrict__ div_out,
                                             double inv_dt, int i, int j, double& dabs, double& ch) {
    const long long k = i * g.pitch + j;
    const int jg = g.jg0 + j;
    const bool interior = i >= 1 && i < g.ny - 1 && jg >= 1 && jg < g.nxg - 1 && j >= 1 && j < g.nx - 1;
    double D = 0.0, p = 0.0;
    if (interior) {
        D = div_at<T>(g, us, vs, i, j);
        if (!(mask && mask[k])) {
            const double rhs = -D * inv_dt;
            if ((i + jg) & 1) p = (g.ax * 0.0 + g.ay * 0.0 - rhs) * g.dinv;
            else {
                const double pe = odd_phi_at<T>(g, us, vs, mask, inv_dt, i, j + 1);
                const double pw = odd_phi_at<T>(g, us, vs, mask, inv_dt, i, j - 1);
                const double pn = odd_phi_at<T>(g, us, vs, mask, inv_dt, i + 1, j);
                const double ps = odd_phi_at<T>(g, us, vs, mask, inv_dt, i - 1, j);
                p = (g.ax * (pe + pw) + g.ay * (pn + ps) - rhs) * g.dinv;
            }
        }
    }
    phi[k] = (T)p;
    if (div_out) div_out[k] = (T)D;
    if (j >= g.own0 && j < g.own1) { dabs = fmax(dabs, fabs(D)); ch = fmax(ch, fabs(p)); }
}

}  // namespace b200cfd
