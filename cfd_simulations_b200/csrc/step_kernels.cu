// step_kernels.cu — momentum predictor, divergence, correction, finalize (BC + IBM + diagnostics) and the
// device-resident step control.  Reference: python/flow_over_cylinder (Fischer)/v5.py:96-200,228-237,316-441.
// All arithmetic is done in double after the load; values the reference materialises as arrays of the
// field dtype are rounded to T at the same points (Rnd<T>), so fp32 storage follows the reference too.
#include "kernels.h"
#include "cellmath.cuh"

namespace b200cfd {

#define BX 64
#define BY 4
static inline dim3 grid2d(const Geom& g) { return dim3((g.nx + BX - 1) / BX, (g.ny + BY - 1) / BY); }
static const dim3 kBlock(BX, BY);

// ------------------------------------------------------------------------------------------------
// unit kernels
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void k_supg_tau(Geom g, const T* __restrict__ u, const T* __restrict__ v, T* __restrict__ tau, double dt,
                           double nu_eff, const T* __restrict__ nu_t) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double out = 0.0;
    if (is_interior(g, i, j)) {
        const double ne = nu_t ? Rnd<T>::r(nu_eff + ld(nu_t, k)) : nu_eff;
        out = tau_cell<T>(g, ld(u, k), ld(v, k), dt, ne);
    }
    tau[k] = (T)out;
}

template <typename T>
__global__ void k_convection(Geom g, const T* __restrict__ u, const T* __restrict__ v, const T* __restrict__ phi,
                             const T* __restrict__ tau, T* __restrict__ conv, int supg) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double out = 0.0;
    if (is_interior(g, i, j)) {
        const double pc = ld(phi, k), pe = ld(phi, k + 1), pw = ld(phi, k - 1), pn = ld(phi, k + g.pitch), ps = ld(phi, k - g.pitch);
        out = supg ? conv_supg_cell<T>(g, ld(u, k), ld(v, k), pc, pe, pw, pn, ps, ld(tau, k))
                   : conv_upwind_cell<T>(g, ld(u, k), ld(v, k), pc, pe, pw, pn, ps);
    }
    conv[k] = (T)out;
}

template <typename T>
__global__ void k_laplacian(Geom g, const T* __restrict__ phi, T* __restrict__ lap, double nu_eff, const T* __restrict__ nu_t) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double out = 0.0;
    if (is_interior(g, i, j)) {
        const double ne = nu_t ? Rnd<T>::r(nu_eff + ld(nu_t, k)) : nu_eff;
        out = lap_cell<T>(g, ne, ld(phi, k), ld(phi, k + 1), ld(phi, k - 1), ld(phi, k + g.pitch), ld(phi, k - g.pitch));
    }
    lap[k] = (T)out;
}

// v5.py:178-187 (+ fused max|div| for the per-step log line v5.py:410)
template <typename T>
__global__ void k_divergence(Geom g, const T* __restrict__ u, const T* __restrict__ v, T* __restrict__ div,
                             unsigned long long* max_out) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    double out = 0.0;
    if (i < g.ny && j < g.nx) {
        const long long k = i * g.pitch + j;
        if (is_interior(g, i, j))
            out = Rnd<T>::r((ld(u, k + 1) - ld(u, k - 1)) * g.hdx + (ld(v, k + g.pitch) - ld(v, k - g.pitch)) * g.hdy);
        div[k] = (T)out;
    }
    if (max_out) block_atomic_max(fabs(out), max_out);
}

// v5.py:189-200
template <typename T>
__global__ void k_gradient(Geom g, const T* __restrict__ phi, T* __restrict__ gx, T* __restrict__ gy) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double ox = 0.0, oy = 0.0;
    if (is_interior(g, i, j)) {
        ox = (ld(phi, k + 1) - ld(phi, k - 1)) * g.hdx;
        oy = (ld(phi, k + g.pitch) - ld(phi, k - g.pitch)) * g.hdy;
    }
    gx[k] = (T)ox;
    gy[k] = (T)oy;
}

// v5.py:96-110 (forward differences)
template <typename T>
__global__ void k_smagorinsky(Geom g, const T* __restrict__ u, const T* __restrict__ v, T* __restrict__ nu_t, double csd2) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double out = 0.0;
    if (is_interior(g, i, j)) {
        const double uc = ld(u, k), vc = ld(v, k);
        const double dudx = (ld(u, k + 1) - uc) / g.dx, dudy = (ld(u, k + g.pitch) - uc) / g.dy;
        const double dvdx = (ld(v, k + 1) - vc) / g.dx, dvdy = (ld(v, k + g.pitch) - vc) / g.dy;
        out = csd2 * sqrt(2 * (dudx * dudx + dvdy * dvdy) + (dudy + dvdx) * (dudy + dvdx));
    }
    nu_t[k] = (T)out;
}

// v5.py:349-360: one thread per border cell; order inflow, outflow, walls (corners end as 0)
template <typename T>
__global__ void k_apply_bc(Geom g, T* __restrict__ u, T* __restrict__ v, const double* __restrict__ y, double V_inf,
                           double y_max, double pert_scale, double phase) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < g.ny) {
        const int i = t;
        const bool wall = (i == 0 || i == g.ny - 1);
        if (g.jg0 == 0) {
            u[i * g.pitch] = wall ? (T)0 : (T)inflow_u(V_inf, y_max, y[i], pert_scale, phase);
            v[i * g.pitch] = (T)0;
        }
        if (g.jg0 + g.nx == g.nxg) {
            u[i * g.pitch + g.nx - 1] = wall ? (T)0 : u[i * g.pitch + g.nx - 2];
            v[i * g.pitch + g.nx - 1] = wall ? (T)0 : v[i * g.pitch + g.nx - 2];
        }
    } else if (t < g.ny + g.nx) {
        const int j = t - g.ny;
        u[j] = (T)0; v[j] = (T)0;
        u[(long long)(g.ny - 1) * g.pitch + j] = (T)0;
        v[(long long)(g.ny - 1) * g.pitch + j] = (T)0;
    }
}

// v5.py:228-237
template <typename T>
__global__ void k_apply_ibm(Geom g, T* __restrict__ u, T* __restrict__ v, const double* __restrict__ ibm, double fs) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    const double m = ibm[k];
    if (m > 0) {
        u[k] = (T)(ld(u, k) * (1.0 - m * fs));
        v[k] = (T)(ld(v, k) * (1.0 - m * fs));
    }
}


// ------------------------------------------------------------------------------------------------
// K1: fused momentum predictor  u*, v* = IBM(BC(u + dt(-C + L)))            v5.py:388-407
// One thread per cell.  The thread of the global outflow column evaluates the predictor at column
// nx-2 (BC copies the already predicted neighbour, v5.py:355-356); wall rows are 0; the global inflow
// column takes the perturbed profile.  Local edge columns of a slab that are not global edges are halo
// columns: they are not written here (the halo exchange fills them).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(BX* BY)
k_predictor(Geom g, Phys p, const StepCtl* __restrict__ ctl, const T* __restrict__ u, const T* __restrict__ v,
            T* __restrict__ us, T* __restrict__ vs, T* __restrict__ tau_out, const T* __restrict__ nu_t,
            const double* __restrict__ ibm, const double* __restrict__ y) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    predictor_point<T>(g, p, ctl, u, v, us, vs, tau_out, nu_t, ibm, y, i, j);
}

// ------------------------------------------------------------------------------------------------
// K3: u = u* - dt grad(phi), fused max|grad phi|                              v5.py:413-417
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(BX* BY)
k_correct(Geom g, StepCtl* __restrict__ ctl, const T* __restrict__ us, const T* __restrict__ vs,
          const T* __restrict__ phi, T* __restrict__ u, T* __restrict__ v) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    double gm = 0.0;
    if (i < g.ny && j < g.nx) {
        const long long k = i * g.pitch + j;
        const double dt = ctl->dt;
        double gx = 0.0, gy = 0.0;
        if (is_interior(g, i, j)) {
            gx = Rnd<T>::r((ld(phi, k + 1) - ld(phi, k - 1)) * g.hdx);
            gy = Rnd<T>::r((ld(phi, k + g.pitch) - ld(phi, k - g.pitch)) * g.hdy);
        }
        u[k] = (T)axmy<T>(ld(us, k), dt, gx);
        v[k] = (T)axmy<T>(ld(vs, k), dt, gy);
        if (j >= g.own0 && j < g.own1) gm = sqrt(Rnd<T>::r(Rnd<T>::r(gx * gx) + Rnd<T>::r(gy * gy)));
    }
    block_atomic_max(gm, &ctl->max_grad);
}

// ------------------------------------------------------------------------------------------------
// K5: finalize.  Reads ua, va (after clean_divergence), writes u, v (different buffers):
//   max|div(ua,va)| (v5.py:421-422) -> BC (v5.py:424) -> IBM (:425) -> nanmax|vorticity| (:427-428)
//   -> nanmean energy (:431-432) -> clip (:437-438) -> post-clip max|u|,|v| for the next adaptive dt.
// BC+IBM are pointwise in (ua, va), so the vorticity stencil re-evaluates them at the 4 neighbours.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(BX* BY)
k_finalize(Geom g, Phys p, StepCtl* __restrict__ ctl, const T* __restrict__ ua, const T* __restrict__ va,
           T* __restrict__ u, T* __restrict__ v, const double* __restrict__ ibm, const uint8_t* __restrict__ mask,
           const double* __restrict__ y, double* __restrict__ ke_part, unsigned long long* __restrict__ cnt_part) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    FinAcc acc;
    const PlainVel<T> vel{ua, va};
    if (i < g.ny && j < g.nx) finalize_point<T>(g, p, ctl, vel, u, v, ibm, mask, y, i, j, acc);
    finalize_block_reduce(acc, ctl, ke_part, cnt_part, blockIdx.y * gridDim.x + blockIdx.x);
}

// max|u|, max|v| of the bound state (set_step / first step: what adaptive_time_step reads, v5.py:322)
template <typename T>
__global__ void k_max_abs2(Geom g, const T* __restrict__ u, const T* __restrict__ v, StepCtl* ctl) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    double mu = 0.0, mv = 0.0;
    if (i < g.ny && j >= g.own0 && j < g.own1) {
        mu = fabs(ld(u, i * g.pitch + j)); mv = fabs(ld(v, i * g.pitch + j));
    }
    block_atomic_max(mu, &ctl->max_u);
    block_atomic_max(mv, &ctl->max_v);
}

// monitor_simulation_health (v5.py:599-613): non-finite flag, max|u|, max|v|, max|div(u,v)| of the state
// out[0..3] = {nonfinite, max|u|, max|v|, max|div|} as order-preserving bit patterns (zeroed by the launcher)
template <typename T>
__global__ void __launch_bounds__(BX* BY)
k_health(Geom g, const T* __restrict__ u, const T* __restrict__ v, unsigned long long* __restrict__ out) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    double mu = 0.0, mv = 0.0, md = 0.0;
    int bad = 0;
    if (i < g.ny && j >= g.own0 && j < g.own1) {
        const long long k = i * g.pitch + j;
        const double a = ld(u, k), b = ld(v, k);
        bad = !isfinite(a) || !isfinite(b);
        mu = fabs(a); mv = fabs(b);
        if (is_interior(g, i, j))
            md = fabs(Rnd<T>::r((ld(u, k + 1) - ld(u, k - 1)) * g.hdx + (ld(v, k + g.pitch) - ld(v, k - g.pitch)) * g.hdy));
    }
    block_atomic_max(mu, out + 1);
    block_atomic_max(mv, out + 2);
    block_atomic_max(md, out + 3);
    if (__syncthreads_or(bad) && threadIdx.x == 0 && threadIdx.y == 0) atomicMax(out, 0x3ff0000000000000ull);  // 1.0
}

// ------------------------------------------------------------------------------------------------
// device-resident step control                                                v5.py:316-326,351-352,406
// ------------------------------------------------------------------------------------------------
__device__ void ctl_prepare(StepCtl* c, const CtlParams& cp) {
    const long long step = c->step;
    double dt;
    if (!cp.adaptive_dt) dt = cp.dt_base;
    else if (step < 1000) dt = cp.scalar_f32 ? (double)(float)0.00002 : 0.00002;
    else {
        double vm = fmax(fmax(c->in_max_u, c->in_max_v), 1e-10);
        const double dt_cfl = cp.cfl_target * cp.hmin / vm;
        const double nu_total = cp.is_f32 ? (double)(float)((float)((float)cp.nu + (float)c->in_mean_nut) + (float)cp.av)
                                          : (cp.nu + c->in_mean_nut) + cp.av;
        const double dt_visc = 0.4 * (cp.hmin * cp.hmin) / nu_total;
        if (cp.is_f32) {
            // float32 fields: vel_max and nu_total are np.float32, so under NumPy 2 promotion the Python-float numerators
            // are cast to float32 and both divisions, the min and the clip are float32 operations (v5.py:322-326)
            const float fc = __fdiv_rn((float)(cp.cfl_target * cp.hmin), (float)vm);
            const float fv = __fdiv_rn((float)(0.4 * (cp.hmin * cp.hmin)), (float)nu_total);
            dt = (double)fminf(fmaxf(fminf(fc, fv), (float)cp.dt_min), (float)cp.dt_max);
        } else {
            dt = fmin(dt_cfl, dt_visc);
            dt = fmin(fmax(dt, cp.dt_min), cp.dt_max);
            if (cp.scalar_f32) dt = (double)(float)dt;           // v5: np.float32(...); v3 / v4: float(...)
        }
    }
    c->dt = dt;
    c->fs = fmin(1.0, (double)step / (double)cp.initial_steps);
    c->pert_scale = fmin(1.0, (double)step / 1000.0) * cp.pert_amplitude;
    c->phase = 0.02 * (double)step;
    c->max_div_pre = c->max_grad = c->max_div_post = c->max_vort = c->max_u = c->max_v = 0ull;
    c->nonfinite = 0u;
    c->pois_change = 0ull; c->pois_last = 0.0; c->pois_iters = 0; c->pois_done = 0; c->pois_blocks_done = 0u;
}

__global__ void k_ctl_init(StepCtl* c, CtlParams cp, long long step, int use_measured_max, const double* __restrict__ nut_part) {
    if (threadIdx.x || blockIdx.x) return;
    c->step = step;
    if (use_measured_max) { c->in_max_u = bits_to_double(c->max_u); c->in_max_v = bits_to_double(c->max_v); }
    double nut = 0.0;                                        // LES: mean of the nu_t attribute as it stands (v5.py:323)
    if (cp.use_les && nut_part) for (int k = 0; k < NUT_PARTS; ++k) nut += nut_part[k];
    c->in_mean_nut = nut / cp.n_cells_global;
    ctl_prepare(c, cp);
}

__global__ void k_ctl_zero_max(StepCtl* c) { if (!threadIdx.x && !blockIdx.x) { c->max_u = 0ull; c->max_v = 0ull; } }

// End of a step, part 1: fold this slab's running maxima and the fixed-order sum of its energy partials into the
// reduction vector red[RED_N] (what the slabs all_gather; on one GPU part 2 reads it directly).
// close_rb1: the "as executed" Poisson bookkeeping (what k_rb1_close does) is folded in here, saving a launch per step.
// Written for latency (it is one block on the critical path between two steps: 7-8 us at 4096x1024 before, 13 % of the
// 600x180 step): the partials are loaded eight at a time before they are added (same order of additions, so the same
// bits), thread 0 requests the control block's maxima before the reduction tree instead of after it, and the vector is
// handed on in `loc` as well as stored, so that the closing part does not read it back from memory.
__device__ __forceinline__ void ctl_reduce_local(StepCtl* c, const double* __restrict__ ke_part,
                                                 const unsigned long long* __restrict__ cnt_part, int n_part,
                                                 double* __restrict__ red, int close_rb1, double tol,
                                                 const double* __restrict__ nut_part, double (&loc)[RED_N]) {
    __shared__ double s_e[256];
    __shared__ unsigned long long s_c[256];
    unsigned long long m_pre = 0, m_grad = 0, m_post = 0, m_vort = 0, m_u = 0, m_v = 0, m_pc = 0;
    unsigned nonfinite = 0;
    if (threadIdx.x == 0) {
        m_pre = c->max_div_pre; m_grad = c->max_grad; m_post = c->max_div_post; m_vort = c->max_vort; m_u = c->max_u; m_v = c->max_v;
        m_pc = c->pois_change; nonfinite = c->nonfinite;
    }
    double e = 0.0; unsigned long long n = 0;
    for (int b0 = threadIdx.x; b0 < n_part; b0 += 256 * 8) {
        double ev[8]; unsigned long long nv[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int b = b0 + q * 256;
            ev[q] = b < n_part ? ke_part[b] : 0.0;           // + 0.0 is exact: the sums keep the bits of the plain loop
            nv[q] = b < n_part ? cnt_part[b] : 0ull;
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) { e += ev[q]; n += nv[q]; }
    }
    s_e[threadIdx.x] = e; s_c[threadIdx.x] = n;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s_e[threadIdx.x] += s_e[threadIdx.x + o]; s_c[threadIdx.x] += s_c[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x) return;
    double pois_last = c->pois_last;
    if (close_rb1) {
        const double ch = bits_to_double(m_pc);
        c->pois_last = ch; c->pois_iters = 1; c->pois_change = 0ull; c->pois_done = ch < tol ? 1 : 0;
        pois_last = ch;
    }
    loc[RED_MAX_DIV_PRE] = bits_to_double(m_pre);
    loc[RED_MAX_GRAD] = bits_to_double(m_grad);
    loc[RED_MAX_DIV_POST] = bits_to_double(m_post);
    loc[RED_MAX_VORT] = bits_to_double(m_vort);
    loc[RED_MAX_U] = bits_to_double(m_u);
    loc[RED_MAX_V] = bits_to_double(m_v);
    loc[RED_POIS_CHANGE] = pois_last;
    loc[RED_NONFINITE] = (double)nonfinite;
    loc[RED_ENERGY_SUM] = s_e[0];
    loc[RED_ENERGY_CNT] = (double)s_c[0];
    double nut = 0.0;                                        // LES: fixed-order sum of the nu_t partials
    if (nut_part) for (int k = 0; k < NUT_PARTS; ++k) nut += nut_part[k];
    loc[RED_NUT_SUM] = nut;
    for (int k = RED_NUT_SUM + 1; k < RED_N; ++k) loc[k] = 0.0;
#pragma unroll
    for (int k = 0; k < RED_N; ++k) red[k] = loc[k];
}

// End of a step, part 2: combine the reduction vectors of all slabs (max is exact and order independent; the energy sum
// is taken in rank order), write the ring record, prepare the next step's dt / ramps (adaptive_time_step, v5.py:316-326).
__device__ __forceinline__ void ctl_end_step(StepCtl* c, const CtlParams& cp, const double* gathered, int nranks,
                                             StepScalars* ring, int cap) {
    double m[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    double es = 0.0, ec = 0.0, nut = 0.0;
    for (int r = 0; r < nranks; ++r) {
        const double* v = gathered + (size_t)r * RED_N;
        for (int k = 0; k < 8; ++k) m[k] = fmax(m[k], v[k]);
        es += v[RED_ENERGY_SUM]; ec += v[RED_ENERGY_CNT]; nut += v[RED_NUT_SUM];
    }
    StepScalars r;
    r.dt = c->dt;
    r.max_div_pre = m[RED_MAX_DIV_PRE];
    r.max_grad_phi = m[RED_MAX_GRAD];
    r.max_div_post = m[RED_MAX_DIV_POST];
    r.max_vorticity = m[RED_MAX_VORT];
    r.mean_energy = ec > 0.0 ? es / ec : __longlong_as_double(0x7ff8000000000000ll);
    r.max_abs_u = m[RED_MAX_U];
    r.max_abs_v = m[RED_MAX_V];
    r.poisson_iters = (double)c->pois_iters;
    r.poisson_change = m[RED_POIS_CHANGE];
    r.nonfinite = m[RED_NONFINITE];
    r.step = (double)c->step;
    ring[(int)(c->step % cap)] = r;
    c->in_max_u = r.max_abs_u; c->in_max_v = r.max_abs_v;
    if (cp.use_les) c->in_mean_nut = nut / cp.n_cells_global;     // np.mean(self.nu_t) of the step just taken (v5.py:323)
    c->step += 1;
    ctl_prepare(c, cp);
}
__global__ void k_ctl_reduce_local(StepCtl* c, const double* __restrict__ ke_part, const unsigned long long* __restrict__ cnt_part,
                                   int n_part, double* __restrict__ red, int close_rb1, double tol,
                                   const double* __restrict__ nut_part) {
    double loc[RED_N];
    ctl_reduce_local(c, ke_part, cnt_part, n_part, red, close_rb1, tol, nut_part, loc);
}
__global__ void k_ctl_end_step(StepCtl* c, CtlParams cp, const double* __restrict__ gathered, int nranks,
                               StepScalars* ring, int cap) {
    if (threadIdx.x || blockIdx.x) return;
    ctl_end_step(c, cp, gathered, nranks, ring, cap);
}
// one GPU: both parts in one launch (thread 0 hands the vector on in registers)
__global__ void k_ctl_reduce_end(StepCtl* c, CtlParams cp, const double* __restrict__ ke_part,
                                 const unsigned long long* __restrict__ cnt_part, int n_part, double* red, int close_rb1,
                                 double tol, const double* __restrict__ nut_part, StepScalars* ring, int cap) {
    pdl_wait(); pdl_release();
    double loc[RED_N];
    ctl_reduce_local(c, ke_part, cnt_part, n_part, red, close_rb1, tol, nut_part, loc);
    if (threadIdx.x) return;
    ctl_end_step(c, cp, loc, 1, ring, cap);
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
#define DISPATCH(dtype, CALL_F64, CALL_F32) do { if ((dtype) == 0) { CALL_F64; } else { CALL_F32; } } while (0)

void launch_supg_tau(int dtype, const Geom& g, const void* u, const void* v, void* tau, double dt, double nu_eff,
                     const void* nu_t, cudaStream_t s) {
    DISPATCH(dtype,
             (k_supg_tau<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, (double*)tau, dt, nu_eff, (const double*)nu_t)),
             (k_supg_tau<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, (float*)tau, dt, nu_eff, (const float*)nu_t)));
}
void launch_convection(int dtype, const Geom& g, const void* u, const void* v, const void* phi, const void* tau,
                       void* conv, int supg, cudaStream_t s) {
    DISPATCH(dtype,
             (k_convection<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, (const double*)phi, (const double*)tau, (double*)conv, supg)),
             (k_convection<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, (const float*)phi, (const float*)tau, (float*)conv, supg)));
}
void launch_laplacian(int dtype, const Geom& g, const void* phi, void* lap, double nu_eff, const void* nu_t, cudaStream_t s) {
    DISPATCH(dtype,
             (k_laplacian<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)phi, (double*)lap, nu_eff, (const double*)nu_t)),
             (k_laplacian<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)phi, (float*)lap, nu_eff, (const float*)nu_t)));
}
void launch_divergence(int dtype, const Geom& g, const void* u, const void* v, void* div, unsigned long long* max_out,
                       cudaStream_t s) {
    DISPATCH(dtype,
             (k_divergence<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, (double*)div, max_out)),
             (k_divergence<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, (float*)div, max_out)));
}
void launch_gradient(int dtype, const Geom& g, const void* phi, void* gx, void* gy, cudaStream_t s) {
    DISPATCH(dtype,
             (k_gradient<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)phi, (double*)gx, (double*)gy)),
             (k_gradient<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)phi, (float*)gx, (float*)gy)));
}
void launch_smagorinsky(int dtype, const Geom& g, const void* u, const void* v, void* nu_t, double cs, cudaStream_t s) {
    const double delta = sqrt(g.dx * g.dy);
    const double csd2 = (cs * delta) * (cs * delta);
    DISPATCH(dtype,
             (k_smagorinsky<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, (double*)nu_t, csd2)),
             (k_smagorinsky<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, (float*)nu_t, csd2)));
}
// Sum of a field over the owned columns, NUT_PARTS fixed partials (block b sums rows b, b + NUT_PARTS, ...): the result
// does not depend on scheduling.  Only used for mean(nu_t) when LES is on.
template <typename T>
__global__ void __launch_bounds__(256)
k_field_sum(Geom g, const T* __restrict__ f, double* __restrict__ partials) {
    double acc = 0.0;
    for (int i = blockIdx.x; i < g.ny; i += NUT_PARTS)
        for (int j = g.own0 + threadIdx.x; j < g.own1; j += 256) acc += (double)f[(long long)i * g.pitch + j];
    const double bs = block_sum(acc);
    if (threadIdx.x == 0) partials[blockIdx.x] = bs;
}
void launch_field_sum(int dtype, const Geom& g, const void* f, double* partials, cudaStream_t s) {
    DISPATCH(dtype, (k_field_sum<double><<<NUT_PARTS, 256, 0, s>>>(g, (const double*)f, partials)),
             (k_field_sum<float><<<NUT_PARTS, 256, 0, s>>>(g, (const float*)f, partials)));
}
void launch_apply_bc(int dtype, const Geom& g, void* u, void* v, const double* y, double V_inf, double y_max,
                     long long step, cudaStream_t s, double amplitude) {
    const double sc = (double)step / 1000.0;
    const double pert_scale = (sc < 1.0 ? sc : 1.0) * amplitude, phase = 0.02 * (double)step;
    const int n = g.ny + g.nx;
    DISPATCH(dtype,
             (k_apply_bc<double><<<(n + 255) / 256, 256, 0, s>>>(g, (double*)u, (double*)v, y, V_inf, y_max, pert_scale, phase)),
             (k_apply_bc<float><<<(n + 255) / 256, 256, 0, s>>>(g, (float*)u, (float*)v, y, V_inf, y_max, pert_scale, phase)));
}
void launch_apply_ibm(int dtype, const Geom& g, void* u, void* v, const double* ibm, double fs, cudaStream_t s) {
    DISPATCH(dtype,
             (k_apply_ibm<double><<<grid2d(g), kBlock, 0, s>>>(g, (double*)u, (double*)v, ibm, fs)),
             (k_apply_ibm<float><<<grid2d(g), kBlock, 0, s>>>(g, (float*)u, (float*)v, ibm, fs)));
}
void launch_predictor(int dtype, const Geom& g, const Phys& p, const StepCtl* ctl, const void* u, const void* v,
                      void* us, void* vs, void* tau, const void* nu_t, const double* ibm, const double* y, cudaStream_t s) {
    DISPATCH(dtype,
             (k_predictor<double><<<grid2d(g), kBlock, 0, s>>>(g, p, ctl, (const double*)u, (const double*)v, (double*)us, (double*)vs, (double*)tau, (const double*)nu_t, ibm, y)),
             (k_predictor<float><<<grid2d(g), kBlock, 0, s>>>(g, p, ctl, (const float*)u, (const float*)v, (float*)us, (float*)vs, (float*)tau, (const float*)nu_t, ibm, y)));
}
void launch_correct(int dtype, const Geom& g, StepCtl* ctl, const void* us, const void* vs, const void* phi,
                    void* u, void* v, cudaStream_t s) {
    DISPATCH(dtype,
             (k_correct<double><<<grid2d(g), kBlock, 0, s>>>(g, ctl, (const double*)us, (const double*)vs, (const double*)phi, (double*)u, (double*)v)),
             (k_correct<float><<<grid2d(g), kBlock, 0, s>>>(g, ctl, (const float*)us, (const float*)vs, (const float*)phi, (float*)u, (float*)v)));
}
int finalize_num_partials(const Geom& g) { dim3 gr = grid2d(g); return (int)(gr.x * gr.y); }
void launch_finalize(int dtype, const Geom& g, const Phys& p, StepCtl* ctl, const void* ua, const void* va,
                     void* u, void* v, const double* ibm, const uint8_t* mask, const double* y,
                     double* ke_partials, unsigned long long* cnt_partials, cudaStream_t s) {
    DISPATCH(dtype,
             (k_finalize<double><<<grid2d(g), kBlock, 0, s>>>(g, p, ctl, (const double*)ua, (const double*)va, (double*)u, (double*)v, ibm, mask, y, ke_partials, cnt_partials)),
             (k_finalize<float><<<grid2d(g), kBlock, 0, s>>>(g, p, ctl, (const float*)ua, (const float*)va, (float*)u, (float*)v, ibm, mask, y, ke_partials, cnt_partials)));
}
// compute_vorticity (v5.py:365-373): central differences on the interior, 0 on the border, NaN inside the cylinder;
// arithmetic in the field dtype (NumPy keeps float32 arrays float32 against Python scalars)
template <typename T>
__global__ void __launch_bounds__(BX* BY)
k_vorticity(Geom g, const T* __restrict__ u, const T* __restrict__ v, const uint8_t* __restrict__ mask, T* __restrict__ w) {
    const int j = blockIdx.x * BX + threadIdx.x, i = blockIdx.y * BY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double o = 0.0;
    if (is_interior(g, i, j))
        o = Rnd<T>::r(Rnd<T>::r(Rnd<T>::r(ld(v, k + 1) - ld(v, k - 1)) / Rnd<T>::r(g.two_dx)) -
                      Rnd<T>::r(Rnd<T>::r(ld(u, k + g.pitch) - ld(u, k - g.pitch)) / Rnd<T>::r(g.two_dy)));
    if (mask && mask[k]) o = __longlong_as_double(0x7ff8000000000000ll);
    w[k] = (T)o;
}
void launch_vorticity(int dtype, const Geom& g, const void* u, const void* v, const uint8_t* mask, void* w, cudaStream_t s) {
    DISPATCH(dtype,
             (k_vorticity<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, mask, (double*)w)),
             (k_vorticity<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, mask, (float*)w)));
}
void launch_ctl_init(StepCtl* ctl, const CtlParams& cp, long long step, int use_measured_max, const double* nut_partials,
                     cudaStream_t s) {
    k_ctl_init<<<1, 32, 0, s>>>(ctl, cp, step, use_measured_max, nut_partials);
}
void launch_ctl_reduce_local(StepCtl* ctl, const double* ke_partials, const unsigned long long* cnt_partials,
                             int n_partials, double* red, int close_rb1, double tol, const double* nut_partials, cudaStream_t s) {
    k_ctl_reduce_local<<<1, 256, 0, s>>>(ctl, ke_partials, cnt_partials, n_partials, red, close_rb1, tol, nut_partials);
}
void launch_ctl_reduce_end(StepCtl* ctl, const CtlParams& cp, const double* ke_partials, const unsigned long long* cnt_partials,
                           int n_partials, double* red, int close_rb1, double tol, const double* nut_partials,
                           StepScalars* ring, int ring_cap, cudaStream_t s) {
    launch_pdl(k_ctl_reduce_end, dim3(1), dim3(256), s, ctl, cp, ke_partials, cnt_partials, n_partials, red, close_rb1, tol, nut_partials, ring, ring_cap);
}
void launch_ctl_end_step(StepCtl* ctl, const CtlParams& cp, const double* gathered, int nranks, StepScalars* ring,
                         int ring_cap, cudaStream_t s) {
    k_ctl_end_step<<<1, 32, 0, s>>>(ctl, cp, gathered, nranks, ring, ring_cap);
}
void launch_health(int dtype, const Geom& g, const void* u, const void* v, unsigned long long* out4, cudaStream_t s) {
    cudaMemsetAsync(out4, 0, 4 * sizeof(unsigned long long), s);
    DISPATCH(dtype,
             (k_health<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, out4)),
             (k_health<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, out4)));
}
void launch_max_abs2(int dtype, const Geom& g, const void* u, const void* v, StepCtl* ctl, cudaStream_t s) {
    k_ctl_zero_max<<<1, 32, 0, s>>>(ctl);
    DISPATCH(dtype,
             (k_max_abs2<double><<<grid2d(g), kBlock, 0, s>>>(g, (const double*)u, (const double*)v, ctl)),
             (k_max_abs2<float><<<grid2d(g), kBlock, 0, s>>>(g, (const float*)u, (const float*)v, ctl)));
}

}  // namespace b200cfd
