/*
 * TEST INFRASTRUCTURE — NOT A PRODUCT PATH.
 *
 * CPU restatement (plain C) of the reference's projection time step, written from the
 * reference's behaviour (python/flow_over_cylinder (Fischer)/v5.py) and pinned against
 * goldens minted from the reference itself (oracle/make_golden.py, NUMBA_NUM_THREADS=1).
 * The reference's own test-suite holds no vectors for this path (it has no tests), so
 * parity is "unpinned by reference tests"; it is pinned by reference *outputs* instead.
 *
 * This header is a poor-man's template: it is included twice by cfd_oracle.c with
 *   T   = storage type (double | float)      SFX(name) = name##_f64 | name##_f32
 * Arithmetic is done in double after loading (Numba promotes f32 (+) f64 -> f64, v5.py
 * Appendix-B of SURVEY.md); values the reference materialises as arrays of the field
 * dtype (tau, conv, lap, div, grad, phi) are rounded to T where the reference stores them.
 *
 * Arrays are C-contiguous (ny, nx); i = row (y), j = column (x).
 */

#define IDX(i, j) ((size_t)(i) * (size_t)nx + (size_t)(j))

/* v5.py:178-187 compute_divergence_fast — central differences, border 0 */
void SFX(orc_divergence)(const T* u, const T* v, T* div, int ny, int nx, double dx, double dy)
{
    const double dxi = 0.5 / dx, dyi = 0.5 / dy;
    memset(div, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j)
            div[IDX(i, j)] = (T)(((double)u[IDX(i, j + 1)] - (double)u[IDX(i, j - 1)]) * dxi +
                                 ((double)v[IDX(i + 1, j)] - (double)v[IDX(i - 1, j)]) * dyi);
}

/* v5.py:189-200 compute_gradient_fast — central differences, border 0 */
void SFX(orc_gradient)(const T* phi, T* gx, T* gy, int ny, int nx, double dx, double dy)
{
    const double dxi = 0.5 / dx, dyi = 0.5 / dy;
    memset(gx, 0, sizeof(T) * (size_t)ny * nx);
    memset(gy, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            gx[IDX(i, j)] = (T)(((double)phi[IDX(i, j + 1)] - (double)phi[IDX(i, j - 1)]) * dxi);
            gy[IDX(i, j)] = (T)(((double)phi[IDX(i + 1, j)] - (double)phi[IDX(i - 1, j)]) * dyi);
        }
}

/* v5.py:149-162 compute_supg_stabilization_fast.  nu_eff is a scalar here unless nu_t != NULL
 * (LES on: nu_eff[i,j] = nu_eff + nu_t[i,j], v5.py:388). */
void SFX(orc_supg_tau)(const T* u, const T* v, T* tau, int ny, int nx, double dx, double dy,
                       double dt, double nu_eff, const T* nu_t)
{
    const double h = dx < dy ? dx : dy;
    memset(tau, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            double uu = u[IDX(i, j)], vv = v[IDX(i, j)];
            double vel = sqrt(uu * uu + vv * vv);
            double ne = nu_t ? (double)(T)(nu_eff + (double)nu_t[IDX(i, j)]) : nu_eff;
            if (vel > 1e-10) {
                double Pe = vel * h / (ne + 1e-10);
                double m = Pe / 2.0 < 1.0 ? Pe / 2.0 : 1.0;
                tau[IDX(i, j)] = (T)(h / (2 * vel) * m);
            } else {
                tau[IDX(i, j)] = (T)(dt / 2);
            }
        }
}

/* v5.py:127-147 compute_convection_supg_fast (note the 0.25/dx and 0.25/dx^2 factors) */
void SFX(orc_convection_supg)(const T* u, const T* v, const T* phi, const T* tau, T* conv,
                              int ny, int nx, double dx, double dy)
{
    const double dxi = 0.5 / dx, dyi = 0.5 / dy;
    memset(conv, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            double uu = u[IDX(i, j)], vv = v[IDX(i, j)];
            double pc = phi[IDX(i, j)], pe = phi[IDX(i, j + 1)], pw = phi[IDX(i, j - 1)];
            double pn = phi[IDX(i + 1, j)], ps = phi[IDX(i - 1, j)];
            double dpx = (pe - pw) * (0.5 * dxi), dpy = (pn - ps) * (0.5 * dyi);
            double c = uu * dpx + vv * dpy;
            double t = tau[IDX(i, j)];
            if (t > 0) {
                double d2x = (pe - 2 * pc + pw) * (dxi * dxi);
                double d2y = (pn - 2 * pc + ps) * (dyi * dyi);
                c -= t * (uu * d2x + vv * d2y);
            }
            conv[IDX(i, j)] = (T)c;
        }
}

/* v5.py:112-125 compute_convection_fast (first-order upwind; use_supg=False branch) */
void SFX(orc_convection_upwind)(const T* u, const T* v, const T* phi, T* conv,
                                int ny, int nx, double dx, double dy)
{
    const double dxi = 1.0 / dx, dyi = 1.0 / dy;
    memset(conv, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            double uu = u[IDX(i, j)], vv = v[IDX(i, j)], pc = phi[IDX(i, j)];
            double dpx = uu > 0 ? (pc - (double)phi[IDX(i, j - 1)]) * dxi : ((double)phi[IDX(i, j + 1)] - pc) * dxi;
            double dpy = vv > 0 ? (pc - (double)phi[IDX(i - 1, j)]) * dyi : ((double)phi[IDX(i + 1, j)] - pc) * dyi;
            conv[IDX(i, j)] = (T)(uu * dpx + vv * dpy);
        }
}

/* v5.py:164-176 compute_laplacian_fast */
void SFX(orc_laplacian)(const T* phi, T* lap, int ny, int nx, double dx, double dy,
                        double nu_eff, const T* nu_t)
{
    const double ax = 1.0 / (dx * dx), ay = 1.0 / (dy * dy);
    memset(lap, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            double pc = phi[IDX(i, j)];
            double ne = nu_t ? (double)(T)(nu_eff + (double)nu_t[IDX(i, j)]) : nu_eff;
            lap[IDX(i, j)] = (T)(ne * (((double)phi[IDX(i, j + 1)] - 2 * pc + (double)phi[IDX(i, j - 1)]) * ax +
                                       ((double)phi[IDX(i + 1, j)] - 2 * pc + (double)phi[IDX(i - 1, j)]) * ay));
        }
}

/* v5.py:96-110 compute_smagorinsky_viscosity_fast (forward differences; off in shipped configs) */
void SFX(orc_smagorinsky)(const T* u, const T* v, T* nu_t, int ny, int nx, double dx, double dy, double cs)
{
    const double delta = sqrt(dx * dy);
    const double csd2 = (cs * delta) * (cs * delta);
    memset(nu_t, 0, sizeof(T) * (size_t)ny * nx);
    #pragma omp parallel for schedule(static)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            double uc = u[IDX(i, j)], vc = v[IDX(i, j)];
            double dudx = ((double)u[IDX(i, j + 1)] - uc) / dx, dudy = ((double)u[IDX(i + 1, j)] - uc) / dy;
            double dvdx = ((double)v[IDX(i, j + 1)] - vc) / dx, dvdy = ((double)v[IDX(i + 1, j)] - vc) / dy;
            double s = sqrt(2 * (dudx * dudx + dvdy * dvdy) + (dudy + dvdx) * (dudy + dvdx));
            nu_t[IDX(i, j)] = (T)(csd2 * s);
        }
}

/* v5.py:349-360 apply_boundary_conditions — order matters: inflow col, outflow col, then walls */
void SFX(orc_apply_bc_amp)(T* u, T* v, int ny, int nx, const double* y, double y_max, double V_inf, long step, double amplitude);
void SFX(orc_apply_bc)(T* u, T* v, int ny, int nx, const double* y, double y_max, double V_inf, long step)
{
    SFX(orc_apply_bc_amp)(u, v, ny, nx, y, y_max, V_inf, step, 0.01);
}
void SFX(orc_apply_bc_amp)(T* u, T* v, int ny, int nx, const double* y, double y_max, double V_inf, long step, double amplitude)
{
    double sc = (double)step / 1000.0;
    const double pert_scale = (sc < 1.0 ? sc : 1.0) * amplitude;
    for (int i = 0; i < ny; ++i) {
        double pert = pert_scale * sin(2 * M_PI * y[i] / y_max + 0.02 * (double)step);
        u[IDX(i, 0)] = (T)(V_inf * (1 + pert));
        v[IDX(i, 0)] = (T)0;
        u[IDX(i, nx - 1)] = u[IDX(i, nx - 2)];
        v[IDX(i, nx - 1)] = v[IDX(i, nx - 2)];
    }
    for (int j = 0; j < nx; ++j) {
        u[IDX(0, j)] = 0; u[IDX(ny - 1, j)] = 0;
        v[IDX(0, j)] = 0; v[IDX(ny - 1, j)] = 0;
    }
}

/* v5.py:228-237 apply_ibm_fast — whole array including the border; ibm is always float64 */
void SFX(orc_apply_ibm)(T* u, T* v, const double* ibm, int ny, int nx, double fs)
{
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < ny; ++i)
        for (int j = 0; j < nx; ++j) {
            double m = ibm[IDX(i, j)];
            if (m > 0) {
                u[IDX(i, j)] = (T)((double)u[IDX(i, j)] * (1.0 - m * fs));
                v[IDX(i, j)] = (T)((double)v[IDX(i, j)] * (1.0 - m * fs));
            }
        }
}

/* v5.py:202-226 solve_pressure_gauss_seidel_fast, AS WRITTEN (true max-reduction, tolerance break).
 * omega = 1 is the reference; omega != 1 gives red-black SOR (phi += omega (phi_gs - phi)).
 * "as executed" under Numba parallel=True is exactly max_iter = 1 (SURVEY finding 3).
 * Returns the number of iterations performed; *last_change = max|dphi| of the last iteration. */
int SFX(orc_rbgs)(T* phi, const T* div, const unsigned char* mask, int ny, int nx, double dx, double dy,
                  double dt, int max_iter, double tol, double omega, double* last_change)
{
    const double a = 1.0 / (dx * dx), b = 1.0 / (dy * dy);
    const double d = 1.0 / (2.0 * (a + b)), dti = 1.0 / dt;
    int it = 0;
    double mc = 0.0;
    for (it = 0; it < max_iter; ++it) {
        mc = 0.0;
        for (int color = 0; color < 2; ++color) {
            #pragma omp parallel for schedule(static) reduction(max : mc)
            for (int i = 1; i < ny - 1; ++i)
                for (int j = 1 + (i + color) % 2; j < nx - 1; j += 2) {
                    if (mask && mask[IDX(i, j)]) continue;
                    double rhs = -(double)div[IDX(i, j)] * dti;
                    double old = phi[IDX(i, j)];
                    double pn = (a * ((double)phi[IDX(i, j + 1)] + (double)phi[IDX(i, j - 1)]) +
                                 b * ((double)phi[IDX(i + 1, j)] + (double)phi[IDX(i - 1, j)]) - rhs) * d;
                    if (omega != 1.0) pn = old + omega * (pn - old);
                    double ch = fabs(pn - old);   /* measured before the store rounds to T (v5.py:220) */
                    if (ch > mc) mc = ch;
                    phi[IDX(i, j)] = (T)pn;
                }
        }
        if (mc < tol) { ++it; break; }
    }
    if (last_change) *last_change = mc;
    return it;
}

/* v5.py:336-346 Jacobi branch of solve_pressure_fast: hard-coded 0.25 and dx^2, mask -> 0, border
 * untouched (copy of old = 0), fixed sweep count, no tolerance.  phi is in/out; tmp is scratch. */
void SFX(orc_jacobi)(T* phi, T* tmp, const T* div, const unsigned char* mask, int ny, int nx,
                     double dx, double dt, int sweeps)
{
    const double dx2 = dx * dx;
    T* a = phi; T* b = tmp;
    for (int s = 0; s < sweeps; ++s) {
        memcpy(b, a, sizeof(T) * (size_t)ny * nx);
        #pragma omp parallel for schedule(static)
        for (int i = 1; i < ny - 1; ++i)
            for (int j = 1; j < nx - 1; ++j) {
                double val = 0.25 * ((double)a[IDX(i, j + 1)] + (double)a[IDX(i, j - 1)] +
                                     (double)a[IDX(i + 1, j)] + (double)a[IDX(i - 1, j)] -
                                     dx2 * (double)div[IDX(i, j)] / dt);
                b[IDX(i, j)] = (T)val;
            }
        if (mask)
            for (size_t k = 0; k < (size_t)ny * nx; ++k) if (mask[k]) b[k] = 0;
        T* t = a; a = b; b = t;
    }
    if (a != phi) memcpy(phi, a, sizeof(T) * (size_t)ny * nx);
}

/* One lexicographic Gauss-Seidel sweep of clean_divergence_fast (v5.py:250-253) as executed
 * with one Numba thread: cell (i,j) sees NEW (i-1,j),(i,j-1) and OLD (i+1,j),(i,j+1). */
void SFX(orc_lexgs_sweep)(T* psi, const T* div, int ny, int nx, double dx, double dy)
{
    const double a = 1.0 / (dx * dx), b = 1.0 / (dy * dy), d = 1.0 / (2.0 * (a + b));
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j)
            psi[IDX(i, j)] = (T)((a * ((double)psi[IDX(i, j + 1)] + (double)psi[IDX(i, j - 1)]) +
                                  b * ((double)psi[IDX(i + 1, j)] + (double)psi[IDX(i - 1, j)]) -
                                  (double)div[IDX(i, j)]) * d);
}

/* v5.py:239-257 clean_divergence_fast(u, v, dx, dy, iterations): psi persists across iterations.
 * work = 4 scratch arrays of ny*nx (psi, div, gx, gy). If psi_out != NULL the final psi is copied out. */
void SFX(orc_clean_divergence)(T* u, T* v, int ny, int nx, double dx, double dy, int iterations, T* work, T* psi_out)
{
    size_t n = (size_t)ny * nx;
    T *psi = work, *div = work + n, *gx = work + 2 * n, *gy = work + 3 * n;
    memset(psi, 0, sizeof(T) * n);
    for (int it = 0; it < iterations; ++it) {
        SFX(orc_divergence)(u, v, div, ny, nx, dx, dy);
        SFX(orc_lexgs_sweep)(psi, div, ny, nx, dx, dy);
        SFX(orc_gradient)(psi, gx, gy, ny, nx, dx, dy);
        #pragma omp parallel for schedule(static)
        for (int i = 1; i < ny - 1; ++i)
            for (int j = 1; j < nx - 1; ++j) {
                u[IDX(i, j)] = (T)(u[IDX(i, j)] - gx[IDX(i, j)]);   /* same-dtype subtraction (NumPy) */
                v[IDX(i, j)] = (T)(v[IDX(i, j)] - gy[IDX(i, j)]);
            }
    }
    if (psi_out) memcpy(psi_out, psi, sizeof(T) * n);
}

/* v5.py:365-373 + :428  nanmax|vorticity| with cylinder cells = NaN (ignored) */
double SFX(orc_max_vorticity)(const T* u, const T* v, const unsigned char* mask, int ny, int nx, double dx, double dy)
{
    double m = 0.0; /* border vorticity is 0, so the max is >= 0 */
    #pragma omp parallel for schedule(static) reduction(max : m)
    for (int i = 1; i < ny - 1; ++i)
        for (int j = 1; j < nx - 1; ++j) {
            if (mask && mask[IDX(i, j)]) continue;
            double w = ((double)v[IDX(i, j + 1)] - (double)v[IDX(i, j - 1)]) / (2 * dx) -
                       ((double)u[IDX(i + 1, j)] - (double)u[IDX(i - 1, j)]) / (2 * dy);
            w = fabs((double)(T)w);
            if (w > m) m = w;   /* NaN compares false -> ignored like nanmax */
        }
    return m;
}

/* v5.py:362-363 + :431-432  nanmean(0.5 (u^2+v^2)) */
double SFX(orc_mean_energy)(const T* u, const T* v, int ny, int nx)
{
    double s = 0.0; long cnt = 0;
    #pragma omp parallel for schedule(static) reduction(+ : s, cnt)
    for (int i = 0; i < ny; ++i) {
        double rs = 0.0; long rc = 0;
        for (int j = 0; j < nx; ++j) {
            double uu = u[IDX(i, j)], vv = v[IDX(i, j)];
            double e = (double)(T)(0.5 * (double)(T)((double)(T)(uu * uu) + (double)(T)(vv * vv)));
            if (e == e) { rs += e; ++rc; }
        }
        s += rs; cnt += rc;
    }
    return cnt ? s / (double)cnt : NAN;
}

double SFX(orc_max_abs)(const T* a, size_t n)
{
    double m = 0.0;
    #pragma omp parallel for schedule(static) reduction(max : m)
    for (size_t k = 0; k < n; ++k) { double x = fabs((double)a[k]); if (x > m) m = x; }
    return m;
}

/* v5.py:316-326 adaptive_time_step.  max|u|,|v| are of the (post-clip) state entering the step. */
double SFX(orc_adaptive_dt)(const orc_config* c, const T* u, const T* v, const T* nu_t, long step)
{
    if (!c->adaptive_dt) return c->dt_base;
    if (step < 1000) return c->scalar_f32 ? (double)(float)0.00002 : 0.00002;
    size_t n = (size_t)c->ny * c->nx;
    double vm = SFX(orc_max_abs)(u, n), vm2 = SFX(orc_max_abs)(v, n);
    if (vm2 > vm) vm = vm2;
    if (vm < 1e-10) vm = 1e-10;
    double h = c->dx < c->dy ? c->dx : c->dy;
    double dt_cfl = c->cfl_target * h / vm;
    double mean_nut = 0.0;
    if (c->use_les && nu_t) { for (size_t k = 0; k < n; ++k) mean_nut += (double)nu_t[k]; mean_nut /= (double)n; }
    /* nu (f32) + mean(nu_t) (field dtype) + av (f32): f64 sum in fp64 mode, f32 sum in fp32 mode */
    double nu_total = sizeof(T) == 8 ? (c->nu + mean_nut) + c->av
                                     : (double)(float)((float)((float)c->nu + (float)mean_nut) + (float)c->av);
    double dt_visc = 0.4 * h * h / nu_total;
    if (sizeof(T) == 4) {
        /* float32 fields: vel_max and nu_total are np.float32, so under NumPy 2 promotion (NEP 50) the Python-float
         * numerators are cast to float32 and both divisions, the min and the clip are float32 operations (v5.py:322-326) */
        float fc = (float)(c->cfl_target * h) / (float)vm, fv = (float)(0.4 * (h * h)) / (float)nu_total;
        float fd = fc < fv ? fc : fv;
        if (fd < (float)c->dt_min) fd = (float)c->dt_min;
        if (fd > (float)c->dt_max) fd = (float)c->dt_max;
        return (double)fd;
    }
    double dt = dt_cfl < dt_visc ? dt_cfl : dt_visc;
    if (dt < c->dt_min) dt = c->dt_min;
    if (dt > c->dt_max) dt = c->dt_max;
    return c->scalar_f32 ? (double)(float)dt : dt;      /* v5: np.float32(...); v3 / v4: float(...) */
}

/* v5.py:375-441 OptimizedTurbulentSolver.time_step(), in execution order.
 * work: 8 scratch arrays of ny*nx.  scal[8] = {dt, max|div*|, max|grad phi|, max|div| post,
 * max|vort|, mean KE, poisson iterations, poisson last change}.  Returns dt. */
double SFX(orc_time_step)(const orc_config* c, T* u, T* v, T* phi, T* u_star, T* v_star, T* div, T* tau, T* nu_t,
                          const double* ibm, const unsigned char* mask, const double* y, long step,
                          T* work, double* scal)
{
    const int ny = c->ny, nx = c->nx;
    const size_t n = (size_t)ny * nx;
    T *conv_u = work, *conv_v = work + n, *lap_u = work + 2 * n, *lap_v = work + 3 * n;
    const double dt = SFX(orc_adaptive_dt)(c, u, v, nu_t, step);
    /* nu_eff = nu + nu_t + av (array of field dtype; a constant when LES is off, v5.py:388).
     * With LES on we evaluate (nu+av)+nu_t instead of (nu+nu_t)+av: last-ulp difference only; LES is
     * off in every shipped configuration (v5.py:60-61,624-625). */
    const double nu_eff = sizeof(T) == 8 ? c->nu + c->av : (double)(float)((float)c->nu + (float)c->av);
    const T* nut_arr = NULL;
    if (c->use_les) { SFX(orc_smagorinsky)(u, v, nu_t, ny, nx, c->dx, c->dy, c->cs); nut_arr = nu_t; }
    else memset(nu_t, 0, sizeof(T) * n);
    if (c->use_supg) {
        SFX(orc_supg_tau)(u, v, tau, ny, nx, c->dx, c->dy, dt, nu_eff, nut_arr);
        SFX(orc_convection_supg)(u, v, u, tau, conv_u, ny, nx, c->dx, c->dy);
        SFX(orc_convection_supg)(u, v, v, tau, conv_v, ny, nx, c->dx, c->dy);
    } else {
        SFX(orc_convection_upwind)(u, v, u, conv_u, ny, nx, c->dx, c->dy);
        SFX(orc_convection_upwind)(u, v, v, conv_v, ny, nx, c->dx, c->dy);
    }
    SFX(orc_laplacian)(u, lap_u, ny, nx, c->dx, c->dy, nu_eff, nut_arr);
    SFX(orc_laplacian)(v, lap_v, ny, nx, c->dx, c->dy, nu_eff, nut_arr);
    /* u* = u_old + dt * (-conv + lap) in the field dtype (NumPy; dt is an f32 scalar) */
    #pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) {
        if (sizeof(T) == 8) {
            u_star[k] = (T)((double)u[k] + dt * (-(double)conv_u[k] + (double)lap_u[k]));
            v_star[k] = (T)((double)v[k] + dt * (-(double)conv_v[k] + (double)lap_v[k]));
        } else {
            float fdt = (float)dt;
            u_star[k] = (T)((float)u[k] + fdt * (-(float)conv_u[k] + (float)lap_u[k]));
            v_star[k] = (T)((float)v[k] + fdt * (-(float)conv_v[k] + (float)lap_v[k]));
        }
    }
    SFX(orc_apply_bc_amp)(u_star, v_star, ny, nx, y, c->y_max, c->V_inf, step, c->pert_amplitude);
    double fs = (double)step / (double)c->initial_steps; if (fs > 1.0) fs = 1.0;
    SFX(orc_apply_ibm)(u_star, v_star, ibm, ny, nx, fs);
    SFX(orc_divergence)(u_star, v_star, div, ny, nx, c->dx, c->dy);
    scal[0] = dt;
    scal[1] = SFX(orc_max_abs)(div, n);
    memset(phi, 0, sizeof(T) * n);
    scal[6] = 0; scal[7] = 0;
    if (c->poisson_mode == ORC_POISSON_JACOBI) {
        SFX(orc_jacobi)(phi, work + 4 * n, div, mask, ny, nx, c->dx, c->cfg_dt, c->pressure_iterations);
        scal[6] = c->pressure_iterations;
    } else {
        int iters = c->poisson_mode == ORC_POISSON_AS_EXECUTED ? 1 : c->pressure_iterations;
        scal[6] = SFX(orc_rbgs)(phi, div, mask, ny, nx, c->dx, c->dy, c->cfg_dt, iters,
                                c->pressure_tolerance, c->omega, &scal[7]);
    }
    T *gx = work + 4 * n, *gy = work + 5 * n;
    SFX(orc_gradient)(phi, gx, gy, ny, nx, c->dx, c->dy);
    double gm = 0.0;
    for (size_t k = 0; k < n; ++k) {
        double g = sqrt((double)(T)((double)gx[k] * gx[k]) + (double)(T)((double)gy[k] * gy[k]));
        if (g > gm) gm = g;
    }
    scal[2] = gm;
    #pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) {
        if (sizeof(T) == 8) {
            u[k] = (T)((double)u_star[k] - dt * (double)gx[k]);
            v[k] = (T)((double)v_star[k] - dt * (double)gy[k]);
        } else {
            float fdt = (float)dt;
            u[k] = (T)((float)u_star[k] - fdt * (float)gx[k]);
            v[k] = (T)((float)v_star[k] - fdt * (float)gy[k]);
        }
    }
    SFX(orc_clean_divergence)(u, v, ny, nx, c->dx, c->dy, c->clean_iters, work, NULL);
    SFX(orc_divergence)(u, v, work, ny, nx, c->dx, c->dy);
    scal[3] = SFX(orc_max_abs)(work, n);
    SFX(orc_apply_bc_amp)(u, v, ny, nx, y, c->y_max, c->V_inf, step, c->pert_amplitude);
    SFX(orc_apply_ibm)(u, v, ibm, ny, nx, fs);
    scal[4] = SFX(orc_max_vorticity)(u, v, mask, ny, nx, c->dx, c->dy);
    scal[5] = SFX(orc_mean_energy)(u, v, ny, nx);
    const double mv = c->max_velocity;
    #pragma omp parallel for schedule(static)
    for (size_t k = 0; k < n; ++k) {
        double a = u[k], b = v[k];
        u[k] = (T)(a < -mv ? -mv : (a > mv ? mv : a));
        v[k] = (T)(b < -mv ? -mv : (b > mv ? mv : b));
    }
    return dt;
}

#undef IDX
