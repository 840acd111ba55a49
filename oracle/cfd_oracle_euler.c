/*
 * TEST INFRASTRUCTURE — CPU oracle of the compressible-Euler finite-volume step of the reference,
 * python/shockwave_analysis/v1_shock.py.  Plain-C restatement; NOT part of the product:
 * only tests/, __graft_entry__.smoke() and bench_euler.py's cpu_baseline / --impl reference legs may use it.
 * Pinned to outputs of the reference itself (oracle/make_golden_euler.py -> tests/golden/euler_*.npz).
 *
 * Layout: U is (nx, ny, 4) float32, y then component fastest (v1_shock.py:252).  Scalar arithmetic is float64 with
 * float32 rounding exactly where the reference materialises float32 arrays (v1_shock.py:92-93,188-193,195-196).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    double gamma, epsilon, max_val, cfl, dx, dy;
    double rho_inf, u_inf, p_inf;
    int32_t nx, ny, solver_type /* 0 hllc, 1 roe */, n_reflect;
} eu_cfg;

static inline double clipd(double x, double a, double b) { return fmax(a, fmin(x, b)); } /* scalar_clip :70-72 */

/* compute_fluxes, v1_shock.py:84-95: primitives in f64, physical fluxes rounded to f32. */
static void eu_phys(const float* U, double g, double eps, double mv, float* F, float* G) {
    double rho = fmax((double)U[0], eps);
    double u = clipd((double)U[1] / rho, -mv, mv);
    double v = clipd((double)U[2] / rho, -mv, mv);
    double E = clipd((double)U[3] / rho, eps, mv);
    double p = fmax((g - 1) * rho * (E - 0.5 * (u * u + v * v)), eps);
    if (F) { F[0] = (float)(rho * u); F[1] = (float)(rho * u * u + p); F[2] = (float)(rho * u * v); F[3] = (float)(rho * u * (E + p / rho)); }
    if (G) { G[0] = (float)(rho * v); G[1] = (float)(rho * v * u); G[2] = (float)(rho * v * v + p); G[3] = (float)(rho * v * (E + p / rho)); }
}

/* hllc_solver, v1_shock.py:147-209.  xnormal != 0: normal = (1,0), else (0,1). */
void eu_hllc(const float* UL, const float* UR, double g, int xnormal, double eps, double mv, float* out) {
    double nxn = xnormal ? 1.0 : 0.0, nyn = xnormal ? 0.0 : 1.0;
    double rL = fmax((double)UL[0], eps);
    if (rL <= eps) { out[0] = out[1] = out[2] = out[3] = 0.f; return; }                      /* :152-153 */
    double uL = clipd((double)UL[1] / rL, -mv, mv), vL = clipd((double)UL[2] / rL, -mv, mv);
    double EL = clipd((double)UL[3] / rL, eps, mv);
    double pL = fmax((g - 1) * rL * (EL - 0.5 * (uL * uL + vL * vL)), eps);
    double rR = fmax((double)UR[0], eps);
    if (rR <= eps) { out[0] = out[1] = out[2] = out[3] = 0.f; return; }                      /* :160-161 */
    double uR = clipd((double)UR[1] / rR, -mv, mv), vR = clipd((double)UR[2] / rR, -mv, mv);
    double ER = clipd((double)UR[3] / rR, eps, mv);
    double pR = fmax((g - 1) * rR * (ER - 0.5 * (uR * uR + vR * vR)), eps);
    double qL = uL * nxn + vL * nyn, qR = uR * nxn + vR * nyn;
    double aL = sqrt(fmax(g * pL / rL, eps)), aR = sqrt(fmax(g * pR / rR, eps));
    double SL = fmin(qL - aL, qR - aR), SR = fmax(qL + aL, qR + aR);                         /* :172-173 */
    double SM = (rR * qR * (SR - qR) - rL * qL * (SL - qL) + pL - pR) / (rR * (SR - qR) - rL * (SL - qL) + eps);
    float fL[4], fR[4];
    if (xnormal) { eu_phys(UL, g, eps, mv, fL, NULL); eu_phys(UR, g, eps, mv, fR, NULL); }
    else         { eu_phys(UL, g, eps, mv, NULL, fL); eu_phys(UR, g, eps, mv, NULL, fR); }
    if (SL >= 0) { memcpy(out, fL, 16); return; }                                            /* :186-187 */
    if (SR <= 0) { memcpy(out, fR, 16); return; }                                            /* :188-189 */
    double pstar = rL * (qL - SL) * (qL - SM) + pL;                                          /* :191 */
    if (SM >= 0) {
        double c = rL * (SL - qL) / (SL - SM + eps);
        float us[4] = { (float)c, (float)(c * (SM * nxn)), (float)(c * (xnormal ? vL : SM)),
                        (float)(c * (EL + (pstar * SM - pL * qL) / (rL * (SL - qL) + eps))) };   /* :192-197 */
        for (int k = 0; k < 4; k++) out[k] = (float)((double)fL[k] + SL * (double)(float)(us[k] - UL[k]));
        return;
    }
    double c = rR * (SR - qR) / (SR - SM + eps);
    float us[4] = { (float)c, (float)(c * (SM * nxn)), (float)(c * (xnormal ? vR : SM)),
                    (float)(c * (ER + (pstar * SM - pR * qR) / (rR * (SR - qR) + eps))) };       /* :202-207 */
    for (int k = 0; k < 4; k++) out[k] = (float)((double)fR[k] + SR * (double)(float)(us[k] - UR[k]));
}

static double superbee(double a, double b) {                                                    /* :77-82 */
    if (a * b <= 0) return 0.0;
    double r = a / (b + 1e-10);
    return fmax(0.0, fmax(fmin(2 * r, 1.0), fmin(r, 2.0))) * b;
}

/* roe_solver, v1_shock.py:97-145 (a Rusanov-like flux with the sum of three Roe wave speeds; the limiter results
 * delta_rho/u/v are computed and never used, :119-121).  The flux is ALWAYS built from F (compute_fluxes()[0]),
 * also for the y-normal (:139-140). */
void eu_roe(const float* UL, const float* UR, double g, int xnormal, double eps, double mv, float* out) {
    double nxn = xnormal ? 1.0 : 0.0, nyn = xnormal ? 0.0 : 1.0;
    double rL = fmax((double)UL[0], eps);
    if (rL <= eps) { out[0] = out[1] = out[2] = out[3] = 0.f; return; }
    double uL = clipd((double)UL[1] / rL, -mv, mv), vL = clipd((double)UL[2] / rL, -mv, mv);
    double EL = clipd((double)UL[3] / rL, eps, mv);
    double pL = fmax((g - 1) * rL * (EL - 0.5 * (uL * uL + vL * vL)), eps);
    double hL = ((double)UL[3] + pL) / rL;
    double rR = fmax((double)UR[0], eps);
    if (rR <= eps) { out[0] = out[1] = out[2] = out[3] = 0.f; return; }
    double uR = clipd((double)UR[1] / rR, -mv, mv), vR = clipd((double)UR[2] / rR, -mv, mv);
    double ER = clipd((double)UR[3] / rR, eps, mv);
    double pR = fmax((g - 1) * rR * (ER - 0.5 * (uR * uR + vR * vR)), eps);
    double hR = ((double)UR[3] + pR) / rR;
    (void)superbee;
    double sL = sqrt(rL), sR = sqrt(rR), den = sL + sR + eps;
    double ur = (uL * sL + uR * sR) / den, vr = (vL * sL + vR * sR) / den, hr = (hL * sL + hR * sR) / den;
    double ar = sqrt(fmax((g - 1) * (hr - 0.5 * (ur * ur + vr * vr)), eps));
    double vn = ur * nxn + vr * nyn;
    double er = 0.05 * ar;
    double l1 = fmax(er, fabs(vn)), l2 = fmax(er, fabs(vn + ar)), l3 = fmax(er, fabs(vn - ar));
    float fL[4], fR[4];
    eu_phys(UL, g, eps, mv, fL, NULL);
    eu_phys(UR, g, eps, mv, fR, NULL);
    for (int k = 0; k < 4; k++) {
        float favg = (float)(0.5 * (double)(float)(fL[k] + fR[k]));          /* 0.5 * f32 array stays f32 */
        float dU = UR[k] - UL[k];
        out[k] = (float)((double)favg - 0.5 * (l1 + l2 + l3) * (double)dU);
    }
}

/* ShockwaveSolver.compute_dt, v1_shock.py:263-275.  NumPy on f32 arrays (Python-float scalars are weak: f32 math),
 * array_clip is @njit with f64 bounds -> its result is f64 holding the f32 quotient. */
double eu_compute_dt(const eu_cfg* c, const float* U, double* max_sx, double* max_sy) {
    long n = (long)c->nx * c->ny;
    float epsf = (float)c->epsilon;
    double msx = -INFINITY, msy = -INFINITY;
    int nan_seen = 0;
#pragma omp parallel for reduction(max : msx, msy) reduction(| : nan_seen)
    for (long k = 0; k < n; k++) {
        const float* q = U + 4 * k;
        float rho = fmaxf(q[0], epsf);
        double u = clipd((double)(float)(q[1] / rho), -c->max_val, c->max_val);
        double v = clipd((double)(float)(q[2] / rho), -c->max_val, c->max_val);
        double E = clipd((double)(float)(q[3] / rho), c->epsilon, c->max_val);
        double p = fmax((double)((float)(c->gamma - 1) * rho) * (E - 0.5 * (u * u + v * v)), c->epsilon); /* (gamma-1)*rho is f32 NumPy math */
        double a = sqrt(clipd(c->gamma * p / (double)rho, c->epsilon, c->max_val));
        double sx = fabs(u) + a, sy = fabs(v) + a;
        if (sx != sx || sy != sy) nan_seen = 1;
        msx = fmax(msx, sx); msy = fmax(msy, sy);
    }
    if (nan_seen) { if (max_sx) *max_sx = NAN; if (max_sy) *max_sy = NAN; return NAN; }
    msx = fmin(msx, c->max_val); msy = fmin(msy, c->max_val);
    if (max_sx) *max_sx = msx;
    if (max_sy) *max_sy = msy;
    double dtx = c->cfl * c->dx / fmax(msx, c->epsilon), dty = c->cfl * c->dy / fmax(msy, c->epsilon);
    return fmin(dtx, dty);
}

/* update_state, v1_shock.py:211-223: border cells of U_new stay 0. */
void eu_update_state(const eu_cfg* c, const float* U, const float* F, const float* G, double dt, float* Un) {
    int nx = c->nx, ny = c->ny;
    memset(Un, 0, sizeof(float) * 4 * (size_t)nx * ny);
    double ax = dt / c->dx, ay = dt / c->dy, eps = c->epsilon, mv = c->max_val;
#pragma omp parallel for
    for (int i = 1; i < nx - 1; i++)
        for (int j = 1; j < ny - 1; j++) {
            size_t o = 4 * ((size_t)i * ny + j), ow = 4 * ((size_t)(i - 1) * ny + j), os = o - 4;
            double up[4];
            for (int k = 0; k < 4; k++)
                up[k] = (double)U[o + k] - ax * (double)(float)(F[o + k] - F[ow + k]) - ay * (double)(float)(G[o + k] - G[os + k]);
            up[0] = fmax(up[0], eps);
            up[1] = clipd(up[1], -mv * up[0], mv * up[0]);
            up[2] = clipd(up[2], -mv * up[0], mv * up[0]);
            up[3] = fmax(up[3], eps * up[0]);
            for (int k = 0; k < 4; k++) Un[o + k] = (float)up[k];
        }
}

/* apply_boundary_conditions, v1_shock.py:277-289, + wedge zeroing :312-313.  The loop over j at :283-289 flips the
 * y-momentum of the columns whose Y[0,j] lies under the wedge line at X[0,j] (a PREFIX j < n_reflect, empty for the
 * shipped geometry) and otherwise copies the y-border columns; the copies are idempotent, so the loop equals
 * "flip the prefix, then copy once if any column was not flagged". */
void eu_apply_bc(const eu_cfg* c, float* U, const uint8_t* wedge) {
    int nx = c->nx, ny = c->ny;
    float E_inf = (float)(c->rho_inf * (c->p_inf / (c->rho_inf * (c->gamma - 1)) + 0.5 * c->u_inf * c->u_inf));
    for (int j = 0; j < ny; j++) {
        float* q = U + 4 * (size_t)j;
        q[0] = (float)c->rho_inf; q[1] = (float)(c->rho_inf * c->u_inf); q[2] = 0.f; q[3] = E_inf;
    }
    memcpy(U + 4 * (size_t)(nx - 1) * ny, U + 4 * (size_t)(nx - 2) * ny, sizeof(float) * 4 * ny);
    for (int i = 0; i < nx; i++)
        for (int j = 0; j < c->n_reflect && j < ny; j++) U[4 * ((size_t)i * ny + j) + 2] = -U[4 * ((size_t)i * ny + j) + 2];
    if (c->n_reflect < ny)
        for (int i = 0; i < nx; i++) {
            float* r = U + 4 * (size_t)i * ny;
            memcpy(r, r + 4, 16);
            memcpy(r + 4 * (size_t)(ny - 1), r + 4 * (size_t)(ny - 2), 16);
        }
    if (wedge)
        for (size_t k = 0; k < (size_t)nx * ny; k++)
            if (wedge[k]) { U[4 * k + 1] = 0.f; U[4 * k + 2] = 0.f; }
}

/* ShockwaveSolver.time_step, v1_shock.py:291-317, n times.  U is updated in place; dts[n] receives each dt
 * (0.0 where the reference returns early on a non-finite dt, :294-296). Returns 0, or -1 on allocation failure. */
int eu_step(const eu_cfg* c, float* U, const uint8_t* wedge, int n, double* dts) {
    int nx = c->nx, ny = c->ny;
    size_t cells = (size_t)nx * ny;
    float* F = (float*)malloc(16 * cells), *G = (float*)malloc(16 * cells), *Un = (float*)malloc(16 * cells);
    if (!F || !G || !Un) { free(F); free(G); free(Un); return -1; }
    void (*flux)(const float*, const float*, double, int, double, double, float*) = c->solver_type ? eu_roe : eu_hllc;
    for (int s = 0; s < n; s++) {
        double dt = eu_compute_dt(c, U, NULL, NULL);
        if (!isfinite(dt)) { if (dts) dts[s] = 0.0; continue; }
        memset(F, 0, 16 * cells);
        memset(G, 0, 16 * cells);
#pragma omp parallel for
        for (int i = 1; i < nx; i++)                                   /* :301-303 */
            for (int j = 0; j < ny; j++)
                flux(U + 4 * ((size_t)(i - 1) * ny + j), U + 4 * ((size_t)i * ny + j), c->gamma, 1, c->epsilon, c->max_val,
                     F + 4 * ((size_t)(i - 1) * ny + j));
#pragma omp parallel for
        for (int i = 0; i < nx; i++)                                   /* :304-306 */
            for (int j = 1; j < ny; j++)
                flux(U + 4 * ((size_t)i * ny + j - 1), U + 4 * ((size_t)i * ny + j), c->gamma, 0, c->epsilon, c->max_val,
                     G + 4 * ((size_t)i * ny + j - 1));
        eu_update_state(c, U, F, G, dt, Un);
        memcpy(U, Un, 16 * cells);
        eu_apply_bc(c, U, wedge);
        if (dts) dts[s] = dt;
    }
    free(F); free(G); free(Un);
    return 0;
}

void eu_phys_fluxes(const float* U, double g, double eps, double mv, float* F, float* G) { eu_phys(U, g, eps, mv, F, G); }
