/*
 * TEST INFRASTRUCTURE — NOT A PRODUCT PATH.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.
 *
 * Plain-C CPU restatement of the hot path of Santhosh-Sathyamurthy/cfd-simulations:
 *   python/flow_over_cylinder (Fischer)/v5.py:96-257,316-441   (projection time step)
 *   python/flow_over_cylinder (Fischer)/v2.py:284-306          (lexicographic SOR fallback)
 *   python/shockwave_analysis/v1_shock.py:84-95,147-223,263-317 (Euler HLLC step; cfd_oracle_euler.c)
 * The reference is Python/Numba (no C sources to compile), so this is a "port" oracle.  It is pinned
 * against outputs of the reference itself run in the build container with NUMBA_NUM_THREADS=1
 * (oracle/make_golden.py -> tests/golden/ npz files).  The reference has no tests of its own, hence no
 * reference-held golden vectors exist for this path.
 *
 * Build: make -C oracle   (gcc -O2 -fopenmp; NO -ffast-math so results are thread-count independent)
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stddef.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

enum { ORC_POISSON_AS_EXECUTED = 0, ORC_POISSON_AS_WRITTEN = 1, ORC_POISSON_JACOBI = 2 };

/* Mirrors OptimizedTurbulentConfig (v5.py:41-94); nu, av, cfg_dt are the float32-rounded scalars. */
typedef struct {
    int ny, nx;
    double dx, dy;
    double nu, av, cfg_dt;
    double V_inf, y_max;
    double cfl_target, dt_min, dt_max, dt_base, max_velocity;
    int initial_steps, use_supg, use_les, adaptive_dt;
    double cs;
    int poisson_mode, pressure_iterations;
    double pressure_tolerance, omega;
    int clean_iters;
    /* v3 / v4 variants of the same step: inflow perturbation amplitude (v5, v4: 0.01, v4.py:349; v3: 0.005, v3.py:329) and
     * whether the scalars nu, dt, artificial_viscosity and the adaptive dt are float32-rounded (v5.py:80-82,321,326) or
     * plain Python floats (v3 / v4: v4.py:79-80,319,324) */
    double pert_amplitude;
    int scalar_f32;
} orc_config;

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)

#define T double
#define SFX(name) CAT(name, _f64)
#include "cfd_oracle_impl.h"
#undef T
#undef SFX

#define T float
#define SFX(name) CAT(name, _f32)
#include "cfd_oracle_impl.h"
#undef T
#undef SFX

/* v2.py:284-306 (v1.py:275-297) solve_pressure_iterative: lexicographic SOR, omega = 1.5, on
 * lap(phi) = +div/dt; after each sweep Neumann left/bottom/top, Dirichlet-0 right, mask -> 0.
 * Returns sweeps done. float64 only (the v1/v2 scripts are float64). */
int orc_sor_v2(double* phi, const double* div, const unsigned char* mask, int ny, int nx,
               double dx, double dy, double dt, double omega, int max_iter, double tol, double* last_change)
{
    const double dx2 = dx * dx, dy2 = dy * dy, denom = 2.0 * (dx2 + dy2);
    size_t n = (size_t)ny * nx;
    double* old = (double*)malloc(sizeof(double) * n);
    int it; double mc = 0.0;
#define P(i, j) phi[(size_t)(i) * nx + (j)]
    for (it = 0; it < max_iter; ++it) {
        memcpy(old, phi, sizeof(double) * n);
        for (int i = 1; i < ny - 1; ++i)
            for (int j = 1; j < nx - 1; ++j) {
                if (mask && mask[(size_t)i * nx + j]) continue;
                double rhs = div[(size_t)i * nx + j] / dt;
                double gs = ((P(i, j + 1) + P(i, j - 1)) * dy2 + (P(i + 1, j) + P(i - 1, j)) * dx2 - rhs * dx2 * dy2) / denom;
                P(i, j) = (1 - omega) * P(i, j) + omega * gs;
            }
        for (int i = 0; i < ny; ++i) { P(i, 0) = P(i, 1); P(i, nx - 1) = 0.0; }
        for (int j = 0; j < nx; ++j) { P(0, j) = P(1, j); P(ny - 1, j) = P(ny - 2, j); }
        if (mask) for (size_t k = 0; k < n; ++k) if (mask[k]) phi[k] = 0.0;
        mc = 0.0;
        for (size_t k = 0; k < n; ++k) { double c = fabs(phi[k] - old[k]); if (c > mc) mc = c; }
        if (mc < tol) { ++it; break; }
    }
#undef P
    free(old);
    if (last_change) *last_change = mc;
    return it;
}

int orc_abi_version(void) { return 1; }

/* OpenMP team control for the timed CPU baseline (bench.py): torchrun exports OMP_NUM_THREADS=1 to its workers, so
 * the bench sets the team size explicitly and reports the size a parallel region REALLY gets. */
#ifdef _OPENMP
#include <omp.h>
void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_get_threads(void)
{
    int n = 1;
#pragma omp parallel
    {
#pragma omp single
        n = omp_get_num_threads();
    }
    return n;
}
#else
void orc_set_threads(int n) { (void)n; }
int orc_get_threads(void) { return 1; }
#endif
