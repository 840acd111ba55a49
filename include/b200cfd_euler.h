/*
 * b200cfd_euler.h — C ABI of the compressible-Euler finite-volume step in libb200cfd.so (sm_100a):
 * the B200 implementation of ShockwaveSolver.time_step of Santhosh-Sathyamurthy/cfd-simulations,
 * python/shockwave_analysis/v1_shock.py:291-317 (BASELINE.json configs[4]).
 *
 * The reference has no FFI: its boundary is the Python method surface of ShockwaveSolver (v1_shock.py:225-328) and
 * the module-level @njit functions (v1_shock.py:69-223); each entry point cites what it replaces.  Conventions are
 * those of b200cfd.h: caller-owned DEVICE pointers, work enqueued on the caller's cudaStream_t (void*), 0 on success /
 * negative B200CFD_E* codes, b200cfd_euler_last_error() for the message, no CPU fallback.
 *
 * Layout: U is (nx, ny, 4) float32, y then component fastest — the reference's own layout (v1_shock.py:252), i.e.
 * one float4 per cell.  Scalar arithmetic is float64 as in the reference (Numba promotes f32*f64 -> f64).
 * The state ping-pongs between two caller-owned buffers; which one is current is device-resident (a step that the
 * reference would skip on a non-finite dt, v1_shock.py:294-296, leaves it unchanged).
 */
#ifndef B200CFD_EULER_H
#define B200CFD_EULER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { B200CFD_EULER_HLLC = 0, B200CFD_EULER_ROE = 1 };   /* ShockwaveConfig.solver_type, v1_shock.py:59 */

/* Mirrors ShockwaveConfig (v1_shock.py:35-67) plus the x-slab decomposition (x is the slow axis: halos are
 * contiguous rows of ny*4 floats). */
typedef struct b200cfd_euler_config {
    int32_t nx, ny;             /* LOCAL array shape: rows (x, incl. halo rows of a slab) and columns (y) */
    int32_t solver_type;        /* B200CFD_EULER_HLLC | B200CFD_EULER_ROE */
    int32_t n_reflect;          /* columns j < n_reflect take the reflective branch of v1_shock.py:284-286 (0 as shipped) */
    int32_t rank, nranks;       /* slab index / count (single GPU: 0 of 1) */
    int32_t halo_lo, halo_hi;   /* halo rows stored below / above the owned rows (0 or 1) */
    int32_t exact_division;     /* 1: an IEEE division wherever the reference divides; 0: one reciprocal per cell (<= 1 ulp f64 apart) */
    int32_t reserved;
    double  gamma, epsilon, max_val, cfl, dx, dy;
    double  rho_inf, u_inf, p_inf;
} b200cfd_euler_config;

typedef struct b200cfd_euler b200cfd_euler;

const char* b200cfd_euler_last_error(const b200cfd_euler* h);   /* h may be NULL: last creation error */

/* ShockwaveSolver.__init__ (v1_shock.py:226-231), device part: control block + dt ring. */
int b200cfd_euler_create(const b200cfd_euler_config* cfg, int device, b200cfd_euler** out);
int b200cfd_euler_destroy(b200cfd_euler* h);

/* Bind the two state buffers (each nx*ny*4 float32) and the wedge mask (uint8 nx*ny, v1_shock.py:239-248).
 * U_a holds the current state.  Enqueues compute_dt (v1_shock.py:263-275) on it and sets self.step = step. */
int b200cfd_euler_bind(b200cfd_euler* h, void* U_a, void* U_b, const uint8_t* wedge_mask, int64_t step, void* stream);

/* ShockwaveSolver.time_step() x n (v1_shock.py:291-317): fused HLLC/Roe fluxes + update_state + boundary conditions
 * + wedge zeroing + the next step's compute_dt reduction; the state and dt stay in HBM.  use_graph != 0 replays a
 * captured CUDA graph per step. */
int b200cfd_euler_step(b200cfd_euler* h, int n_steps, int use_graph, void* stream);

/* Same through a HOST array as the reference's time_step() sees self.U: upload, n steps, download (synchronises).
 * dt_last (host, may be NULL) receives the last step's dt. */
int b200cfd_euler_step_host(b200cfd_euler* h, void* U_host, int n_steps, int use_graph, double* dt_last, void* stream);

/* x-slab stepping: step_begin = the fused kernel; the caller then all-reduces (MAX) the 4 doubles at
 * b200cfd_euler_reduction_ptr() over the slabs and exchanges the halo rows of the NEW current buffer;
 * step_end(reduced) closes the step (next dt, ring record, buffer flip). */
int   b200cfd_euler_step_begin(b200cfd_euler* h, void* stream);
int   b200cfd_euler_step_end(b200cfd_euler* h, const double* reduced_dev, void* stream);
void* b200cfd_euler_reduction_ptr(b200cfd_euler* h);
int   b200cfd_euler_reduction_len(void);
/* compute_dt over the slabs at start-up: local maxima -> reduction vector (then all-reduce, then set_dt_from). */
int   b200cfd_euler_local_speeds(b200cfd_euler* h, void* stream);
int   b200cfd_euler_set_dt_from(b200cfd_euler* h, const double* reduced_dev, int64_t step, void* stream);

/* Host view of the device-resident control state {step, index of the current buffer (0 = U_a), dt of the NEXT step,
 * steps skipped so far}. Synchronises. */
int b200cfd_euler_state(b200cfd_euler* h, int64_t* step, int32_t* current, double* next_dt, int64_t* skipped, void* stream);

/* dt returned by time_step() for steps [first, first+n) (0.0 where the reference returns early). Synchronises. */
int b200cfd_euler_fetch_dts(b200cfd_euler* h, int64_t first, int n, double* out, void* stream);

int b200cfd_euler_launches_per_step(b200cfd_euler* h);

/* ---- one export per reference function, for unit parity (device pointers) ---- */
/* hllc_solver / roe_solver (v1_shock.py:147-209 / 97-145) on n face pairs: UL, UR, out are (n, 4) float32. */
int b200cfd_euler_k_riemann(const void* UL, const void* UR, int n, int solver_type, int x_normal,
                            double gamma, double epsilon, double max_val, void* out, void* stream);
/* compute_fluxes (v1_shock.py:84-95): F, G (n, 4) float32 */
int b200cfd_euler_k_fluxes(const void* U, int n, double gamma, double epsilon, double max_val, void* F, void* G, void* stream);
/* update_state (v1_shock.py:211-223) with explicit flux arrays F, G of U's shape; border cells of U_new are 0. */
int b200cfd_euler_k_update_state(b200cfd_euler* h, const void* U, const void* F, const void* G, double dt, void* U_new, void* stream);
/* compute_dt (v1_shock.py:263-275) of an arbitrary state of the handle's shape; out3 (host) = {dt, max|u|+a, max|v|+a}. Synchronises. */
int b200cfd_euler_k_compute_dt(b200cfd_euler* h, const void* U, double* out3, void* stream);
/* apply_boundary_conditions + wedge zeroing (v1_shock.py:277-289,312-313) in place */
int b200cfd_euler_k_apply_bc(b200cfd_euler* h, void* U, const uint8_t* wedge_mask, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200CFD_EULER_H */
