/*
 * b200cfd.h — C ABI of libb200cfd.so: the B200 (sm_100a) implementation of the projection time step of
 * Santhosh-Sathyamurthy/cfd-simulations, python/flow_over_cylinder (Fischer)/v5.py.
 *
 * The reference has no FFI/plugin interface: its boundary is the Python method surface of
 * OptimizedTurbulentSolver (v5.py:259-441) and the module-level @njit kernels (v5.py:96-257).  Each
 * entry point below cites the reference function it replaces.  Conventions:
 *   - every field pointer is a DEVICE pointer owned by the caller (e.g. a torch tensor's data_ptr());
 *     the library never frees caller memory and owns only its scratch inside the opaque handle;
 *   - arrays are row-major (ny, nx), x fastest, dtype = B200CFD_F64 | B200CFD_F32 (the reference's
 *     memory_efficient switch, v5.py:287); ibm_mask is always float64, cylinder mask is uint8 (bool);
 *   - all work is enqueued on the caller's cudaStream_t (passed as void*); no host synchronisation
 *     unless the function's comment says "synchronises";
 *   - return 0 on success, negative B200CFD_E* otherwise; b200cfd_last_error() gives a message;
 *   - one handle per GPU, single caller per handle (the reference is single-threaded Python).
 * There is no CPU fallback: without a CUDA device every compute entry point returns B200CFD_ECUDA.
 */
#ifndef B200CFD_H
#define B200CFD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200CFD_ABI_VERSION 3

enum { B200CFD_F64 = 0, B200CFD_F32 = 1 };
enum { B200CFD_OK = 0, B200CFD_EINVAL = -1, B200CFD_ECUDA = -2, B200CFD_ENOMEM = -3, B200CFD_ESTATE = -4 };

/* How "the reference's pressure solve" is interpreted (SURVEY.md finding 3). */
enum {
    B200CFD_POISSON_AS_EXECUTED = 0, /* v5.py:202-226 under Numba parallel=True: exactly 1 red-black iteration */
    B200CFD_POISSON_AS_WRITTEN  = 1, /* same source, true max|dphi| reduction, tolerance break, <= N iterations */
    B200CFD_POISSON_JACOBI      = 2, /* v5.py:336-346 (use_fast_pressure=False): fixed sweep count, 0.25 / dx^2 */
    B200CFD_POISSON_SOR_V2      = 3  /* v2.py:284-306 / v1.py:275-297 solve_pressure_iterative: SOR on lap(phi) = +div/dt,
                                        Neumann left/bottom/top, Dirichlet-0 right after every sweep, max|dphi| < tol stop;
                                        red-black ordering (same converged field as the reference's lexicographic sweep);
                                        b200cfd_poisson only, not a time-step mode */
};
enum {
    B200CFD_CLEANDIV_LEXICOGRAPHIC = 0, /* v5.py:250-253 at one Numba thread: lexicographic Gauss-Seidel (parity mode) */
    B200CFD_CLEANDIV_REDBLACK      = 1  /* NOT reference parity: red-black ordering of the same sweep */
};

/* Mirrors OptimizedTurbulentConfig (v5.py:41-94). nu, av, cfg_dt are the float32-rounded values the
 * reference stores (v5.py:80-82) widened to double. */
typedef struct b200cfd_config {
    int32_t ny, nx;              /* LOCAL array shape (rows, columns incl. halo columns of a slab) */
    int32_t dtype;               /* B200CFD_F64 | B200CFD_F32 */
    int32_t poisson_mode, pressure_iterations, cleandiv_mode, clean_iters;
    int32_t initial_steps, use_supg, use_les, adaptive_dt;
    double  dx, dy;
    double  nu, av, cfg_dt;
    double  V_inf, y_max;
    double  cfl_target, dt_min, dt_max, dt_base, max_velocity;
    double  cs;                  /* smagorinsky_constant */
    double  pressure_tolerance, omega;
    /* x-slab decomposition (single GPU: rank 0 of 1, x_offset 0, nx_global = nx, halos 0).
     * x_offset = global column of local column 0; local columns [halo_left, nx - halo_right) are owned,
     * the others are halo columns (recomputed redundantly, refreshed by the caller once per step). */
    int32_t rank, nranks, nx_global, x_offset, halo_left, halo_right;
    /* v3 / v4 variants of the same step, honoured when variant_set != 0 (else the v5 values 0.01 / 1):
     * pert_amplitude = inflow perturbation amplitude (v5.py:351, v4.py:349: 0.01; v3.py:329: 0.005);
     * scalar_f32 = the warm-up dt and the adaptive dt are float32-rounded (v5.py:321,326) or plain doubles (v4.py:319,324). */
    double  pert_amplitude;
    int32_t scalar_f32, variant_set;
} b200cfd_config;

/* Per-step diagnostics, the five values v5.py logs each step (:410,:415,:422,:429,:435) plus dt. */
typedef struct b200cfd_step_scalars {
    double dt;              /* adaptive_time_step() result used by this step (v5.py:316-326) */
    double max_div_pre;     /* max|div u*|   (v5.py:410) */
    double max_grad_phi;    /* max|grad phi| (v5.py:414-415) */
    double max_div_post;    /* max|div u| after clean_divergence, before BC (v5.py:421-422) */
    double max_vorticity;   /* nanmax|omega| (v5.py:427-429) */
    double mean_energy;     /* nanmean(0.5(u^2+v^2)) -> energy_history (v5.py:431-433) */
    double max_abs_u, max_abs_v; /* post-clip maxima: what adaptive_time_step / monitor_simulation_health see */
    double poisson_iters, poisson_change;
    double nonfinite;       /* 1.0 if any non-finite u,v was seen (monitor_simulation_health, v5.py:601) */
    double step;            /* step index this record belongs to */
} b200cfd_step_scalars;

typedef struct b200cfd_handle b200cfd_handle;

int         b200cfd_abi_version(void);
const char* b200cfd_last_error(const b200cfd_handle* h);   /* h may be NULL: last creation error */
int         b200cfd_device_count(void);

/* OptimizedTurbulentSolver.__init__ (v5.py:260-267) device part: scratch + control block. */
int b200cfd_create(const b200cfd_config* cfg, int device, b200cfd_handle** out);
int b200cfd_destroy(b200cfd_handle* h);

/* Bind caller-owned state arrays (names = the reference's attributes, v5.py:285-297).
 * tau_supg / div_u_star may be NULL (then they are not materialised). y = self.y (float64[ny]). */
int b200cfd_bind_fields(b200cfd_handle* h, void* u, void* v, void* phi, void* u_star, void* v_star,
                        void* div_u_star, void* tau_supg, void* nu_t,
                        const double* ibm_mask, const uint8_t* cylinder_mask, const double* y);

/* self.step = step (and re-derive what adaptive_time_step needs from the bound u, v). */
int b200cfd_set_step(b200cfd_handle* h, int64_t step, void* stream);
int b200cfd_get_step(b200cfd_handle* h, int64_t* step);

/* OptimizedTurbulentSolver.time_step() x n_steps (v5.py:375-441); state stays in HBM, dt and
 * diagnostics go to a device ring buffer. use_graph != 0 replays one captured CUDA graph per step. */
int b200cfd_step(b200cfd_handle* h, int n_steps, int use_graph, void* stream);

/* Same, but through HOST arrays as the reference's time_step() sees them: uploads u,v (pinned or
 * pageable host memory), runs n_steps, downloads u,v and the last step's scalars. Synchronises. */
int b200cfd_step_host(b200cfd_handle* h, void* u_host, void* v_host, int n_steps, int use_graph,
                      b200cfd_step_scalars* last, void* stream);

/* ---- x-slab (multi-GPU) stepping: one process per GPU, one handle per slab -------------------------------
 * time_step() = step_begin (all kernels of the step except the closing one; fills this slab's reduction vector)
 *   -> caller all_gathers the b200cfd_reduction_len() doubles at b200cfd_reduction_ptr() over the slabs (NCCL)
 *   -> step_end(gathered, nranks): global maxima / energy, ring record, next dt (adaptive_time_step v5.py:316-326)
 *   -> caller refreshes the halo columns of u and v with pack_columns / NCCL send-recv / unpack_columns.
 * b200cfd_required_halo gives the halo widths for which that single exchange per step is sufficient. */
int   b200cfd_required_halo(const b200cfd_config* cfg, int* left, int* right);
int   b200cfd_step_begin(b200cfd_handle* h, int use_graph, void* stream);
int   b200cfd_step_end(b200cfd_handle* h, const double* gathered_dev, int nranks, void* stream);
/* Overlap of the halo exchange with the tail of the step: when b200cfd_step_splits(h) is 1, step_begin
 * can be issued in two parts.  part 1 = the step up to the finalize of the column blocks at the slab's edges (its halo
 * columns and the owned columns its neighbours need); after it the caller starts b200cfd_halo_rows(pack) -> NCCL
 * send/recv -> b200cfd_halo_rows(unpack) on a communication stream, while part 2 (the finalize of every other column
 * block + this slab's reduction vector) runs on the compute stream. */
int   b200cfd_step_splits(b200cfd_handle* h);
int   b200cfd_step_part(b200cfd_handle* h, int part, int use_graph, void* stream);
/* All fields (1..4) and both sides of a halo refresh in ONE launch (replaces 2 x nfields pack_columns calls). */
int   b200cfd_halo_rows(b200cfd_handle* h, int pack, void* const* fields, int nfields, int j_left, int n_left, void* buf_left,
                        int j_right, int n_right, void* buf_right, void* stream);
void* b200cfd_reduction_ptr(b200cfd_handle* h);
/* Device address of one double: max|dphi| over the OWNED cells of the last completed red-black iteration of
 * b200cfd_poisson (the reference's max_change, v5.py:212-225).  The slab Poisson solve all-reduces (MAX) it over the ranks
 * (NCCL) instead of reading it back per iteration: see OptimizedTurbulentSolver.solve_pressure_slabs. */
void* b200cfd_poisson_change_ptr(b200cfd_handle* h);
int   b200cfd_reduction_len(void);
/* self.step = step with the GLOBAL max|u|, max|v| of the state (what adaptive_time_step reads). Synchronises. */
int   b200cfd_set_step_global(b200cfd_handle* h, int64_t step, double max_abs_u, double max_abs_v, void* stream);
/* field[:, j0:j0+ncols] -> buf[ny][ncols] and back (device pointers; dtype of the handle). */
int   b200cfd_pack_columns(b200cfd_handle* h, const void* field, int j0, int ncols, void* buf, void* stream);
int   b200cfd_unpack_columns(b200cfd_handle* h, void* field, int j0, int ncols, const void* buf, void* stream);

/* monitor_simulation_health (v5.py:599-613) reductions over the bound u, v:
 * out4 = {nonfinite (0/1), max|u|, max|v|, max|div(u,v)|}. Synchronises. */
int b200cfd_health(b200cfd_handle* h, double* out4, void* stream);

/* Per-phase CUDA-event timing of b200cfd_step(use_graph = 0): begin, run steps, end.  end synchronises, writes the
 * ';'-joined phase names and the mean milliseconds per step of each phase, returns the number of phases. */
int b200cfd_profile_begin(b200cfd_handle* h);
int b200cfd_profile_end(b200cfd_handle* h, char* names, int names_cap, double* ms, int cap);

/* Kernels one time step enqueues (as counted by the last b200cfd_step call). */
int b200cfd_launches_per_step(b200cfd_handle* h);
/* Tiling of the lexicographic clean-divergence sweep: {strip_w, tile_h, apron_h, apron_v, n_strips, n_tiles,
 * n_sm, n_energy_partials}. */
void b200cfd_get_clean_plan(b200cfd_handle* h, int* out8);

/* Copy the scalars of steps [first_step, first_step+n) out of the ring. Synchronises the stream. */
int b200cfd_fetch_scalars(b200cfd_handle* h, int64_t first_step, int n, b200cfd_step_scalars* out, void* stream);

/* solve_pressure_fast / solve_pressure_gauss_seidel_fast (v5.py:328-347, 202-226) on arbitrary
 * device arrays: phi (in/out, NOT zeroed here), rhs_div = div_u_star, mask may be NULL.
 * mode: AS_EXECUTED (1 iteration), AS_WRITTEN (<= max_iter, tolerance break, exact per-iteration max),
 * JACOBI (max_iter sweeps; uses dx only), SOR_V2 (v1/v2 fallback solver, omega = 1.5 there).  omega = 1 is the v5
 * reference; omega != 1 is red-black SOR.
 * temporal_depth T >= 1: iterations fused per HBM pass by the shared-memory (TMA) kernel; 0 = auto.
 * iters_done / last_change (host pointers, may be NULL) are written after a stream sync when non-NULL. */
int b200cfd_poisson(b200cfd_handle* h, void* phi, const void* rhs_div, const uint8_t* mask,
                    double dx, double dy, double dt, int mode, int max_iter, double tol, double omega,
                    int temporal_depth, int* iters_done, double* last_change, void* stream);
/* What the last b200cfd_poisson call did: out4 = {temporal depth used (1 = the one-sweep-per-pass kernels), HBM passes
 * of the temporally blocked kernel (0 if it did not run), bytes per cell one such pass moves, reserved}. */
int b200cfd_poisson_info(b200cfd_handle* h, int* out4);
/* keep != 0: the following b200cfd_poisson calls use the SAME, unchanged rhs_div and mask arrays as the previous call (one
 * solve done in chunks around halo refreshes: the slab solver); the temporally blocked kernel then does not rebuild its
 * classified copy of the right-hand side per call.  Call with keep = 0 before the right-hand side changes. */
int b200cfd_poisson_keep_rhs(b200cfd_handle* h, int keep);

/* ---- one export per reference kernel, for unit parity (all device pointers, shape of the handle) ---- */
/* compute_supg_stabilization_fast (v5.py:149-162) */
int b200cfd_k_supg_tau(b200cfd_handle* h, const void* u, const void* v, void* tau, double dt, double nu_eff, void* stream);
/* compute_convection_supg_fast (v5.py:127-147) / compute_convection_fast (v5.py:112-125) */
int b200cfd_k_convection(b200cfd_handle* h, const void* u, const void* v, const void* phi, const void* tau,
                         void* conv, int supg, void* stream);
/* compute_laplacian_fast (v5.py:164-176) */
int b200cfd_k_laplacian(b200cfd_handle* h, const void* phi, void* lap, double nu_eff, void* stream);
/* compute_divergence_fast (v5.py:178-187) */
int b200cfd_k_divergence(b200cfd_handle* h, const void* u, const void* v, void* div, void* stream);
/* compute_gradient_fast (v5.py:189-200) */
int b200cfd_k_gradient(b200cfd_handle* h, const void* phi, void* gx, void* gy, void* stream);
/* compute_smagorinsky_viscosity_fast (v5.py:96-110) */
int b200cfd_k_smagorinsky(b200cfd_handle* h, const void* u, const void* v, void* nu_t, double cs, void* stream);
/* OptimizedTurbulentSolver.compute_vorticity (v5.py:365-373): 0 on the border, NaN where mask != 0 (mask may be NULL);
 * used by the snapshot writer (v5.py:454-470) so that only finished fields cross PCIe */
int b200cfd_k_vorticity(b200cfd_handle* h, const void* u, const void* v, const uint8_t* mask, void* w, void* stream);
/* apply_boundary_conditions (v5.py:349-360) for an explicit step index */
int b200cfd_k_apply_bc(b200cfd_handle* h, void* u, void* v, const double* y, int64_t step, void* stream);
/* apply_ibm_fast (v5.py:228-237) */
int b200cfd_k_apply_ibm(b200cfd_handle* h, void* u, void* v, const double* ibm, double force_strength, void* stream);
/* clean_divergence_fast (v5.py:239-257); psi_out may be NULL */
int b200cfd_k_clean_divergence(b200cfd_handle* h, void* u, void* v, int iterations, int mode, void* psi_out, void* stream);
/* fused predictor u*, v* = BC(IBM(u + dt(-C + L))) (v5.py:388-407) for an explicit step / dt */
int b200cfd_k_predictor(b200cfd_handle* h, const void* u, const void* v, void* u_star, void* v_star,
                        void* tau_or_null, const double* ibm, const double* y, int64_t step, double dt, void* stream);

/* ---- allocation-free entry point ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * `sweeps` Jacobi sweeps of solve_pressure_fast (the arithmetic of b200cfd_poisson with mode = JACOBI, temporal_depth = 1)
 * on dense row-major (ny, nx) device arrays of the CURRENT device: phi in/out (not zeroed here), rhs_div = div_u_star,
 * mask may be NULL.  dtype = B200CFD_F64 / B200CFD_F32.  Needs no handle and calls no cudaMalloc / cudaFree / allocating
 * library routine: the ping-pong buffer is a file-scope device array of B200CFD_STATIC_SCRATCH_BYTES that every GPU which
 * loads the library holds until the process exits.  ny * nx * sizeof(element) beyond that size is rejected with
 * B200CFD_EINVAL (b200cfd_last_error(NULL) says so); use b200cfd_poisson there.  All calls on a device share the array and
 * it keeps the last call's contents (nothing in it has to start at zero: a call's first sweep writes every cell it later
 * reads); calls that overlap on the GPU (different streams, no ordering between them) race on it.  Asynchronous on
 * `stream`. */
#define B200CFD_STATIC_SCRATCH_BYTES (8u << 20)   /* 8 MiB: 1024 x 1024 fp64 cells, 2048 x 1024 fp32 cells */
int b200cfd_jacobi_static(int dtype, int ny, int nx, void* phi, const void* rhs_div, const uint8_t* mask,
                          double dx, double dt, int sweeps, void* stream);

/* ---- test hook (host only, no CUDA call) ----
 * The cost-balanced tile plan of the streaming Poisson kernel (csrc/poisson_stream_plan.h), run on the host with the code
 * the device runs.  No counterpart in the reference (v5.py:202-226 and :336-346 sweep the whole array in one prange); it
 * exists so that the CPU test-suite can check the plan's invariants: every strip's rows covered exactly once, cuts on
 * multiples of 3, at most `ntile` tiles.  cflags: [n_strips][NC] bytes (1 = the chunk runs the class selects);
 * tiles_out: ntile x 4 ints (strip, first row, one past the last row, 0; first == last: unused entry). */
int b200cfd_stream_plan_host(const uint8_t* cflags, int n_strips, int NC, int ny, int shift, int ntile, int alpha_pct,
                             int min_groups, int* tiles_out);

#ifdef __cplusplus
}
#endif
#endif /* B200CFD_H */
