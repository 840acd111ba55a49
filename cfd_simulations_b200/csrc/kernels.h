// kernels.h — host-side launchers of the CUDA kernels (one translation unit per kernel family).
#pragma once
#include "common.cuh"

namespace b200cfd {

struct Phys {
    double nu_eff;        // nu + av in the field dtype (v5.py:388, LES off) ; base value when LES on
    double V_inf, y_max, max_velocity;
    double inv_cfg_dt;    // 1/cfg.dt  (Poisson RHS uses the CONFIG dt, v5.py:210,333)
    int use_supg;
    int pad;
};

// control: prepare ctl for `step` from in_max_u/v; close a step into the scalar ring
struct CtlParams {
    double cfl_target, dt_min, dt_max, dt_base, hmin, nu, av;
    int adaptive_dt, initial_steps, is_f32, use_les;
    double n_cells_global;      // ny * nx_global: denominator of mean(nu_t) (v5.py:323)
    double pert_amplitude;      // inflow perturbation amplitude (v5 / v4: 0.01, v3: 0.005)
    int scalar_f32, pad;        // warm-up / adaptive dt rounded to float32 (v5) or kept as doubles (v3 / v4)
};
#define NUT_PARTS 256           // fixed number of partial sums of nu_t (deterministic two-stage reduction)
// sum of nu_t over the OWNED columns into NUT_PARTS partials (LES only; feeds adaptive_time_step's mean(nu_t))
void launch_field_sum(int dtype, const Geom& g, const void* f, double* partials, cudaStream_t s);

// ---- unit kernels (one per reference @njit function) -------------------------------------------
void launch_supg_tau(int dtype, const Geom& g, const void* u, const void* v, void* tau, double dt, double nu_eff,
                     const void* nu_t, cudaStream_t s);
void launch_convection(int dtype, const Geom& g, const void* u, const void* v, const void* phi, const void* tau,
                       void* conv, int supg, cudaStream_t s);
void launch_laplacian(int dtype, const Geom& g, const void* phi, void* lap, double nu_eff, const void* nu_t, cudaStream_t s);
void launch_divergence(int dtype, const Geom& g, const void* u, const void* v, void* div, unsigned long long* max_out,
                       cudaStream_t s);
void launch_gradient(int dtype, const Geom& g, const void* phi, void* gx, void* gy, cudaStream_t s);
void launch_smagorinsky(int dtype, const Geom& g, const void* u, const void* v, void* nu_t, double cs, cudaStream_t s);
void launch_apply_bc(int dtype, const Geom& g, void* u, void* v, const double* y, double V_inf, double y_max,
                     long long step, cudaStream_t s, double amplitude = 0.01);
void launch_apply_ibm(int dtype, const Geom& g, void* u, void* v, const double* ibm, double fs, cudaStream_t s);

// ---- fused step kernels ------------------------------------------------------------------------
// K1: u*, v* = IBM(BC(u + dt(-C + L)))        reads ctl->dt/fs/pert (device)   v5.py:388-407
void launch_predictor(int dtype, const Geom& g, const Phys& p, const StepCtl* ctl, const void* u, const void* v,
                      void* us, void* vs, void* tau_or_null, const void* nu_t_or_null, const double* ibm,
                      const double* y, cudaStream_t s);
// K3: u = u* - dt grad(phi), max|grad phi|                                     v5.py:413-417
void launch_correct(int dtype, const Geom& g, StepCtl* ctl, const void* us, const void* vs, const void* phi,
                    void* u, void* v, cudaStream_t s);
// K5: post-div max, BC, IBM, vorticity max, energy sum, clip, post-clip maxima  v5.py:421-438
void launch_finalize(int dtype, const Geom& g, const Phys& p, StepCtl* ctl, const void* ua, const void* va,
                     void* u, void* v, const double* ibm, const uint8_t* mask, const double* y,
                     double* ke_partials, unsigned long long* cnt_partials, cudaStream_t s);
int finalize_num_partials(const Geom& g);
// compute_vorticity (v5.py:365-373), for the snapshot writer
void launch_vorticity(int dtype, const Geom& g, const void* u, const void* v, const uint8_t* mask, void* w, cudaStream_t s);
void launch_ctl_init(StepCtl* ctl, const CtlParams& cp, long long step, int use_measured_max, const double* nut_partials,
                     cudaStream_t s);
void launch_ctl_reduce_local(StepCtl* ctl, const double* ke_partials, const unsigned long long* cnt_partials,
                             int n_partials, double* red, int close_rb1, double tol, const double* nut_partials, cudaStream_t s);
// one GPU: reduce_local + end_step in one launch
void launch_ctl_reduce_end(StepCtl* ctl, const CtlParams& cp, const double* ke_partials, const unsigned long long* cnt_partials,
                           int n_partials, double* red, int close_rb1, double tol, const double* nut_partials,
                           StepScalars* ring, int ring_cap, cudaStream_t s);
void launch_ctl_end_step(StepCtl* ctl, const CtlParams& cp, const double* gathered, int nranks, StepScalars* ring,
                         int ring_cap, cudaStream_t s);
// ---- x-slab halos: strided columns <-> contiguous buffer [ny][ncols] (halo.cu) -------------------
void launch_pack_columns(int es, const Geom& g, const void* field, int j0, int ncols, void* buf, cudaStream_t s);
void launch_unpack_columns(int es, const Geom& g, void* field, int j0, int ncols, const void* buf, cudaStream_t s);
// all fields (nf <= 4) and both sides in ONE launch: columns [jl, jl+nl) <-> bl[nf][ny][nl], [jr, jr+nr) <-> br[nf][ny][nr]
void launch_halo_rows(int es, bool pack, const Geom& g, void* const* fields, int nf, int jl, int nl, void* bl, int jr, int nr,
                      void* br, cudaStream_t s);
void launch_health(int dtype, const Geom& g, const void* u, const void* v, unsigned long long* out4, cudaStream_t s);
void launch_max_abs2(int dtype, const Geom& g, const void* u, const void* v, StepCtl* ctl, cudaStream_t s);

// ---- row-marching fast kernels (step_fast.cu): SUPG on, LES off, even nx, 16-byte aligned fields ----
int tile_flags_dims(const Geom& g, int* ntx, int* nty);     // returns ntx * nty (bytes of the flag map)
void launch_tile_flags(const Geom& g, const double* ibm, const uint8_t* mask, uint8_t* flags, cudaStream_t s);
bool fast_path_ok(const Geom& g, const void* const* ptrs, int n);
// after_ctl: the kernel enqueued right before this one on `s` is this handle's k_ctl_reduce_end (the previous step of the
// same b200cfd_step call): the launch may then overlap it (programmatic dependent launch, see pdl_release)
void launch_predictor_fast(int dtype, const Geom& g, const Phys& p, const StepCtl* ctl, const void* u, const void* v,
                           void* us, void* vs, void* tau_or_null, const double* ibm, const double* y,
                           const uint8_t* flags, int n_sm, cudaStream_t s, bool after_ctl = false);
void launch_div_rb1_fast(int dtype, const Geom& g, const void* us, const void* vs, const uint8_t* mask, void* phi,
                         void* div_out, double inv_dt, StepCtl* ctl, const uint8_t* flags, int n_sm, cudaStream_t s);
// K2 + K3 fused (as executed, omega = 1): div, one red-black iteration, velocity correction, three maxima
void launch_div_rb1_correct_fast(int dtype, const Geom& g, const void* us, const void* vs, const uint8_t* mask, void* phi,
                                 void* div_out, void* ua, void* va, double inv_dt, StepCtl* ctl, const uint8_t* flags,
                                 int n_sm, cudaStream_t s);
int finalize_fast_num_partials(const Geom& g, int n_sm);
// part 0: every column block.  x-slabs split the launch so that the halo exchange can start while the bulk of the kernel
// still runs: part 1 = the column blocks that overlap [0, c_left) or [c_right, nx) (the halo columns and the owned columns
// the neighbours need), part 2 = all the others.
void launch_finalize_fast(int dtype, const Geom& g, const Phys& p, StepCtl* ctl, const void* ua, const void* va,
                          const void* psi_or_null, const void* psi1_or_null, void* u, void* v, const double* ibm, const uint8_t* mask,
                          const double* y, double* ke_partials, unsigned long long* cnt_partials, const uint8_t* flags,
                          int n_sm, cudaStream_t s, int part = 0, int c_left = 0, int c_right = 0);
// K5 with its rows staged by TMA (finalize_tma.cu); psi (the last pending correction) is required
bool finalize_tma_ok(int dtype, const Geom& g, const void* ua, const void* va, const void* psi, const void* psi1);
int finalize_tma_num_partials(const Geom& g, int n_sm);
bool launch_finalize_tma(int dtype, const Geom& g, const Phys& p, StepCtl* ctl, const void* ua, const void* va, const void* psi,
                         const void* psi1, void* u, void* v, const double* ibm, const uint8_t* mask, const double* y,
                         double* ke_partials, unsigned long long* cnt_partials, const uint8_t* flags, int n_sm, cudaStream_t s);
void launch_rb1_close(StepCtl* ctl, double tol, cudaStream_t s);

// ---- Poisson -----------------------------------------------------------------------------------
// one red-black colour half-sweep (global colour parity), fused max|dphi|, last block closes the iteration
void launch_rbgs_color(int dtype, const Geom& g, void* phi, const void* div, const uint8_t* mask, double inv_dt,
                       double omega, int color, StepCtl* ctl, double tol, int close_iteration, cudaStream_t s);
void launch_jacobi(int dtype, const Geom& g, const void* phi_in, void* phi_out, const void* div, const uint8_t* mask,
                   double dx2_over_dt, cudaStream_t s);
// Neumann left/bottom/top, Dirichlet-0 right of v1/v2 solve_pressure_iterative (v2.py:299-303)
void launch_sor_v2_bc(int dtype, const Geom& g, void* phi, const StepCtl* ctl, cudaStream_t s);
void launch_poisson_reset(StepCtl* ctl, cudaStream_t s);
// fused div(u*,v*) + exactly one red-black iteration from phi = 0 ("as executed"), + max|div|
void launch_div_rb1(int dtype, const Geom& g, const void* us, const void* vs, const uint8_t* mask, void* phi,
                    void* div_out, double inv_dt, StepCtl* ctl, double tol, cudaStream_t s);

// ---- divergence clean-up (lexicographic GS by independent overlapped warp tiles) -----------------
struct CleanPlan {
    int strip_w, tile_h, apron_h, apron_v; int n_strips, n_tiles; int j_first, j_end;
    int pfd;        // L2 prefetch distance of the lean sweep, rows
    int two_pass;   // 1: tiles sweep only their own rows from zero inflow, then k_clean_corr adds the homogeneous response
                    //    to the true inflow over the first apron_v rows of every tile (cleandiv.cu); 0: apron rows re-swept
};
CleanPlan clean_plan(const Geom& g, int dtype, int n_sm);
void launch_clean_sweep(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v,
                        const void* psi_old /*NULL = zeros*/, void* psi_new, void* uo /*NULL = no fused correction*/, void* vo,
                        cudaStream_t s);
// the same sweep with its rows staged in shared memory by TMA (cleandiv_tma.cu).  mode 0: psi_old = 0; 1: psi_old streamed;
// 2: psi_old = psi_1 streamed and the correction u,v -= grad(psi_1) applied on the load side.  false = not applicable / failed
bool clean_sweep_tma_ok(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v, const void* psi_old,
                        const void* psi_new);
bool launch_clean_sweep_tma(int dtype, const Geom& g, const CleanPlan& cp, const void* u, const void* v, const void* psi_old,
                            void* psi_new, int mode, cudaStream_t s);
int clean_sweep_launches(const CleanPlan& cp);   // kernels one launch_clean_sweep enqueues (2 with the two-pass plan)
void launch_clean_apply(int dtype, const Geom& g, void* u, void* v, const void* psi, cudaStream_t s);

// message returned by b200cfd_last_error(NULL) on this thread (c_api.cu), for entry points that have no handle
void set_global_error(const char* msg);

}  // namespace b200cfd
