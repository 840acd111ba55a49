// c_api.cu — the C ABI of libb200cfd.so (include/b200cfd.h).  Host orchestration only: every numerical
// operation is a CUDA kernel in step_kernels.cu / poisson.cu / poisson_tb.cu / cleandiv.cu.
// There is no CPU fallback anywhere in this library.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <new>
#include <string>
#include <vector>

#include "../../include/b200cfd.h"
#include "kernels.h"
#include "poisson_tb.h"

using namespace b200cfd;

static_assert(sizeof(StepScalars) == sizeof(b200cfd_step_scalars), "ring record layout");

static thread_local std::string g_create_error;

struct b200cfd_handle {
    b200cfd_config cfg;
    int device = 0, n_sm = 148;
    int es = 8;                       // element size of the field dtype
    Geom g;
    Phys phys;
    CtlParams cp;
    CleanPlan clean;
    // library-owned device memory
    StepCtl* ctl = nullptr;
    StepScalars* ring = nullptr;
    int ring_cap = 8192;
    void *ua = nullptr, *va = nullptr, *psi[2] = {nullptr, nullptr}, *tmp = nullptr;
    double* nut_part = nullptr;           // LES: NUT_PARTS partial sums of nu_t (mean(nu_t) of adaptive_time_step)
    void *ub = nullptr, *vb = nullptr;   // second velocity pair: output of the clean sweep with the correction fused in
    double* ke_part = nullptr;
    double* red = nullptr;            // this slab's reduction vector (RED_N doubles)
    unsigned long long* cnt_part = nullptr;
    int n_part = 0;                   // capacity of the partial arrays
    int n_part_used = 0;              // blocks of the finalize kernel in use
    uint8_t* flags = nullptr;         // tile-flag map of ibm / mask (step_fast.cu)
    bool fast = false;                // row-marching fast kernels applicable (decided when the fields are bound)
    bool fuse_apply = false;          // clean sweep writes the corrected velocity itself: parity-green but measured SLOWER on B200 (the sweep is issue-bound), off unless B200CFD_FUSE_APPLY is set
    bool fuse_k23 = true;             // K2 + K3 in one kernel on the fast path (B200CFD_NO_FUSE_K23 turns it off)
    bool fin_tma = false;             // finalize through the TMA-staged kernel (finalize_tma.cu); decided when the fields are bound
    bool fuse_corr = true;            // 2 clean iterations on the fast path: the second sweep applies the first correction on its load side and finalize applies both (no k_clean_apply pass); B200CFD_FUSE_CORR=0 turns it off
    // caller-owned (bound) device memory
    void *u = nullptr, *v = nullptr, *phi = nullptr, *us = nullptr, *vs = nullptr, *div = nullptr, *tau = nullptr, *nu_t = nullptr;
    const double* ibm = nullptr;
    const uint8_t* mask = nullptr;
    const double* y = nullptr;
    bool bound = false;
    long long step = 0;               // host mirror of ctl->step
    // cached CUDA graph of one time step
    cudaStream_t cap_stream = nullptr;
    cudaGraphExec_t step_graph = nullptr;
    cudaGraphExec_t multi_graph = nullptr;   // graph_steps consecutive steps in one graph (the step is device-driven, so k steps are k copies)
    int graph_steps = 8;                     // B200CFD_GRAPH_STEPS; 1 = one graph launch per step
    cudaGraphExec_t begin_graph = nullptr;   // the step without its closing kernel (multi-GPU: all_gather in between)
    cudaGraphExec_t part_graph[2] = {nullptr, nullptr};   // x-slabs: the step up to the finalize of the edge columns / the rest
    int req_halo_l = 0, req_halo_r = 0;      // the halo widths every slab of this decomposition uses (b200cfd_required_halo)
    PoissonTB* tb = nullptr;          // temporally blocked Poisson state (tensor maps, scratch)
    bool last_poisson_tb = false;     // the last b200cfd_poisson call ran through the temporally blocked kernel
    int gpu_launches = 0;             // kernels enqueued by the last b200cfd_step call (per step)
    // optional per-phase profiling (CUDA events between the kernels of a step; non-graph path only)
    bool profiling = false;
    std::vector<std::string> prof_names;
    std::vector<cudaEvent_t> prof_events;     // prof_names.size() + 1 events per profiled step, appended
    std::vector<double> prof_ms;              // accumulated per phase
    int prof_steps = 0;
    std::string err;
};

namespace b200cfd { void set_global_error(const char* msg) { g_create_error = msg; } }

static int fail(b200cfd_handle* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf; else g_create_error = buf;
    return code;
}
#define CU(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(h, B200CFD_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)
#define CHECK_LAUNCH(h) do { cudaError_t e_ = cudaPeekAtLastError(); if (e_ != cudaSuccess) return fail(h, B200CFD_ECUDA, "kernel launch: %s", cudaGetErrorString(e_)); } while (0)

static Geom make_geom(int ny, int nx, long long pitch, double dx, double dy, int jg0, int nxg, int own0 = 0, int own1 = -1) {
    Geom g;
    g.ny = ny; g.nx = nx; g.pitch = pitch; g.jg0 = jg0; g.nxg = nxg;
    g.own0 = own0; g.own1 = own1 < 0 ? nx : own1;
    g.dx = dx; g.dy = dy;
    g.hdx = 0.5 / dx; g.hdy = 0.5 / dy;                 // v5.py:131-132,182-183,193-194
    g.qdx = 0.5 * g.hdx; g.qdy = 0.5 * g.hdy;           // v5.py:137-138
    g.hdx2 = g.hdx * g.hdx; g.hdy2 = g.hdy * g.hdy;     // v5.py:141-142
    g.ax = 1.0 / (dx * dx); g.ay = 1.0 / (dy * dy);     // v5.py:168-169,207-208
    g.dinv = 1.0 / (2.0 * (g.ax + g.ay));               // v5.py:209
    g.rdx = 1.0 / dx; g.rdy = 1.0 / dy;                 // v5.py:116-117
    g.hmin = dx < dy ? dx : dy;
    g.two_dx = 2 * dx; g.two_dy = 2 * dy;
    return g;
}

extern "C" {

int b200cfd_abi_version(void) { return B200CFD_ABI_VERSION; }
const char* b200cfd_last_error(const b200cfd_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }
int b200cfd_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int b200cfd_create(const b200cfd_config* cfg, int device, b200cfd_handle** out) {
    if (!cfg || !out) return fail(nullptr, B200CFD_EINVAL, "null argument");
    if (cfg->ny < 4 || cfg->nx < 4) return fail(nullptr, B200CFD_EINVAL, "grid must be at least 4x4 (got %dx%d)", cfg->ny, cfg->nx);
    if (cfg->dtype != B200CFD_F64 && cfg->dtype != B200CFD_F32) return fail(nullptr, B200CFD_EINVAL, "bad dtype %d", cfg->dtype);
    if (!(cfg->dx > 0) || !(cfg->dy > 0)) return fail(nullptr, B200CFD_EINVAL, "dx, dy must be positive");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(nullptr, B200CFD_ECUDA, "no CUDA device (this library has no CPU fallback): %s", cudaGetErrorString(e)); }
    if (device < 0 || device >= ndev) return fail(nullptr, B200CFD_EINVAL, "device %d out of range (%d devices)", device, ndev);
    b200cfd_handle* h = new (std::nothrow) b200cfd_handle();
    if (!h) return fail(nullptr, B200CFD_ENOMEM, "host allocation failed");
    h->cfg = *cfg;
    h->device = device;
    h->es = cfg->dtype == B200CFD_F64 ? 8 : 4;
#define CUC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(nullptr, B200CFD_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); b200cfd_destroy(h); return B200CFD_ECUDA; } } while (0)
    CUC(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUC(cudaGetDeviceProperties(&prop, device));
    h->n_sm = prop.multiProcessorCount;
    const int nxg = cfg->nranks > 1 ? cfg->nx_global : cfg->nx;
    const int jg0 = cfg->nranks > 1 ? cfg->x_offset : 0;
    if (cfg->halo_left < 0 || cfg->halo_right < 0 || cfg->halo_left + cfg->halo_right >= cfg->nx) { delete h; return fail(nullptr, B200CFD_EINVAL, "bad halo widths %d/%d for nx=%d", cfg->halo_left, cfg->halo_right, cfg->nx); }
    if (jg0 < 0 || jg0 + cfg->nx > nxg) { delete h; return fail(nullptr, B200CFD_EINVAL, "slab [%d,%d) outside the global grid of %d columns", jg0, jg0 + cfg->nx, nxg); }
    h->g = make_geom(cfg->ny, cfg->nx, cfg->nx, cfg->dx, cfg->dy, jg0, nxg, cfg->halo_left, cfg->nx - cfg->halo_right);
    const bool f32 = cfg->dtype == B200CFD_F32;
    h->phys.nu_eff = f32 ? (double)(float)((float)cfg->nu + (float)cfg->av) : cfg->nu + cfg->av;   // v5.py:388 in the field dtype
    h->phys.V_inf = cfg->V_inf; h->phys.y_max = cfg->y_max; h->phys.max_velocity = cfg->max_velocity;
    h->phys.inv_cfg_dt = 1.0 / cfg->cfg_dt;
    h->phys.use_supg = cfg->use_supg; h->phys.pad = 0;
    h->cp.cfl_target = cfg->cfl_target; h->cp.dt_min = cfg->dt_min; h->cp.dt_max = cfg->dt_max; h->cp.dt_base = cfg->dt_base;
    h->cp.hmin = h->g.hmin; h->cp.nu = cfg->nu; h->cp.av = cfg->av;
    h->cp.adaptive_dt = cfg->adaptive_dt; h->cp.initial_steps = cfg->initial_steps > 0 ? cfg->initial_steps : 1;
    h->cp.is_f32 = f32; h->cp.use_les = cfg->use_les;
    h->cp.n_cells_global = (double)cfg->ny * (double)nxg;
    h->cp.pert_amplitude = cfg->variant_set ? cfg->pert_amplitude : 0.01;
    h->cp.scalar_f32 = cfg->variant_set ? cfg->scalar_f32 : 1;
    h->cp.pad = 0;
    if (cfg->use_les) { CUC(cudaMalloc(&h->nut_part, NUT_PARTS * sizeof(double))); CUC(cudaMemset(h->nut_part, 0, NUT_PARTS * sizeof(double))); }
    h->clean = clean_plan(h->g, cfg->dtype, h->n_sm);
    if (cfg->nranks > 1) b200cfd_required_halo(cfg, &h->req_halo_l, &h->req_halo_r);
    const size_t n = (size_t)cfg->ny * cfg->nx, fb = n * h->es;
    CUC(cudaMalloc(&h->ctl, sizeof(StepCtl)));
    CUC(cudaMemset(h->ctl, 0, sizeof(StepCtl)));
    CUC(cudaMalloc(&h->ring, sizeof(StepScalars) * h->ring_cap));
    CUC(cudaMemset(h->ring, 0, sizeof(StepScalars) * h->ring_cap));
    CUC(cudaMalloc(&h->ua, fb)); CUC(cudaMalloc(&h->va, fb));
    CUC(cudaMalloc(&h->ub, fb)); CUC(cudaMalloc(&h->vb, fb));
    CUC(cudaMemset(h->ub, 0, fb)); CUC(cudaMemset(h->vb, 0, fb));
    CUC(cudaMalloc(&h->psi[0], fb)); CUC(cudaMalloc(&h->psi[1], fb));
    CUC(cudaMemset(h->ua, 0, fb)); CUC(cudaMemset(h->va, 0, fb));
    CUC(cudaMemset(h->psi[0], 0, fb)); CUC(cudaMemset(h->psi[1], 0, fb));   // psi borders must be 0 and are never written
    h->n_part = finalize_num_partials(h->g);
    { const int nf = finalize_fast_num_partials(h->g, h->n_sm); if (nf > h->n_part) h->n_part = nf; }
    { const int nf = finalize_tma_num_partials(h->g, h->n_sm); if (nf > h->n_part) h->n_part = nf; }
    h->n_part_used = finalize_num_partials(h->g);
    { int ntx, nty; const int nb = tile_flags_dims(h->g, &ntx, &nty); CUC(cudaMalloc(&h->flags, nb)); CUC(cudaMemset(h->flags, 3, nb)); }
    CUC(cudaMalloc(&h->ke_part, sizeof(double) * h->n_part));
    CUC(cudaMalloc(&h->cnt_part, sizeof(unsigned long long) * (h->n_part > 4 ? h->n_part : 4)));
    CUC(cudaMalloc(&h->red, sizeof(double) * RED_N));
    CUC(cudaMemset(h->red, 0, sizeof(double) * RED_N));
    CUC(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
    launch_ctl_init(h->ctl, h->cp, 0, 0, nullptr, nullptr);
    CUC(cudaDeviceSynchronize());
#undef CUC
    *out = h;
    return B200CFD_OK;
}

int b200cfd_destroy(b200cfd_handle* h) {
    if (!h) return B200CFD_OK;
    cudaSetDevice(h->device);
    if (h->step_graph) cudaGraphExecDestroy(h->step_graph);
    if (h->multi_graph) cudaGraphExecDestroy(h->multi_graph);
    if (h->begin_graph) cudaGraphExecDestroy(h->begin_graph);
    for (cudaGraphExec_t ge : h->part_graph) if (ge) cudaGraphExecDestroy(ge);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    if (h->tb) poisson_tb_destroy(h->tb);
    cudaFree(h->ctl); cudaFree(h->ring); cudaFree(h->ua); cudaFree(h->va); cudaFree(h->ub); cudaFree(h->vb); cudaFree(h->nut_part); cudaFree(h->psi[0]); cudaFree(h->psi[1]);
    cudaFree(h->tmp); cudaFree(h->ke_part); cudaFree(h->cnt_part); cudaFree(h->red); cudaFree(h->flags);
    delete h;
    return B200CFD_OK;
}

int b200cfd_bind_fields(b200cfd_handle* h, void* u, void* v, void* phi, void* u_star, void* v_star,
                        void* div_u_star, void* tau_supg, void* nu_t,
                        const double* ibm_mask, const uint8_t* cylinder_mask, const double* y) {
    if (!h) return B200CFD_EINVAL;
    if (!u || !v || !phi || !u_star || !v_star || !ibm_mask || !y)
        return fail(h, B200CFD_EINVAL, "u, v, phi, u_star, v_star, ibm_mask and y are required");
    if (h->cfg.use_les && !nu_t) return fail(h, B200CFD_EINVAL, "use_les needs nu_t");
    h->u = u; h->v = v; h->phi = phi; h->us = u_star; h->vs = v_star; h->div = div_u_star; h->tau = tau_supg; h->nu_t = nu_t;
    h->ibm = ibm_mask; h->mask = cylinder_mask; h->y = y;
    h->bound = true;
    {   // fast kernels: the default configuration on even, 16-byte aligned grids; tile flags of ibm / mask
        const void* ptrs[] = {u, v, phi, u_star, v_star, div_u_star, tau_supg, h->ua, h->va, h->psi[0], h->psi[1]};
        h->fast = h->cfg.use_supg && !h->cfg.use_les && fast_path_ok(h->g, ptrs, 11) && !getenv("B200CFD_NO_FAST");
        h->fuse_k23 = !getenv("B200CFD_NO_FUSE_K23");
        h->fuse_apply = getenv("B200CFD_FUSE_APPLY") != nullptr;
        { const char* e = getenv("B200CFD_FUSE_CORR"); h->fuse_corr = !(e && *e == '0'); }
        { const char* e = getenv("B200CFD_GRAPH_STEPS"); if (e && atoi(e) >= 1) h->graph_steps = atoi(e); }
        CU(h, cudaSetDevice(h->device));
        launch_tile_flags(h->g, ibm_mask, cylinder_mask, h->flags, nullptr);
        CU(h, cudaDeviceSynchronize());
        h->fin_tma = h->fast && h->cfg.cleandiv_mode == B200CFD_CLEANDIV_LEXICOGRAPHIC && h->cfg.clean_iters > 0 &&
                     finalize_tma_ok(h->cfg.dtype, h->g, h->ua, h->va, h->psi[0], h->psi[1]);
        h->n_part_used = h->fin_tma ? finalize_tma_num_partials(h->g, h->n_sm)
                                    : (h->fast ? finalize_fast_num_partials(h->g, h->n_sm) : finalize_num_partials(h->g));
    }
    if (h->step_graph) { cudaGraphExecDestroy(h->step_graph); h->step_graph = nullptr; }
    if (h->multi_graph) { cudaGraphExecDestroy(h->multi_graph); h->multi_graph = nullptr; }
    if (h->begin_graph) { cudaGraphExecDestroy(h->begin_graph); h->begin_graph = nullptr; }
    for (cudaGraphExec_t& ge : h->part_graph) if (ge) { cudaGraphExecDestroy(ge); ge = nullptr; }
    return B200CFD_OK;
}

int b200cfd_set_step(b200cfd_handle* h, int64_t step, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    launch_max_abs2(h->cfg.dtype, h->g, h->u, h->v, h->ctl, s);
    if (h->cfg.use_les && h->nu_t) launch_field_sum(h->cfg.dtype, h->g, h->nu_t, h->nut_part, s);
    launch_ctl_init(h->ctl, h->cp, (long long)step, 1, h->cfg.use_les ? h->nut_part : nullptr, s);
    CHECK_LAUNCH(h);
    h->step = step;
    return B200CFD_OK;
}
int b200cfd_get_step(b200cfd_handle* h, int64_t* step) {
    if (!h || !step) return B200CFD_EINVAL;
    *step = h->step;
    return B200CFD_OK;
}

// Enqueue the Poisson solve of a time step (phi zeroed first, v5.py:331). Returns launches enqueued.
static int enqueue_step_poisson(b200cfd_handle* h, cudaStream_t s) {
    const b200cfd_config& c = h->cfg;
    const int dt = c.dtype;
    const size_t fb = (size_t)c.ny * c.nx * h->es;
    int nl = 0;
    if (c.poisson_mode == B200CFD_POISSON_AS_EXECUTED && c.omega == 1.0) {
        if (h->fast) {
            launch_div_rb1_fast(dt, h->g, h->us, h->vs, h->mask, h->phi, h->div, h->phys.inv_cfg_dt, h->ctl, h->flags, h->n_sm, s);
            launch_rb1_close(h->ctl, c.pressure_tolerance, s);
        } else {
            launch_div_rb1(dt, h->g, h->us, h->vs, h->mask, h->phi, h->div, h->phys.inv_cfg_dt, h->ctl, c.pressure_tolerance, s);
        }
        return 2;
    }
    void* div = h->div ? h->div : h->tmp;   // tmp doubles as div scratch when the caller does not want div_u_star
    launch_divergence(dt, h->g, h->us, h->vs, div, &h->ctl->max_div_pre, s); ++nl;
    cudaMemsetAsync(h->phi, 0, fb, s);
    if (c.poisson_mode == B200CFD_POISSON_JACOBI) {
        // v5.py:336-346 uses dx only; ping-pong phi <-> ua (ua is free until the correction kernel)
        const double k = c.dx * c.dx / c.cfg_dt;
        cudaMemsetAsync(h->ua, 0, fb, s);
        void* a = h->phi; void* b = h->ua;
        for (int it = 0; it < c.pressure_iterations; ++it) {
            launch_jacobi(dt, h->g, a, b, div, h->mask, k, s); ++nl;
            void* t = a; a = b; b = t;
        }
        if (a != h->phi) cudaMemcpyAsync(h->phi, a, fb, cudaMemcpyDeviceToDevice, s);
    } else {
        const int iters = c.poisson_mode == B200CFD_POISSON_AS_EXECUTED ? 1 : c.pressure_iterations;
        for (int it = 0; it < iters; ++it) {
            launch_rbgs_color(dt, h->g, h->phi, div, h->mask, h->phys.inv_cfg_dt, c.omega, 0, h->ctl, c.pressure_tolerance, 0, s);
            launch_rbgs_color(dt, h->g, h->phi, div, h->mask, h->phys.inv_cfg_dt, c.omega, 1, h->ctl, c.pressure_tolerance, 1, s);
            nl += 2;
        }
    }
    return nl;
}

static void prof_mark(b200cfd_handle* h, const char* name, cudaStream_t s, int& idx);
// out_u / out_v (in-step fast path only): where the velocity that still needs the deferred last correction lives
static int enqueue_clean(b200cfd_handle* h, void* u, void* v, int iterations, int mode, cudaStream_t s, void** last_psi, bool in_step = false,
                         bool defer_last_apply = false, void** out_u = nullptr, void** out_v = nullptr, void** first_psi = nullptr) {
    const int dt = h->cfg.dtype;
    const size_t fb = (size_t)h->cfg.ny * h->cfg.nx * h->es;
    int nl = 0;
    void* last = nullptr;
    if (first_psi) *first_psi = nullptr;
    if (mode == B200CFD_CLEANDIV_LEXICOGRAPHIC && in_step && defer_last_apply && first_psi && iterations == 2 && h->fuse_corr &&
        clean_sweep_tma_ok(dt, h->g, h->clean, u, v, h->psi[0], h->psi[1])) {
        // both corrections deferred: sweep 1 (psi_old = 0) -> psi[0]; sweep 2 streams psi[0] as psi_old AND applies
        // u,v -= grad(psi[0]) on its load side -> psi[1]; the finalize kernel applies both to the stored u, v
        int pi = 0;
        if (launch_clean_sweep_tma(dt, h->g, h->clean, u, v, nullptr, h->psi[0], 0, s)) {
            prof_mark(h, "clean_sweep_1", s, pi);
            if (launch_clean_sweep_tma(dt, h->g, h->clean, u, v, h->psi[0], h->psi[1], 2, s)) {
                prof_mark(h, "clean_sweep_2", s, pi);
                *first_psi = h->psi[0];
                if (last_psi) *last_psi = h->psi[1];
                if (out_u) *out_u = u;
                if (out_v) *out_v = v;
                return 2;
            }
        }
        cudaGetLastError();   // a failed launch falls through to the unfused path below, which redoes both sweeps
    }
    if (mode == B200CFD_CLEANDIV_REDBLACK) {
        // NOT reference parity: same update, red-black ordering, in place on psi[0] (inv_dt = -1 turns
        // the Poisson colour kernel's "+ div/dt" into "- div")
        cudaMemsetAsync(h->psi[0], 0, fb, s);
        for (int it = 0; it < iterations; ++it) {
            launch_divergence(dt, h->g, u, v, h->psi[1], nullptr, s);
            launch_rbgs_color(dt, h->g, h->psi[0], h->psi[1], nullptr, -1.0, 1.0, 0, nullptr, 0.0, 0, s);
            launch_rbgs_color(dt, h->g, h->psi[0], h->psi[1], nullptr, -1.0, 1.0, 1, nullptr, 0.0, 0, s);
            launch_clean_apply(dt, h->g, u, v, h->psi[0], s);
            nl += 4;
        }
        last = h->psi[0];
    } else {
        for (int it = 0; it < iterations; ++it) {
            void* pn = h->psi[it & 1];
            const void* po = it == 0 ? nullptr : h->psi[(it - 1) & 1];
            int pi = 0;
            const bool last_deferred = defer_last_apply && it == iterations - 1;
            if (out_u && out_v && h->fuse_apply && !h->clean.two_pass && !last_deferred) {
                // sweep + correction in one kernel, written to the other velocity pair
                void* uo = (u == h->ua) ? h->ub : h->ua;
                void* vo = (v == h->va) ? h->vb : h->va;
                launch_clean_sweep(dt, h->g, h->clean, u, v, po, pn, uo, vo, s);
                if (in_step) prof_mark(h, it == 0 ? "clean_sweep_apply_1" : "clean_sweep_apply_2", s, pi);
                u = uo; v = vo;
                ++nl; last = pn;
                continue;
            }
            launch_clean_sweep(dt, h->g, h->clean, u, v, po, pn, nullptr, nullptr, s);
            nl += clean_sweep_launches(h->clean) - 1;
            if (in_step) prof_mark(h, it == 0 ? "clean_sweep_1" : "clean_sweep_2", s, pi);
            if (!last_deferred) {
                launch_clean_apply(dt, h->g, u, v, pn, s); ++nl;
                if (in_step) prof_mark(h, it == 0 ? "clean_apply_1" : "clean_apply_2", s, pi);
            }
            ++nl;
            last = pn;
        }
    }
    if (last_psi) *last_psi = last;
    if (out_u) *out_u = u;
    if (out_v) *out_v = v;
    return nl;
}

// profiling: an event after each phase of a step (first profiled step defines the phase names)
static void prof_mark(b200cfd_handle* h, const char* name, cudaStream_t s, int& idx) {
    if (!h->profiling) return;
    if (h->prof_steps == 0 && name) h->prof_names.push_back(name);
    cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, s);
    h->prof_events.push_back(e);
    ++idx;
}

// does the step of this handle split into "up to the finalize of the edge columns" + "the rest" (x-slab overlap)?
static bool step_splits(const b200cfd_handle* h) {
    const b200cfd_config& c = h->cfg;
    return h->fast && !h->fin_tma && !h->fuse_apply && c.cleandiv_mode == B200CFD_CLEANDIV_LEXICOGRAPHIC && c.clean_iters > 0 && c.nranks > 1 &&
           c.poisson_mode == B200CFD_POISSON_AS_EXECUTED && c.omega == 1.0;
}

// part 0: the whole step.  part 1 / 2 (step_splits only): the step up to and including the finalize of the column blocks
// at the slab's edges / the finalize of the other column blocks + the closing reduction.
static int enqueue_step(b200cfd_handle* h, cudaStream_t s, bool close = true, int part = 0, bool after_ctl = false) {
    const b200cfd_config& c = h->cfg;
    const int dt = c.dtype;
    int nl = 0, pi = 0, close_rb1 = 0;
    const void* nut = nullptr;
    // columns whose finalize result the halo exchange needs (sent to a neighbour) or overwrites (this slab's halos)
    const int c_left = c.rank > 0 ? h->g.own0 + h->req_halo_r : 0;
    const int c_right = c.rank < c.nranks - 1 ? h->g.own1 - h->req_halo_l : h->g.nx;
    if (part == 2) {
        // the psi pointers of the fused-correction path are fixed: sweep 1 -> psi[0], sweep 2 -> psi[1]
        const bool two = h->fuse_corr && c.clean_iters == 2 && clean_sweep_tma_ok(dt, h->g, h->clean, h->ua, h->va, h->psi[0], h->psi[1]);
        void* last_psi = two ? h->psi[1] : h->psi[(c.clean_iters - 1) & 1];
        launch_finalize_fast(dt, h->g, h->phys, h->ctl, h->ua, h->va, last_psi, two ? h->psi[0] : nullptr, h->u, h->v, h->ibm, h->mask,
                             h->y, h->ke_part, h->cnt_part, h->flags, h->n_sm, s, 2, c_left, c_right);
        close_rb1 = (h->fuse_k23 ? 1 : 0);
        launch_ctl_reduce_local(h->ctl, h->ke_part, h->cnt_part, h->n_part_used, h->red, close_rb1, c.pressure_tolerance, h->nut_part, s);
        return 2;
    }
    prof_mark(h, nullptr, s, pi);
    if (c.use_les) {
        launch_smagorinsky(dt, h->g, h->u, h->v, h->nu_t, c.cs, s);
        launch_field_sum(dt, h->g, h->nu_t, h->nut_part, s);      // mean(nu_t) for the NEXT step's adaptive dt (v5.py:323)
        nut = h->nu_t; nl += 2;
    }
    if (h->fast) launch_predictor_fast(dt, h->g, h->phys, h->ctl, h->u, h->v, h->us, h->vs, h->tau, h->ibm, h->y, h->flags, h->n_sm, s, after_ctl);
    else launch_predictor(dt, h->g, h->phys, h->ctl, h->u, h->v, h->us, h->vs, h->tau, nut, h->ibm, h->y, s);
    ++nl;
    prof_mark(h, "predictor", s, pi);
    if (h->fast && h->fuse_k23 && c.poisson_mode == B200CFD_POISSON_AS_EXECUTED && c.omega == 1.0) {
        // K2 + K3 in one pass: u*, v* are read once and phi is never re-read
        launch_div_rb1_correct_fast(dt, h->g, h->us, h->vs, h->mask, h->phi, h->div, h->ua, h->va, h->phys.inv_cfg_dt, h->ctl,
                                    h->flags, h->n_sm, s);
        close_rb1 = 1;               // the Poisson bookkeeping of k_rb1_close is folded into the end-of-step control kernel
        nl += 1;
        prof_mark(h, "poisson_correct", s, pi);
    } else {
        nl += enqueue_step_poisson(h, s);
        prof_mark(h, "poisson", s, pi);
        launch_correct(dt, h->g, h->ctl, h->us, h->vs, h->phi, h->ua, h->va, s); ++nl;
        prof_mark(h, "correct", s, pi);
    }
    // fast path: the last clean-divergence correction u,v -= grad(psi) is applied on the fly by the finalize kernel
    const bool defer = h->fast && c.cleandiv_mode == B200CFD_CLEANDIV_LEXICOGRAPHIC && c.clean_iters > 0;
    void *last_psi = nullptr, *first_psi = nullptr;
    void *fu = h->ua, *fv = h->va;   // velocity entering the finalize kernel
    nl += enqueue_clean(h, h->ua, h->va, c.clean_iters, c.cleandiv_mode, s, &last_psi, true, defer, defer ? &fu : nullptr, defer ? &fv : nullptr,
                        defer ? &first_psi : nullptr);
    if (c.cleandiv_mode != B200CFD_CLEANDIV_LEXICOGRAPHIC) prof_mark(h, "clean_divergence", s, pi);
    if (part == 1) {
        // edge column blocks only; the rest of the finalize and the closing reduction are part 2
        launch_finalize_fast(dt, h->g, h->phys, h->ctl, fu, fv, defer ? last_psi : nullptr, defer ? first_psi : nullptr, h->u, h->v, h->ibm, h->mask,
                             h->y, h->ke_part, h->cnt_part, h->flags, h->n_sm, s, 1, c_left, c_right);
        return nl + 1;
    }
    if (h->fin_tma && defer && last_psi) {
        if (!launch_finalize_tma(dt, h->g, h->phys, h->ctl, fu, fv, last_psi, first_psi, h->u, h->v, h->ibm, h->mask, h->y, h->ke_part, h->cnt_part, h->flags, h->n_sm, s))
            h->err = "launch of the TMA finalize kernel failed";
    } else if (h->fast) launch_finalize_fast(dt, h->g, h->phys, h->ctl, fu, fv, defer ? last_psi : nullptr, defer ? first_psi : nullptr, h->u, h->v, h->ibm, h->mask, h->y, h->ke_part, h->cnt_part, h->flags, h->n_sm, s);
    else launch_finalize(dt, h->g, h->phys, h->ctl, h->ua, h->va, h->u, h->v, h->ibm, h->mask, h->y, h->ke_part, h->cnt_part, s);
    ++nl;
    prof_mark(h, (defer && first_psi) ? "finalize2" : "finalize", s, pi);
    if (close) launch_ctl_reduce_end(h->ctl, h->cp, h->ke_part, h->cnt_part, h->n_part_used, h->red, close_rb1, c.pressure_tolerance, h->nut_part, h->ring, h->ring_cap, s);
    else launch_ctl_reduce_local(h->ctl, h->ke_part, h->cnt_part, h->n_part_used, h->red, close_rb1, c.pressure_tolerance, h->nut_part, s);
    ++nl;
    prof_mark(h, "ctl", s, pi);
    if (h->profiling) ++h->prof_steps;
    return nl;
}

static int ensure_tmp(b200cfd_handle* h) {
    if (h->tmp) return B200CFD_OK;
    const size_t fb = (size_t)h->cfg.ny * h->cfg.nx * h->es;
    CU(h, cudaMalloc(&h->tmp, fb));
    CU(h, cudaMemset(h->tmp, 0, fb));
    return B200CFD_OK;
}

int b200cfd_step(b200cfd_handle* h, int n_steps, int use_graph, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    if (n_steps < 0) return fail(h, B200CFD_EINVAL, "n_steps < 0");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    if (!h->div && !(h->cfg.poisson_mode == B200CFD_POISSON_AS_EXECUTED && h->cfg.omega == 1.0)) {
        int rc = ensure_tmp(h); if (rc) return rc;
    }
    if (use_graph) {
        if (!h->step_graph) {
            cudaGraph_t graph = nullptr;
            CU(h, cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
            h->gpu_launches = enqueue_step(h, h->cap_stream);
            cudaError_t e = cudaStreamEndCapture(h->cap_stream, &graph);
            if (e != cudaSuccess) return fail(h, B200CFD_ECUDA, "graph capture: %s", cudaGetErrorString(e));
            e = cudaGraphInstantiate(&h->step_graph, graph, 0);
            cudaGraphDestroy(graph);
            if (e != cudaSuccess) { h->step_graph = nullptr; return fail(h, B200CFD_ECUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
        }
        int left = n_steps;
        // several steps per graph launch: the kernel-to-kernel (programmatic) edges then also span the step boundaries, which
        // between two graph launches are a full serialisation plus a launch latency
        const int K = h->graph_steps;
        // (built on the first graph call whatever its length: a caller that warms up with a few steps must not find the
        // capture and instantiation of the long graph inside its timed region)
        if (K > 1) {
            if (!h->multi_graph) {
                cudaGraph_t graph = nullptr;
                CU(h, cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
                for (int k = 0; k < K; ++k) enqueue_step(h, h->cap_stream, true, 0, k > 0);   // k > 0: right behind the control kernel of step k - 1
                cudaError_t e = cudaStreamEndCapture(h->cap_stream, &graph);
                if (e != cudaSuccess) return fail(h, B200CFD_ECUDA, "graph capture (%d steps): %s", K, cudaGetErrorString(e));
                e = cudaGraphInstantiate(&h->multi_graph, graph, 0);
                cudaGraphDestroy(graph);
                if (e != cudaSuccess) { h->multi_graph = nullptr; return fail(h, B200CFD_ECUDA, "graph instantiate (%d steps): %s", K, cudaGetErrorString(e)); }
                cudaGraphUpload(h->multi_graph, s);          // not on the first timed launch either
            }
            for (; left >= K; left -= K) CU(h, cudaGraphLaunch(h->multi_graph, s));
        }
        for (int k = 0; k < left; ++k) CU(h, cudaGraphLaunch(h->step_graph, s));
    } else {
        for (int k = 0; k < n_steps; ++k) h->gpu_launches = enqueue_step(h, s, true, 0, k > 0);
        CHECK_LAUNCH(h);
    }
    h->step += n_steps;
    return B200CFD_OK;
}

// ---- multi-GPU (x-slab) stepping: begin = the whole step except its closing kernel; the caller all_gathers the
// RED_N-double reduction vectors of all slabs; end = closing kernel on the gathered vectors.
int b200cfd_step_begin(b200cfd_handle* h, int use_graph, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    if (h->cfg.nranks > 1 && !(h->cfg.poisson_mode == B200CFD_POISSON_AS_EXECUTED && h->cfg.omega == 1.0))
        return fail(h, B200CFD_EINVAL, "slab-decomposed stepping supports poisson_mode as_executed only (iterated solves need a halo exchange per iteration: use b200cfd_poisson with pack/unpack_columns)");
    if (use_graph) {
        if (!h->begin_graph) {
            cudaGraph_t graph = nullptr;
            CU(h, cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
            h->gpu_launches = enqueue_step(h, h->cap_stream, false) + 1;
            cudaError_t e = cudaStreamEndCapture(h->cap_stream, &graph);
            if (e != cudaSuccess) return fail(h, B200CFD_ECUDA, "graph capture: %s", cudaGetErrorString(e));
            e = cudaGraphInstantiate(&h->begin_graph, graph, 0);
            cudaGraphDestroy(graph);
            if (e != cudaSuccess) { h->begin_graph = nullptr; return fail(h, B200CFD_ECUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
        }
        CU(h, cudaGraphLaunch(h->begin_graph, s));
    } else {
        h->gpu_launches = enqueue_step(h, s, false) + 1;
        CHECK_LAUNCH(h);
    }
    return B200CFD_OK;
}
// x-slab overlap: part 1 = the step up to the finalize of the column blocks at the slab's edges (the halo columns and the
// owned columns the neighbours need); the caller can then start the halo exchange on another stream while part 2 (the
// finalize of all the other column blocks + this slab's reduction vector) runs.  Returns B200CFD_ESTATE when the step
// of this handle does not split (general kernels): use b200cfd_step_begin then.
int b200cfd_step_part(b200cfd_handle* h, int part, int use_graph, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    if (part != 1 && part != 2) return fail(h, B200CFD_EINVAL, "part must be 1 or 2");
    if (!step_splits(h)) return fail(h, B200CFD_ESTATE, "the step of this handle does not split (b200cfd_step_splits)");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    cudaGraphExec_t& ge = h->part_graph[part - 1];
    if (use_graph) {
        if (!ge) {
            cudaGraph_t graph = nullptr;
            CU(h, cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
            const int n = enqueue_step(h, h->cap_stream, false, part);
            if (part == 1) h->gpu_launches = n + 3;          // + finalize of the other blocks, the local reduction, the closing kernel
            cudaError_t e = cudaStreamEndCapture(h->cap_stream, &graph);
            if (e != cudaSuccess) return fail(h, B200CFD_ECUDA, "graph capture: %s", cudaGetErrorString(e));
            e = cudaGraphInstantiate(&ge, graph, 0);
            cudaGraphDestroy(graph);
            if (e != cudaSuccess) { ge = nullptr; return fail(h, B200CFD_ECUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
        }
        CU(h, cudaGraphLaunch(ge, s));
    } else {
        const int n = enqueue_step(h, s, false, part);
        if (part == 1) h->gpu_launches = n + 3;
        CHECK_LAUNCH(h);
    }
    return B200CFD_OK;
}
int b200cfd_step_splits(b200cfd_handle* h) { return (h && h->bound && step_splits(h)) ? 1 : 0; }

// All fields (nfields <= 4) and both sides of a halo refresh in one launch.  pack != 0: columns [j_left, j_left+n_left) of
// every field -> buf_left[nfields][ny][n_left], [j_right, j_right+n_right) -> buf_right[nfields][ny][n_right]; else the inverse.
int b200cfd_halo_rows(b200cfd_handle* h, int pack, void* const* fields, int nfields, int j_left, int n_left, void* buf_left,
                      int j_right, int n_right, void* buf_right, void* stream) {
    if (!h || !fields || nfields < 1 || nfields > 4) return fail(h, B200CFD_EINVAL, "1..4 fields");
    if (n_left < 0 || n_right < 0 || (n_left > 0 && (!buf_left || j_left < 0 || j_left + n_left > h->g.nx)) ||
        (n_right > 0 && (!buf_right || j_right < 0 || j_right + n_right > h->g.nx)))
        return fail(h, B200CFD_EINVAL, "columns outside the slab");
    CU(h, cudaSetDevice(h->device));
    launch_halo_rows(h->es, pack != 0, h->g, fields, nfields, j_left, n_left, buf_left, j_right, n_right, buf_right, (cudaStream_t)stream);
    CHECK_LAUNCH(h);
    return B200CFD_OK;
}

int b200cfd_step_end(b200cfd_handle* h, const double* gathered, int nranks, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    if (nranks < 1) return fail(h, B200CFD_EINVAL, "nranks < 1");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    launch_ctl_end_step(h->ctl, h->cp, gathered ? gathered : h->red, gathered ? nranks : 1, h->ring, h->ring_cap, s);
    CHECK_LAUNCH(h);
    h->step += 1;
    return B200CFD_OK;
}
void* b200cfd_reduction_ptr(b200cfd_handle* h) { return h ? (void*)h->red : nullptr; }
int b200cfd_reduction_len(void) { return RED_N; }
void* b200cfd_poisson_change_ptr(b200cfd_handle* h) { return h ? (void*)&h->ctl->pois_last : nullptr; }
// set_step for slabs: the caller supplies the GLOBAL post-clip maxima adaptive_time_step needs (v5.py:322)
int b200cfd_set_step_global(b200cfd_handle* h, int64_t step, double max_abs_u, double max_abs_v, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    unsigned long long bits[2];
    memcpy(&bits[0], &max_abs_u, 8); memcpy(&bits[1], &max_abs_v, 8);
    CU(h, cudaMemcpyAsync(&h->ctl->max_u, &bits[0], 8, cudaMemcpyHostToDevice, s));
    CU(h, cudaMemcpyAsync(&h->ctl->max_v, &bits[1], 8, cudaMemcpyHostToDevice, s));
    CU(h, cudaStreamSynchronize(s));
    // (LES on slabs: mean(nu_t) restarts from 0 here and is exact again from the first step taken, which all_gathers it)
    launch_ctl_init(h->ctl, h->cp, (long long)step, 1, nullptr, s);
    CHECK_LAUNCH(h);
    h->step = step;
    return B200CFD_OK;
}
int b200cfd_pack_columns(b200cfd_handle* h, const void* field, int j0, int ncols, void* buf, void* stream) {
    if (!h || !field || !buf) return fail(h, B200CFD_EINVAL, "null argument");
    if (j0 < 0 || ncols < 0 || j0 + ncols > h->g.nx) return fail(h, B200CFD_EINVAL, "columns [%d,%d) outside the slab", j0, j0 + ncols);
    CU(h, cudaSetDevice(h->device));
    launch_pack_columns(h->es, h->g, field, j0, ncols, buf, (cudaStream_t)stream);
    CHECK_LAUNCH(h);
    return B200CFD_OK;
}
int b200cfd_unpack_columns(b200cfd_handle* h, void* field, int j0, int ncols, const void* buf, void* stream) {
    if (!h || !field || !buf) return fail(h, B200CFD_EINVAL, "null argument");
    if (j0 < 0 || ncols < 0 || j0 + ncols > h->g.nx) return fail(h, B200CFD_EINVAL, "columns [%d,%d) outside the slab", j0, j0 + ncols);
    CU(h, cudaSetDevice(h->device));
    launch_unpack_columns(h->es, h->g, field, j0, ncols, buf, (cudaStream_t)stream);
    CHECK_LAUNCH(h);
    return B200CFD_OK;
}
// Halo widths (columns) a slab needs so that ONE exchange of u, v per time step suffices: everything
// else of the step is recomputed redundantly in the halo.  Left = 2 x the horizontal apron of the lexicographic
// sweep + the stencil radii of the other phases; right = the stencil radii.  No device needed.
int b200cfd_required_halo(const b200cfd_config* cfg, int* left, int* right) {
    if (!cfg || !left || !right || !(cfg->dx > 0) || !(cfg->dy > 0)) return B200CFD_EINVAL;
    Geom g = make_geom(cfg->ny, cfg->nx, cfg->nx, cfg->dx, cfg->dy, 0, cfg->nx);
    CleanPlan cp = clean_plan(g, cfg->dtype, 148);
    const int iters = cfg->clean_iters > 0 ? cfg->clean_iters : 0;
    *left = iters * (cp.apron_h + 3) + 8;
    *right = 3 * iters + 8;
    // round both up to 16 columns (128 B in fp64): with a slab width that is a multiple of 16 every row of the local
    // array then starts on a 128-byte boundary (an unaligned pitch costs one extra 32-byte sector per warp access)
    *left = (*left + 15) / 16 * 16;
    *right = (*right + 15) / 16 * 16;
    return B200CFD_OK;
}

int b200cfd_health(b200cfd_handle* h, double* out4, void* stream) {
    if (!h || !h->bound || !out4) return fail(h, B200CFD_ESTATE, "bind_fields first");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    unsigned long long* d = h->cnt_part;   // free between steps (holds >= 4 words)
    launch_health(h->cfg.dtype, h->g, h->u, h->v, d, s);
    CHECK_LAUNCH(h);
    CU(h, cudaMemcpyAsync(out4, d, 4 * sizeof(double), cudaMemcpyDeviceToHost, s));   // bit patterns of non-negative doubles
    CU(h, cudaStreamSynchronize(s));
    return B200CFD_OK;
}

/* Per-phase profiling of b200cfd_step (use_graph = 0): CUDA events between the kernels of each step. */
int b200cfd_profile_begin(b200cfd_handle* h) {
    if (!h) return B200CFD_EINVAL;
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    h->prof_events.clear(); h->prof_names.clear(); h->prof_ms.clear(); h->prof_steps = 0;
    h->profiling = true;
    return B200CFD_OK;
}
/* Synchronises; writes up to cap phase names (';'-joined into names, NUL-terminated) and the mean ms per step
 * of each phase; returns the number of phases (or a negative error). */
int b200cfd_profile_end(b200cfd_handle* h, char* names, int names_cap, double* ms, int cap) {
    if (!h || !h->profiling) return fail(h, B200CFD_ESTATE, "profile_begin first");
    h->profiling = false;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaDeviceSynchronize());
    const int np = (int)h->prof_names.size();
    std::vector<double> acc(np, 0.0);
    const int per = np + 1;
    for (int st = 0; st < h->prof_steps; ++st)
        for (int k = 0; k < np; ++k) {
            float t = 0.f;
            cudaEventElapsedTime(&t, h->prof_events[(size_t)st * per + k], h->prof_events[(size_t)st * per + k + 1]);
            acc[k] += t;
        }
    std::string joined;
    for (int k = 0; k < np; ++k) { if (k) joined += ';'; joined += h->prof_names[k]; if (k < cap && ms) ms[k] = h->prof_steps ? acc[k] / h->prof_steps : 0.0; }
    if (names && names_cap > 0) { strncpy(names, joined.c_str(), names_cap - 1); names[names_cap - 1] = 0; }
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    h->prof_events.clear();
    return np;
}

int b200cfd_launches_per_step(b200cfd_handle* h) { return h ? h->gpu_launches : 0; }

int b200cfd_fetch_scalars(b200cfd_handle* h, int64_t first_step, int n, b200cfd_step_scalars* out, void* stream) {
    if (!h || !out || n < 0) return fail(h, B200CFD_EINVAL, "bad argument");
    if (n > h->ring_cap) return fail(h, B200CFD_EINVAL, "at most %d records are kept", h->ring_cap);
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    for (int k = 0; k < n;) {
        const int slot = (int)((first_step + k) % h->ring_cap);
        int run = h->ring_cap - slot; if (run > n - k) run = n - k;
        CU(h, cudaMemcpyAsync(out + k, h->ring + slot, sizeof(StepScalars) * run, cudaMemcpyDeviceToHost, s));
        k += run;
    }
    CU(h, cudaStreamSynchronize(s));
    for (int k = 0; k < n; ++k)
        if ((int64_t)out[k].step != first_step + k) return fail(h, B200CFD_ESTATE, "step %lld is not in the ring (found %lld)", (long long)(first_step + k), (long long)out[k].step);
    return B200CFD_OK;
}

int b200cfd_step_host(b200cfd_handle* h, void* u_host, void* v_host, int n_steps, int use_graph,
                      b200cfd_step_scalars* last, void* stream) {
    if (!h || !h->bound) return fail(h, B200CFD_ESTATE, "bind_fields first");
    if (!u_host || !v_host) return fail(h, B200CFD_EINVAL, "null host array");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t fb = (size_t)h->cfg.ny * h->cfg.nx * h->es;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaMemcpyAsync(h->u, u_host, fb, cudaMemcpyHostToDevice, s));
    CU(h, cudaMemcpyAsync(h->v, v_host, fb, cudaMemcpyHostToDevice, s));
    int rc = b200cfd_set_step(h, h->step, stream);   // adaptive_time_step reads the arrays it is given (v5.py:322)
    if (rc) return rc;
    rc = b200cfd_step(h, n_steps, use_graph, stream);
    if (rc) return rc;
    CU(h, cudaMemcpyAsync(u_host, h->u, fb, cudaMemcpyDeviceToHost, s));
    CU(h, cudaMemcpyAsync(v_host, h->v, fb, cudaMemcpyDeviceToHost, s));
    if (last && n_steps > 0) return b200cfd_fetch_scalars(h, h->step - 1, 1, last, stream);
    CU(h, cudaStreamSynchronize(s));
    return B200CFD_OK;
}

int b200cfd_poisson(b200cfd_handle* h, void* phi, const void* rhs_div, const uint8_t* mask,
                    double dx, double dy, double dt, int mode, int max_iter, double tol, double omega,
                    int temporal_depth, int* iters_done, double* last_change, void* stream) {
    if (!h || !phi || !rhs_div) return fail(h, B200CFD_EINVAL, "null argument");
    if (!(dx > 0) || !(dy > 0) || !(dt != 0) || max_iter < 0) return fail(h, B200CFD_EINVAL, "bad dx/dy/dt/max_iter");
    cudaStream_t s = (cudaStream_t)stream;
    CU(h, cudaSetDevice(h->device));
    const int dtp = h->cfg.dtype;
    // slabs: max|dphi| is taken over the OWNED columns only (halo columns lose validity between refreshes)
    Geom g = make_geom(h->g.ny, h->g.nx, h->g.pitch, dx, dy, h->g.jg0, h->g.nxg, h->g.own0, h->g.own1);
    const size_t fb = (size_t)g.ny * g.nx * h->es;
    launch_poisson_reset(h->ctl, s);
    h->last_poisson_tb = false;
    int done_iters = 0, tb_done_iters = -1;
    double tb_last_change = 0.0;
    if (mode == B200CFD_POISSON_JACOBI) {
        const double k = dx * dx / dt;
        int it = 0;
        if (temporal_depth != 1 && max_iter >= 2) {
            int rc = poisson_tb_jacobi(&h->tb, dtp, g, h->n_sm, phi, rhs_div, mask, k, max_iter, temporal_depth, &it, s);
            if (rc < 0) return fail(h, B200CFD_ECUDA, "temporally blocked Jacobi: %s", poisson_tb_error());
            h->last_poisson_tb = it > 0;
        }
        if (it < max_iter) {
            int rc = ensure_tmp(h); if (rc) return rc;
            void* a = phi; void* b = h->tmp;
            for (; it < max_iter; ++it) { launch_jacobi(dtp, g, a, b, rhs_div, mask, k, s); void* t = a; a = b; b = t; }
            if (a != phi) CU(h, cudaMemcpyAsync(phi, a, fb, cudaMemcpyDeviceToDevice, s));
        }
        done_iters = max_iter;
    } else if (mode == B200CFD_POISSON_SOR_V2) {
        // v2.py:284-306: SOR on lap(phi) = +div/dt (sign flipped w.r.t. v5: inv_dt < 0), wall conditions after every
        // sweep, stop when max|dphi| < tol.  Red-black ordering (north_star): same fixed point as the reference's
        // lexicographic sweep, reached along a different path.
        if (h->cfg.nranks > 1) return fail(h, B200CFD_EINVAL, "SOR_V2 is single-slab only");
        const double inv_dt = -1.0 / dt;
        const bool check = (tol > 0.0);
        const int chunk = 32;
        int it = 0;
        while (it < max_iter) {
            const int m = (max_iter - it < chunk) ? max_iter - it : chunk;
            for (int q = 0; q < m; ++q) {
                launch_rbgs_color(dtp, g, phi, rhs_div, mask, inv_dt, omega, 0, h->ctl, tol, 0, s);
                launch_rbgs_color(dtp, g, phi, rhs_div, mask, inv_dt, omega, 1, h->ctl, tol, 1, s);
                launch_sor_v2_bc(dtp, g, phi, h->ctl, s);
            }
            it += m;
            if (check) {
                int done = 0;
                CU(h, cudaMemcpyAsync(&done, &h->ctl->pois_done, sizeof(int), cudaMemcpyDeviceToHost, s));
                CU(h, cudaStreamSynchronize(s));
                if (done) break;
            }
        }
        done_iters = it;
    } else {
        const int iters = mode == B200CFD_POISSON_AS_EXECUTED ? (max_iter < 1 ? max_iter : 1) : max_iter;
        const double inv_dt = 1.0 / dt;
        const bool check = (tol > 0.0);
        int it = 0;
        double tb_last = 0.0;
        bool tb_ran = false;
        if (temporal_depth != 1 && iters >= 2) {
            // TMA temporally blocked passes; in tolerance mode they stop at exactly the iteration the sequential
            // algorithm stops at (per-iteration max|dphi| slots), synchronising once per pass
            int rc = poisson_tb_rbsor(&h->tb, dtp, g, h->n_sm, phi, rhs_div, mask, inv_dt, omega, iters, temporal_depth, &it, s,
                                      check ? tol : 0.0, &tb_last);
            if (rc < 0) return fail(h, B200CFD_ECUDA, "temporally blocked red-black: %s", poisson_tb_error());
            tb_ran = it > 0;
            h->last_poisson_tb = tb_ran;
        }
        const int chunk = 32;
        while (!tb_ran && it < iters) {
            const int m = (iters - it < chunk) ? iters - it : chunk;
            for (int q = 0; q < m; ++q) {
                launch_rbgs_color(dtp, g, phi, rhs_div, mask, inv_dt, omega, 0, h->ctl, tol, 0, s);
                launch_rbgs_color(dtp, g, phi, rhs_div, mask, inv_dt, omega, 1, h->ctl, tol, 1, s);
            }
            it += m;
            if (check) {   // tolerance mode: look at the device flag once per chunk (the kernels stop themselves exactly)
                int done = 0;
                CU(h, cudaMemcpyAsync(&done, &h->ctl->pois_done, sizeof(int), cudaMemcpyDeviceToHost, s));
                CU(h, cudaStreamSynchronize(s));
                if (done) break;
            }
        }
        done_iters = it;
        if (tb_ran) { tb_done_iters = it; tb_last_change = tb_last; }
    }
    CHECK_LAUNCH(h);
    if (mode != B200CFD_POISSON_JACOBI && tb_done_iters >= 0) {
        if (iters_done) *iters_done = tb_done_iters;
        if (last_change) *last_change = tb_last_change;
        if (iters_done || last_change) CU(h, cudaStreamSynchronize(s));
        return B200CFD_OK;
    }
    if (iters_done || last_change) {
        StepCtl c;
        CU(h, cudaMemcpyAsync(&c, h->ctl, sizeof(StepCtl), cudaMemcpyDeviceToHost, s));
        CU(h, cudaStreamSynchronize(s));
        const bool counted = (mode != B200CFD_POISSON_JACOBI) && c.pois_iters > 0;
        if (iters_done) *iters_done = counted ? c.pois_iters : done_iters;
        if (last_change) *last_change = c.pois_last;
    }
    return B200CFD_OK;
}

// keep != 0: the following b200cfd_poisson calls solve on the SAME, unchanged rhs_div and mask arrays as the previous call
// (one solve done in chunks around halo refreshes): the temporally blocked kernel's classified copy of the right-hand side is
// not rebuilt per call.  Reset with keep = 0 before the right-hand side changes.
int b200cfd_poisson_keep_rhs(b200cfd_handle* h, int keep) {
    if (!h) return B200CFD_EINVAL;
    poisson_tb_keep_rhs(&h->tb, keep);
    return B200CFD_OK;
}

// What the last b200cfd_poisson call did: out4 = {temporal depth used (1 = one-sweep kernels), HBM passes of the
// temporally blocked kernel (0 when it did not run), bytes per cell per pass of that kernel, reserved}.
int b200cfd_poisson_info(b200cfd_handle* h, int* out4) {
    if (!h || !out4) return B200CFD_EINVAL;
    int depth = 0, passes = 0;
    poisson_tb_info(h->tb, &depth, &passes);
    const bool tb = h->last_poisson_tb;
    out4[0] = tb ? depth : 1; out4[1] = tb ? passes : 0; out4[2] = 3 * h->es; out4[3] = 0;
    return B200CFD_OK;
}

// ---- unit exports ---------------------------------------------------------------------------------
#define UNIT_PRE() if (!h) return B200CFD_EINVAL; cudaStream_t s = (cudaStream_t)stream; CU(h, cudaSetDevice(h->device));
#define UNIT_POST() CHECK_LAUNCH(h); return B200CFD_OK;

int b200cfd_k_supg_tau(b200cfd_handle* h, const void* u, const void* v, void* tau, double dt, double nu_eff, void* stream) {
    UNIT_PRE(); launch_supg_tau(h->cfg.dtype, h->g, u, v, tau, dt, nu_eff, nullptr, s); UNIT_POST();
}
int b200cfd_k_convection(b200cfd_handle* h, const void* u, const void* v, const void* phi, const void* tau,
                         void* conv, int supg, void* stream) {
    UNIT_PRE(); launch_convection(h->cfg.dtype, h->g, u, v, phi, tau, conv, supg, s); UNIT_POST();
}
int b200cfd_k_laplacian(b200cfd_handle* h, const void* phi, void* lap, double nu_eff, void* stream) {
    UNIT_PRE(); launch_laplacian(h->cfg.dtype, h->g, phi, lap, nu_eff, nullptr, s); UNIT_POST();
}
int b200cfd_k_divergence(b200cfd_handle* h, const void* u, const void* v, void* div, void* stream) {
    UNIT_PRE(); launch_divergence(h->cfg.dtype, h->g, u, v, div, nullptr, s); UNIT_POST();
}
int b200cfd_k_gradient(b200cfd_handle* h, const void* phi, void* gx, void* gy, void* stream) {
    UNIT_PRE(); launch_gradient(h->cfg.dtype, h->g, phi, gx, gy, s); UNIT_POST();
}
int b200cfd_k_smagorinsky(b200cfd_handle* h, const void* u, const void* v, void* nu_t, double cs, void* stream) {
    UNIT_PRE(); launch_smagorinsky(h->cfg.dtype, h->g, u, v, nu_t, cs, s); UNIT_POST();
}
int b200cfd_k_vorticity(b200cfd_handle* h, const void* u, const void* v, const uint8_t* mask, void* w, void* stream) {
    UNIT_PRE(); launch_vorticity(h->cfg.dtype, h->g, u, v, mask, w, s); UNIT_POST();
}
int b200cfd_k_apply_bc(b200cfd_handle* h, void* u, void* v, const double* y, int64_t step, void* stream) {
    UNIT_PRE(); launch_apply_bc(h->cfg.dtype, h->g, u, v, y, h->cfg.V_inf, h->cfg.y_max, (long long)step, s, h->cp.pert_amplitude); UNIT_POST();
}
int b200cfd_k_apply_ibm(b200cfd_handle* h, void* u, void* v, const double* ibm, double force_strength, void* stream) {
    UNIT_PRE(); launch_apply_ibm(h->cfg.dtype, h->g, u, v, ibm, force_strength, s); UNIT_POST();
}
int b200cfd_k_clean_divergence(b200cfd_handle* h, void* u, void* v, int iterations, int mode, void* psi_out, void* stream) {
    UNIT_PRE();
    void* last = nullptr;
    enqueue_clean(h, u, v, iterations, mode, s, &last);
    if (psi_out && last) CU(h, cudaMemcpyAsync(psi_out, last, (size_t)h->cfg.ny * h->cfg.nx * h->es, cudaMemcpyDeviceToDevice, s));
    UNIT_POST();
}
int b200cfd_k_predictor(b200cfd_handle* h, const void* u, const void* v, void* u_star, void* v_star,
                        void* tau_or_null, const double* ibm, const double* y, int64_t step, double dt, void* stream) {
    UNIT_PRE();
    // a private control block carrying the explicit step / dt
    StepCtl c; memset(&c, 0, sizeof c);
    c.dt = dt; c.step = step;
    c.fs = (double)step / (double)h->cp.initial_steps; if (c.fs > 1.0) c.fs = 1.0;
    double sc = (double)step / 1000.0; if (sc > 1.0) sc = 1.0;
    c.pert_scale = sc * h->cp.pert_amplitude; c.phase = 0.02 * (double)step;
    StepCtl* d = nullptr;
    CU(h, cudaMalloc(&d, sizeof(StepCtl)));
    CU(h, cudaMemcpyAsync(d, &c, sizeof c, cudaMemcpyHostToDevice, s));
    launch_predictor(h->cfg.dtype, h->g, h->phys, d, u, v, u_star, v_star, tau_or_null, nullptr, ibm, y, s);
    CU(h, cudaStreamSynchronize(s));
    cudaFree(d);
    UNIT_POST();
}

void b200cfd_get_clean_plan(b200cfd_handle* h, int* out8) {
    if (!h || !out8) return;
    out8[0] = h->clean.strip_w; out8[1] = h->clean.tile_h; out8[2] = h->clean.apron_h; out8[3] = h->clean.apron_v;
    out8[4] = h->clean.n_strips; out8[5] = h->clean.n_tiles; out8[6] = h->n_sm; out8[7] = h->n_part;
}

}  // extern "C"
