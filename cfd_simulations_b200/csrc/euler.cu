// euler.cu — fused compressible-Euler finite-volume step for sm_100a + its C ABI (include/b200cfd_euler.h).
// Replaces ShockwaveSolver.time_step (python/shockwave_analysis/v1_shock.py:291-317): both flux loops (:301-306),
// update_state (:211-223), apply_boundary_conditions (:277-289), the wedge zeroing (:312-313) and the NEXT step's
// compute_dt reduction (:263-275) in ONE pass over the state: 16 B read + 16 B written + 1 B of mask per cell.
//
// Design (B200-first, not a translation of the reference's three array sweeps):
//   * state = one float4 per cell, (nx, ny) cells, y fastest -> a warp reads 32 consecutive cells of a row with one
//     128-bit load per lane (512 B, coalesced);
//   * a warp owns a window of 32 columns (30 written: one ghost lane each side supplies the y-neighbour) and MARCHES
//     along x over a chunk of rows: the x-face flux of the previous row is carried in registers, the y-face flux of the
//     neighbour column comes from a warp shuffle, so every face flux is evaluated once (+1/15 ghost redundancy) and
//     every cell's primitives once — no F/G arrays, no shared memory, no block barrier;
//   * the next row's load is issued before the current row's arithmetic (software pipelining);
//   * dt lives in a device-resident control block; per-warp maxima go through atomicMax on the bit pattern of
//     non-negative doubles, so n steps are n graph replays with no host round trip.
// The arithmetic is float64 (as the reference's), so the kernel is bound by the FP64 pipe, not by HBM.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>

#include "../../include/b200cfd.h"
#include "../../include/b200cfd_euler.h"
#include "euler.cuh"

namespace b200cfd { namespace eu {

#ifndef EU_UNROLL2
#define EU_UNROLL2 0                // 1: march unrolled by two rows (Cell variables swap roles instead of being copied): 138 registers instead of 124, measured slower (41.1 vs 43.4 Gcell/s)
#endif
#ifndef EU_PF
#define EU_PF 6                     // L2 prefetch distance of the state rows (0 = off)
#endif
constexpr int WARPS = 4;            // warps per block (independent: no barrier in the step kernel)
constexpr int OUT_LANES = 30;       // written columns per warp window
constexpr int RING = 1 << 16;

struct Ctl {
    double dt;                              // dt of the step about to run (compute_dt of the current state)
    unsigned long long msx, msy, nanflag;   // accumulators: bits of non-negative doubles (= the reduction vector)
    unsigned long long pad;
    long long step;                         // self.step: steps actually taken
    long long calls;                        // time_step() calls recorded in the ring
    long long skipped;
    int skip;                               // dt is not finite: time_step() returns 0.0 and leaves U alone (:294-296)
    int pad2;
};

struct StepParams {
    int nx, ny;
    int r_first, r_last;     // local rows that update_state writes (interior rows of the global domain owned here)
    int inflow_row;          // local index of global row 0 (free stream, :279-282) or -1
    int outflow_row;         // local index of global row nx-1 (copy of the row before it, :283) or -1
    int n_reflect;
    int ch, nwin, nchunk;
    double dx, dy;
    float4 fs;
    Phys phys;
};

struct Acc { double sx, sy; unsigned nan; };

template <bool EXACT>
__device__ __forceinline__ void put(float4* __restrict__ out, const unsigned wedge_byte, size_t idx, float4 q,
                                    const Phys& P, Acc& acc) {
    if (wedge_byte) { q.y = 0.f; q.z = 0.f; }                       // :312-313
    out[idx] = q;
    double sx, sy;
    cell_speeds<EXACT>(q, P, sx, sy);
    acc.nan |= (q.x != q.x) | (q.y != q.y) | (q.z != q.z) | (q.w != q.w) | (sx != sx) | (sy != sy);
    acc.sx = vmax<EXACT>(sx, acc.sx);        // a NaN speed never wins (it is flagged above)
    acc.sy = vmax<EXACT>(sy, acc.sy);
}

// One cell of row i after update_state, through the y-boundary part of apply_boundary_conditions (:284-289):
// flip the reflective prefix, then (if any column is not reflective) copy into the two border columns.
// wb: the wedge-mask byte of cell (i, c), loaded by the caller ahead of the arithmetic (a load issued here, right before
// its use, left a full global-load latency exposed on every row: 21 % of the kernel's stall samples).
template <bool EXACT>
__device__ __forceinline__ void emit(float4* __restrict__ out, const uint8_t* __restrict__ wedge, const StepParams& S,
                                     int i, int c, float4 q, Acc& acc, const unsigned wb) {
    if (c < S.n_reflect) q.z = -q.z;
    const size_t row = (size_t)i * S.ny;
    put<EXACT>(out, wb, row + c, q, S.phys, acc);
    if (S.n_reflect < S.ny) {
        if (c == 1) put<EXACT>(out, wedge[row], row, q, S.phys, acc);
        if (c == S.ny - 2) put<EXACT>(out, wedge[row + S.ny - 1], row + S.ny - 1, q, S.phys, acc);
    }
}

__device__ __forceinline__ void passthrough(const float4* __restrict__ in, float4* __restrict__ out, const StepParams& S, int i, int c) {
    const size_t row = (size_t)i * S.ny;
    out[row + c] = in[row + c];
    if (c == 1) out[row] = in[row];
    if (c == S.ny - 2) out[row + S.ny - 1] = in[row + S.ny - 1];
}

template <int SOLVER, bool EXACT>
__global__ void __launch_bounds__(32 * WARPS)
k_euler_step(const StepParams S, const float4* __restrict__ in, float4* __restrict__ out, const uint8_t* __restrict__ wedge,
             Ctl* __restrict__ ctl) {
    const int lane = threadIdx.x & 31;
    const int wid = blockIdx.x * WARPS + (threadIdx.x >> 5);
    const int win = wid % S.nwin, chunk = wid / S.nwin;
    if (chunk >= S.nchunk) return;
    const int c = win * OUT_LANES + lane;
    const int cc = min(c, S.ny - 1);
    const bool writes = lane >= 1 && lane <= OUT_LANES && c <= S.ny - 2;
    const int r0 = S.r_first + chunk * S.ch;
    const int r1 = min(r0 + S.ch - 1, S.r_last);
    const Phys P = S.phys;

    if (ctl->skip) {                       // non-finite dt: the state passes through unchanged
        if (writes) {
            if (chunk == 0 && S.inflow_row >= 0) passthrough(in, out, S, S.inflow_row, c);
            for (int i = r0; i <= r1; ++i) passthrough(in, out, S, i, c);
            if (r1 == S.r_last && S.outflow_row >= 0) passthrough(in, out, S, S.outflow_row, c);
        }
        return;
    }
    const double dt = ctl->dt;
    const double ax = dt / S.dx, ay = dt / S.dy;
    Acc acc = {0.0, 0.0, 0u};

    if (writes && chunk == 0 && S.inflow_row >= 0) emit<EXACT>(out, wedge, S, S.inflow_row, c, S.fs, acc, wedge[(size_t)S.inflow_row * S.ny + c]);   // :279-282

    const float4* col = in + cc;
    float4 qN = __ldg(col + (size_t)(r0 + 1) * S.ny);
    Cell C = make_cell<EXACT>(__ldg(col + (size_t)r0 * S.ny), P);
    float4 Fw;
    {
        const Cell W = make_cell<EXACT>(__ldg(col + (size_t)(r0 - 1) * S.ny), P);
        Fw = face_flux<SOLVER, true, EXACT>(W, C, P);
    }
    const uint8_t* wcol = wedge + cc;
    unsigned wbN = __ldg(wcol + (size_t)r0 * S.ny);
    // one row: Cc = cell (i, c), Nn receives cell (i+1, c).  Called with the two Cell variables swapped on alternate rows, so
    // that the 24 registers of a Cell are renamed instead of copied.
    auto row = [&](const int i, const Cell& Cc, Cell& Nn) {
        const float4 qNN = __ldg(col + (size_t)min(i + 2, S.nx - 1) * S.ny);   // next iteration's row, in flight during the math
        if (EU_PF > 0 && i + EU_PF < S.nx) asm volatile("prefetch.global.L2 [%0];" ::"l"(col + (size_t)(i + EU_PF) * S.ny));
        const unsigned wb = wbN;                                               // wedge bytes: this row's arrived, next row's requested
        wbN = __ldg(wcol + (size_t)min(i + 1, S.nx - 1) * S.ny);
        unsigned wbO = 0;
        if (i + 1 == S.outflow_row) wbO = __ldg(wcol + (size_t)S.outflow_row * S.ny);
        Nn = make_cell<EXACT>(qN, P);
        const float4 Fe = face_flux<SOLVER, true, EXACT>(Cc, Nn, P);            // x-face (i | i+1)
        const Cell Y = shfl_down_cell<SOLVER>(Cc);                              // cell (i, c+1)
        const float4 Gn = face_flux<SOLVER, false, EXACT>(Cc, Y, P);            // y-face (c | c+1)
        const float4 Gs = shfl_up4(Gn);                                         // y-face (c-1 | c)
        if (writes) {
            const float4 q = update_cell<EXACT>(Cc.U, Fw, Fe, Gs, Gn, ax, ay, P);
            emit<EXACT>(out, wedge, S, i, c, q, acc, wb);
            if (i + 1 == S.outflow_row) emit<EXACT>(out, wedge, S, S.outflow_row, c, q, acc, wbO);       // :283
        }
        Fw = Fe;
        qN = qNN;
    };
    Cell D;
    int i = r0;
#if EU_UNROLL2
    for (; i + 1 <= r1; i += 2) { row(i, C, D); row(i + 1, D, C); }
    if (i <= r1) row(i, C, D);
#else
    for (; i <= r1; ++i) { row(i, C, D); C = D; }
#endif
    // block-free reduction: warp shuffle, one atomic per warp and quantity
    double sx = acc.sx, sy = acc.sy;
    unsigned nan = acc.nan;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx = fmax(sx, __shfl_xor_sync(0xffffffffu, sx, o));
        sy = fmax(sy, __shfl_xor_sync(0xffffffffu, sy, o));
        nan |= __shfl_xor_sync(0xffffffffu, nan, o);
    }
    if (lane == 0) {
        atomicMax(&ctl->msx, (unsigned long long)__double_as_longlong(sx));
        atomicMax(&ctl->msy, (unsigned long long)__double_as_longlong(sy));
        if (nan) atomicMax(&ctl->nanflag, (unsigned long long)__double_as_longlong(1.0));
    }
}

// compute_dt's reduction over an arbitrary state (start-up, unit parity): grid-stride, one atomic per warp.
__global__ void __launch_bounds__(256)
k_euler_speeds(const float4* __restrict__ U, long long n, Phys P, unsigned long long* __restrict__ red) {
    double sx = 0.0, sy = 0.0;
    unsigned nan = 0;
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
        const float4 q = __ldg(U + k);
        double a, b;
        cell_speeds(q, P, a, b);
        nan |= (q.x != q.x) | (q.y != q.y) | (q.z != q.z) | (q.w != q.w) | (a != a) | (b != b);
        sx = fmax(sx, a);
        sy = fmax(sy, b);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx = fmax(sx, __shfl_xor_sync(0xffffffffu, sx, o));
        sy = fmax(sy, __shfl_xor_sync(0xffffffffu, sy, o));
        nan |= __shfl_xor_sync(0xffffffffu, nan, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(red + 0, (unsigned long long)__double_as_longlong(sx));
        atomicMax(red + 1, (unsigned long long)__double_as_longlong(sy));
        if (nan) atomicMax(red + 2, (unsigned long long)__double_as_longlong(1.0));
    }
}

struct DtParams { double cfl, dx, dy, eps, mv; };

// dt = min(cfl dx / max(min(max_sx, max_val), eps), cfl dy / ...) (v1_shock.py:271-275); NaN if any cell was NaN
__device__ __forceinline__ double dt_from(const double* red, const DtParams& D) {
    if (red[2] != 0.0) return nan("");
    const double msx = fmin(red[0], D.mv), msy = fmin(red[1], D.mv);
    return fmin(D.cfl * D.dx / fmax(msx, D.eps), D.cfl * D.dy / fmax(msy, D.eps));
}

// closes a step: ring record, counters, next dt from the (all-reduced) maxima, accumulators cleared
__global__ void k_euler_ctl_end(Ctl* c, const double* __restrict__ reduced, double* __restrict__ ring, DtParams D) {
    if (threadIdx.x || blockIdx.x) return;
    ring[c->calls % RING] = c->skip ? 0.0 : c->dt;
    c->calls += 1;
    if (c->skip) c->skipped += 1;
    else {
        c->step += 1;
        const double next = dt_from(reduced, D);
        c->dt = next;
        c->skip = !isfinite(next);
    }
    c->msx = 0ull; c->msy = 0ull; c->nanflag = 0ull;
}

__global__ void k_euler_ctl_init(Ctl* c, const double* __restrict__ reduced, DtParams D, long long step) {
    if (threadIdx.x || blockIdx.x) return;
    const double dt = dt_from(reduced, D);
    c->dt = dt;
    c->skip = !isfinite(dt);
    if (step >= 0) { c->step = step; c->calls = step; c->skipped = 0; }   // step < 0: only dt is re-derived
    c->msx = 0ull; c->msy = 0ull; c->nanflag = 0ull;
}

__global__ void k_euler_zero_red(unsigned long long* red) { if (!threadIdx.x && !blockIdx.x) { red[0] = red[1] = red[2] = red[3] = 0ull; } }

// ---- unit-parity kernels (one per reference function) ----
template <int SOLVER>
__global__ void k_riemann(const float4* __restrict__ UL, const float4* __restrict__ UR, int n, int xn, Phys P, float4* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const Cell L = make_cell<true>(UL[k], P), R = make_cell<true>(UR[k], P);
    out[k] = xn ? face_flux<SOLVER, true>(L, R, P) : face_flux<SOLVER, false>(L, R, P);
}

__global__ void k_fluxes(const float4* __restrict__ U, int n, Phys P, float4* __restrict__ F, float4* __restrict__ G) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const Cell c = make_cell<true>(U[k], P);
    F[k] = c.Fp; G[k] = c.Gp;
}

__global__ void k_update_state(const float4* __restrict__ U, const float4* __restrict__ F, const float4* __restrict__ G,
                               int nx, int ny, double ax, double ay, Phys P, float4* __restrict__ Un) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= ny) return;
    const size_t o = (size_t)i * ny + j;
    if (i == 0 || i == nx - 1 || j == 0 || j == ny - 1) { Un[o] = f4zero(); return; }
    Un[o] = update_cell(U[o], F[o - ny], F[o], G[o - 1], G[o], ax, ay, P);
}

__global__ void k_bc_rows(float4* U, int nx, int ny, float4 fs) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ny) return;
    U[j] = fs;
    U[(size_t)(nx - 1) * ny + j] = U[(size_t)(nx - 2) * ny + j];
}
__global__ void k_bc_flip(float4* U, int nx, int ny, int n_reflect) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= n_reflect || j >= ny) return;
    U[(size_t)i * ny + j].z = -U[(size_t)i * ny + j].z;
}
__global__ void k_bc_cols(float4* U, int nx, int ny) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nx) return;
    U[(size_t)i * ny] = U[(size_t)i * ny + 1];
    U[(size_t)i * ny + ny - 1] = U[(size_t)i * ny + ny - 2];
}
__global__ void k_bc_wedge(float4* U, const uint8_t* __restrict__ w, long long n) {
    const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (k >= n || !w[k]) return;
    U[k].y = 0.f; U[k].z = 0.f;
}

}}  // namespace b200cfd::eu

// =====================================================================================================================
// C ABI (include/b200cfd_euler.h).  Host orchestration only; there is no CPU fallback.
// =====================================================================================================================
using namespace b200cfd::eu;

static thread_local std::string g_eu_create_error;

struct b200cfd_euler {
    b200cfd_euler_config cfg;
    int device = 0, n_sm = 148;
    StepParams sp;
    DtParams dp;
    Ctl* ctl = nullptr;             // library-owned
    double* ring = nullptr;
    unsigned long long* red_tmp = nullptr;   // scratch reduction vector for k_compute_dt
    float4* U[2] = {nullptr, nullptr};       // caller-owned
    const uint8_t* wedge = nullptr;
    bool bound = false;
    long long calls = 0;            // host mirror of ctl->calls (buffer parity: current = U[(calls - calls0) & 1])
    long long calls0 = 0;
    cudaGraphExec_t graph[2] = {nullptr, nullptr};
    cudaStream_t cap_stream = nullptr;
    int grid = 0;
    std::string err;
};

static int efail(b200cfd_euler* h, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) h->err = buf; else g_eu_create_error = buf;
    return code;
}
#define ECU(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return efail(h, B200CFD_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); } while (0)
#define ELAUNCH(h) do { cudaError_t e_ = cudaPeekAtLastError(); if (e_ != cudaSuccess) return efail(h, B200CFD_ECUDA, "kernel launch: %s", cudaGetErrorString(e_)); } while (0)

static Phys make_phys(double gamma, double eps, double mv) {
    Phys P;
    P.gamma = gamma; P.gm1 = gamma - 1; P.eps = eps; P.mv = mv;
    P.epsf = (float)eps; P.gm1f = (float)(gamma - 1);
    return P;
}

static int cur_index(const b200cfd_euler* h) { return (int)((h->calls - h->calls0) & 1); }

static void launch_step(b200cfd_euler* h, int src, cudaStream_t s) {
    const float4* in = h->U[src];
    float4* out = h->U[src ^ 1];
    const bool exact = h->cfg.exact_division != 0;
    const int threads = 32 * WARPS;
    if (h->cfg.solver_type == B200CFD_EULER_HLLC) {
        if (exact) k_euler_step<0, true><<<h->grid, threads, 0, s>>>(h->sp, in, out, h->wedge, h->ctl);
        else       k_euler_step<0, false><<<h->grid, threads, 0, s>>>(h->sp, in, out, h->wedge, h->ctl);
    } else {
        if (exact) k_euler_step<1, true><<<h->grid, threads, 0, s>>>(h->sp, in, out, h->wedge, h->ctl);
        else       k_euler_step<1, false><<<h->grid, threads, 0, s>>>(h->sp, in, out, h->wedge, h->ctl);
    }
}

static void drop_graphs(b200cfd_euler* h) {
    for (int k = 0; k < 2; ++k) if (h->graph[k]) { cudaGraphExecDestroy(h->graph[k]); h->graph[k] = nullptr; }
}

extern "C" {

const char* b200cfd_euler_last_error(const b200cfd_euler* h) { return h ? h->err.c_str() : g_eu_create_error.c_str(); }

int b200cfd_euler_create(const b200cfd_euler_config* cfg, int device, b200cfd_euler** out) {
    if (!cfg || !out) return efail(nullptr, B200CFD_EINVAL, "null argument");
    if (cfg->nx < 3 || cfg->ny < 3) return efail(nullptr, B200CFD_EINVAL, "grid must be at least 3x3 (got %dx%d)", cfg->nx, cfg->ny);
    if (!(cfg->dx > 0) || !(cfg->dy > 0)) return efail(nullptr, B200CFD_EINVAL, "dx, dy must be positive");
    if (cfg->solver_type != B200CFD_EULER_HLLC && cfg->solver_type != B200CFD_EULER_ROE) return efail(nullptr, B200CFD_EINVAL, "bad solver_type %d", cfg->solver_type);
    if (cfg->nranks < 1 || cfg->rank < 0 || cfg->rank >= cfg->nranks) return efail(nullptr, B200CFD_EINVAL, "bad rank %d of %d", cfg->rank, cfg->nranks);
    if (cfg->halo_lo < 0 || cfg->halo_lo > 1 || cfg->halo_hi < 0 || cfg->halo_hi > 1) return efail(nullptr, B200CFD_EINVAL, "halo rows must be 0 or 1");
    if ((cfg->rank > 0) != (cfg->halo_lo == 1) || (cfg->rank < cfg->nranks - 1) != (cfg->halo_hi == 1))
        return efail(nullptr, B200CFD_EINVAL, "interior slab sides need exactly one halo row");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) { cudaGetLastError(); return efail(nullptr, B200CFD_ECUDA, "no CUDA device (this library has no CPU fallback): %s", cudaGetErrorString(e)); }
    if (device < 0 || device >= ndev) return efail(nullptr, B200CFD_EINVAL, "device %d out of range (%d devices)", device, ndev);
    b200cfd_euler* h = new (std::nothrow) b200cfd_euler();
    if (!h) return efail(nullptr, B200CFD_ENOMEM, "host allocation failed");
    h->cfg = *cfg;
    h->device = device;
#define ECUC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { efail(nullptr, B200CFD_ECUDA, "%s: %s", #call, cudaGetErrorString(e_)); b200cfd_euler_destroy(h); return B200CFD_ECUDA; } } while (0)
    ECUC(cudaSetDevice(device));
    cudaDeviceProp prop;
    ECUC(cudaGetDeviceProperties(&prop, device));
    h->n_sm = prop.multiProcessorCount;
    ECUC(cudaMalloc(&h->ctl, sizeof(Ctl)));
    ECUC(cudaMemset(h->ctl, 0, sizeof(Ctl)));
    ECUC(cudaMalloc(&h->ring, sizeof(double) * RING));
    ECUC(cudaMemset(h->ring, 0, sizeof(double) * RING));
    ECUC(cudaMalloc(&h->red_tmp, 4 * sizeof(unsigned long long)));
    ECUC(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));   // graphs are captured here (the caller's stream may be the legacy default stream)
#undef ECUC
    StepParams& S = h->sp;
    S.nx = cfg->nx; S.ny = cfg->ny;
    const bool first = cfg->rank == 0, last = cfg->rank == cfg->nranks - 1;
    S.r_first = first ? 1 : cfg->halo_lo;                       // global row 0 is never updated (update_state :214)
    S.r_last = last ? cfg->nx - 2 : cfg->nx - 1 - cfg->halo_hi;
    S.inflow_row = first ? 0 : -1;
    S.outflow_row = last ? cfg->nx - 1 : -1;
    S.n_reflect = cfg->n_reflect < 0 ? 0 : (cfg->n_reflect > cfg->ny ? cfg->ny : cfg->n_reflect);
    S.dx = cfg->dx; S.dy = cfg->dy;
    S.phys = make_phys(cfg->gamma, cfg->epsilon, cfg->max_val);
    // free stream exactly as apply_boundary_conditions assigns it into the f32 array (:279-282)
    S.fs = make_float4((float)cfg->rho_inf, (float)(cfg->rho_inf * cfg->u_inf), 0.0f,
                       (float)(cfg->rho_inf * (cfg->p_inf / (cfg->rho_inf * (cfg->gamma - 1)) + 0.5 * (cfg->u_inf * cfg->u_inf))));
    S.nwin = (cfg->ny - 2 + OUT_LANES - 1) / OUT_LANES;
    const int rows = S.r_last - S.r_first + 1;
    if (rows < 1) { b200cfd_euler_destroy(h); return efail(nullptr, B200CFD_EINVAL, "slab has no interior rows"); }
    // chunk of rows per warp: enough warps to fill the machine ~16 deep, chunks long enough to amortise the 2 extra rows
    long long want = (long long)h->n_sm * 16 * WARPS;
    long long ch = (long long)rows * S.nwin / want;
    if (ch < 8) ch = 8;
    if (ch > 128) ch = 128;
    if (ch > rows) ch = rows;
    S.ch = (int)ch;
    S.nchunk = (rows + S.ch - 1) / S.ch;
    h->grid = (int)(((long long)S.nwin * S.nchunk + WARPS - 1) / WARPS);
    h->dp = DtParams{cfg->cfl, cfg->dx, cfg->dy, cfg->epsilon, cfg->max_val};
    *out = h;
    return B200CFD_OK;
}

int b200cfd_euler_destroy(b200cfd_euler* h) {
    if (!h) return B200CFD_OK;
    cudaSetDevice(h->device);
    drop_graphs(h);
    cudaFree(h->ctl); cudaFree(h->ring); cudaFree(h->red_tmp);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    delete h;
    return B200CFD_OK;
}

int b200cfd_euler_local_speeds(b200cfd_euler* h, void* stream) {
    if (!h || !h->bound) return efail(h, B200CFD_ESTATE, "bind first");
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    unsigned long long* red = &h->ctl->msx;
    k_euler_zero_red<<<1, 32, 0, s>>>(red);
    const long long n = (long long)h->cfg.nx * h->cfg.ny;
    k_euler_speeds<<<h->n_sm * 8, 256, 0, s>>>(h->U[cur_index(h)], n, h->sp.phys, red);
    ELAUNCH(h);
    return B200CFD_OK;
}

int b200cfd_euler_set_dt_from(b200cfd_euler* h, const double* reduced_dev, int64_t step, void* stream) {
    if (!h || !h->bound) return efail(h, B200CFD_ESTATE, "bind first");
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    k_euler_ctl_init<<<1, 32, 0, s>>>(h->ctl, reduced_dev ? reduced_dev : (const double*)&h->ctl->msx, h->dp, (long long)step);
    ELAUNCH(h);
    // the current buffer keeps its place: re-base the parity so that cur_index() is unchanged
    const int cur = cur_index(h);
    h->calls = step; h->calls0 = step - cur;
    return B200CFD_OK;
}

int b200cfd_euler_bind(b200cfd_euler* h, void* U_a, void* U_b, const uint8_t* wedge_mask, int64_t step, void* stream) {
    if (!h) return B200CFD_EINVAL;
    if (!U_a || !U_b || !wedge_mask) return efail(h, B200CFD_EINVAL, "null device pointer");
    if (((uintptr_t)U_a | (uintptr_t)U_b) & 15) return efail(h, B200CFD_EINVAL, "state buffers must be 16-byte aligned");
    drop_graphs(h);
    h->U[0] = (float4*)U_a; h->U[1] = (float4*)U_b; h->wedge = wedge_mask;
    h->bound = true;
    h->calls = h->calls0 = step;
    int rc = b200cfd_euler_local_speeds(h, stream);
    if (rc) return rc;
    if (h->cfg.nranks == 1) return b200cfd_euler_set_dt_from(h, nullptr, step, stream);
    return B200CFD_OK;    // slabs: the caller all-reduces the reduction vector, then calls set_dt_from
}

void* b200cfd_euler_reduction_ptr(b200cfd_euler* h) { return h ? (void*)&h->ctl->msx : nullptr; }
int   b200cfd_euler_reduction_len(void) { return 4; }

int b200cfd_euler_step_begin(b200cfd_euler* h, void* stream) {
    if (!h || !h->bound) return efail(h, B200CFD_ESTATE, "bind first");
    ECU(h, cudaSetDevice(h->device));
    launch_step(h, cur_index(h), (cudaStream_t)stream);
    ELAUNCH(h);
    return B200CFD_OK;
}

int b200cfd_euler_step_end(b200cfd_euler* h, const double* reduced_dev, void* stream) {
    if (!h || !h->bound) return efail(h, B200CFD_ESTATE, "bind first");
    ECU(h, cudaSetDevice(h->device));
    k_euler_ctl_end<<<1, 32, 0, (cudaStream_t)stream>>>(h->ctl, reduced_dev ? reduced_dev : (const double*)&h->ctl->msx, h->ring, h->dp);
    ELAUNCH(h);
    h->calls += 1;
    return B200CFD_OK;
}

int b200cfd_euler_launches_per_step(b200cfd_euler*) { return 2; }

int b200cfd_euler_step(b200cfd_euler* h, int n_steps, int use_graph, void* stream) {
    if (!h || !h->bound) return efail(h, B200CFD_ESTATE, "bind first");
    if (n_steps < 0) return efail(h, B200CFD_EINVAL, "n_steps < 0");
    if (h->cfg.nranks != 1) return efail(h, B200CFD_ESTATE, "slabs step with step_begin / step_end");
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    if (use_graph && !h->graph[0]) {
        for (int src = 0; src < 2; ++src) {
            cudaGraph_t g = nullptr;
            ECU(h, cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
            launch_step(h, src, h->cap_stream);
            k_euler_ctl_end<<<1, 32, 0, h->cap_stream>>>(h->ctl, (const double*)&h->ctl->msx, h->ring, h->dp);
            cudaError_t e = cudaStreamEndCapture(h->cap_stream, &g);
            if (e != cudaSuccess) return efail(h, B200CFD_ECUDA, "graph capture: %s", cudaGetErrorString(e));
            e = cudaGraphInstantiate(&h->graph[src], g, 0);
            cudaGraphDestroy(g);
            if (e != cudaSuccess) { h->graph[src] = nullptr; drop_graphs(h); return efail(h, B200CFD_ECUDA, "graph instantiate: %s", cudaGetErrorString(e)); }
        }
    }
    for (int k = 0; k < n_steps; ++k) {
        const int src = cur_index(h);
        if (use_graph) ECU(h, cudaGraphLaunch(h->graph[src], s));
        else {
            launch_step(h, src, s);
            k_euler_ctl_end<<<1, 32, 0, s>>>(h->ctl, (const double*)&h->ctl->msx, h->ring, h->dp);
            ELAUNCH(h);
        }
        h->calls += 1;
    }
    return B200CFD_OK;
}

int b200cfd_euler_step_host(b200cfd_euler* h, void* U_host, int n_steps, int use_graph, double* dt_last, void* stream) {
    if (!h || !h->bound) return efail(h, B200CFD_ESTATE, "bind first");
    if (!U_host) return efail(h, B200CFD_EINVAL, "null host array");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t bytes = (size_t)h->cfg.nx * h->cfg.ny * sizeof(float4);
    ECU(h, cudaSetDevice(h->device));
    ECU(h, cudaMemcpyAsync(h->U[cur_index(h)], U_host, bytes, cudaMemcpyHostToDevice, s));
    // compute_dt reads the array it is given (v1_shock.py:293)
    int rc = b200cfd_euler_local_speeds(h, stream);
    if (rc) return rc;
    k_euler_ctl_init<<<1, 32, 0, s>>>(h->ctl, (const double*)&h->ctl->msx, h->dp, -1LL);
    rc = b200cfd_euler_step(h, n_steps, use_graph, stream);
    if (rc) return rc;
    ECU(h, cudaMemcpyAsync(U_host, h->U[cur_index(h)], bytes, cudaMemcpyDeviceToHost, s));
    if (dt_last && n_steps > 0) ECU(h, cudaMemcpyAsync(dt_last, h->ring + ((h->calls - 1) % RING), sizeof(double), cudaMemcpyDeviceToHost, s));
    ECU(h, cudaStreamSynchronize(s));
    return B200CFD_OK;
}

int b200cfd_euler_state(b200cfd_euler* h, int64_t* step, int32_t* current, double* next_dt, int64_t* skipped, void* stream) {
    if (!h) return B200CFD_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    Ctl c;
    ECU(h, cudaMemcpyAsync(&c, h->ctl, sizeof c, cudaMemcpyDeviceToHost, s));
    ECU(h, cudaStreamSynchronize(s));
    if (step) *step = c.step;
    if (current) *current = cur_index(h);
    if (next_dt) *next_dt = c.dt;
    if (skipped) *skipped = c.skipped;
    return B200CFD_OK;
}

int b200cfd_euler_fetch_dts(b200cfd_euler* h, int64_t first, int n, double* out, void* stream) {
    if (!h || !out || n < 0) return efail(h, B200CFD_EINVAL, "bad argument");
    if (first < h->calls - RING || first + n > h->calls || first < 0) return efail(h, B200CFD_EINVAL, "calls [%lld, %lld) are not in the ring (have up to %lld)", (long long)first, (long long)first + n, h->calls);
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    for (int k = 0; k < n;) {
        const long long idx = (first + k) % RING;
        const int run = (int)((RING - idx) < (long long)(n - k) ? (RING - idx) : (n - k));
        ECU(h, cudaMemcpyAsync(out + k, h->ring + idx, sizeof(double) * run, cudaMemcpyDeviceToHost, s));
        k += run;
    }
    ECU(h, cudaStreamSynchronize(s));
    return B200CFD_OK;
}

int b200cfd_euler_k_riemann(const void* UL, const void* UR, int n, int solver_type, int x_normal,
                            double gamma, double epsilon, double max_val, void* out, void* stream) {
    if (!UL || !UR || !out || n < 0) return efail(nullptr, B200CFD_EINVAL, "bad argument");
    const Phys P = make_phys(gamma, epsilon, max_val);
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) return B200CFD_OK;
    if (solver_type == B200CFD_EULER_HLLC) k_riemann<0><<<(n + 127) / 128, 128, 0, s>>>((const float4*)UL, (const float4*)UR, n, x_normal, P, (float4*)out);
    else k_riemann<1><<<(n + 127) / 128, 128, 0, s>>>((const float4*)UL, (const float4*)UR, n, x_normal, P, (float4*)out);
    ELAUNCH(nullptr);
    return B200CFD_OK;
}

int b200cfd_euler_k_fluxes(const void* U, int n, double gamma, double epsilon, double max_val, void* F, void* G, void* stream) {
    if (!U || !F || !G || n < 0) return efail(nullptr, B200CFD_EINVAL, "bad argument");
    if (n == 0) return B200CFD_OK;
    k_fluxes<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>((const float4*)U, n, make_phys(gamma, epsilon, max_val), (float4*)F, (float4*)G);
    ELAUNCH(nullptr);
    return B200CFD_OK;
}

int b200cfd_euler_k_update_state(b200cfd_euler* h, const void* U, const void* F, const void* G, double dt, void* U_new, void* stream) {
    if (!h || !U || !F || !G || !U_new) return efail(h, B200CFD_EINVAL, "null argument");
    ECU(h, cudaSetDevice(h->device));
    dim3 grid((h->cfg.ny + 127) / 128, h->cfg.nx);
    k_update_state<<<grid, 128, 0, (cudaStream_t)stream>>>((const float4*)U, (const float4*)F, (const float4*)G, h->cfg.nx, h->cfg.ny,
                                                           dt / h->cfg.dx, dt / h->cfg.dy, h->sp.phys, (float4*)U_new);
    ELAUNCH(h);
    return B200CFD_OK;
}

int b200cfd_euler_k_compute_dt(b200cfd_euler* h, const void* U, double* out3, void* stream) {
    if (!h || !U || !out3) return efail(h, B200CFD_EINVAL, "null argument");
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    k_euler_zero_red<<<1, 32, 0, s>>>(h->red_tmp);
    k_euler_speeds<<<h->n_sm * 8, 256, 0, s>>>((const float4*)U, (long long)h->cfg.nx * h->cfg.ny, h->sp.phys, h->red_tmp);
    ELAUNCH(h);
    double red[4];
    ECU(h, cudaMemcpyAsync(red, h->red_tmp, sizeof red, cudaMemcpyDeviceToHost, s));
    ECU(h, cudaStreamSynchronize(s));
    const DtParams& D = h->dp;
    const double msx = red[0] < D.mv ? red[0] : D.mv, msy = red[1] < D.mv ? red[1] : D.mv;
    const double dtx = D.cfl * D.dx / (msx > D.eps ? msx : D.eps), dty = D.cfl * D.dy / (msy > D.eps ? msy : D.eps);
    out3[0] = red[2] != 0.0 ? nan("") : (dtx < dty ? dtx : dty);
    out3[1] = msx; out3[2] = msy;
    return B200CFD_OK;
}

int b200cfd_euler_k_apply_bc(b200cfd_euler* h, void* U, const uint8_t* wedge_mask, void* stream) {
    if (!h || !U) return efail(h, B200CFD_EINVAL, "null argument");
    if (h->cfg.nranks != 1) return efail(h, B200CFD_ESTATE, "unit kernel: single slab only");
    cudaStream_t s = (cudaStream_t)stream;
    ECU(h, cudaSetDevice(h->device));
    const int nx = h->cfg.nx, ny = h->cfg.ny;
    k_bc_rows<<<(ny + 127) / 128, 128, 0, s>>>((float4*)U, nx, ny, h->sp.fs);
    if (h->sp.n_reflect > 0) k_bc_flip<<<dim3((h->sp.n_reflect + 127) / 128, nx), 128, 0, s>>>((float4*)U, nx, ny, h->sp.n_reflect);
    if (h->sp.n_reflect < ny) k_bc_cols<<<(nx + 127) / 128, 128, 0, s>>>((float4*)U, nx, ny);
    if (wedge_mask) k_bc_wedge<<<(unsigned)(((long long)nx * ny + 255) / 256), 256, 0, s>>>((float4*)U, wedge_mask, (long long)nx * ny);
    ELAUNCH(h);
    return B200CFD_OK;
}

}  // extern "C"
