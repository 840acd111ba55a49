// poisson.cu — pressure-Poisson kernels, baseline (one HBM pass per sweep) versions.
// Reference: python/flow_over_cylinder (Fischer)/v5.py:202-226 (red-black Gauss-Seidel) and :336-346 (Jacobi branch).
// The temporally blocked shared-memory/TMA versions live in poisson_tb.cu; these are the general
// fallbacks (any shape, any pitch) and the building blocks of the tolerance-terminated mode.
#include "kernels.h"
#include "cellmath.cuh"

namespace b200cfd {

template <typename T> __device__ __forceinline__ double ldp(const T* p, long long k) { return (double)p[k]; }

// One colour half-sweep.  colour 0 <=> (i + j_global) odd (v5.py:215: j starts at 1 + (i+color)%2).
// Each thread owns one cell of the colour.  max|phi_new - phi| is reduced per block and folded into
// ctl->pois_change with one atomicMax; when `close_iteration` is set the last block to finish closes the
// iteration (pois_last, pois_iters, pois_done = change < tol) so a tolerance-terminated solve needs no
// host round trip: once pois_done is set every later launch returns immediately.
#define PBX 64
#define PBY 4
template <typename T>
__global__ void __launch_bounds__(PBX* PBY)
k_rbgs_color(Geom g, T* __restrict__ phi, const T* __restrict__ div, const uint8_t* __restrict__ mask, double inv_dt,
             double omega, int color, StepCtl* __restrict__ ctl, double tol, int close_iteration) {
    if (ctl && ctl->pois_done) return;
    const int i = blockIdx.y * PBY + threadIdx.y;
    // first local column of this colour in row i: global (i + jg) parity == 1 - color ... colour 0 -> odd
    const int par = (i + g.jg0 + color + 1) & 1;         // parity of local j for cells of this colour in row i
    const int j = 2 * (blockIdx.x * PBX + threadIdx.x) + par;
    double ch = 0.0;
    if (i >= 1 && i < g.ny - 1 && j >= 1 && j < g.nx - 1 && g.jg0 + j >= 1 && g.jg0 + j < g.nxg - 1) {
        const long long k = i * g.pitch + j;
        if (!(mask && mask[k])) {
            const double rhs = -ldp(div, k) * inv_dt;
            const double old = ldp(phi, k);
            double pn = (g.ax * (ldp(phi, k + 1) + ldp(phi, k - 1)) + g.ay * (ldp(phi, k + g.pitch) + ldp(phi, k - g.pitch)) - rhs) * g.dinv;
            if (omega != 1.0) pn = old + omega * (pn - old);
            if (j >= g.own0 && j < g.own1) ch = fabs(pn - old);
            phi[k] = (T)pn;
        }
    }
    if (!ctl) return;
    block_atomic_max(ch, &ctl->pois_change);
    if (close_iteration) {
        __shared__ bool s_last;
        if (threadIdx.x == 0 && threadIdx.y == 0) {
            __threadfence();
            const unsigned total = gridDim.x * gridDim.y;
            s_last = (atomicAdd(&ctl->pois_blocks_done, 1u) == total - 1);
        }
        __syncthreads();
        if (s_last && threadIdx.x == 0 && threadIdx.y == 0) {
            __threadfence();
            const double c = bits_to_double(atomicExch(&ctl->pois_change, 0ull));
            ctl->pois_last = c;
            ctl->pois_iters += 1;
            ctl->pois_blocks_done = 0u;
            if (c < tol) ctl->pois_done = 1;
            __threadfence();
        }
    }
}

// v5.py:338-346: phi_new = copy(phi); interior = 0.25 (E + W + N + S - dx^2 div / dt); masked -> 0.
template <typename T>
__global__ void __launch_bounds__(PBX* PBY)
k_jacobi(Geom g, const T* __restrict__ in, T* __restrict__ out, const T* __restrict__ div, const uint8_t* __restrict__ mask,
         double dx2_over_dt) {
    const int j = blockIdx.x * PBX + threadIdx.x, i = blockIdx.y * PBY + threadIdx.y;
    if (i >= g.ny || j >= g.nx) return;
    const long long k = i * g.pitch + j;
    double o;
    if (mask && mask[k]) o = 0.0;
    else if (i >= 1 && i < g.ny - 1 && j >= 1 && j < g.nx - 1 && g.jg0 + j >= 1 && g.jg0 + j < g.nxg - 1)
        o = 0.25 * (ldp(in, k + 1) + ldp(in, k - 1) + ldp(in, k + g.pitch) + ldp(in, k - g.pitch) - dx2_over_dt * ldp(div, k));
    else o = ldp(in, k);
    out[k] = (T)o;
}

__global__ void k_poisson_reset(StepCtl* c) {
    if (threadIdx.x || blockIdx.x) return;
    c->pois_change = 0ull; c->pois_last = 0.0; c->pois_iters = 0; c->pois_done = 0; c->pois_blocks_done = 0u;
}

// K2 fused "as executed" pressure solve (SURVEY finding 3): div(u*,v*) + exactly ONE red-black iteration from
// phi = 0, in one pass.  Odd cells:  phi = (D/dt_c) d.  Even cells: phi = (a(pE+pW) + b(pN+pS) + D/dt_c) d with
// the four (odd) neighbours evaluated in place from D at those cells.  Masked / border cells stay 0.
// Also materialises div_u_star (the reference keeps it as an attribute, v5.py:409) and max|div| (v5.py:410).
template <typename T>
__global__ void __launch_bounds__(PBX* PBY)
k_div_rb1(Geom g, const T* __restrict__ us, const T* __restrict__ vs, const uint8_t* __restrict__ mask,
          T* __restrict__ phi, T* __restrict__ div_out, double inv_dt, StepCtl* __restrict__ ctl) {
    const int j = blockIdx.x * PBX + threadIdx.x, i = blockIdx.y * PBY + threadIdx.y;
    double dabs = 0.0, ch = 0.0;
    if (i < g.ny && j < g.nx) divrb1_point<T>(g, us, vs, mask, phi, div_out, inv_dt, i, j, dabs, ch);
    if (ctl) {
        block_atomic_max(dabs, &ctl->max_div_pre);
        block_atomic_max(ch, &ctl->pois_change);
    }
}
__global__ void k_rb1_close(StepCtl* c, double tol) {
    if (threadIdx.x || blockIdx.x) return;
    const double ch = bits_to_double(c->pois_change);
    c->pois_last = ch; c->pois_iters = 1; c->pois_change = 0ull; c->pois_done = ch < tol ? 1 : 0;
}

void launch_rb1_close(StepCtl* ctl, double tol, cudaStream_t s) { k_rb1_close<<<1, 32, 0, s>>>(ctl, tol); }

static inline dim3 pgrid(const Geom& g) { return dim3((g.nx + PBX - 1) / PBX, (g.ny + PBY - 1) / PBY); }
static inline dim3 pgrid_color(const Geom& g) { return dim3((g.nx / 2 + 1 + PBX - 1) / PBX, (g.ny + PBY - 1) / PBY); }
static const dim3 kPBlock(PBX, PBY);

void launch_rbgs_color(int dtype, const Geom& g, void* phi, const void* div, const uint8_t* mask, double inv_dt,
                       double omega, int color, StepCtl* ctl, double tol, int close_iteration, cudaStream_t s) {
    if (dtype == 0)
        k_rbgs_color<double><<<pgrid_color(g), kPBlock, 0, s>>>(g, (double*)phi, (const double*)div, mask, inv_dt, omega, color, ctl, tol, close_iteration);
    else
        k_rbgs_color<float><<<pgrid_color(g), kPBlock, 0, s>>>(g, (float*)phi, (const float*)div, mask, inv_dt, omega, color, ctl, tol, close_iteration);
}
void launch_jacobi(int dtype, const Geom& g, const void* in, void* out, const void* div, const uint8_t* mask,
                   double dx2_over_dt, cudaStream_t s) {
    if (dtype == 0)
        k_jacobi<double><<<pgrid(g), kPBlock, 0, s>>>(g, (const double*)in, (double*)out, (const double*)div, mask, dx2_over_dt);
    else
        k_jacobi<float><<<pgrid(g), kPBlock, 0, s>>>(g, (const float*)in, (float*)out, (const float*)div, mask, dx2_over_dt);
}
// v2.py:299-303 (v1.py:290-294): after every SOR sweep  phi[:,0] = phi[:,1]; phi[:,-1] = 0; phi[0,:] = phi[1,:];
// phi[-1,:] = phi[-2,:]  (in this order: the wall rows copy the already updated side columns, so the corners end as
// phi[1,1], 0, phi[-2,1], 0).  One launch: the column part skips the wall rows, the row part reproduces the corners.
template <typename T>
__global__ void __launch_bounds__(256)
k_sor_v2_bc(Geom g, T* __restrict__ phi, const StepCtl* __restrict__ ctl) {
    (void)ctl;   // always applied: idempotent once the colour kernels have stopped, and the stopping sweep needs it
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= 1 && t < g.ny - 1) {
        const long long r = (long long)t * g.pitch;
        phi[r] = phi[r + 1];
        phi[r + g.nx - 1] = (T)0;
    }
    if (t < g.nx) {
        const int jr = t == 0 ? 1 : t;                                  // column whose value the wall rows copy
        const long long r1 = g.pitch, rm = (long long)(g.ny - 2) * g.pitch, rl = (long long)(g.ny - 1) * g.pitch;
        const bool right = t == g.nx - 1;
        phi[t] = right ? (T)0 : phi[r1 + jr];
        phi[rl + t] = right ? (T)0 : phi[rm + jr];
    }
}
void launch_sor_v2_bc(int dtype, const Geom& g, void* phi, const StepCtl* ctl, cudaStream_t s) {
    const int n = g.nx > g.ny ? g.nx : g.ny;
    if (dtype == 0) k_sor_v2_bc<double><<<(n + 255) / 256, 256, 0, s>>>(g, (double*)phi, ctl);
    else k_sor_v2_bc<float><<<(n + 255) / 256, 256, 0, s>>>(g, (float*)phi, ctl);
}
void launch_poisson_reset(StepCtl* ctl, cudaStream_t s) { k_poisson_reset<<<1, 32, 0, s>>>(ctl); }

void launch_div_rb1(int dtype, const Geom& g, const void* us, const void* vs, const uint8_t* mask, void* phi,
                    void* div_out, double inv_dt, StepCtl* ctl, double tol, cudaStream_t s) {
    if (dtype == 0)
        k_div_rb1<double><<<pgrid(g), kPBlock, 0, s>>>(g, (const double*)us, (const double*)vs, mask, (double*)phi, (double*)div_out, inv_dt, ctl);
    else
        k_div_rb1<float><<<pgrid(g), kPBlock, 0, s>>>(g, (const float*)us, (const float*)vs, mask, (float*)phi, (float*)div_out, inv_dt, ctl);
    if (ctl) k_rb1_close<<<1, 32, 0, s>>>(ctl, tol);
}

}  // namespace b200cfd
