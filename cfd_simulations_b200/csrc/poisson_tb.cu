// poisson_tb.cu — temporally blocked pressure-Poisson sweeps: several Jacobi sweeps / red-black SOR iterations per
// HBM pass.  Reference: python/flow_over_cylinder (Fischer)/v5.py:336-346 (Jacobi branch) and :202-226 (red-black
// Gauss-Seidel; omega != 1 gives SOR, the converged-solution counterpart of v2.py:284-306).
//
// One CTA owns a TX x TY tile of the output.  It stages the tile plus a halo of H cells (H = depth for Jacobi,
// 2*depth for red-black) of phi and of the pre-scaled right-hand side in shared memory with TMA
// (cp.async.bulk.tensor.2d, completion on an mbarrier; out-of-bounds cells are zero-filled by the hardware, which is
// exactly the Dirichlet-0 exterior), runs `depth` sweeps on the shrinking trapezoid entirely in shared memory and
// writes the tile back.  Passes ping-pong between two global arrays because neighbouring CTAs read each other's halos.
// The cylinder mask is folded into the right-hand side once per solve (masked cell = NaN sentinel), so a pass moves
// 8 (phi in) + 8 (rhs) + 8 (phi out) B/cell instead of 25 B/cell PER SWEEP for the one-sweep kernels.
// In tolerance mode every CTA also reduces max|dphi| of each of its iterations over the cells it owns (exact values),
// warp shuffle -> block -> one atomicMax per iteration slot, so the caller can stop at exactly the iteration at which
// the sequential algorithm stops.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <cstdlib>
#include <string>

#include "poisson_tb.h"
#include "tma.cuh"

namespace b200cfd {

#define TB_RW 128          // staged region width (cells): one thread column each, TMA box inner dimension
#define TB_THREADS 256     // 2 threads per column (upper / lower half of the rows)
#define TB_MAX_DEPTH 8

static std::string g_tb_err;
const char* poisson_tb_error() { return g_tb_err.c_str(); }

struct PoissonTB {
    PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
    void* rhs_s = nullptr;      // pre-scaled right-hand side with the mask folded in (NaN = masked)
    void* alt = nullptr;        // ping-pong partner of phi
    size_t bytes = 0;
    unsigned long long* slots = nullptr;   // per-iteration max|dphi| of the last pass (TB_MAX_DEPTH entries)
    int last_depth = 0, last_passes = 0;   // what the last solve did (b200cfd_poisson_info)
    PoissonStream* stream = nullptr;       // state of the streaming-pipeline kernel (poisson_stream.cu)
    bool last_stream = false;
};

void poisson_tb_keep_rhs(PoissonTB** ptb, int keep) {
    if (!*ptb) *ptb = new PoissonTB();
    poisson_stream_keep_rhs(&(*ptb)->stream, keep);
}
void poisson_tb_info(const PoissonTB* tb, int* depth, int* passes) {
    if (tb && tb->last_stream) { poisson_stream_info(tb->stream, depth, passes); return; }
    *depth = tb ? tb->last_depth : 0;
    *passes = tb ? tb->last_passes : 0;
}

void poisson_tb_destroy(PoissonTB* tb) {
    if (!tb) return;
    cudaFree(tb->rhs_s); cudaFree(tb->alt); cudaFree(tb->slots);
    poisson_stream_destroy(tb->stream);
    delete tb;
}

struct TBParams {
    int ny, nx; long long pitch; int jg0, nxg, own0, own1;
    int TX, TY, H, RH, depth;
    double ax, ay, dinv, omega, scale;
};

// rhs_s = masked ? NaN : div.  The scale (Jacobi: dx^2/dt, v5.py:341; red-black: -1/dt, v5.py:217-219) is applied in
// double inside the sweep, exactly as the one-sweep kernels do, so both paths round identically in fp32 storage too.
template <typename T>
__global__ void k_tb_make_rhs(long long n, const T* __restrict__ div, const uint8_t* __restrict__ mask, double scale,
                              T* __restrict__ out) {
    const T nan_t = (T)__longlong_as_double(0x7ff8000000000000ll);
    for (long long k = blockIdx.x * 256ll + threadIdx.x; k < n; k += (long long)gridDim.x * 256ll)
        out[k] = (mask && mask[k]) ? nan_t : div[k];
}

// MODE 0: Jacobi (ping-pong in shared memory), MODE 1: red-black Gauss-Seidel / SOR (in place in shared memory)
template <typename T, int MODE>
__global__ void __launch_bounds__(TB_THREADS)
k_poisson_tb(const __grid_constant__ CUtensorMap tm_phi, const __grid_constant__ CUtensorMap tm_rhs, T* __restrict__ out,
             TBParams P, unsigned long long* __restrict__ slots) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int RH = P.RH, H = P.H;
    T* sA = reinterpret_cast<T*>(smem_raw);
    T* sR = sA + RH * TB_RW;
    T* sB = sR + RH * TB_RW;                                   // Jacobi only
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw + (size_t)(MODE == 0 ? 3 : 2) * RH * TB_RW * sizeof(T));
    __shared__ double s_red[TB_THREADS / 32];
    const int tid = threadIdx.x, col = tid & (TB_RW - 1), half = tid >> 7;
    const int gx0 = blockIdx.x * P.TX - H, gy0 = blockIdx.y * P.TY - H;
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(bar, 2u * (uint32_t)(RH * TB_RW * sizeof(T)));
        tma_load_2d(sA, &tm_phi, gx0, gy0, bar);
        tma_load_2d(sR, &tm_rhs, gx0, gy0, bar);
    }
    mbar_wait(bar, 0);
    const int jl = gx0 + col, jg = P.jg0 + jl;                  // local / global column of this thread
    const bool col_in = jl >= 1 && jl < P.nx - 1 && jg >= 1 && jg < P.nxg - 1;
    const int rlo = half ? RH / 2 : 0, rhi = half ? RH : RH / 2;
    if (MODE == 0) {
        T* src = sA; T* dst = sB;
        for (int s = 1; s <= P.depth; ++s) {
            const int a = max(rlo, s), b = min(rhi, RH - s);
            if (col >= s && col < TB_RW - s && a < b) {
                double cm = (double)src[(a - 1) * TB_RW + col], cc = (double)src[a * TB_RW + col];
                for (int i = a; i < b; ++i) {
                    const double cp = (double)src[(i + 1) * TB_RW + col];
                    const double r = P.scale * (double)sR[i * TB_RW + col];
                    const int gi = gy0 + i;
                    double o = cc;
                    if (r != r) o = 0.0;                                                  // masked -> 0 (v5.py:345)
                    else if (col_in && gi >= 1 && gi < P.ny - 1)
                        o = 0.25 * ((double)src[i * TB_RW + col + 1] + (double)src[i * TB_RW + col - 1] + cp + cm - r);
                    dst[i * TB_RW + col] = (T)o;
                    cm = cc; cc = cp;
                }
            }
            __syncthreads();
            T* t = src; src = dst; dst = t;
        }
        sA = src;                                               // result buffer
    } else {
        const bool own_col = jl >= P.own0 && jl < P.own1 && col >= H && col < H + P.TX;
        for (int t = 1; t <= P.depth; ++t) {
            double ch = 0.0;
            for (int color = 0; color < 2; ++color) {
                const int hs = 2 * (t - 1) + color + 1;
                const int a = max(rlo, hs), b = min(rhi, RH - hs);
                if (col_in && col >= hs && col < TB_RW - hs) {
                    // colour 0 <=> (i + j) odd in global indices (v5.py:215)
                    const int want = color == 0 ? 1 : 0;
                    int i = a + ((((gy0 + a + jg) & 1) != want) ? 1 : 0);
                    for (; i < b; i += 2) {
                        const int gi = gy0 + i;
                        const double r = P.scale * (double)sR[i * TB_RW + col];
                        if (gi < 1 || gi >= P.ny - 1 || r != r) continue;                // border / masked: untouched
                        const double old = (double)sA[i * TB_RW + col];
                        double pn = (P.ax * ((double)sA[i * TB_RW + col + 1] + (double)sA[i * TB_RW + col - 1]) +
                                     P.ay * ((double)sA[(i + 1) * TB_RW + col] + (double)sA[(i - 1) * TB_RW + col]) - r) * P.dinv;
                        if (P.omega != 1.0) pn = old + P.omega * (pn - old);
                        sA[i * TB_RW + col] = (T)pn;
                        if (own_col && i >= H && i < H + P.TY) ch = fmax(ch, fabs(pn - old));
                    }
                }
                __syncthreads();
            }
            if (slots) {                                        // exact max|dphi| of iteration t over the cells this CTA owns
                ch = warp_max(ch);
                if ((tid & 31) == 0) s_red[tid >> 5] = ch;
                __syncthreads();
                if (tid < 32) {
                    double w = tid < TB_THREADS / 32 ? s_red[tid] : 0.0;
                    w = warp_max(w);
                    if (tid == 0 && w > 0.0) atomicMax(&slots[t - 1], (unsigned long long)__double_as_longlong(w));
                }
                __syncthreads();
            }
        }
    }
    // write the owned tile back (coalesced along x)
    if (col >= H && col < H + P.TX && jl >= 0 && jl < P.nx) {
        const int a = max(rlo, H), b = min(rhi, H + P.TY);
        for (int i = a; i < b; ++i) {
            const int gi = gy0 + i;
            if (gi >= 0 && gi < P.ny) out[(long long)gi * P.pitch + jl] = sA[i * TB_RW + col];
        }
    }
}

// ---- host side -----------------------------------------------------------------------------------
static int tb_init(PoissonTB** ptb, size_t bytes) {
    if (!*ptb || !(*ptb)->encode) {
        PoissonTB* tb = *ptb ? *ptb : new PoissonTB();
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) {
            cudaGetLastError(); g_tb_err = "cuTensorMapEncodeTiled is not available from this driver"; if (!*ptb) delete tb; return -1;
        }
        tb->encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
        if (cudaMalloc(&tb->slots, TB_MAX_DEPTH * sizeof(unsigned long long)) != cudaSuccess) { g_tb_err = "cudaMalloc(slots)"; if (!*ptb) delete tb; return -1; }
        *ptb = tb;
    }
    PoissonTB* tb = *ptb;
    if (tb->bytes < bytes) {
        cudaFree(tb->rhs_s); cudaFree(tb->alt); tb->rhs_s = tb->alt = nullptr; tb->bytes = 0;
        if (cudaMalloc(&tb->rhs_s, bytes) != cudaSuccess || cudaMalloc(&tb->alt, bytes) != cudaSuccess) {
            cudaGetLastError(); g_tb_err = "cudaMalloc of the temporal-blocking scratch failed"; return -1;
        }
        tb->bytes = bytes;
    }
    return 0;
}

static bool tb_supported(const Geom& g, int es, const void* a, const void* b) {
    // TMA: 16-byte aligned base and row pitch; the staged region must not dwarf the grid
    return ((uintptr_t)a % 16 == 0) && ((uintptr_t)b % 16 == 0) && ((g.pitch * es) % 16 == 0) && g.nx >= 32 && g.ny >= 16;
}

static int tb_map(PoissonTB* tb, CUtensorMap* m, int dtype, const Geom& g, const void* base, int RH) {
    const int es = dtype == 0 ? 8 : 4;
    cuuint64_t dims[2] = {(cuuint64_t)g.nx, (cuuint64_t)g.ny};
    cuuint64_t strides[1] = {(cuuint64_t)(g.pitch * es)};
    cuuint32_t box[2] = {TB_RW, (cuuint32_t)RH};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = tb->encode(m, dtype == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                            const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { g_tb_err = "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")"; return -1; }
    return 0;
}

// geometry of a pass: staged rows RH (shared-memory budget: 2 CTAs per SM), halo H, owned tile TX x TY
static void tb_geometry(int dtype, int mode, int depth, TBParams& P) {
    P.depth = depth;
    const int nbuf = mode == 0 ? 3 : 2;
    const int es = dtype == 0 ? 8 : 4;
    // the box origin (tile origin - H) must stay 16-byte aligned in x: odd element offsets fault on sm_100a
    const int al = 16 / es;
    P.H = ((mode == 0 ? depth : 2 * depth) + al - 1) / al * al;
    int RH = (int)((110 * 1024) / ((size_t)nbuf * TB_RW * es));      // <= 110 KB per CTA
    if (RH > 64) RH = 64;
    RH &= ~1;
    P.RH = RH;
    P.TX = TB_RW - 2 * P.H;
    P.TY = RH - 2 * P.H;
}

template <typename T, int MODE>
static int tb_launch(const CUtensorMap& mp, const CUtensorMap& mr, void* out, const TBParams& P, unsigned long long* slots,
                     cudaStream_t s) {
    const size_t smem = (size_t)(MODE == 0 ? 3 : 2) * P.RH * TB_RW * sizeof(T) + 16;
    static size_t configured = 0;
    if (configured < smem) {
        if (cudaFuncSetAttribute(k_poisson_tb<T, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            g_tb_err = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)"; return -1;
        }
        configured = smem;
    }
    const dim3 grid((P.nx + P.TX - 1) / P.TX, (P.ny + P.TY - 1) / P.TY);
    k_poisson_tb<T, MODE><<<grid, TB_THREADS, smem, s>>>(mp, mr, (T*)out, P, slots);
    return cudaPeekAtLastError() == cudaSuccess ? 0 : -1;
}

// runs `iters` sweeps / iterations in passes of `depth`; phi holds the result on return.  slots_out (device,
// TB_MAX_DEPTH entries, may be NULL) receives the per-iteration max|dphi| of the LAST pass (red-black only).
static int tb_run(PoissonTB** ptb, int dtype, int mode, const Geom& g, void* phi, const void* div, const uint8_t* mask,
                  double scale, double omega, int iters, int depth, int* iters_done, cudaStream_t s, double tol = 0.0,
                  double* last_change = nullptr) {
    *iters_done = 0;
    const int es = dtype == 0 ? 8 : 4;
    if (iters <= 0) return 0;
    if (poisson_stream_enabled()) {
        // default: the register-resident streaming pipeline; falls through to the tile kernel below when not applicable
        if (!*ptb) *ptb = new PoissonTB();
        (*ptb)->last_stream = false;
        const int rc = poisson_stream_run(&(*ptb)->stream, dtype, mode, g, 148, phi, div, mask, scale, omega, iters, depth, iters_done, s,
                                          tol, last_change);
        if (rc < 0) { g_tb_err = poisson_stream_error(); return -1; }
        if (*iters_done > 0) { (*ptb)->last_stream = true; return 0; }
    }
    if (depth <= 0) depth = mode == 0 ? 4 : 3;
    if (depth > TB_MAX_DEPTH) depth = TB_MAX_DEPTH;
    if (depth > iters) depth = iters;
    const size_t bytes = (size_t)g.ny * g.pitch * es;
    if (tb_init(ptb, bytes)) return -1;
    PoissonTB* tb = *ptb;
    if (!tb_supported(g, es, phi, tb->alt)) return 0;
    TBParams P;
    P.ny = g.ny; P.nx = g.nx; P.pitch = g.pitch; P.jg0 = g.jg0; P.nxg = g.nxg; P.own0 = g.own0; P.own1 = g.own1;
    P.ax = g.ax; P.ay = g.ay; P.dinv = g.dinv; P.omega = omega; P.scale = scale;
    tb_geometry(dtype, mode, depth, P);
    if (P.TY < 4 || P.TX < 16) return 0;
    const long long n = (long long)g.ny * g.pitch;
    const int nb = (int)((n + 255) / 256 < 148 * 32 ? (n + 255) / 256 : 148 * 32);
    if (dtype == 0) k_tb_make_rhs<double><<<nb, 256, 0, s>>>(n, (const double*)div, mask, scale, (double*)tb->rhs_s);
    else k_tb_make_rhs<float><<<nb, 256, 0, s>>>(n, (const float*)div, mask, scale, (float*)tb->rhs_s);
    CUtensorMap m_phi, m_alt, m_rhs;
    if (tb_map(tb, &m_phi, dtype, g, phi, P.RH) || tb_map(tb, &m_alt, dtype, g, tb->alt, P.RH) || tb_map(tb, &m_rhs, dtype, g, tb->rhs_s, P.RH)) return -1;
    void* cur = phi; void* nxt = tb->alt;
    int done = 0;
    const bool check = tol > 0.0 && mode == 1;
    tb->last_depth = depth; tb->last_passes = 0;
    auto pass = [&](int d) -> int {
        P.depth = d;
        ++tb->last_passes;
        const CUtensorMap& mc = cur == phi ? m_phi : m_alt;
        unsigned long long* sl = check ? tb->slots : nullptr;
        if (check && cudaMemsetAsync(tb->slots, 0, TB_MAX_DEPTH * sizeof(unsigned long long), s) != cudaSuccess) return -1;
        int rc;
        if (dtype == 0) rc = mode == 0 ? tb_launch<double, 0>(mc, m_rhs, nxt, P, sl, s) : tb_launch<double, 1>(mc, m_rhs, nxt, P, sl, s);
        else rc = mode == 0 ? tb_launch<float, 0>(mc, m_rhs, nxt, P, sl, s) : tb_launch<float, 1>(mc, m_rhs, nxt, P, sl, s);
        if (rc && g_tb_err.empty()) g_tb_err = "kernel launch failed";
        return rc;
    };
    while (done < iters) {
        const int d = iters - done < depth ? iters - done : depth;  // last pass may be shallower (the halo is wide enough)
        if (pass(d)) return -1;
        if (check) {
            // exact per-iteration max|dphi| of this pass: stop at the first iteration below the tolerance, as the
            // sequential algorithm does (v5.py:224-225).  The pass input is still intact (ping-pong), so a pass that
            // overshoots is redone with exactly the right depth.
            double ch[TB_MAX_DEPTH];
            if (cudaMemcpyAsync(ch, tb->slots, sizeof ch, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) {
                g_tb_err = "reading the convergence slots failed"; return -1;
            }
            int stop = 0;
            for (int t = 0; t < d; ++t) if (ch[t] < tol) { stop = t + 1; break; }
            if (stop) {
                if (stop < d && pass(stop)) return -1;
                void* t2 = cur; cur = nxt; nxt = t2;
                done += stop;
                if (last_change) *last_change = ch[stop - 1];
                break;
            }
            if (last_change) *last_change = ch[d - 1];
        }
        void* t = cur; cur = nxt; nxt = t;
        done += d;
    }
    if (cur != phi && cudaMemcpyAsync(phi, cur, bytes, cudaMemcpyDeviceToDevice, s) != cudaSuccess) { g_tb_err = "copy back"; return -1; }
    *iters_done = done;
    return 0;
}

int poisson_tb_jacobi(PoissonTB** tb, int dtype, const Geom& g, int n_sm, void* phi, const void* div, const uint8_t* mask,
                      double dx2_over_dt, int sweeps, int depth, int* iters_done, cudaStream_t s) {
    (void)n_sm;
    return tb_run(tb, dtype, 0, g, phi, div, mask, dx2_over_dt, 1.0, sweeps, depth, iters_done, s);
}
int poisson_tb_rbsor(PoissonTB** tb, int dtype, const Geom& g, int n_sm, void* phi, const void* div, const uint8_t* mask,
                     double inv_dt, double omega, int iters, int depth, int* iters_done, cudaStream_t s, double tol,
                     double* last_change) {
    (void)n_sm;
    return tb_run(tb, dtype, 1, g, phi, div, mask, -inv_dt, omega, iters, depth, iters_done, s, tol, last_change);
}

}  // namespace b200cfd
