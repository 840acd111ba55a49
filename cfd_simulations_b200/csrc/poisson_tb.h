// poisson_tb.h — temporally blocked (several sweeps per HBM pass) Poisson kernels: TMA-staged tile + halo in
// shared memory.  Each entry runs as many of the requested sweeps as it can in full passes of depth T and
// reports how many it did in *iters_done (the caller finishes the remainder with the one-sweep kernels).
#pragma once
#include "common.cuh"

namespace b200cfd {
struct PoissonTB;
const char* poisson_tb_error();
void poisson_tb_destroy(PoissonTB* tb);
void poisson_tb_info(const PoissonTB* tb, int* depth, int* passes);   // depth and HBM passes of the last solve
// returns 0 on success (possibly with *iters_done == 0 when the shape is not supported), < 0 on CUDA errors
int poisson_tb_jacobi(PoissonTB** tb, int dtype, const Geom& g, int n_sm, void* phi, const void* div, const uint8_t* mask,
                      double dx2_over_dt, int sweeps, int depth, int* iters_done, cudaStream_t s);
int poisson_tb_rbsor(PoissonTB** tb, int dtype, const Geom& g, int n_sm, void* phi, const void* div, const uint8_t* mask,
                     double inv_dt, double omega, int iters, int depth, int* iters_done, cudaStream_t s, double tol = 0.0,
                     double* last_change = nullptr);   // tol > 0: stops at the first iteration with max|dphi| < tol (synchronises per pass)

// ---- poisson_stream.cu: the register-resident streaming pipeline (default); same contract as the entries above ----
struct PoissonStream;
bool poisson_stream_enabled();                       // B200CFD_POISSON_STREAM=0 selects the shared-memory tile kernel above
const char* poisson_stream_error();
void poisson_stream_destroy(PoissonStream* ps);
void poisson_stream_info(const PoissonStream* ps, int* depth, int* passes);
// keep != 0: the following solves use the SAME, unchanged right-hand side and mask arrays as the previous one (a solve done
// in chunks around halo refreshes): the classified copy of the rhs and the tile flags are not rebuilt
void poisson_stream_keep_rhs(PoissonStream** ps, int keep);
void poisson_tb_keep_rhs(PoissonTB** tb, int keep);
// mode 0: Jacobi (scale = dx^2/dt), mode 1: red-black (scale = -1/dt).  tol > 0 (mode 1): device-side stop at exactly the
// iteration the sequential algorithm stops at; the host polls once per batch of passes.
int poisson_stream_run(PoissonStream** ps, int dtype, int mode, const Geom& g, int n_sm, void* phi, const void* div,
                       const uint8_t* mask, double scale, double omega, int iters, int depth, int* iters_done,
                       cudaStream_t s, double tol, double* last_change);
}  // namespace b200cfd
