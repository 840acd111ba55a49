// poisson_stream.cu — temporally blocked pressure-Poisson sweeps as a register-resident streaming pipeline.
// Reference: python/flow_over_cylinder (Fischer)/v5.py:336-346 (Jacobi branch) and :202-226 (red-black Gauss-Seidel;
// omega != 1 gives SOR, the converged-solution counterpart of v2.py:284-306).
//
// One warp owns a 128-column window (4 columns per lane) and streams DOWN a chunk of rows.  Rows of phi and of the
// right-hand side arrive through a per-warp ring of TMA boxes (cp.async.bulk.tensor + mbarrier, 3 rows per slot,
// several slots ahead, zero fill outside the array; fp64 rows land with the 32-byte swizzle so that the lanes' 128-bit row
// loads are free of bank conflicts, see STREAM_SWZ).  Each row then runs through a pipeline of L "levels" held in
// registers: level k computes row q-k of sweep k from the rows q-k-1, q-k, q-k+1 of sweep k-1 (two of them kept in
// registers, the third just produced), x-neighbours by warp shuffle, right-hand side of row q-k re-read from the ring.
// Jacobi: one level per sweep; red-black: two levels (red half-sweep, black half-sweep) per iteration.  The row that
// leaves the last level is staged in shared memory and written back with a TMA store, 3 rows per box.
// One HBM pass = L sweeps (L/2 red-black iterations) at 8 (phi in) + 8 (rhs) + 8 (phi out) B/cell; the only redundancy is
// the 2 x H halo columns of the window (H = L rounded up to 4) and 2 x L rows per row chunk, so unlike the tile kernel
// of poisson_tb.cu (4 shared-memory loads + 1 store per cell and sweep, 1.39 x staged area) depth costs almost nothing:
// shared memory carries 8 B per cell and sweep (the right-hand side), everything else lives in registers.
// Cell classes are data, not coordinates: k_stream_make_rhs folds the cylinder mask AND the non-updated border cells into
// the right-hand side as sentinels (NaN = masked, +Inf = keep), so the kernel has no position tests at all; the selects
// that honour them run only in the 3-row chunks whose levels read one (per-chunk flags, also from k_stream_make_rhs).
// Every tile is resident at once, so a pass lasts as long as its most expensive tile: the tiles come from a cost-balanced
// plan built on the device from those flags (poisson_stream_plan.h, k_stream_plan).
// Tolerance mode (red-black): per-iteration max|dphi| over the owned cells goes to its own slot (exact, order free); the
// last block of a pass decides ON THE DEVICE whether the sequential algorithm would have stopped inside this pass
// (v5.py:224-225) and, if so, with which depth the pass has to be redone from its intact input; every later launch
// exits at once.  The host only polls the flag once per batch of passes.
#include <cstdlib>
#include <string>
#include <type_traits>

#include "poisson_tb.h"
#include "poisson_stream_plan.h"
#include "tma.cuh"

namespace b200cfd {

#define SW 128        // window width = TMA box width
#define SPL 4         // columns per lane
#ifndef SWARPS
#define SWARPS 4      // warps (independent tiles) per block
#endif
#ifndef STREAM_MINB
#define STREAM_MINB 2 // blocks per SM the kernel is compiled for (registers) and the plan fills (SWARPS * STREAM_MINB tiles per SM)
#endif
#define SKR 3         // rows per ring slot = unroll of the row loop
#define FULLM 0xffffffffu
// A/B switches (tools/build_variant.sh): shuffle the north row one row step early (see NbW / NbE) instead of the centre row
// at the point of use, per mode; red-black levels load only the two rhs values they use (two LDS.64) instead of the whole
// 4-cell row.  Measured at 8192^2 (profiles/r04b_alpha.txt): early shuffles cost Jacobi 4 % (964 -> 924 GLUPS, 8 more
// registers) and gain SOR 4 % (461 -> 480); LDS.64 is within the noise (+2 % SOR, -1 % GS).
#ifndef STREAM_EARLY_SHFL_J
#define STREAM_EARLY_SHFL_J 0
#endif
#ifndef STREAM_EARLY_SHFL_RB
#define STREAM_EARLY_SHFL_RB 1
#endif
#ifndef STREAM_STAGE1
#define STREAM_STAGE1 0   // one staging buffer for the output rows instead of two (3 KB less shared memory per warp)
#endif
#ifndef STREAM_ST_SWZ
#define STREAM_ST_SWZ 0   // fp64: the staged output rows are written swizzled too and stored through a swizzled 3-D map
#endif
#ifndef STREAM_PLAN_STRIP_MAJOR
#define STREAM_PLAN_STRIP_MAJOR 0   // tile table ordered strip by strip instead of row-major over (tile-in-strip, strip)
#endif
#ifndef STREAM_SWZ
#define STREAM_SWZ 1      // fp64: phi and rhs rows land in shared memory with TMA's 32-byte swizzle (conflict-free 128-bit row loads)
#endif
#ifndef STREAM_RB_LDS64
#define STREAM_RB_LDS64 0
#endif

struct StreamCtl {            // device-resident control of a tolerance-terminated solve
    int done;                 // tolerance reached: every later launch exits at once
    int redo_depth;           // > 0: the stopping pass must be redone with this many iterations (from its intact input)
    int stop_pass;            // index of the pass in which the solve stopped
    int iters;                // iterations done when it stopped
    double last_change;       // max|dphi| of the last iteration done
    unsigned int blocks_done; // last-block election
    int pad;
};

struct StreamParams {
    int ny, nx;
    int OW, H;                // owned columns per strip, halo columns per side
    int tile_h;               // owned rows per row chunk (multiple of 3)
    int n_strips, n_rchunks;
    int ntile;                // entries of the tile plan (k_stream_plan)
    int NC;                   // chunk flags per strip (global 3-row chunk ids, see k_stream_make_rhs)
    int own0, own1;           // slab-owned local columns (max|dphi| is taken over them only)
    int jg0;                  // global column of local column 0 (colour parity)
    int nact;                 // active levels of this pass (<= L); the rest pass their input through
    int iter0;                // iterations done before this pass (bookkeeping of the stop)
    double ax, ay, dinv, omega, scale, tol;
};

template <typename T> struct SRow;
template <> struct SRow<double> {
    static __device__ __forceinline__ void ld(const double* p, double (&o)[SPL]) {
        const double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
        o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
    }
    // row written by a swizzled TMA box (tma_make_3d_swz32): sw = 2 for the lanes whose 32-byte unit lies in an odd 128-byte
    // line (lane bit 2), where the unit's two 16-byte halves are exchanged; 0 otherwise
    static __device__ __forceinline__ void ld_swz(const double* p, int sw, double (&o)[SPL]) {
        const double2 a = *reinterpret_cast<const double2*>(p + sw), b = *reinterpret_cast<const double2*>(p + 2 - sw);
        o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
    }
    static __device__ __forceinline__ void st(double* p, const double (&x)[SPL]) {
        *reinterpret_cast<double2*>(p) = make_double2(x[0], x[1]);
        *reinterpret_cast<double2*>(p + 2) = make_double2(x[2], x[3]);
    }
    static __device__ __forceinline__ void st_swz(double* p, int sw, const double (&x)[SPL]) {
        *reinterpret_cast<double2*>(p + sw) = make_double2(x[0], x[1]);
        *reinterpret_cast<double2*>(p + 2 - sw) = make_double2(x[2], x[3]);
    }
};
template <> struct SRow<float> {
    static __device__ __forceinline__ void ld(const float* p, double (&o)[SPL]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
    }
    static __device__ __forceinline__ void ld_swz(const float* p, int, double (&o)[SPL]) { ld(p, o); }   // (one 128-bit load per lane: no conflicts)
    static __device__ __forceinline__ void st_swz(float* p, int, const double (&x)[SPL]) { st(p, x); }
    static __device__ __forceinline__ void st(float* p, const double (&x)[SPL]) {
        *reinterpret_cast<float4*>(p) = make_float4((float)x[0], (float)x[1], (float)x[2], (float)x[3]);
    }
};

// Branch-free selects.  nvcc turns a ternary whose arms are long fp64 expressions into a (divergent, per-cell) branch;
// here every arm is needed anyway (the classes are data), so the select is spelled in PTX.
__device__ __forceinline__ double sel_lt(double a, double b, double x, double y) {       // a < b ? x : y   (false on NaN)
    double o;
    asm("{\n .reg .pred p;\n setp.lt.f64 p, %1, %2;\n selp.f64 %0, %3, %4, p;\n}" : "=d"(o) : "d"(a), "d"(b), "d"(x), "d"(y));
    return o;
}
__device__ __forceinline__ double sel_nan(double r, double x, double y) {                 // isnan(r) ? x : y
    double o;
    asm("{\n .reg .pred p;\n setp.nan.f64 p, %1, %1;\n selp.f64 %0, %2, %3, p;\n}" : "=d"(o) : "d"(r), "d"(x), "d"(y));
    return o;
}
__device__ __forceinline__ double sel_ge(int a, int b, double x, double y) {                  // a >= b ? x : y   (integers)
    double o;
    asm("{\n .reg .pred p;\n setp.ge.s32 p, %1, %2;\n selp.f64 %0, %3, %4, p;\n}" : "=d"(o) : "r"(a), "r"(b), "d"(x), "d"(y));
    return o;
}
__device__ __forceinline__ double sel_gt(int a, int b, double x, double y) {                  // a > b ? x : y    (integers)
    double o;
    asm("{\n .reg .pred p;\n setp.gt.s32 p, %1, %2;\n selp.f64 %0, %3, %4, p;\n}" : "=d"(o) : "r"(a), "r"(b), "d"(x), "d"(y));
    return o;
}
__device__ __forceinline__ double sel_int(int c, double x, double y) {                    // c != 0 ? x : y
    double o;
    asm("{\n .reg .pred p;\n setp.ne.s32 p, %1, 0;\n selp.f64 %0, %2, %3, p;\n}" : "=d"(o) : "r"(c), "d"(x), "d"(y));
    return o;
}

// rhs with the cell classes folded in: NaN = masked cell, +Inf = cell that is never updated (array / global border),
// anything else = div.  The scale (Jacobi: dx^2/dt, v5.py:341; red-black: -1/dt, v5.py:217-219) is applied in double
// inside the sweep, as the one-sweep kernels do, so both paths round identically in fp32 storage too.
// cflags[strip * NC + c] = 1 for every 3-row chunk of a strip's window in which some level of the pipeline reads the rhs of a
// special cell: only those chunks run the version of the chunk body with the class selects.  Chunks are numbered globally:
// every tile starts ingesting at q0 = r0 - (L + E) with r0 a multiple of 3, so chunk c always holds the phi rows
// 3c - (L + E) .. + 2, and its levels read the rhs of rows 3c - (L + E) - L .. 3c - (L + E) + 1.  A special cell in row i
// therefore marks the chunks ceil((i - 1 + L + E) / 3) .. floor((i + 2L + E) / 3).  (Plain stores of the same value: no atomics.)
// The flag used to be one per TILE; all tiles are resident at once, so the kernel lasted as long as its slowest warp, and
// a flagged tile (every border tile, the disc) ran 1.7 x the instructions of a plain one over ALL its rows: the r02z capture
// shows the SMs active for only 69 % of the Jacobi kernel's duration.
template <typename T>
__global__ void k_stream_make_rhs(int ny, int nx, long long pitch, int jg0, int nxg, const T* __restrict__ div,
                                  const uint8_t* __restrict__ mask, T* __restrict__ out, uint8_t* __restrict__ cflags,
                                  int OW, int H, int L, int LE, int n_strips, int NC) {
    const T nan_t = (T)__longlong_as_double(0x7ff8000000000000ll), inf_t = (T)__longlong_as_double(0x7ff0000000000000ll);
    // four columns (256 apart: coalesced) per thread and trip, loads before stores: the one-element-per-thread form ran at a
    // third of the HBM rate (476 us for 8192^2, as long as a whole 6-sweep pass)
    for (int i = blockIdx.y; i < ny; i += gridDim.y)
    for (int j0 = blockIdx.x * 1024 + threadIdx.x; j0 < (int)pitch; j0 += gridDim.x * 1024) {
        const long long k0 = (long long)i * pitch + j0;
        T d[4]; bool msk[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const bool in = j0 + 256 * q < (int)pitch;
            d[q] = in ? div[k0 + 256 * q] : (T)0;
            msk[q] = in && mask && mask[k0 + 256 * q];
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = j0 + 256 * q;
            if (j >= (int)pitch) break;
            const int jg = jg0 + j;
            const bool upd = i >= 1 && i < ny - 1 && j >= 1 && j < nx - 1 && jg >= 1 && jg < nxg - 1;
            out[k0 + 256 * q] = !upd ? inf_t : (msk[q] ? nan_t : d[q]);
            if ((!upd || msk[q]) && j < nx) {
                const int s0 = max((j - H) / OW, 0), s1 = min((j + H) / OW, n_strips - 1);
                const int c0 = max((i - 1 + LE + 2) / 3, 0), c1 = min((i + L + LE) / 3, NC - 1);
                for (int c = c0; c <= c1; ++c)
                    for (int s = s0; s <= s1; ++s) cflags[s * NC + c] = 1;
            }
        }
    }
}

// MODE 0: Jacobi, L sweeps per pass.  MODE 1: red-black Gauss-Seidel / SOR, L/2 iterations per pass (L even).
// role 0: normal pass (exits at once when ctl->done); role 1: redo of pass `pass_index` with ctl->redo_depth iterations.
// SOR: omega != 1 (the relaxation step is compiled in only then: omega = 1 must stay the exact Gauss-Seidel update).
// CHECK: accumulate the per-iteration max|dphi| (tolerance mode).
template <typename T, int MODE, int L, int NS, bool SOR, bool CHECK>
__global__ void __launch_bounds__(32 * SWARPS, STREAM_MINB)
k_poisson_stream(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_rhs,
                 const __grid_constant__ CUtensorMap tm_out, StreamParams P, unsigned long long* __restrict__ slots,
                 StreamCtl* __restrict__ ctl, int role, int pass_index, const uint8_t* __restrict__ cflags,
                 const StreamTile* __restrict__ tiles) {
    constexpr int ARR = SKR * SW;                 // elements per slot (3 rows)
    constexpr int KEEP = (L + SKR - 1) / SKR;     // rhs slots behind the current one whose rows are still needed
    constexpr int NSR = NS + KEEP;                // rhs ring: NS slots of lead + the ones the pipeline still reads
    constexpr int NSTG = STREAM_STAGE1 ? 1 : 2;
    constexpr bool SWZ = STREAM_SWZ && sizeof(T) == 8;
    constexpr bool STSWZ = STREAM_ST_SWZ && sizeof(T) == 8;
    constexpr int WARP_ELEMS = (NS + NSR + NSTG) * ARR;   // phi ring, rhs ring, staging buffers of 3 output rows
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    int nact = P.nact;
    if (ctl) {
        const int done = *reinterpret_cast<volatile int*>(&ctl->done);
        if (role == 0) { if (done) return; }
        else {
            const int rd = *reinterpret_cast<volatile int*>(&ctl->redo_depth), sp = *reinterpret_cast<volatile int*>(&ctl->stop_pass);
            if (!done || rd <= 0 || sp != pass_index) return;
            nact = MODE == 1 ? 2 * rd : rd;
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.x * SWARPS + warp;
    StreamTile td;                                                // this warp's tile: strip and row range from the cost-balanced plan
    td.s = 0; td.r0 = 0; td.r1 = 0; td.pad = 0;
    if (tile < P.ntile) td = tiles[tile];
    const bool active = td.r1 > td.r0;
    T* ring_p = reinterpret_cast<T*>(smem_raw) + (size_t)warp * WARP_ELEMS;      // phi ring: NS slots
    T* ring_r = ring_p + NS * ARR;                                               // rhs ring: NSR slots
    T* stage = ring_r + NSR * ARR;                                               // 2 x 3 output rows (dense, OW per row)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw + (size_t)SWARPS * WARP_ELEMS * sizeof(T)) + warp * NS;
    double ch[MODE == 1 ? L / 2 : 1];
#pragma unroll
    for (int t = 0; t < (MODE == 1 ? L / 2 : 1); ++t) ch[t] = 0.0;
    if (active) {
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < NS; ++k) mbar_init(&bars[k], 1);
            mbar_fence_init();
        }
        __syncwarp();
        const int s = td.s;
        const int x0 = s * P.OW;                                  // first owned column
        const int jw0 = x0 - P.H;                                 // window origin (multiple of 4; may be negative)
        const int c0 = jw0 + SPL * lane;
        const int r0 = td.r0, r1 = td.r1;                        // r0 is a multiple of 3; r1 is one or ny
        // ingest rows q0 .. ; the row leaving level L at ingest step q is q - L.  q0 is chosen so that (i) level L's first
        // valid row (q0 + L) is <= r0 and (ii) row r0 leaves the pipeline (at ingest step r0 + L) as the FIRST row of a chunk
        // (stores are whole 3-row boxes): q0 = r0 - L - e with (2L + e) % 3 == 0
        constexpr int E = (3 - (2 * L) % 3) % 3;
        const int q0 = r0 - L - E;
        const int NCH = (r1 + L - q0 + SKR - 1) / SKR;            // chunks until row r1 - 1 has left the pipeline
        // chunk k: phi rows into slot k % NS of the phi ring, rhs rows into slot k % NSR of the rhs ring (one mbarrier per
        // phi slot covers both boxes)
        auto issue = [&](int k) {
            const int slot = k % NS;
            mbar_expect_tx(&bars[slot], (uint32_t)(2 * ARR * sizeof(T)));
            if (SWZ) {                                            // 3-D swizzled view: (0, x / 4, y); jw0 is a multiple of 4
                tma_load_3d(ring_p + slot * ARR, &tm_in, 0, jw0 / 4, q0 + SKR * k, &bars[slot]);
                tma_load_3d(ring_r + (k % NSR) * ARR, &tm_rhs, 0, jw0 / 4, q0 + SKR * k, &bars[slot]);
            } else {
                tma_load_2d(ring_p + slot * ARR, &tm_in, jw0, q0 + SKR * k, &bars[slot]);
                tma_load_2d(ring_r + (k % NSR) * ARR, &tm_rhs, jw0, q0 + SKR * k, &bars[slot]);
            }
        };
        if (lane == 0) {
            const int n0 = NCH < NS ? NCH : NS;
            for (int k = 0; k < n0; ++k) issue(k);
        }
        // column flags as multiplicative masks: owned by this tile AND by this slab (max|dphi| only)
        double ownm[SPL];
#pragma unroll
        for (int m = 0; m < SPL; ++m) {
            const int c = c0 + m;
            ownm[m] = (c >= x0 && c < x0 + P.OW && c >= P.own0 && c < P.own1 && c < P.nx) ? 1.0 : 0.0;
        }
        // colour of this lane's columns: c0 is a multiple of 4, so cells 0, 2 have the parity of (row + jg0), cells 1, 3 the other
        const int colpar = P.jg0 & 1;
        const double ax = P.ax, ay = P.ay, dinv = P.dinv, omega = P.omega, nscale = -P.scale;
        const double INF = __longlong_as_double(0x7ff0000000000000ll);

        // pipeline registers: R[k] = the two newest rows that entered level k+1 (k = 0: phi as loaded)
        double R[L][3][SPL];
#pragma unroll
        for (int k = 0; k < L; ++k)
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int m = 0; m < SPL; ++m) R[k][j][m] = 0.0;

        // x-neighbours of each level's CENTRE row, optionally shuffled one row step early (STREAM_EARLY_SHFL_*): a level's north
        // row becomes its centre row in the next row step, so the level can shuffle the end cells of its north row now and use
        // the result a whole row step later.  Same values, same operations: results are bit-identical.  It pays for the
        // red-black levels only (see the switches at the top); the r04a capture has 2.5 of the 5.3 stall cycles per
        // instruction on the short scoreboard, but the SASS already had 17+ instructions between a shuffle and its first use.
        // (Jacobi needs both neighbours; a red-black level updates cells 0, 2 or cells 1, 3 and needs only the west or
        // only the east one, and the parity alternates from row to row, so it keeps the one the NEXT row step uses.)
        double NbW[L], NbE[L];
#pragma unroll
        for (int k = 0; k < L; ++k) { NbW[k] = 0.0; NbE[k] = 0.0; }

        const int scol = SPL * lane - P.H;                          // this lane's first column inside the staged (owned) row
        const bool st_lane = scol >= 0 && scol + SPL <= P.OW;
        // swizzled staging (STSWZ): row j of the dense OW-column box starts at byte 8 OW j of a 256-byte aligned buffer; this
        // lane's 32-byte unit has its halves exchanged when bit 7 of its byte offset is set
        int stsw[SKR];
#pragma unroll
        for (int j = 0; j < SKR; ++j) stsw[j] = (STSWZ && st_lane) ? (((8 * P.OW * j + 8 * scol) >> 7) & 1) * 2 : 0;
        const T* ring_rl = ring_r + SPL * lane;
        const int swz = SWZ ? ((lane >> 2) & 1) * 2 : 0;            // element offset of this lane's first 16-byte half (SRow::ld_swz)
        // folded coefficients (the same in both code paths below, so a cell's value does not depend on which tile - plain or
        // special - computed it: redundant halo copies stay bit-identical to the owner's value)
        const double q4 = 0.25 * nscale, axd = ax * dinv, ayd = ay * dinv, nsd = nscale * dinv;

        // The chunk body exists in three versions, chosen per 3-row chunk (warp-uniform): VAR 0, plain chunks (almost all of
        // them), pure arithmetic; VAR 1, chunks in which some level reads the rhs of a masked or never-updated cell, with the
        // class selects; VAR 2, passes with inactive levels (the last, shallower pass of a solve and redo passes), with the
        // class selects and the pass-through of the inactive levels.  All three compute an updated cell's value with the same
        // operations in the same order, so a value does not depend on which version produced it.
        // red-black: which cells of a row have the colour of a level is a compile-time fact once the parity PB of
        // (first row of the chunk + colour offset of the window) is fixed; the chunk body exists in both parities and a
        // warp-uniform branch picks one per chunk (a branch per LEVEL gets if-converted: both arms issue, predicated)
        // one level: row `row` of sweep k+1 from rows S (row-1), C (row), N (row+1) of sweep k and the rhs of `row`.
        // Straight-line code: the shuffles (of the NORTH row, for the next row step: see NbW / NbE) are unconditional.
        auto level = [&](auto pb_tag, auto var_tag, int k, int jrow, int row, const double (&S)[SPL], const double (&C)[SPL],
                         const double (&N)[SPL], const T* rhs_row, double (&out)[SPL]) {
            constexpr int PB = decltype(pb_tag)::value;
            constexpr int VAR = decltype(var_tag)::value;
            constexpr int SENT = 0x7ff00000;                        // high word of +Inf.  The sentinels are +Inf and the positive quiet NaN, so a
            // SIGNED compare of the high word tells them from every finite value (negative values are negative integers)
            double r[SPL] = {0.0, 0.0, 0.0, 0.0};
            if (MODE == 0 || !STREAM_RB_LDS64) SRow<T>::ld_swz(rhs_row, swz, r);
            const int act = k < nact;                               // warp-uniform (always true unless VAR == 2)
            if (MODE == 0) {
#if !STREAM_EARLY_SHFL_J
                const double Wn = __shfl_up_sync(FULLM, C[3], 1), En = __shfl_down_sync(FULLM, C[0], 1);
#else
                const double Wn = NbW[k], En = NbE[k];             // neighbours of C, shuffled during the previous row step
                NbW[k] = __shfl_up_sync(FULLM, N[3], 1);           // neighbours of N = the next row step's C
                NbE[k] = __shfl_down_sync(FULLM, N[0], 1);
#endif
#pragma unroll
                for (int m = 0; m < SPL; ++m) {
                    const double ns = N[m] + S[m], rq = r[m] * q4;
                    const double e = m < SPL - 1 ? C[(m + 1) % SPL] : En, w = m > 0 ? C[(m + SPL - 1) % SPL] : Wn;
                    const double v = Rnd<T>::r(fma(0.25, (e + w) + ns, rq));
                    if (VAR == 2) {
                        // updated cell -> v, masked (NaN) -> 0 (v5.py:345), never-updated border (+Inf) -> kept
                        const double o = sel_lt(fabs(r[m]), INF, v, sel_nan(r[m], 0.0, C[m]));
                        out[m] = sel_int(act, o, C[m]);
                    } else if (VAR == 1) {
                        // the same classes, told apart on the high word with integer compares (off the fp64 pipe and ready
                        // long before v).  Level 1 zeroes the masked cells; in the levels behind it a masked cell already
                        // holds 0, so "keep" covers both classes there.
                        const int a = __double2hiint(r[m]);
                        out[m] = k == 0 ? sel_ge(a, SENT, sel_gt(a, SENT, 0.0, C[m]), v) : sel_ge(a, SENT, C[m], v);
                    } else out[m] = v;
                }
            } else {
                // colour of level k: even k = first colour ((i + j) odd, v5.py:215), odd k = second.  c0 is a multiple
                // of 4, so cells 0, 2 have the parity of (row + jg0) and cells 1, 3 the other one (warp-uniform branch).
                // row = q + jrow with (q + colpar) & 1 == PB, jrow = j - 1 - k a compile-time offset
                const int want = (k & 1) ? 0 : 1;
                const bool even_cells = (((PB + jrow) % 2 + 2) % 2) == want;       // compile time
                const int rown = row >= r0 && row < r1;
                // this row step uses the west neighbour (cells 0, 2) or the east one (cells 1, 3); the next one the other
#if !STREAM_EARLY_SHFL_RB
                const double Wn = __shfl_up_sync(FULLM, C[3], 1), En = __shfl_down_sync(FULLM, C[0], 1);
#else
                const double Wn = NbW[k], En = NbE[k];
                if (even_cells) NbE[k] = __shfl_down_sync(FULLM, N[0], 1);
                else NbW[k] = __shfl_up_sync(FULLM, N[3], 1);
#endif
                if (STREAM_RB_LDS64) {
                    if (even_cells) { r[0] = (double)rhs_row[0 ^ swz]; r[2] = (double)rhs_row[2 ^ swz]; }
                    else { r[1] = (double)rhs_row[1 ^ swz]; r[3] = (double)rhs_row[3 ^ swz]; }
                }
                auto cell = [&](int m, double e, double w) {
                    const double t = fma(ayd, N[m] + S[m], nsd * r[m]);
                    double pn = fma(axd, e + w, t);
                    if (SOR) pn = fma(omega, pn - C[m], C[m]);
                    pn = Rnd<T>::r(pn);
                    if (VAR == 2) pn = sel_int(act, sel_lt(fabs(r[m]), INF, pn, C[m]), C[m]);   // masked / border / inactive: kept
                    else if (VAR == 1) pn = sel_ge(__double2hiint(r[m]), SENT, C[m], pn);   // masked / border: kept
                    out[m] = pn;
                    // branch-free on purpose: a per-lane `if (owned)` here diverges inside the pipeline body and cost 25% of the
                    // tolerance-mode rate when tried (229 -> 172 GLUPS at 8192^2)
                    // (and a compare-and-select instead of fmax(): the library form quiets NaNs in 7-8 instructions; a NaN
                    // candidate loses the comparison, which is the same outcome)
                    if (CHECK) {
                        const double cand = fabs(pn - C[m]) * ownm[m];
                        ch[k / 2] = (rown && cand > ch[k / 2]) ? cand : ch[k / 2];
                    }
                };
                if (even_cells) { out[1] = C[1]; out[3] = C[3]; cell(0, C[1], Wn); cell(2, C[3], C[1]); }
                else { out[0] = C[0]; out[2] = C[2]; cell(1, C[2], C[0]); cell(3, En, C[2]); }
            }
        };
        const bool inactive = nact < L;
        const uint8_t* cf = cflags + (size_t)s * P.NC + r0 / SKR;   // chunk k of this tile is global chunk r0 / 3 + k
        int cf_cur = cf[0];
        int nstore = 0;
        for (int k = 0; k < NCH; ++k) {
            const int slot = k % NS;
            const int cf_next = k + 1 < NCH ? (int)cf[k + 1] : 0;        // a chunk ahead: the load has a whole chunk to arrive
            mbar_wait(&bars[slot], (uint32_t)((k / NS) & 1));
            const T* sp = ring_p + slot * ARR + SPL * lane;
            const int q = q0 + SKR * k;
            T* sg = stage + (NSTG == 2 ? (nstore & 1) : 0) * ARR;
            // rhs rows of the chunks k, k-1, ... (this lane's columns): the level that computes row q + j - 1 - lv reads
            // row (j - 1 - lv) mod 3 of chunk k + floor((j - 1 - lv) / 3), all compile-time offsets
            const T* rb[KEEP + 2];
            {
                const int km = k % NSR;
#pragma unroll
                for (int o = 0; o < KEEP + 2; ++o) { int sl = km - o; if (sl < 0) sl += NSR; rb[o] = ring_rl + sl * ARR; }
            }
            const int orow = q - L;                                  // first row leaving the pipeline in this chunk
            const bool do_store = orow >= r0 && orow < r1;          // whole chunks by construction
            if (NSTG == 2 && do_store && nstore >= 2) {              // the staging buffer's previous TMA store has read it
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
            }
            auto body = [&](auto pb_tag, auto var_tag) {
#pragma unroll
                for (int j = 0; j < SKR; ++j) {
                    // ring row j of this slot = phi row q + j enters level 1 as its north row
                    SRow<T>::ld_swz(sp + j * SW, swz, R[0][(j + 2) % 3]);
#pragma unroll
                    for (int lv = 0; lv < L; ++lv) {
                        const int row = q + j - 1 - lv;                  // row computed by level lv + 1
                        const int off = j - 1 - lv;                      // compile time after unrolling
                        const int back = off >= 0 ? 0 : (-off + SKR - 1) / SKR, jr = off + back * SKR;
                        if (lv + 1 < L) {
                            // the level writes straight into the next level's north row (no copy: the ncu source page
                            // of the first version showed a quarter of the issue slots going to register moves)
                            level(pb_tag, var_tag, lv, off, row, R[lv][j % 3], R[lv][(j + 1) % 3], R[lv][(j + 2) % 3], rb[back] + jr * SW,
                                  R[lv + 1 < L ? lv + 1 : lv][(j + 2) % 3]);
                        } else {
                            double out[SPL];
                            level(pb_tag, var_tag, lv, off, row, R[lv][j % 3], R[lv][(j + 1) % 3], R[lv][(j + 2) % 3], rb[back] + jr * SW, out);
                            if (NSTG == 1 && j == 0 && do_store && nstore >= 1) {   // single buffer: the previous chunk's store has read it
                                if (lane == 0) tma_store_wait_read<0>();             // (issued a third of a chunk ago)
                                __syncwarp();
                            }
                            if (do_store && st_lane) SRow<T>::st_swz(sg + j * P.OW + scol, stsw[j], out);
                        }
                    }
                }
            };
            auto variants = [&](auto pb_tag) {
                if (inactive) body(pb_tag, std::integral_constant<int, 2>{});
                else if (cf_cur) body(pb_tag, std::integral_constant<int, 1>{});
                else body(pb_tag, std::integral_constant<int, 0>{});
            };
            if (MODE == 1 && ((q + colpar) & 1)) variants(std::integral_constant<int, 1>{});
            else variants(std::integral_constant<int, 0>{});
            cf_cur = cf_next;
            if (do_store) {
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if (STSWZ) tma_store_3d(&tm_out, 0, x0 / 4, orow, sg);
                    else tma_store_2d(&tm_out, x0, orow, sg);        // 3 dense rows of the OW owned columns
                    tma_store_commit();
                }
                ++nstore;
            }
            __syncwarp();
            // the phi slot of chunk k is free now; the rhs of chunk k + NS overwrites that of chunk k - KEEP, which no
            // level reads any more (level L of the next chunk reaches back to chunk k + 1 - KEEP)
            if (lane == 0 && k + NS < NCH) issue(k + NS);
        }
        if (lane == 0) tma_store_wait<0>();
    }
    // ---- per-iteration max|dphi| over the owned cells -> slots; the last block closes the pass
    if (MODE == 1 && CHECK) {
        __shared__ double s_red[SWARPS][L / 2];
#pragma unroll
        for (int t = 0; t < L / 2; ++t) {
            const double w = warp_max(ch[t]);
            if (lane == 0) s_red[warp][t] = w;
        }
        __syncthreads();
        if (threadIdx.x < L / 2) {
            double w = 0.0;
            for (int k = 0; k < SWARPS; ++k) w = fmax(w, s_red[k][threadIdx.x]);
            if (w > 0.0) atomicMax(&slots[threadIdx.x], (unsigned long long)__double_as_longlong(w));
        }
        if (ctl && role == 0) {
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) {
                const unsigned int prev = atomicAdd(&ctl->blocks_done, 1u);
                if (prev == gridDim.x - 1) {
                    __threadfence();
                    const int niter = nact / 2;
                    int stop = 0;
                    double last = 0.0;
                    for (int t = 0; t < niter; ++t) {
                        last = __longlong_as_double((long long)atomicAdd(&slots[t], 0ull));
                        if (last < P.tol) { stop = t + 1; break; }
                    }
                    ctl->last_change = last;
                    if (stop) {
                        ctl->iters = P.iter0 + stop;
                        ctl->stop_pass = pass_index;
                        ctl->redo_depth = stop < niter ? stop : 0;
                        __threadfence();
                        ctl->done = 1;
                    } else {
                        ctl->iters = P.iter0 + niter;
                    }
                    for (int t = 0; t < L / 2; ++t) slots[t] = 0ull;
                    ctl->blocks_done = 0u;
                }
            }
        }
    }
}

// The tile plan (poisson_stream_plan.h): one block, the items of every phase spread over its threads.
// seg: [n_strips][nseg] segment sums -> exclusive prefixes; cost, n, first: per strip.
__global__ void __launch_bounds__(1024)
k_stream_plan(StreamPlanArgs a, int* __restrict__ seg, int* __restrict__ cost, int* __restrict__ n, int* __restrict__ first,
              StreamTile* __restrict__ tiles) {
    __shared__ int s_lo, s_hi, s_sum;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int nseg = plan_nseg(a), cap = plan_cap(a);
    for (int it = tid; it < a.n_strips * nseg; it += nt) seg[it] = plan_seg_sum(a, it / nseg, it % nseg);
    if (tid == 0) { s_lo = 1; s_hi = 1; }
    __syncthreads();
    for (int s = tid; s < a.n_strips; s += nt) {
        const int c = plan_strip_prefix(a, seg + (size_t)s * nseg);
        cost[s] = c;
        atomicMax(&s_hi, c);
    }
    __syncthreads();
    // the smallest T with sum_s tiles_at(cost[s], T) <= ntile (bisection; every strip evaluates its own term)
    while (true) {
        const int lo = s_lo, hi = s_hi;
        if (lo >= hi) break;
        const int mid = lo + (hi - lo) / 2;
        __syncthreads();
        if (tid == 0) s_sum = 0;
        __syncthreads();
        int local = 0;
        for (int s = tid; s < a.n_strips; s += nt) local += plan_tiles_at(cost[s], mid, cap);
        if (local) atomicAdd(&s_sum, local);
        __syncthreads();
        if (tid == 0) { if (s_sum <= a.ntile) s_hi = mid; else s_lo = mid + 1; }
        __syncthreads();
    }
    const int T = s_lo;
    for (int s = tid; s < a.n_strips; s += nt) n[s] = plan_tiles_at(cost[s], T, cap);
    __syncthreads();
    if (tid == 0) { int f = 0; for (int s = 0; s < a.n_strips; ++s) { first[s] = f; f += n[s]; } }
    __syncthreads();
    // one warp per strip, one lane per tile
    for (int s = tid / 32; s < a.n_strips; s += nt / 32)
        for (int j = tid % 32; j < n[s]; j += 32) {
            StreamTile t;
            t.s = s; t.pad = 0;
            t.r0 = plan_cut_row(a, s, seg + (size_t)s * nseg, cost[s], n[s], j);
            t.r1 = plan_cut_row(a, s, seg + (size_t)s * nseg, cost[s], n[s], j + 1);
#if STREAM_PLAN_STRIP_MAJOR
            const int idx = first[s] + j;
#else
            const int idx = plan_tile_index(n, a.n_strips, s, j);
#endif
            if (idx < a.ntile) tiles[idx] = t;
        }
}

// ---- host side -----------------------------------------------------------------------------------------------------
struct PoissonStream {
    void* rhs_s = nullptr;
    void* alt = nullptr;
    size_t bytes = 0;
    unsigned long long* slots = nullptr;   // 8 entries
    StreamCtl* ctl = nullptr;
    uint8_t* cflags = nullptr;             // one byte per (strip, global 3-row chunk): its levels read masked / never-updated cells
    size_t cflags_cap = 0;
    StreamTile* tiles = nullptr;           // the cost-balanced tile plan built from cflags
    size_t tiles_cap = 0;
    int* plan_i = nullptr;                 // plan scratch: per strip cost, tiles, first tile; then [n_strips][nseg] segment sums
    size_t plan_cap = 0;
    int plan_ntile = 0, plan_alpha = -1;
    // what rhs_s / cflags were built from: a caller that solves in chunks on an UNCHANGED right-hand side (the slab solver:
    // a few iterations, halo refresh, a few more) declares it with poisson_stream_keep_rhs and skips the rebuild
    const void* rhs_src = nullptr; const void* mask_src = nullptr;
    int rhs_ny = 0, rhs_nx = 0, rhs_L = 0, rhs_dtype = -1;
    bool keep_rhs = false;
    int last_depth = 0, last_passes = 0;
};
static std::string g_stream_err;
const char* poisson_stream_error() { return g_stream_err.c_str(); }
void poisson_stream_destroy(PoissonStream* ps) {
    if (!ps) return;
    cudaFree(ps->rhs_s); cudaFree(ps->alt); cudaFree(ps->slots); cudaFree(ps->ctl); cudaFree(ps->cflags); cudaFree(ps->tiles); cudaFree(ps->plan_i);
    delete ps;
}
void poisson_stream_keep_rhs(PoissonStream** pps, int keep) {
    if (!*pps) *pps = new PoissonStream();
    (*pps)->keep_rhs = keep != 0;
}
void poisson_stream_info(const PoissonStream* ps, int* depth, int* passes) {
    *depth = ps ? ps->last_depth : 0;
    *passes = ps ? ps->last_passes : 0;
}

static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

template <typename T, int MODE, int L, int NS, bool SOR, bool CHECK>
static int stream_launch(const CUtensorMap& mi, const CUtensorMap& mr, const CUtensorMap& mo, const StreamParams& P,
                         unsigned long long* slots, StreamCtl* ctl, int role, int pass_index, const uint8_t* cflags, const StreamTile* tiles, cudaStream_t s) {
    constexpr int KEEP = (L + SKR - 1) / SKR;
    const size_t smem = (size_t)SWARPS * (NS + NS + KEEP + (STREAM_STAGE1 ? 1 : 2)) * SKR * SW * sizeof(T) + SWARPS * NS * sizeof(uint64_t);
    auto kern = k_poisson_stream<T, MODE, L, NS, SOR, CHECK>;
    static bool configured = false;
    if (!configured) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            cudaGetLastError(); g_stream_err = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)"; return -1;
        }
        // STREAM_MINB blocks per SM only fit with the largest shared-memory carve-out
        if (cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared) != cudaSuccess) cudaGetLastError();
        configured = true;
    }
    const int ntile = P.ntile;
    kern<<<(ntile + SWARPS - 1) / SWARPS, 32 * SWARPS, smem, s>>>(mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles);
    if (cudaPeekAtLastError() != cudaSuccess) { g_stream_err = cudaGetErrorString(cudaGetLastError()); return -1; }
    return 0;
}

// pipeline depths instantiated: 2 and 6 levels (a shallower pass runs in the deeper pipeline with inactive levels)
#ifndef STREAM_NS
#define STREAM_NS 2
#endif
template <typename T, int MODE, bool SOR, bool CHECK>
static int stream_depth(int L, const CUtensorMap& mi, const CUtensorMap& mr, const CUtensorMap& mo, const StreamParams& P,
                        unsigned long long* slots, StreamCtl* ctl, int role, int pass_index, const uint8_t* cflags, const StreamTile* tiles, cudaStream_t s) {
    if (L <= 2) return stream_launch<T, MODE, 2, STREAM_NS, SOR, CHECK>(mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s);
    return stream_launch<T, MODE, 6, STREAM_NS, SOR, CHECK>(mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s);
}
template <typename T, int MODE>
static int stream_dispatch(int L, const CUtensorMap& mi, const CUtensorMap& mr, const CUtensorMap& mo, const StreamParams& P,
                           unsigned long long* slots, StreamCtl* ctl, int role, int pass_index, const uint8_t* cflags, const StreamTile* tiles, cudaStream_t s) {
    if constexpr (MODE == 0) {
        return stream_depth<T, MODE, false, false>(L, mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s);
    } else {
        const bool sor = P.omega != 1.0, chk = slots != nullptr;
        if (sor) return chk ? stream_depth<T, MODE, true, true>(L, mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s)
                            : stream_depth<T, MODE, true, false>(L, mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s);
        return chk ? stream_depth<T, MODE, false, true>(L, mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s)
                   : stream_depth<T, MODE, false, false>(L, mi, mr, mo, P, slots, ctl, role, pass_index, cflags, tiles, s);
    }
}

bool poisson_stream_enabled() {
    static const int on = env_int("B200CFD_POISSON_STREAM", 1);
    return on != 0 && tma_encoder() != nullptr;
}

// Runs `iters` sweeps (mode 0) / iterations (mode 1) on (phi, div, mask).  depth: iterations or sweeps per pass (0 = auto).
// tol > 0 (mode 1): stop at the first iteration with max|dphi| < tol, exactly as the sequential algorithm; the host polls
// the device flag once per `poll` passes.  Returns 0 and *iters_done = 0 when the shape is not supported.
int poisson_stream_run(PoissonStream** pps, int dtype, int mode, const Geom& g, int n_sm, void* phi, const void* div,
                       const uint8_t* mask, double scale, double omega, int iters, int depth, int* iters_done,
                       cudaStream_t s, double tol, double* last_change) {
    *iters_done = 0;
    const int es = dtype == 0 ? 8 : 4;
    if (iters <= 0) return 0;
    if (((uintptr_t)phi % 16) || ((g.pitch * es) % 16) || g.nx < 64 || g.ny < 16 || g.nx % 4) return 0;
    if (!*pps) *pps = new PoissonStream();
    PoissonStream* ps = *pps;
    if (!ps->slots) {
        if (cudaMalloc(&ps->slots, 8 * sizeof(unsigned long long)) != cudaSuccess || cudaMalloc(&ps->ctl, sizeof(StreamCtl)) != cudaSuccess) {
            cudaGetLastError(); g_stream_err = "cudaMalloc(control)"; return -1;
        }
    }
    const size_t bytes = (size_t)g.ny * g.pitch * es;
    if (ps->bytes < bytes) {
        cudaFree(ps->rhs_s); cudaFree(ps->alt); ps->rhs_s = ps->alt = nullptr; ps->bytes = 0; ps->rhs_src = nullptr;
        if (cudaMalloc(&ps->rhs_s, bytes) != cudaSuccess || cudaMalloc(&ps->alt, bytes) != cudaSuccess) {
            cudaGetLastError(); g_stream_err = "cudaMalloc of the temporal-blocking scratch failed"; return -1;
        }
        ps->bytes = bytes;
    }
    // levels per pass
    int per = depth > 0 ? depth : (mode == 0 ? 6 : 3);             // sweeps / iterations per pass
    if (mode == 0) { if (per > 6) per = 6; }
    else if (per > 3) per = 3;
    if (per > iters) per = iters;
    int L = mode == 0 ? per : 2 * per;
    L = L <= 2 ? 2 : 6;                                             // instantiated pipeline depths
    const int cap = mode == 0 ? L : L / 2;                          // sweeps / iterations one pass of this depth can hold
    StreamParams P;
    P.ny = g.ny; P.nx = g.nx; P.own0 = g.own0; P.own1 = g.own1; P.jg0 = g.jg0;
    P.ax = g.ax; P.ay = g.ay; P.dinv = g.dinv; P.omega = omega; P.scale = scale; P.tol = tol;
    P.H = (L + 3) / 4 * 4;
    P.OW = SW - 2 * P.H;
    P.n_strips = (g.nx + P.OW - 1) / P.OW;
    // tiles: every tile resident at once (one warp per tile), 8 warps per SM; the plan kernel below cuts the strips into
    // ntile row ranges of equal cost (at least 4 groups of 3 rows each)
    const int want = env_int("B200CFD_STREAM_WARPS_PER_SM", SWARPS * STREAM_MINB) * n_sm;
    const int G = (g.ny + SKR - 1) / SKR, min_groups = 4;
    int per_strip_max = G / min_groups;                            // (the plan's own cap, which knows alpha, is tighter)
    if (per_strip_max < 1) per_strip_max = 1;
    long long nt = want;
    if (nt > (long long)P.n_strips * per_strip_max) nt = (long long)P.n_strips * per_strip_max;
    if (nt < P.n_strips) nt = P.n_strips;
    P.ntile = (int)nt;
    P.n_rchunks = (P.ntile + P.n_strips - 1) / P.n_strips;        // informational: average tiles per strip
    P.tile_h = (g.ny + P.n_rchunks - 1) / P.n_rchunks;             // informational
    const long long n = (long long)g.ny * g.pitch;
    const int nb = (int)((n + 255) / 256 < 148 * 32 ? (n + 255) / 256 : 148 * 32);
    const int E = (3 - (2 * L) % 3) % 3;                            // the kernel's E: every tile starts ingesting at r0 - (L + E)
    P.NC = (g.ny + 2 * L + E) / SKR + 2;                            // global chunk ids 0 .. NC - 1 cover every tile's last chunk
    const size_t nflag = (size_t)P.n_strips * P.NC;
    if (ps->cflags_cap < nflag) {
        cudaFree(ps->cflags); ps->cflags = nullptr; ps->cflags_cap = 0; ps->rhs_src = nullptr;
        if (cudaMalloc(&ps->cflags, nflag) != cudaSuccess) { cudaGetLastError(); g_stream_err = "cudaMalloc(chunk flags)"; return -1; }
        ps->cflags_cap = nflag;
    }
    // (the flags depend on the strip geometry and the chunk phase, both functions of L, not on the row chunking)
    if (ps->tiles_cap < (size_t)P.ntile) {
        cudaFree(ps->tiles); ps->tiles = nullptr; ps->tiles_cap = 0; ps->rhs_src = nullptr;
        if (cudaMalloc(&ps->tiles, (size_t)P.ntile * sizeof(StreamTile)) != cudaSuccess) { cudaGetLastError(); g_stream_err = "cudaMalloc(tile plan)"; return -1; }
        ps->tiles_cap = (size_t)P.ntile;
    }
    const size_t plan_ints = (size_t)P.n_strips * (3 + (size_t)((G + PLAN_SEG - 1) / PLAN_SEG));
    if (ps->plan_cap < plan_ints) {
        cudaFree(ps->plan_i); ps->plan_i = nullptr; ps->plan_cap = 0; ps->rhs_src = nullptr;
        if (cudaMalloc(&ps->plan_i, plan_ints * sizeof(int)) != cudaSuccess) { cudaGetLastError(); g_stream_err = "cudaMalloc(plan scratch)"; return -1; }
        ps->plan_cap = plan_ints;
    }
    // extra cost of a chunk with class selects, percent of a plain chunk.  Measured, not derived from instruction counts (the
    // select body has 1.4 x the instructions): at 8192^2 Jacobi gives 814 / 867 / 924 GLUPS at 0 / 35 / 50 before the swizzle
    // and 1091 / 1211 / 1217 / 1139 at 50 / 100 / 150 / 250 after it; red-black is flat from 35 to 150
    // (profiles/r04b_alpha.txt, r04d_alpha.txt, r04e_alpha.txt; tools/alpha_sweep.py)
    int alpha = env_int("B200CFD_STREAM_ALPHA", 100);
    if (alpha < 0) alpha = 0;
    if (alpha > 400) alpha = 400;
    const bool reuse = ps->keep_rhs && ps->rhs_src == div && ps->mask_src == (const void*)mask && ps->rhs_ny == g.ny &&
                       ps->rhs_nx == g.nx && ps->rhs_L == L && ps->rhs_dtype == dtype && ps->plan_ntile == P.ntile && ps->plan_alpha == alpha;
    const dim3 gnb((unsigned)((g.pitch + 1023) / 1024), (unsigned)(g.ny < 2048 ? g.ny : 2048));
    (void)nb;
    if (reuse) { /* rhs_s and cflags still describe this right-hand side */ }
    else if (cudaMemsetAsync(ps->cflags, 0, nflag, s) != cudaSuccess) { g_stream_err = "cudaMemsetAsync(chunk flags)"; return -1; }
    else if (dtype == 0) k_stream_make_rhs<double><<<gnb, 256, 0, s>>>(g.ny, g.nx, g.pitch, g.jg0, g.nxg, (const double*)div, mask, (double*)ps->rhs_s,
                                                                  ps->cflags, P.OW, P.H, L, L + E, P.n_strips, P.NC);
    else k_stream_make_rhs<float><<<gnb, 256, 0, s>>>(g.ny, g.nx, g.pitch, g.jg0, g.nxg, (const float*)div, mask, (float*)ps->rhs_s,
                                                    ps->cflags, P.OW, P.H, L, L + E, P.n_strips, P.NC);
    if (!reuse) {
        StreamPlanArgs pa;
        pa.cflags = ps->cflags; pa.n_strips = P.n_strips; pa.NC = P.NC; pa.ny = g.ny; pa.shift = (2 * L + E) / SKR;
        pa.ntile = P.ntile; pa.alpha_pct = alpha; pa.min_groups = min_groups;
        if (cudaMemsetAsync(ps->tiles, 0, (size_t)P.ntile * sizeof(StreamTile), s) != cudaSuccess) { g_stream_err = "cudaMemsetAsync(tile plan)"; return -1; }
        k_stream_plan<<<1, 1024, 0, s>>>(pa, ps->plan_i + 3 * (size_t)P.n_strips, ps->plan_i, ps->plan_i + P.n_strips, ps->plan_i + 2 * (size_t)P.n_strips,
                                         ps->tiles);
        if (cudaPeekAtLastError() != cudaSuccess) { g_stream_err = cudaGetErrorString(cudaGetLastError()); return -1; }
    }
    ps->rhs_src = div; ps->mask_src = mask; ps->rhs_ny = g.ny; ps->rhs_nx = g.nx; ps->rhs_L = L; ps->rhs_dtype = dtype;
    ps->plan_ntile = P.ntile; ps->plan_alpha = alpha;
    CUtensorMap m_phi, m_alt, m_rhs, o_phi, o_alt;
    const bool swz = STREAM_SWZ && dtype == 0;                      // fp64 rows land swizzled (tma_make_3d_swz32)
    auto make_in = [&](CUtensorMap* m, const void* base) {
        return swz ? tma_make_3d_swz32(m, base, g.nx, g.ny, g.pitch, SW, SKR) : tma_make_2d(m, dtype, base, g.nx, g.ny, g.pitch, SW, SKR);
    };
    if (!make_in(&m_phi, phi) || !make_in(&m_alt, ps->alt) || !make_in(&m_rhs, ps->rhs_s) ||
        !((STREAM_ST_SWZ && dtype == 0) ? tma_make_3d_swz32(&o_phi, phi, g.nx, g.ny, g.pitch, P.OW, SKR) : tma_make_2d(&o_phi, dtype, phi, g.nx, g.ny, g.pitch, P.OW, SKR)) ||
        !((STREAM_ST_SWZ && dtype == 0) ? tma_make_3d_swz32(&o_alt, ps->alt, g.nx, g.ny, g.pitch, P.OW, SKR) : tma_make_2d(&o_alt, dtype, ps->alt, g.nx, g.ny, g.pitch, P.OW, SKR))) {
        g_stream_err = "cuTensorMapEncodeTiled failed"; return -1;
    }
    const bool check = tol > 0.0 && mode == 1;
    if (check) {
        if (cudaMemsetAsync(ps->ctl, 0, sizeof(StreamCtl), s) != cudaSuccess || cudaMemsetAsync(ps->slots, 0, 8 * sizeof(unsigned long long), s) != cudaSuccess) {
            g_stream_err = "cudaMemsetAsync(control)"; return -1;
        }
    }
    void* cur = phi; void* nxt = ps->alt;
    int done = 0, pass = 0;
    ps->last_depth = cap; ps->last_passes = 0;
    const int poll = env_int("B200CFD_STREAM_POLL", 32);
    auto launch = [&](int role, int d) -> int {
        P.nact = mode == 0 ? d : 2 * d;
        P.iter0 = done;
        const CUtensorMap& mi = cur == phi ? m_phi : m_alt;
        const CUtensorMap& mo = cur == phi ? o_alt : o_phi;
        unsigned long long* sl = check ? ps->slots : nullptr;
        StreamCtl* ct = check ? ps->ctl : nullptr;
        if (dtype == 0) return mode == 0 ? stream_dispatch<double, 0>(L, mi, m_rhs, mo, P, sl, ct, role, pass, ps->cflags, ps->tiles, s)
                                         : stream_dispatch<double, 1>(L, mi, m_rhs, mo, P, sl, ct, role, pass, ps->cflags, ps->tiles, s);
        return mode == 0 ? stream_dispatch<float, 0>(L, mi, m_rhs, mo, P, sl, ct, role, pass, ps->cflags, ps->tiles, s)
                         : stream_dispatch<float, 1>(L, mi, m_rhs, mo, P, sl, ct, role, pass, ps->cflags, ps->tiles, s);
    };
    StreamCtl hc;
    bool stopped = false;
    while (done < iters) {
        const int d = iters - done < cap ? iters - done : cap;     // the last pass may be shallower
        if (launch(0, d)) return -1;
        if (check && launch(1, d)) return -1;                       // redo of this pass if the stop fell inside it (else exits at once)
        ++ps->last_passes;
        void* t = cur; cur = nxt; nxt = t;
        done += d; ++pass;
        if (check && (pass % poll == 0 || done >= iters)) {
            if (cudaMemcpyAsync(&hc, ps->ctl, sizeof hc, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess) {
                g_stream_err = "reading the convergence flag failed"; return -1;
            }
            if (hc.done) { stopped = true; break; }
        }
    }
    if (check) {
        if (stopped) {
            // the result is what pass hc.stop_pass wrote: pass p reads (p even ? phi : alt) and writes the other one
            done = hc.iters;
            cur = (hc.stop_pass & 1) ? phi : ps->alt;
            ps->last_passes = hc.stop_pass + 1;
        }
        if (last_change) *last_change = hc.last_change;
    }
    if (cur != phi && cudaMemcpyAsync(phi, cur, bytes, cudaMemcpyDeviceToDevice, s) != cudaSuccess) { g_stream_err = "copy back"; return -1; }
    *iters_done = done;
    return 0;
}

}  // namespace b200cfd

// Host run of the tile plan (the same functions k_stream_plan runs, poisson_stream_plan.h), for the CPU test-suite: no CUDA call.
// cflags: [n_strips][NC] host bytes; tiles_out: ntile x 4 ints (strip, r0, r1, 0).  Returns 0, or -1 on bad arguments.
extern "C" int b200cfd_stream_plan_host(const uint8_t* cflags, int n_strips, int NC, int ny, int shift, int ntile, int alpha_pct,
                                        int min_groups, int* tiles_out) {
    using namespace b200cfd;
    if (!cflags || !tiles_out || n_strips < 1 || NC < 1 || ny < 1 || ntile < n_strips || min_groups < 1 || shift < 0 ||
        alpha_pct < 0 || alpha_pct > 400) return -1;
    StreamPlanArgs a;
    a.cflags = cflags; a.n_strips = n_strips; a.NC = NC; a.ny = ny; a.shift = shift; a.ntile = ntile; a.alpha_pct = alpha_pct;
    a.min_groups = min_groups;
    const int nseg = plan_nseg(a), cap = plan_cap(a);
    int* seg = new int[(size_t)n_strips * nseg];
    int* cost = new int[3 * (size_t)n_strips];
    int* n = cost + n_strips;
    int* first = n + n_strips;
    StreamTile* t = reinterpret_cast<StreamTile*>(tiles_out);
    for (int i = 0; i < 4 * ntile; ++i) tiles_out[i] = 0;
    for (int it = 0; it < n_strips * nseg; ++it) seg[it] = plan_seg_sum(a, it / nseg, it % nseg);
    int lo = 1, hi = 1;
    for (int s = 0; s < n_strips; ++s) { cost[s] = plan_strip_prefix(a, seg + (size_t)s * nseg); if (cost[s] > hi) hi = cost[s]; }
    while (lo < hi) {
        const int mid = lo + (hi - lo) / 2;
        int sum = 0;
        for (int s = 0; s < n_strips; ++s) sum += plan_tiles_at(cost[s], mid, cap);
        if (sum <= ntile) hi = mid; else lo = mid + 1;
    }
    int f = 0;
    for (int s = 0; s < n_strips; ++s) { n[s] = plan_tiles_at(cost[s], lo, cap); first[s] = f; f += n[s]; }
    for (int s = 0; s < n_strips; ++s)
        for (int j = 0; j < n[s]; ++j) {
#if STREAM_PLAN_STRIP_MAJOR
            StreamTile& d = t[first[s] + j];
#else
            StreamTile& d = t[plan_tile_index(n, n_strips, s, j)];
#endif
            d.s = s; d.pad = 0;
            d.r0 = plan_cut_row(a, s, seg + (size_t)s * nseg, cost[s], n[s], j);
            d.r1 = plan_cut_row(a, s, seg + (size_t)s * nseg, cost[s], n[s], j + 1);
        }
    delete[] seg; delete[] cost;
    return 0;
}

