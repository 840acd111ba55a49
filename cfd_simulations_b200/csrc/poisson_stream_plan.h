// poisson_stream_plan.h — cost-balanced tile plan of the streaming Poisson kernel (poisson_stream.cu).
//
// Every tile (one warp: a 128-column window of one strip x a range of rows) is resident at once, so a pass lasts as long as
// its most expensive tile.  A 3-row chunk whose levels read masked or never-updated cells runs the chunk body with the class
// selects (~1.4 x the instructions of the plain body), and such chunks are not spread evenly: the first and last strip have
// one in EVERY chunk (the border columns), the strips under the disc have a run of them.  With equal row ranges the r02z
// capture showed the SMs active for 69 % of the Jacobi kernel's duration.  The plan therefore gives every strip a number of
// tiles in proportion to its cost and cuts each strip where the running cost crosses equal shares.
//
// The functions are __host__ __device__ and sequential per strip: the device runs them with one thread per strip
// (k_stream_plan), the CPU test-suite runs the same code through b200cfd_stream_plan_host.
// No counterpart in the reference (v5.py:202-226, :336-346 sweep the whole array in one Numba prange).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define PLAN_HD __host__ __device__
#else
#define PLAN_HD
#endif

namespace b200cfd {

struct StreamTile { int s, r0, r1, pad; };   // strip, first owned row (multiple of 3), one past the last owned row; r1 <= r0: empty

struct StreamPlanArgs {
    const uint8_t* cflags;   // [n_strips][NC] chunk flags written by k_stream_make_rhs
    int n_strips, NC;
    int ny;
    int shift;               // rows 3g .. 3g + 2 leave the pipeline in global chunk g + shift  (shift = (2L + E) / 3)
    int ntile;               // tiles to hand out (= warps resident at once)
    int alpha_pct;           // extra cost of a flagged chunk, in percent of a plain one
    int min_groups;          // fewest 3-row groups a tile may own
};

PLAN_HD inline int plan_groups(const StreamPlanArgs& a) { return (a.ny + 2) / 3; }

PLAN_HD inline int plan_group_cost(const StreamPlanArgs& a, int s, int g) {
    int c = g + a.shift;
    if (c > a.NC - 1) c = a.NC - 1;
    return 100 + (a.cflags[(size_t)s * a.NC + c] ? a.alpha_pct : 0);
}

// cost of strip s (percent units)
PLAN_HD inline long long plan_strip_cost(const StreamPlanArgs& a, int s) {
    const int G = plan_groups(a);
    long long t = 0;
    for (int g = 0; g < G; ++g) t += plan_group_cost(a, s, g);
    return t;
}

// tiles per strip: the allocation that minimises the most expensive tile, max_s cost[s] / n[s], subject to sum n[s] <= ntile
// (bisection on that maximum T: n[s] = ceil(cost[s] / T)), every strip at least one tile, none more than G / min_groups;
// tiles left over go to the strips with the highest cost per tile.  first[s] = index of strip s's first tile.
// Serial; n_strips is small (74 at 8192 columns).
PLAN_HD inline void plan_allocate(const StreamPlanArgs& a, const long long* cost, int* n, int* first) {
    const int G = plan_groups(a);
    int cap = G / a.min_groups;
    if (cap < 1) cap = 1;
    long long lo = 1, hi = 1;
    for (int s = 0; s < a.n_strips; ++s) if (cost[s] > hi) hi = cost[s];
    while (lo < hi) {
        const long long mid = lo + (hi - lo) / 2;
        long long need = 0;
        for (int s = 0; s < a.n_strips; ++s) {
            long long k = (cost[s] + mid - 1) / mid;
            if (k < 1) k = 1;
            if (k > cap) k = cap;
            need += k;
        }
        if (need <= a.ntile) hi = mid; else lo = mid + 1;
    }
    int used = 0;
    for (int s = 0; s < a.n_strips; ++s) {
        long long k = (cost[s] + lo - 1) / lo;
        if (k < 1) k = 1;
        if (k > cap) k = cap;
        n[s] = (int)k;
        used += n[s];
    }
    for (int spare = a.ntile - used; spare > 0; --spare) {
        int best = -1;
        for (int s = 0; s < a.n_strips; ++s)
            if (n[s] < cap && (best < 0 || cost[s] * n[best] > cost[best] * n[s])) best = s;
        if (best < 0) break;
        ++n[best];
    }
    int f = 0;
    for (int s = 0; s < a.n_strips; ++s) { first[s] = f; f += n[s]; }
}

// cuts strip s into n tiles of (nearly) equal cost; entries that cannot be filled (short grids) are left empty
PLAN_HD inline void plan_cut(const StreamPlanArgs& a, int s, long long cost, int n, int first, StreamTile* tiles) {
    const int G = plan_groups(a);
    // the fixed cost of a tile (filling its pipeline) is the same for all and does not move the cuts
    long long acc = 0;
    int j = 0, start = 0;
    for (int g = 0; g < G && j < n - 1; ++g) {
        acc += plan_group_cost(a, s, g);
        const int len = g + 1 - start, left = G - (g + 1);
        if (acc * n >= (long long)(j + 1) * cost && len >= a.min_groups && left >= (n - 1 - j) * a.min_groups) {
            tiles[first + j].s = s; tiles[first + j].r0 = 3 * start; tiles[first + j].r1 = 3 * (g + 1); tiles[first + j].pad = 0;
            start = g + 1;
            ++j;
        }
    }
    tiles[first + j].s = s; tiles[first + j].r0 = 3 * start; tiles[first + j].r1 = a.ny; tiles[first + j].pad = 0;
    for (++j; j < n; ++j) { tiles[first + j].s = s; tiles[first + j].r0 = 0; tiles[first + j].r1 = 0; tiles[first + j].pad = 0; }
}

}  // namespace b200cfd
