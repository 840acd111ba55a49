// poisson_stream_plan.h — cost-balanced tile plan of the streaming Poisson kernel (poisson_stream.cu).
//
// Every tile (one warp: a 128-column window of one strip x a range of rows) is resident at once, so a pass lasts as long as
// its most expensive tile.  A 3-row chunk whose levels read masked or never-updated cells runs the chunk body with the class
// selects (~1.4 x the instructions of the plain body), and such chunks are not spread evenly: the first and last strip have
// one in EVERY chunk (the border columns), the strips under the disc have a run of them.  With equal row ranges the r02z
// capture showed the SMs active for 69 % of the Jacobi kernel's duration.  The plan therefore gives every strip a number of
// tiles in proportion to its cost and cuts each strip where the running cost crosses equal shares.
//
// The functions are __host__ __device__ and each handles one item (a segment of a strip, a strip, a tile): the device runs
// them with the items spread over the threads of one block (k_stream_plan), the CPU test-suite runs the same functions in
// loops through b200cfd_stream_plan_host.  (A first version ran serially per strip, with the allocation on one thread in
// 64-bit arithmetic: 0.8 ms at 8192^2, as long as two passes of the kernel it plans.)
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
    int alpha_pct;           // extra cost of a flagged chunk, in percent of a plain one (0 .. 400)
    int min_groups;          // fewest 3-row groups a tile may own (guaranteed through plan_cap)
};

PLAN_HD inline int plan_groups(const StreamPlanArgs& a) { return (a.ny + 2) / 3; }
constexpr int PLAN_SEG = 64;                                          // groups per segment of a strip
PLAN_HD inline int plan_nseg(const StreamPlanArgs& a) { return (plan_groups(a) + PLAN_SEG - 1) / PLAN_SEG; }
// most tiles a strip may get: a tile's share of its strip's cost is at least 100 G / cap, a group costs at most
// 100 + alpha_pct, so with cap <= 100 G / ((100 + alpha_pct) min_groups) every tile owns at least min_groups groups
// without any further rule
PLAN_HD inline int plan_cap(const StreamPlanArgs& a) {
    const long long cap = 100ll * plan_groups(a) / ((100ll + a.alpha_pct) * a.min_groups);
    return cap < 1 ? 1 : (int)cap;
}

PLAN_HD inline int plan_group_cost(const StreamPlanArgs& a, int s, int g) {
    int c = g + a.shift;
    if (c > a.NC - 1) c = a.NC - 1;
    return 100 + (a.cflags[(size_t)s * a.NC + c] ? a.alpha_pct : 0);
}

// item (strip, segment): cost of the segment's groups (percent units)
PLAN_HD inline int plan_seg_sum(const StreamPlanArgs& a, int s, int seg) {
    const int G = plan_groups(a);
    int g1 = (seg + 1) * PLAN_SEG;
    if (g1 > G) g1 = G;
    int t = 0;
    for (int g = seg * PLAN_SEG; g < g1; ++g) t += plan_group_cost(a, s, g);
    return t;
}

// item strip: turns the strip's segment sums into exclusive prefix sums in place; returns the strip's cost
PLAN_HD inline int plan_strip_prefix(const StreamPlanArgs& a, int* seg) {
    const int n = plan_nseg(a);
    int acc = 0;
    for (int i = 0; i < n; ++i) { const int v = seg[i]; seg[i] = acc; acc += v; }
    return acc;
}

// tiles a strip of this cost gets when no tile may cost more than T
PLAN_HD inline int plan_tiles_at(int cost, int T, int cap) {
    int k = (int)(((unsigned)cost + (unsigned)T - 1u) / (unsigned)T);
    if (k < 1) k = 1;
    if (k > cap) k = cap;
    return k;
}

// item (strip, tile j of n): first owned row of the tile = 3 x (number of leading groups whose cost stays below j / n of the
// strip's); pre = the strip's exclusive segment prefix sums.  j == n gives ny (one past the last tile's rows).
PLAN_HD inline int plan_cut_row(const StreamPlanArgs& a, int s, const int* pre, int cost, int n, int j) {
    if (j <= 0) return 0;
    if (j >= n) return a.ny;
    const int G = plan_groups(a), nseg = plan_nseg(a);
    const int thr = (int)(((long long)j * cost + n - 1) / n);         // first g with acc(g) >= thr closes tile j - 1
    int seg = 0;
    while (seg + 1 < nseg && pre[seg + 1] < thr) ++seg;
    int acc = pre[seg], g = seg * PLAN_SEG;
    for (; g < G; ++g) {
        acc += plan_group_cost(a, s, g);
        if (acc >= thr) break;
    }
    int row = 3 * (g + 1);
    if (row > a.ny) row = a.ny;
    return row;
}

// item (strip, tile j): position of the tile in the table.  Row-major over (j, strip): the j-th tiles of all strips are
// neighbours in the table, so the 4 warps of a block (4 consecutive entries) and the blocks next to it work on ADJACENT strips
// at (nearly) the same rows, and what the machine reads at any moment is runs of neighbouring 1 KB row segments rather
// than 1184 unrelated ones (strips of equal cost have identical cuts).  n[] = tiles per strip.
PLAN_HD inline int plan_tile_index(const int* n, int n_strips, int s, int j) {
    int idx = 0;
    for (int t = 0; t < n_strips; ++t) idx += n[t] < j ? n[t] : j;        // tiles 0 .. j-1 of every strip
    for (int t = 0; t < s; ++t) idx += n[t] > j ? 1 : 0;                  // tile j of the strips before s
    return idx;
}

}  // namespace b200cfd
