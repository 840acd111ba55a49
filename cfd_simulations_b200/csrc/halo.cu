// halo.cu — x-slab halo plumbing: a slab's halo is a set of COLUMNS of a row-major (ny, nx) field, i.e. strided in
// memory, while NCCL send/recv wants one contiguous message.  pack: field[:, j0:j0+ncols] -> buf[ny][ncols];
// unpack: the inverse.  One thread per element, x fastest on both sides (ncols is 10..100, so a warp covers whole
// short rows: reads are 8*ncols-byte contiguous runs, writes fully coalesced).
#include "kernels.h"

namespace b200cfd {

template <typename T, bool PACK>
__global__ void __launch_bounds__(256)
k_columns(long long pitch, int ny, int j0, int ncols, T* __restrict__ field, T* __restrict__ buf) {
    const long long n = (long long)ny * ncols;
    for (long long t = blockIdx.x * 256ll + threadIdx.x; t < n; t += (long long)gridDim.x * 256ll) {
        const int i = (int)(t / ncols), c = (int)(t - (long long)i * ncols);
        if (PACK) buf[t] = field[i * pitch + j0 + c];
        else field[i * pitch + j0 + c] = buf[t];
    }
}

static inline int col_blocks(int ny, int ncols) {
    long long n = ((long long)ny * ncols + 255) / 256;
    return (int)(n < 1 ? 1 : (n > 148 * 16 ? 148 * 16 : n));
}

void launch_pack_columns(int es, const Geom& g, const void* field, int j0, int ncols, void* buf, cudaStream_t s) {
    if (ncols <= 0) return;
    if (es == 8) k_columns<double, true><<<col_blocks(g.ny, ncols), 256, 0, s>>>(g.pitch, g.ny, j0, ncols, (double*)field, (double*)buf);
    else k_columns<float, true><<<col_blocks(g.ny, ncols), 256, 0, s>>>(g.pitch, g.ny, j0, ncols, (float*)field, (float*)buf);
}
void launch_unpack_columns(int es, const Geom& g, void* field, int j0, int ncols, const void* buf, cudaStream_t s) {
    if (ncols <= 0) return;
    if (es == 8) k_columns<double, false><<<col_blocks(g.ny, ncols), 256, 0, s>>>(g.pitch, g.ny, j0, ncols, (double*)field, (double*)buf);
    else k_columns<float, false><<<col_blocks(g.ny, ncols), 256, 0, s>>>(g.pitch, g.ny, j0, ncols, (float*)field, (float*)buf);
}

}  // namespace b200cfd
