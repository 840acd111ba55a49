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


// ---- all fields, both sides, one launch ---------------------------------------------------------------------------
// A halo refresh moves the columns [jl, jl+nl) (towards the left neighbour) and [jr, jr+nr) (towards the right one) of
// NF fields.  One warp per row: lanes run along the columns, so the strided side is read / written in contiguous runs
// of 8*n bytes and the buffer side ([field][row][column], what the NCCL message is) is fully coalesced; no index
// division.  PACK: fields -> buffers; else buffers -> fields.
struct HaloFields { void* f[4]; };
template <typename T, bool PACK>
__global__ void __launch_bounds__(256)
k_halo_rows(HaloFields hf, int nf, long long pitch, int ny, int jl, int nl, T* __restrict__ bl, int jr, int nr, T* __restrict__ br) {
    const int lane = threadIdx.x & 31;
    for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < ny; row += gridDim.x * 8) {
        for (int k = 0; k < nf; ++k) {
            T* fr = (T*)hf.f[k] + (long long)row * pitch;
            if (nl > 0) {
                T* b = bl + ((long long)k * ny + row) * nl;
                for (int c = lane; c < nl; c += 32) { if (PACK) b[c] = fr[jl + c]; else fr[jl + c] = b[c]; }
            }
            if (nr > 0) {
                T* b = br + ((long long)k * ny + row) * nr;
                for (int c = lane; c < nr; c += 32) { if (PACK) b[c] = fr[jr + c]; else fr[jr + c] = b[c]; }
            }
        }
    }
}

void launch_halo_rows(int es, bool pack, const Geom& g, void* const* fields, int nf, int jl, int nl, void* bl, int jr, int nr,
                      void* br, cudaStream_t s) {
    if (nf <= 0 || (nl <= 0 && nr <= 0)) return;
    HaloFields hf = {};
    for (int k = 0; k < nf && k < 4; ++k) hf.f[k] = fields[k];
    int nb = (g.ny + 7) / 8;
    if (nb > 148 * 8) nb = 148 * 8;
    if (es == 8) {
        if (pack) k_halo_rows<double, true><<<nb, 256, 0, s>>>(hf, nf, g.pitch, g.ny, jl, nl, (double*)bl, jr, nr, (double*)br);
        else k_halo_rows<double, false><<<nb, 256, 0, s>>>(hf, nf, g.pitch, g.ny, jl, nl, (double*)bl, jr, nr, (double*)br);
    } else {
        if (pack) k_halo_rows<float, true><<<nb, 256, 0, s>>>(hf, nf, g.pitch, g.ny, jl, nl, (float*)bl, jr, nr, (float*)br);
        else k_halo_rows<float, false><<<nb, 256, 0, s>>>(hf, nf, g.pitch, g.ny, jl, nl, (float*)bl, jr, nr, (float*)br);
    }
}

}  // namespace b200cfd
