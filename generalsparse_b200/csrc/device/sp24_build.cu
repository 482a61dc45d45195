// The 2:4 image (host/sp24_format.hpp) built on the device: the plan-time counterpart of transform.cu.  The host builder takes ~0.5 s for the C2
// weight (download of the chain's arrays, a sort per row, byte-wise writes, upload of 56 MB); these passes take the chain's DEVICE arrays
// -- column ids, fp16 values, the row level's first_nz_indices -- and leave the image where the kernel reads it.  Same layout, bit for bit, for
// the compressed operand, the metadata, the unit table of the stream without empty tiles and the row table; the tile-local remainder lists
// a row's entries in column order (the host orders them against shared-memory bank conflicts: a different order of the same sum).
// One thread per stored entry: its rank inside its group of four columns comes from a walk over its (at most three) left neighbours, what
// the group's first element needs to know about the second from a walk to the right; padding zeros (modify_*_by_col_pad_in_sub_matrix.cc)
// are skipped as on the host.  Unsorted or duplicate columns inside a row raise a flag and the plan falls back to the host builder,
// which refuses them with its own message.
#include "common.cuh"
#include "sp24_build.hpp"

#include <algorithm>

namespace gsd {
namespace {

#define TRY(x)                                                                                                                                       \
    do {                                                                                                                                             \
        cudaError_t e_ = (x);                                                                                                                        \
        if (e_ != cudaSuccess) return e_;                                                                                                            \
    } while (0)

constexpr uint32_t ROWS = 128, COLS = 128, A_BYTES = ROWS * (COLS / 2) * 2, E_BYTES = ROWS * 16, TILE_BYTES = A_BYTES + E_BYTES, HEADER = 272;

// host/sp24_format.hpp: sp24_a_offset, sp24_e_byte, sp24_e_shift, sp24_b_row_offset
__device__ __forceinline__ uint32_t a_offset(uint32_t r, uint32_t slot) { return r * 128u + ((((slot >> 3) ^ (r & 7u)) & 7u) << 4) + (slot & 7u) * 2u; }
__device__ __forceinline__ uint32_t e_byte(uint32_t r, uint32_t k)
{
    const uint32_t L = (r & 7u) + 8u * ((k >> 4) & 1u) + 16u * (r >> 4);
    return A_BYTES + 16u * L + 4u * (k >> 5) + 2u * ((r >> 3) & 1u) + ((k >> 3) & 1u);
}
__device__ __forceinline__ uint32_t e_shift(uint32_t k) { return ((k >> 2) & 1u) * 4u; }
__device__ __forceinline__ uint32_t b_row_offset(uint32_t k, uint32_t row_bytes) { return k * row_bytes + ((((k * row_bytes) >> 7) & (row_bytes / 16u - 1u)) << 4); }

struct build_args {
    const void *col;
    const uint16_t *val;
    const uint32_t *rowptr;
    uint32_t n_rows, n_cols, nnz, k_tiles, col_is_u16;
    uint32_t *tile_img, *rem_cnt, *bad; // counts: non-zeros the image of a tile holds, remainder entries per (tile, tile row); bad input flag
    // write pass
    unsigned char *img;
    const uint32_t *unit_of_tile; // null: every tile is a unit
    uint32_t unit_bytes, row_bytes;
};

__device__ __forceinline__ uint32_t col_at(const build_args &a, uint32_t i) { return a.col_is_u16 ? (uint32_t)((const uint16_t *)a.col)[i] : ((const uint32_t *)a.col)[i]; }

template <bool WRITE> __global__ void __launch_bounds__(256) k_sp24_entries(build_args a)
{
    for (uint64_t e64 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e64 < a.nnz; e64 += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t e = (uint32_t)e64;
        const uint16_t v = a.val[e];
        if ((v & 0x7fffu) == 0) continue; // +-0: the chain's padding
        uint32_t lo = 0, hi = a.n_rows; // the row: the last one that starts at or before e
        while (hi - lo > 1) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (a.rowptr[mid] <= e) lo = mid;
            else hi = mid;
        }
        const uint32_t r = lo, rs = a.rowptr[r], re = a.rowptr[r + 1];
        if (e >= re) continue; // beyond the row level (cannot happen for a chain's arrays)
        const uint32_t c = col_at(a, e), g = c >> 2, kt = c / COLS;
        if (c >= a.n_cols) {
            *a.bad = 1;
            continue;
        }
        // left neighbours: the rank inside the group; for a remainder entry also what the earlier groups of this tile row overflow by
        uint32_t rank = 0, extra = 0, cnt = 0, gcur = 0xffffffffu, prev = c;
        for (uint32_t j = e; j > rs;) {
            j--;
            if ((a.val[j] & 0x7fffu) == 0) continue;
            const uint32_t cj = col_at(a, j);
            if (cj >= prev) *a.bad = 1; // not ascending (or a duplicate): the host builder refuses it
            prev = cj;
            if ((cj >> 2) == g) {
                rank++;
                continue;
            }
            if (rank < 2 || cj / COLS != kt) break; // image entries need nothing further left; the tile's window ends here
            if ((cj >> 2) != gcur) extra += cnt > 2 ? cnt - 2 : 0, gcur = cj >> 2, cnt = 1;
            else cnt++;
        }
        extra += cnt > 2 ? cnt - 2 : 0;
        const uint32_t tile = (r / ROWS) * a.k_tiles + kt, rr = r % ROWS;
        if (!WRITE) {
            if (rank >= 2) atomicAdd(a.rem_cnt + (size_t)tile * ROWS + rr, 1u);
            else atomicAdd(a.tile_img + tile, 1u);
            continue;
        }
        unsigned char *unit = a.img + (size_t)(a.unit_of_tile ? a.unit_of_tile[tile] : tile) * a.unit_bytes;
        if (rank >= 2) { // entries of tile row rr: start[rr] .. start[rr + 1], column order
            const uint32_t start = ((const uint16_t *)(unit + TILE_BYTES))[rr];
            ((uint32_t *)(unit + TILE_BYTES + HEADER))[start + extra + rank - 2] = (b_row_offset(c % COLS, a.row_bytes) << 16) | v;
            continue;
        }
        const uint32_t k = (g * 4u) % COLS, slot = k / 2u, idx = c & 3u;
        if (rank == 1) {
            *(uint16_t *)(unit + a_offset(rr, slot + 1)) = v;
            continue;
        }
        // the group's first element writes the metadata: it needs the index of the second, if there is one
        bool has2 = false;
        uint32_t idx2 = 3;
        for (uint32_t j = e + 1; j < re; j++) {
            if ((a.val[j] & 0x7fffu) == 0) continue;
            const uint32_t cj = col_at(a, j);
            if ((cj >> 2) == g) has2 = true, idx2 = cj & 3u;
            break;
        }
        uint32_t nib;
        if (has2) *(uint16_t *)(unit + a_offset(rr, slot)) = v, nib = idx | (idx2 << 2);
        else if (idx == 3u) *(uint16_t *)(unit + a_offset(rr, slot + 1)) = v, nib = 0u | (3u << 2); // a lone element in the last column: (0 holding zero, 3)
        else *(uint16_t *)(unit + a_offset(rr, slot)) = v, nib = idx | (3u << 2);                    // a lone element elsewhere: (itself, 3 holding zero)
        // every group starts as 0x4 ("keeps elements 0 and 1, both zero"); two groups share a byte, four bytes a word
        const uint32_t eb = e_byte(rr, k);
        atomicXor((uint32_t *)(unit + (eb & ~3u)), ((nib ^ 0x4u) << e_shift(k)) << ((eb & 3u) * 8u));
    }
}

// one warp per tile: entries per tile (sum of its 128 row counts) in pass 0; the u16 start[129] header of the tile's unit in pass 1
__global__ void __launch_bounds__(256) k_sp24_tile_rows(const uint32_t *rem_cnt, uint32_t T, uint32_t *tile_rem, unsigned char *img, const uint32_t *unit_of_tile, const uint32_t *keep,
                                                        uint32_t unit_bytes)
{
    const int lane = threadIdx.x & 31;
    for (uint64_t t = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < T; t += ((uint64_t)gridDim.x * blockDim.x) >> 5) {
        uint32_t c[4], mine = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) c[j] = rem_cnt[t * ROWS + lane * 4 + j], mine += c[j];
        uint32_t inc = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += x;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
        if (!img) {
            if (lane == 0) tile_rem[t] = total;
            continue;
        }
        if (keep && !keep[t]) continue; // the tile left the stream (it holds nothing)
        uint16_t *start = (uint16_t *)(img + (size_t)(unit_of_tile ? unit_of_tile[t] : t) * unit_bytes + TILE_BYTES);
        uint32_t run = inc - mine;
#pragma unroll
        for (int j = 0; j < 4; j++) start[lane * 4 + j] = (uint16_t)run, run += c[j];
        if (lane == 31) start[ROWS] = (uint16_t)total;
    }
}

// which tiles stay in the stream (compact_sp24_units: a tile with a non-zero in its image or an entry in its block; a panel with none keeps its first)
__global__ void k_sp24_keep(const uint32_t *tile_img, const uint32_t *tile_rem, uint32_t n_panels, uint32_t k_tiles, uint32_t *keep)
{
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_panels; p += (uint64_t)gridDim.x * blockDim.x) {
        bool any = false;
        for (uint32_t kt = 0; kt < k_tiles; kt++) {
            const bool k = tile_img[p * k_tiles + kt] != 0 || tile_rem[p * k_tiles + kt] != 0;
            keep[p * k_tiles + kt] = k, any = any || k;
        }
        if (!any) keep[p * k_tiles] = 1;
    }
}

// exclusive scan of keep[] (one CTA walks the array: a few thousand tiles), the unit table (k tile of every unit) and the panel table
__global__ void __launch_bounds__(1024) k_sp24_units(const uint32_t *keep, uint32_t T, uint32_t k_tiles, uint32_t n_panels, uint32_t *unit_of_tile, uint32_t *unit_kt, uint32_t *panel_ptr,
                                                      uint32_t *n_units)
{
    __shared__ uint32_t warp_sum[32], carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t base = 0; base < T; base += 1024) {
        const uint32_t t = base + threadIdx.x, k = t < T ? keep[t] : 0u;
        uint32_t inc = k;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += x;
        }
        if (lane == 31) warp_sum[w] = inc;
        __syncthreads();
        if (w == 0) {
            uint32_t s = warp_sum[lane], si = s;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, si, d);
                if (lane >= d) si += x;
            }
            warp_sum[lane] = si - s;
        }
        __syncthreads();
        const uint32_t at = carry + warp_sum[w] + inc - k;
        if (t < T) {
            unit_of_tile[t] = at;
            if (k && unit_kt) unit_kt[at] = t % k_tiles;
            if (panel_ptr && t % k_tiles == 0) panel_ptr[t / k_tiles] = at;
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = at + k;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *n_units = carry;
        if (panel_ptr) panel_ptr[n_panels] = carry;
    }
}

__global__ void k_sp24_sums(const uint32_t *tile_img, const uint32_t *tile_rem, uint32_t T, unsigned long long *sums /* image nnz, remainder nnz */, uint32_t *max_entries)
{
    unsigned long long si = 0, sr = 0;
    uint32_t mx = 0;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < T; t += (uint64_t)gridDim.x * blockDim.x) si += tile_img[t], sr += tile_rem[t], mx = max(mx, tile_rem[t]);
    atomicAdd(sums, si), atomicAdd(sums + 1, sr), atomicMax(max_entries, mx);
}

// a unit before its entries arrive: compressed operand zero, every metadata nibble 0x4, header and entry list zero
__global__ void __launch_bounds__(256) k_sp24_clear(unsigned char *img, uint32_t n_units, uint32_t unit_bytes)
{
    const uint32_t words = unit_bytes / 16;
    for (uint64_t u = blockIdx.x; u < n_units; u += gridDim.x) {
        uint4 *p = (uint4 *)(img + u * unit_bytes);
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) {
            const bool meta = i >= A_BYTES / 16 && i < TILE_BYTES / 16;
            p[i] = meta ? make_uint4(0x44444444u, 0x44444444u, 0x44444444u, 0x44444444u) : make_uint4(0, 0, 0, 0);
        }
    }
}

__global__ void k_sp24_out_row(const uint32_t *rowid, const uint32_t *perm, uint32_t begin_row, uint32_t n_rows, uint32_t n_slots, uint32_t perm_len, uint32_t *out_row, uint32_t *bad)
{
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t v = 0xffffffffu;
        if (i < n_rows) {
            const uint32_t r = rowid[i];
            if (perm && r >= perm_len) *bad = 1;
            else v = (perm ? perm[r] : r) + begin_row;
        }
        out_row[i] = v;
    }
}

// the 128 x 128 tiles that hold a stored entry, for the cost model (engine.cu: auto_candidates)
__global__ void k_tiles_seen(const void *col, uint32_t col_is_u16, const uint32_t *rowptr, uint32_t n_rows, uint32_t nnz, uint32_t k_tiles, unsigned char *seen)
{
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = n_rows;
        while (hi - lo > 1) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (rowptr[mid] <= e) lo = mid;
            else hi = mid;
        }
        if (e >= rowptr[lo + 1]) continue;
        const uint32_t c = col_is_u16 ? (uint32_t)((const uint16_t *)col)[e] : ((const uint32_t *)col)[e];
        if (c / COLS < k_tiles) seen[(size_t)(lo / ROWS) * k_tiles + c / COLS] = 1;
    }
}
__global__ void k_tiles_count(const unsigned char *seen, uint32_t n_panels, uint32_t k_tiles, unsigned long long *out)
{
    unsigned long long n = 0;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_panels; p += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t here = 0;
        for (uint32_t k = 0; k < k_tiles; k++) here += seen[p * k_tiles + k];
        n += here ? here : 1; // every row panel keeps one tile
    }
    if (n) atomicAdd(out, n);
}

inline unsigned grid_for(uint64_t n, unsigned per) { return (unsigned)std::min<uint64_t>(std::max<uint64_t>(1, (n + per - 1) / per), 148u * 32u); }

} // namespace

void sp24_dev_image::release()
{
    cudaFree(tile_img), cudaFree(rem_cnt), cudaFree(tile_rem), cudaFree(keep), cudaFree(unit_of_tile), cudaFree(scal);
    tile_img = rem_cnt = tile_rem = keep = unit_of_tile = nullptr, scal = nullptr;
}

cudaError_t count_tiles128_dev(const void *col, bool col_is_u16, const uint32_t *rowptr, uint32_t n_rows, uint32_t n_cols, uint64_t nnz, cudaStream_t s, uint64_t *out)
{
    const uint32_t n_panels = (n_rows + ROWS - 1) / ROWS, k_tiles = (n_cols + COLS - 1) / COLS;
    const uint64_t T = (uint64_t)n_panels * k_tiles;
    if (T == 0 || nnz > 0xfffffff0ull) return cudaErrorInvalidValue;
    unsigned char *seen = nullptr;
    TRY(cudaMalloc((void **)&seen, T + 16));
    cudaError_t e = cudaMemsetAsync(seen, 0, T + 16, s);
    unsigned long long *cnt = (unsigned long long *)(seen + ((T + 7) & ~7ull)); // the spare bytes behind the flags
    unsigned long long h = 0;
    if (e == cudaSuccess) {
        k_tiles_seen<<<grid_for(nnz, 256 * 4), 256, 0, s>>>(col, col_is_u16 ? 1u : 0u, rowptr, n_rows, (uint32_t)nnz, k_tiles, seen);
        k_tiles_count<<<grid_for(n_panels, 128), 128, 0, s>>>(seen, n_panels, k_tiles, cnt);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h, cnt, 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(seen);
    *out = h;
    return e;
}

cudaError_t sp24_dev_count(const launch_params &p, const uint32_t *rowptr, uint32_t n_rows, cudaStream_t s, sp24_dev_image &d)
{
    d.n_panels = (n_rows + ROWS - 1) / ROWS, d.k_tiles = (p.a.cols + COLS - 1) / COLS;
    const uint64_t T64 = (uint64_t)d.n_panels * d.k_tiles;
    if (T64 == 0 || T64 > 0x3fffffffull || p.a.nnz > 0xfffffff0ull) return cudaErrorInvalidValue;
    d.T = (uint32_t)T64;
    TRY(cudaMalloc((void **)&d.tile_img, (size_t)d.T * 4));
    TRY(cudaMalloc((void **)&d.rem_cnt, (size_t)d.T * ROWS * 4));
    TRY(cudaMalloc((void **)&d.tile_rem, (size_t)d.T * 4));
    TRY(cudaMalloc((void **)&d.scal, 64));
    TRY(cudaMemsetAsync(d.tile_img, 0, (size_t)d.T * 4, s));
    TRY(cudaMemsetAsync(d.rem_cnt, 0, (size_t)d.T * ROWS * 4, s));
    TRY(cudaMemsetAsync(d.scal, 0, 64, s));
    build_args a = {};
    a.col = p.a.col, a.val = (const uint16_t *)p.a.val, a.rowptr = rowptr, a.n_rows = n_rows, a.n_cols = p.a.cols, a.nnz = (uint32_t)p.a.nnz, a.k_tiles = d.k_tiles;
    a.col_is_u16 = p.a.col_is_u16, a.tile_img = d.tile_img, a.rem_cnt = d.rem_cnt, a.bad = (uint32_t *)d.scal + 6;
    k_sp24_entries<false><<<grid_for(p.a.nnz, 256 * 4), 256, 0, s>>>(a);
    k_sp24_tile_rows<<<grid_for((uint64_t)d.T * 32, 256), 256, 0, s>>>(d.rem_cnt, d.T, d.tile_rem, nullptr, nullptr, nullptr, 0);
    k_sp24_sums<<<grid_for(d.T, 256 * 4), 256, 0, s>>>(d.tile_img, d.tile_rem, d.T, d.scal, (uint32_t *)d.scal + 4);
    TRY(cudaGetLastError());
    unsigned long long h[8];
    TRY(cudaMemcpyAsync(h, d.scal, 64, cudaMemcpyDeviceToHost, s));
    TRY(cudaStreamSynchronize(s));
    d.img_nnz = h[0], d.rem_nnz = h[1], d.max_entries = ((uint32_t *)h)[4], d.bad = ((uint32_t *)h)[6] != 0;
    return cudaSuccess;
}

cudaError_t sp24_dev_fill(const launch_params &p, const uint32_t *rowptr, uint32_t n_rows, uint32_t unit_bytes, uint32_t row_bytes, bool compact, double keep_frac, cudaStream_t s,
                          sp24_dev_image &d, unsigned char **img, uint32_t **panel_ptr, uint32_t **unit_kt, uint32_t *n_units)
{
    *img = nullptr, *panel_ptr = nullptr, *unit_kt = nullptr, *n_units = d.T;
    const bool tiled = unit_bytes != TILE_BYTES;
    if (compact) {
        uint32_t kept = 0;
        TRY(cudaMalloc((void **)&d.keep, (size_t)d.T * 4));
        TRY(cudaMalloc((void **)&d.unit_of_tile, (size_t)d.T * 4 + 4));
        TRY(cudaMalloc((void **)unit_kt, (size_t)d.T * 4 + 16));
        TRY(cudaMalloc((void **)panel_ptr, ((size_t)d.n_panels + 1) * 4 + 16));
        k_sp24_keep<<<grid_for(d.n_panels, 128), 128, 0, s>>>(d.tile_img, d.tile_rem, d.n_panels, d.k_tiles, d.keep);
        k_sp24_units<<<1, 1024, 0, s>>>(d.keep, d.T, d.k_tiles, d.n_panels, d.unit_of_tile, *unit_kt, *panel_ptr, (uint32_t *)d.scal + 8);
        TRY(cudaGetLastError());
        TRY(cudaMemcpyAsync(&kept, (uint32_t *)d.scal + 8, 4, cudaMemcpyDeviceToHost, s));
        TRY(cudaStreamSynchronize(s));
        if ((double)kept > keep_frac * (double)d.T) { // not worth it: the stream keeps every tile
            cudaFree(*unit_kt), cudaFree(*panel_ptr), cudaFree(d.unit_of_tile), cudaFree(d.keep);
            *unit_kt = *panel_ptr = nullptr, d.unit_of_tile = d.keep = nullptr;
        } else *n_units = kept;
    }
    TRY(cudaMalloc((void **)img, (size_t)*n_units * unit_bytes + 16));
    k_sp24_clear<<<std::min(*n_units, 148u * 8u), 256, 0, s>>>(*img, *n_units, unit_bytes);
    if (tiled) k_sp24_tile_rows<<<grid_for((uint64_t)d.T * 32, 256), 256, 0, s>>>(d.rem_cnt, d.T, nullptr, *img, d.unit_of_tile, d.keep, unit_bytes);
    TRY(cudaGetLastError());
    build_args a = {};
    a.col = p.a.col, a.val = (const uint16_t *)p.a.val, a.rowptr = rowptr, a.n_rows = n_rows, a.n_cols = p.a.cols, a.nnz = (uint32_t)p.a.nnz, a.k_tiles = d.k_tiles;
    a.col_is_u16 = p.a.col_is_u16, a.tile_img = d.tile_img, a.rem_cnt = d.rem_cnt, a.bad = (uint32_t *)d.scal + 6;
    a.img = *img, a.unit_of_tile = d.unit_of_tile, a.unit_bytes = unit_bytes, a.row_bytes = row_bytes;
    k_sp24_entries<true><<<grid_for(p.a.nnz, 256 * 4), 256, 0, s>>>(a);
    TRY(cudaGetLastError());
    return cudaStreamSynchronize(s);
}

cudaError_t sp24_dev_out_row(const launch_params &p, const uint32_t *rowid, uint32_t n_rows, uint32_t n_panels, cudaStream_t s, sp24_dev_image &d, uint32_t **out_row)
{
    const uint32_t n_slots = n_panels * ROWS;
    TRY(cudaMalloc((void **)out_row, (size_t)n_slots * 4 + 16));
    k_sp24_out_row<<<grid_for(n_slots, 256), 256, 0, s>>>(rowid, p.a.perm, p.a.begin_row, n_rows, n_slots, p.a.rows, *out_row, (uint32_t *)d.scal + 6);
    TRY(cudaGetLastError());
    uint32_t bad = 0;
    TRY(cudaMemcpyAsync(&bad, (uint32_t *)d.scal + 6, 4, cudaMemcpyDeviceToHost, s));
    TRY(cudaStreamSynchronize(s));
    return bad ? cudaErrorInvalidValue : cudaSuccess;
}

} // namespace gsd
