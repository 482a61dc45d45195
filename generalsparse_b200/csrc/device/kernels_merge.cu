// B200 range kernel (GS_FAMILY_B200_MERGE): the device consumer of the chains whose blocks are RANGES OF THE NON-ZERO STREAM that do
// not line up with rows -- merge-path parts (merge_path_{thread,warp,tblock}_operator: get_begin_{rows,nzs}_of_level_after_merge_path.cc),
// nnz-balanced row groups (balanced_interval_row_direction_*: data_transform_common.cc:934-988), column-direction row chunks
// (fixed_interval_col_direction_*_blocking_operator) and row blocks of more than one row.  The reference has no kernel token for the
// first two (its seven drivers never reach them); this is the B200 one for all four.
//
// Work unit: a SEGMENT [a, b) of the stream, walked by a GROUP of LG lanes (LG = min(N, 32)); lane j of the group owns dense columns
// j, j + 32, ... so a gathered row of B is one coalesced request of N values (whole 128-byte lines from N = 32 up, one 32-byte sector
// at N = 8 fp32).  Segments are the chain's own blocks; a block longer than a cap is cut into equal slices at plan time (the hub row of
// a power-law matrix becomes thousands of slices), the row each slice starts in found by binary search in the row pointer -- host
// side, once.  The group loads LG column ids / values per step with one coalesced request and hands them round by shuffles; row
// ends come from the row pointer (a warp-uniform scalar per group).  A row that lies wholly inside a segment is stored plainly; the
// (at most two) rows a segment shares with its neighbours go to C through red.global.add, and C is zeroed per launch, which also
// covers empty rows.  Bound: HBM (A streamed once) while B's working set sits in L1/L2.
#include "common.cuh"

#include "../../../include/gsparse_b200.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <new>
#include <vector>

namespace gsd {

struct b200_merge_plan {
    uint32_t *seg_nz = nullptr, *seg_row = nullptr, *rowptr = nullptr;
    uint32_t n_seg = 0, n_rows = 0;
    uint64_t part_bytes = 0; // the chain's own range array, as the kernel's format bytes count it
    int N = 0;
    char name[128] = {0};
};

namespace {

constexpr int MERGE_WARPS = 8, BATCH = 8;

struct merge_args {
    const void *col, *val, *B;
    void *C;
    const uint32_t *seg_nz, *seg_row, *rowptr, *perm;
    uint32_t n_seg, begin_row;
    int N;
};

template <typename ValT, int NV, int LG> __device__ __forceinline__ void flush(ValT *C, int N, uint32_t out_row, int j, float (&acc)[NV], bool edge)
{
    ValT *dst = C + (size_t)out_row * N + j;
#pragma unroll
    for (int v = 0; v < NV; v++) {
        if (edge) {
            if (acc[v] != 0.0f) atomic_add_out(dst, (size_t)(LG * v), acc[v]);
        } else dst[LG * v] = from_f32<ValT>(acc[v]);
        acc[v] = 0.0f;
    }
}

template <typename ColT, typename ValT, int LG, int NV> __global__ void __launch_bounds__(MERGE_WARPS * 32) k_b200_merge(merge_args p)
{
    const uint32_t tid = blockIdx.x * (MERGE_WARPS * 32) + threadIdx.x;
    const uint32_t g = tid / LG;
    const int j = tid % LG, lane = threadIdx.x & 31;
    if (g >= p.n_seg) return; // whole groups leave together
    const unsigned mask = LG == 32 ? 0xffffffffu : (((1u << LG) - 1u) << (lane & ~(LG - 1)));
    const ColT *col = (const ColT *)p.col;
    const ValT *val = (const ValT *)p.val, *B = (const ValT *)p.B;
    ValT *C = (ValT *)p.C;
    const int N = p.N;
    const uint32_t a = p.seg_nz[g], b = p.seg_nz[g + 1];
    uint32_t r = p.seg_row[g];
    uint32_t row_begin = p.rowptr[r], e = p.rowptr[r + 1];
    bool edge = a != row_begin; // the first row began before this segment: others add to it too
    float acc[NV];
#pragma unroll
    for (int v = 0; v < NV; v++) acc[v] = 0.0f;
    auto out = [&](uint32_t row) { return (p.perm ? p.perm[row] : row) + p.begin_row; };

    for (uint32_t k = a; k < b; k += LG) {
        uint32_t c_mine = 0;
        float v_mine = 0.0f;
        if (k + j < b) c_mine = (uint32_t)col[k + j], v_mine = to_f32(val[k + j]);
        const uint32_t cnt = min((uint32_t)LG, b - k);
#pragma unroll 1
        for (uint32_t j0 = 0; j0 < cnt; j0 += BATCH) {
            float x[BATCH][NV];
            // the gathers of a batch are in flight together; entries past the end repeat the lane's own id with value 0
#pragma unroll
            for (int q = 0; q < BATCH; q++) {
                const uint32_t c = __shfl_sync(mask, c_mine, (int)((j0 + q) & (LG - 1)), LG);
                const ValT *src = B + (size_t)c * N + j;
#pragma unroll
                for (int v = 0; v < NV; v++) x[q][v] = (LG >= BATCH || j0 + q < cnt) ? to_f32(__ldg(src + LG * v)) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < BATCH; q++) {
                if (LG < BATCH && q >= LG) break;
                const uint32_t kk = k + j0 + q;
                const float vv = __shfl_sync(mask, v_mine, (int)((j0 + q) & (LG - 1)), LG);
                if (j0 + q < cnt) {
                    while (kk == e) { // the row ended (possibly followed by empty rows: C keeps its zeros)
                        if (e != row_begin) flush<ValT, NV, LG>(C, N, out(r), j, acc, edge);
                        edge = false;
                        r++, row_begin = e, e = p.rowptr[r + 1];
                    }
#pragma unroll
                    for (int v = 0; v < NV; v++) acc[v] = fmaf(vv, x[q][v], acc[v]);
                }
            }
        }
    }
    if (b > a) flush<ValT, NV, LG>(C, N, out(r), j, acc, edge || b != e); // the last row may go on in the next segment
}

template <typename ColT, typename ValT> const void *pick(int N)
{
    switch (N) {
    case 1: return (const void *)k_b200_merge<ColT, ValT, 1, 1>;
    case 2: return (const void *)k_b200_merge<ColT, ValT, 2, 1>;
    case 4: return (const void *)k_b200_merge<ColT, ValT, 4, 1>;
    case 8: return (const void *)k_b200_merge<ColT, ValT, 8, 1>;
    case 16: return (const void *)k_b200_merge<ColT, ValT, 16, 1>;
    case 32: return (const void *)k_b200_merge<ColT, ValT, 32, 1>;
    case 64: return (const void *)k_b200_merge<ColT, ValT, 32, 2>;
    case 128: return (const void *)k_b200_merge<ColT, ValT, 32, 4>;
    case 256: return (const void *)k_b200_merge<ColT, ValT, 32, 8>;
    case 512: return (const void *)k_b200_merge<ColT, ValT, 32, 16>;
    default: return nullptr;
    }
}

int env_int(const char *name, int dflt)
{
    const char *v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

} // namespace

bool b200_merge_supported(int N) { return pick<uint32_t, float>(N) != nullptr; }

// parts: the chain's range array on the DEVICE (n_parts + 1 offsets into the stream); rows_dev: nz_row_indices on the device (u32, one per
// stored non-zero) or null, in which case chunk_rows (row of every part, col-direction chains) or the parts themselves (one row per part)
// give the row pointer.
cudaError_t b200_merge_prepare(const launch_params &p, const uint32_t *parts_dev, uint32_t n_parts, const uint32_t *nz_rows_dev, const uint32_t *part_rows_dev, bool parts_are_rows,
                               b200_merge_plan **out, cudaStream_t s)
{
    *out = nullptr;
    if (!b200_merge_supported(p.N) || n_parts == 0 || p.a.nnz == 0 || p.a.nnz > 0xfffffff0ull) return cudaErrorInvalidValue;
    b200_merge_plan *pl = new (std::nothrow) b200_merge_plan;
    if (!pl) return cudaErrorMemoryAllocation;
    auto fail = [&](cudaError_t e) {
        b200_merge_free(pl);
        return e;
    };
    cudaError_t e;
    std::vector<uint32_t> parts(n_parts + 1), rowptr;
    if ((e = cudaMemcpyAsync(parts.data(), parts_dev, parts.size() * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
    const uint32_t nnz = (uint32_t)p.a.nnz;
    uint32_t n_rows = 0;
    if (nz_rows_dev) { // COO row ids of the (padded, possibly sorted) stream -> row pointer
        std::vector<uint32_t> rows(nnz);
        if ((e = cudaMemcpyAsync(rows.data(), nz_rows_dev, (size_t)nnz * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        n_rows = rows[nnz - 1] + 1;
        rowptr.assign((size_t)n_rows + 1, 0);
        for (uint32_t i = 0; i < nnz; i++) {
            if (rows[i] >= n_rows || (i && rows[i] < rows[i - 1])) return fail(cudaErrorInvalidValue);
            rowptr[rows[i] + 1]++;
        }
        for (uint32_t r = 0; r < n_rows; r++) rowptr[r + 1] += rowptr[r];
    } else if (part_rows_dev) { // column-direction chunks: part i lies in row part_rows[i]
        std::vector<uint32_t> pr(n_parts);
        if ((e = cudaMemcpyAsync(pr.data(), part_rows_dev, (size_t)n_parts * 4, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(e);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        n_rows = pr[n_parts - 1] + 1;
        rowptr.assign((size_t)n_rows + 1, 0);
        for (uint32_t i = 0; i < n_parts; i++) {
            if (pr[i] >= n_rows || (i && pr[i] < pr[i - 1]) || parts[i + 1] < parts[i]) return fail(cudaErrorInvalidValue);
            rowptr[pr[i] + 1] += parts[i + 1] - parts[i];
        }
        for (uint32_t r = 0; r < n_rows; r++) rowptr[r + 1] += rowptr[r];
    } else if (parts_are_rows) {
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
        n_rows = n_parts, rowptr = parts;
    } else return fail(cudaErrorInvalidValue);
    if (parts[0] != 0 || parts[n_parts] > nnz || rowptr[n_rows] < parts[n_parts]) return fail(cudaErrorInvalidValue);

    // segments: the chain's parts, long ones cut into equal slices (GS_MERGE_CAP non-zeros at most)
    const int LG = p.N < 32 ? p.N : 32;
    const uint32_t cap = (uint32_t)std::max(LG, env_int("GS_MERGE_CAP", LG == 32 ? 512 : 32 * LG));
    std::vector<uint32_t> seg_nz, seg_row;
    for (uint32_t i = 0; i < n_parts; i++) {
        const uint32_t a = parts[i], b = parts[i + 1];
        if (b < a) return fail(cudaErrorInvalidValue);
        if (b == a) continue;
        const uint32_t pieces = (b - a + cap - 1) / cap;
        for (uint32_t q = 0; q < pieces; q++) {
            const uint32_t lo = a + (uint32_t)((uint64_t)(b - a) * q / pieces);
            seg_nz.push_back(lo);
            seg_row.push_back((uint32_t)(std::upper_bound(rowptr.begin(), rowptr.end(), lo) - rowptr.begin()) - 1);
        }
    }
    seg_nz.push_back(parts[n_parts]);
    pl->n_seg = (uint32_t)seg_row.size(), pl->n_rows = n_rows, pl->N = p.N, pl->part_bytes = ((uint64_t)n_parts + 1) * 4;
    if (pl->n_seg == 0) return fail(cudaErrorInvalidValue);
    auto up = [&](uint32_t **dst, const std::vector<uint32_t> &v) {
        cudaError_t e2 = cudaMalloc((void **)dst, v.size() * 4 + 16);
        return e2 != cudaSuccess ? e2 : cudaMemcpyAsync(*dst, v.data(), v.size() * 4, cudaMemcpyHostToDevice, s);
    };
    if ((e = up(&pl->seg_nz, seg_nz)) != cudaSuccess || (e = up(&pl->seg_row, seg_row)) != cudaSuccess || (e = up(&pl->rowptr, rowptr)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
    snprintf(pl->name, sizeof pl->name, "k_b200_merge<N%d,%s,%s,%d lanes/segment,%u segments of <=%u nnz from %u parts>", p.N, p.a.col_is_u16 ? "u16" : "u32", p.a.val_is_f16 ? "f16" : "f32",
             LG, pl->n_seg, cap, n_parts);
    *out = pl;
    return cudaSuccess;
}

cudaError_t b200_merge_launch(const b200_merge_plan *pl, const launch_params &p, cudaStream_t s)
{
    merge_args a;
    a.col = p.a.col, a.val = p.a.val, a.B = p.B, a.C = p.C, a.seg_nz = pl->seg_nz, a.seg_row = pl->seg_row, a.rowptr = pl->rowptr, a.perm = p.a.perm;
    a.n_seg = pl->n_seg, a.begin_row = p.a.begin_row, a.N = pl->N;
    const void *fn = p.a.col_is_u16 ? (p.a.val_is_f16 ? pick<uint16_t, __half>(pl->N) : pick<uint16_t, float>(pl->N)) : (p.a.val_is_f16 ? pick<uint32_t, __half>(pl->N) : pick<uint32_t, float>(pl->N));
    if (!fn) return cudaErrorInvalidValue;
    const int LG = pl->N < 32 ? pl->N : 32;
    const uint64_t threads = (uint64_t)pl->n_seg * LG;
    const unsigned grid = (unsigned)((threads + MERGE_WARPS * 32 - 1) / (MERGE_WARPS * 32));
    void *args[] = {&a};
    return cudaLaunchKernel(fn, dim3(grid), dim3(MERGE_WARPS * 32), args, 0, s);
}

void b200_merge_free(b200_merge_plan *pl)
{
    if (!pl) return;
    cudaFree(pl->seg_nz), cudaFree(pl->seg_row), cudaFree(pl->rowptr);
    delete pl;
}

const char *b200_merge_kernel_name(const b200_merge_plan *pl) { return pl ? pl->name : ""; }

// what one launch streams: A once, the segment table and the row pointer, B and C once each
uint64_t b200_merge_bytes(const b200_merge_plan *pl, uint64_t nnz, uint64_t idx_bytes, uint64_t val_bytes, uint64_t rows, uint64_t cols)
{
    return nnz * (idx_bytes + val_bytes) + ((uint64_t)pl->n_seg * 2 + 1 + pl->n_rows + 1) * 4 + (cols + rows) * pl->N * val_bytes;
}

} // namespace gsd
