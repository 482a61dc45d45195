// gs_* entry points of include/gsparse_b200.h: device context, array upload, kernel plans,
// launches and timing.  Replaces the reference's executor (executor.cc:6-105: write text files,
// fork nvcc, fork a.out, read perf_result) with in-process launches on one CUDA stream.
#include "../capi_internal.hpp"
#include "common.cuh"
#include "nccl_dl.hpp"
#include "transform.hpp"
#include "sp24_build.hpp"
#include "../host/arr_optimization.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <string>
#include <vector>

using namespace gs;

#define CUDA_TRY(expr)                                                                                         \
    do {                                                                                                       \
        cudaError_t e__ = (expr);                                                                              \
        if (e__ != cudaSuccess) throw gs::Error(std::string("CUDA: ") + cudaGetErrorString(e__) + " at " #expr); \
    } while (0)

#define GS_API_BEGIN try {
#define GS_API_END                                                                                             \
    return GS_OK;                                                                                              \
    }                                                                                                          \
    catch (...) { return gs::translate_exception(); }

struct dev_array {
    void *ptr = nullptr;
    uint64_t len = 0;
    int dtype = GS_U32;   // device dtype
    uint64_t max_value = 0;
    uint64_t align = 0;   // gcd of all entries (offset arrays): the row alignment vector loads may rely on
    int model = 0;        // gs::arr_compress_type of the closed form the array follows (0 = none), fitted at upload
    uint64_t model_coefficient = 0, model_intercept = 0;
    bool iota = false;    // entry i == i for every i: a first_row_indices array of a row level with one row per block
};

struct gs_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_in = nullptr, copy_out = nullptr; // gs_spmm_host_async: H2D and D2H beside the launch stream (lazy)
    // gs_spmm_host_async rotates its launches over `stream` and `launch_extra[1 ...]`: consecutive steps use different plans, and a B200
    // launch alone fills one CTA slot per SM of two, so the kernels of two steps run side by side instead of one after the other
    static constexpr int MAX_LAUNCH = 4;
    cudaStream_t launch_extra[MAX_LAUNCH] = {nullptr, nullptr, nullptr, nullptr}; // [0] unused: index 0 is `stream`
    cudaEvent_t ev_join[MAX_LAUNCH] = {nullptr, nullptr, nullptr, nullptr};
    unsigned async_turn = 0;
    bool chain_extra[MAX_LAUNCH] = {false, false, false, false}, dirty_extra[MAX_LAUNCH] = {false, false, false, false};
    // everything this library puts on `stream` comes after what gs_spmm_host_async has put on the other launch streams
    void join_async()
    {
        for (int i = 1; i < MAX_LAUNCH; i++) {
            if (!dirty_extra[i]) continue;
            dirty_extra[i] = false;
            if (cudaEventRecord(ev_join[i], launch_extra[i]) == cudaSuccess) cudaStreamWaitEvent(stream, ev_join[i], 0);
        }
    }
    // true while the last operation this library put on `stream` was a B200 row kernel launch: only then may the
    // next row launch use programmatic dependent launch (it reads A and the panel before griddepcontrol.wait)
    bool row_chain = false;
    cudaStream_t mem_stream()
    {
        row_chain = false;
        join_async();
        return stream;
    }
    std::map<std::string, dev_array> arrays;
    gsd::nccl::comm_t comm = nullptr; // gs_comm_init: this rank's communicator for the optional all-gather of C
    int comm_rank = -1, comm_size = 0;
};

struct gs_plan {
    gs_ctx *ctx = nullptr;
    gs_plan_desc desc{};
    int N = 0;
    int half = 0;
    int kernel = 0; // gs_family actually launched
    gsd::launch_params lp;
    gsd::b200_row_plan *row_plan = nullptr;
    gsd::b200_tile_plan *tile_plan = nullptr;
    gsd::b200_sp24_plan *sp24_plan = nullptr;
    gsd::b200_merge_plan *merge_plan = nullptr;
    gsd::b200_nnz_plan *nnz_plan = nullptr;
    void *B_dev = nullptr;  // owned
    const void *B_use = nullptr;
    void *C_dev = nullptr;
    void *C_ext = nullptr;  // gs_plan_set_output_device: the caller's buffer (not owned); launches write here instead of C_dev
    void *C_out() const { return C_ext ? C_ext : C_dev; }
    void *B_pinned = nullptr, *C_pinned = nullptr; // staging for pageable host buffers is NOT used; these stay null
    uint64_t offset_bytes = 0; // bytes of the offset / bitmap arrays the launched kernel reads
    uint64_t model_bytes = 0; // of chain_offset_bytes: arrays a closed-form model reproduces
    uint64_t chain_offset_bytes = 0; // the same for the chain's own family (desc.family): SURVEY.md 8(d)'s offset term, independent of the kernel chosen
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_b = nullptr, ev_k = nullptr, ev_c = nullptr; // async round trip: B landed, kernel done, C read back
    int last_async = -1;                                        // the launch stream the plan's previous gs_spmm_host_async used
    size_t esz() const { return half ? 2 : 4; }
};

static uint64_t gcd64(uint64_t a, uint64_t b)
{
    while (b) {
        uint64_t t = a % b;
        a = b, b = t;
    }
    return a;
}

static bool key_has(const std::string &key, const char *s) { return key.find(s) != std::string::npos; }

// host array (any unsigned width, f32, f64 or f16) -> the device's canonical dtype for that key
static void upload(gs_ctx *ctx, const std::string &key, const void *host, uint64_t len, int dtype, int half)
{
    GS_REQUIRE(host && len > 0, "empty array " + key);
    CUDA_TRY(cudaSetDevice(ctx->device));
    dev_array d;
    d.len = len;
    std::vector<unsigned char> staged;
    const bool is_val = key_has(key, "nz_vals");
    if (is_val) {
        GS_REQUIRE(dtype == GS_F32 || dtype == GS_F64 || dtype == GS_F16, "values must be f16, f32 or f64");
        d.dtype = half ? GS_F16 : GS_F32;
        staged.resize(len * (half ? 2 : 4));
        for (uint64_t i = 0; i < len; i++) {
            float v = dtype == GS_F32 ? ((const float *)host)[i] : dtype == GS_F64 ? (float)((const double *)host)[i] : __half2float(((const __half *)host)[i]);
            if (half) ((__half *)staged.data())[i] = __float2half_rn(v);
            else ((float *)staged.data())[i] = v;
        }
    } else {
        GS_REQUIRE(dtype == GS_U8 || dtype == GS_U16 || dtype == GS_U32 || dtype == GS_U64, "index arrays must be unsigned integers");
        std::vector<uint64_t> w(len);
        uint64_t mx = 0, g = 0;
        bool iota = true;
        for (uint64_t i = 0; i < len; i++) {
            uint64_t v = dtype == GS_U8 ? ((const uint8_t *)host)[i] : dtype == GS_U16 ? ((const uint16_t *)host)[i] : dtype == GS_U32 ? ((const uint32_t *)host)[i] : ((const uint64_t *)host)[i];
            w[i] = v;
            mx = v > mx ? v : mx;
            g = gcd64(g, v);
            iota = iota && v == i;
        }
        d.max_value = mx, d.align = g, d.iota = iota;
        if (!key_has(key, "nz_col_indices") && !key_has(key, "nz_row_indices")) { // offset arrays only: the streams never follow a formula
            const gs::arr_model m = gs::fit_arr_model(w, (uint64_t)gs::get_config().as_integer("BRANCH_COMPRESS_MAX_SIZE"));
            d.model = (int)m.type, d.model_coefficient = m.coefficient, d.model_intercept = m.intercept;
        }
        const bool is_col = key_has(key, "nz_col_indices");
        const bool is_bits = key_has(key, "thread_bit_map") || key_has(key, "segment_empty_flag") || (key_has(key, "bit_map_of_thread") && key_has(key, "WARP_META"));
        if (is_bits) d.dtype = GS_U64;
        else if (is_col && mx <= 65535ull) d.dtype = GS_U16;
        else {
            GS_REQUIRE(mx <= 4294967295ull, key + " does not fit 32-bit device offsets");
            d.dtype = GS_U32;
        }
        size_t es = d.dtype == GS_U64 ? 8 : d.dtype == GS_U16 ? 2 : 4;
        staged.resize(len * es);
        for (uint64_t i = 0; i < len; i++) {
            if (es == 8) ((uint64_t *)staged.data())[i] = w[i];
            else if (es == 2) ((uint16_t *)staged.data())[i] = (uint16_t)w[i];
            else ((uint32_t *)staged.data())[i] = (uint32_t)w[i];
        }
    }
    auto it = ctx->arrays.find(key);
    if (it != ctx->arrays.end()) {
        cudaFree(it->second.ptr);
        ctx->arrays.erase(it);
    }
    // 16 spare bytes so that 128-bit loads of the last partial vector stay inside the allocation
    CUDA_TRY(cudaMalloc(&d.ptr, staged.size() + 16));
    CUDA_TRY(cudaMemcpyAsync(d.ptr, staged.data(), staged.size(), cudaMemcpyHostToDevice, ctx->mem_stream()));
    CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->arrays[key] = d;
}

static const dev_array *find_arr(const gs_ctx *ctx, const char *pos, const char *name, int id)
{
    auto it = ctx->arrays.find(std::string(pos) + "_" + name + "_" + std::to_string(id));
    return it == ctx->arrays.end() ? nullptr : &it->second;
}

static size_t dsize(int dt) { return dt == GS_U64 || dt == GS_F64 ? 8 : dt == GS_U16 || dt == GS_F16 ? 2 : dt == GS_U8 ? 1 : 4; }

static void bind_arrays(gs_plan *p)
{
    gs_ctx *ctx = p->ctx;
    const int id = p->desc.sub_matrix_id;
    gsd::dev_arrays &a = p->lp.a;
    a = gsd::dev_arrays();
    const dev_array *col = find_arr(ctx, "GLOBAL_META", "nz_col_indices", id), *val = find_arr(ctx, "GLOBAL_META", "nz_vals", id);
    GS_REQUIRE(col && val, "nz_col_indices / nz_vals of the sub-matrix are not uploaded");
    GS_REQUIRE(col->len == val->len, "column and value arrays differ in length");
    GS_REQUIRE((val->dtype == GS_F16) == (p->half != 0), "values were uploaded with a different HALF setting than the plan");
    a.col = col->ptr, a.val = val->ptr, a.col_is_u16 = col->dtype == GS_U16, a.val_is_f16 = val->dtype == GS_F16;
    a.nnz = col->len;
    GS_REQUIRE(col->max_value < p->desc.cols, "a column index exceeds the matrix width");
    a.rows = (uint32_t)p->desc.rows, a.cols = (uint32_t)p->desc.cols;
    auto u32 = [&](const char *pos, const char *name, uint64_t *len = nullptr) -> const uint32_t * {
        const dev_array *d = find_arr(ctx, pos, name, id);
        if (!d) return nullptr;
        GS_REQUIRE(d->dtype == GS_U32, std::string(name) + " has an unexpected device type");
        if (len) *len = d->len;
        return (const uint32_t *)d->ptr;
    };
    auto u64 = [&](const char *pos, const char *name) -> const uint64_t * {
        const dev_array *d = find_arr(ctx, pos, name, id);
        if (!d) return nullptr;
        GS_REQUIRE(d->dtype == GS_U64, std::string(name) + " has an unexpected device type");
        return (const uint64_t *)d->ptr;
    };
    uint64_t n = 0;
    if (p->desc.has_row_permutation) {
        a.perm = u32("GLOBAL_META", "original_nz_row_indices");
        GS_REQUIRE(a.perm, "original_nz_row_indices is not uploaded");
    }
    a.tb_first_row = u32("TBLOCK_META", "first_row_indices");
    a.tb_first_nz = u32("TBLOCK_META", "first_nz_indices", &n);
    if (a.tb_first_nz) a.n_tb = (uint32_t)(n - 1);
    a.tb_first_bmt = u32("TBLOCK_META", "first_BMT_indices", &n);
    if (a.tb_first_bmt) a.n_tb = (uint32_t)(n - 1);
    a.w_first_row = u32("WARP_META", "first_row_indices");
    a.w_first_nz = u32("WARP_META", "first_nz_indices", &n);
    if (a.w_first_nz) a.n_w = (uint32_t)(n - 1);
    a.w_first_bmt = u32("WARP_META", "first_BMT_indices");
    a.w_bit_map = u64("WARP_META", "bit_map_of_thread");
    a.t_first_row = u32("THREAD_META", "first_row_indices");
    a.t_first_row_we = u32("THREAD_META", "first_row_indices_without_ending", &n);
    if (a.t_first_row_we) a.n_t = (uint32_t)n;
    a.t_first_nz = u32("THREAD_META", "first_nz_indices", &n);
    if (a.t_first_nz) a.n_t = (uint32_t)(n - 1);
    else {
        a.t_first_nz = u32("THREAD_META", "first_nz_indices_relative_to_BMW", &n);
        if (a.t_first_nz) a.t_nz_relative = 1, a.n_t = (uint32_t)n;
    }
    a.t_bit_map = u64("THREAD_META", "thread_bit_map");
    a.t_head = u32("THREAD_META", "bit_map_of_thread");
    a.t_seg_ptr = u32("THREAD_META", "segment_ptr");
    a.t_seg_offset = u32("THREAD_META", "segment_offset");
    a.t_seg_empty_flag = u64("THREAD_META", "segment_empty_flag");
    a.t_seg_empty_rows = u32("THREAD_META", "segment_empty_row_indices");
    const dev_array *br = find_arr(ctx, "GLOBAL_META", "begin_row_index", id);
    (void)br; // sub-matrix 0 of an undivided matrix starts at row 0; divided matrices upload begin_row_index
}

static void require(bool ok, const char *what)
{
    if (!ok) throw gs::Error(std::string("the uploaded arrays do not include ") + what + " needed by this kernel family");
}

static uint64_t arr_bytes(const gs_ctx *ctx, const char *pos, const char *name, int id)
{
    const dev_array *d = find_arr(ctx, pos, name, id);
    return d ? d->len * dsize(d->dtype) : 0;
}

// the row level the B200 row / tile / 2:4 kernels read (the same choice their prepare functions make), or null
static const char *row_level(const gsd::dev_arrays &a)
{
    return (a.t_first_nz && a.t_first_row && !a.t_nz_relative) ? "THREAD_META" : (a.w_first_nz && a.w_first_row) ? "WARP_META" : (a.tb_first_nz && a.tb_first_row) ? "TBLOCK_META" : nullptr;
}

// Those kernels treat block i of the row level as row first_row_indices[i] and STORE its sum: that is only SpMM when every block is one
// whole row, i.e. first_row_indices = 0, 1, 2, ... (row-direction blocking with block size 1).  An nnz-direction chain (several blocks per
// row, which the reference refuses to pair with a total reduction) or a row block > 1 would silently overwrite rows.
static void require_one_row_per_block(const gs_plan *p, const char *who)
{
    const char *lvl = row_level(p->lp.a);
    if (!lvl) throw gs::Error(std::string(who) + ": needs a row-direction chain (first_row_indices / first_nz_indices at one level)");
    const dev_array *fr = find_arr(p->ctx, lvl, "first_row_indices", p->desc.sub_matrix_id);
    if (!fr || !fr->iota) throw gs::Error(std::string(who) + ": " + lvl + " first_row_indices is not 0, 1, 2, ...: the kernel needs one whole row per block (row-direction blocking, block size 1)");
}

// the reference-shaped total families reduce a block to one register per dense column and store it as ONE row (kernels_ref.cu): a block of
// several rows (balanced blocking, row block > 1) needs the range kernel instead
static void require_iota(const gs_plan *p, const char *lvl, const char *who)
{
    const dev_array *fr = find_arr(p->ctx, lvl, "first_row_indices", p->desc.sub_matrix_id);
    if (fr && !fr->iota) throw gs::Error(std::string(who) + ": " + lvl + " first_row_indices is not 0, 1, 2, ...: a block of several rows cannot be stored as one row; use b200_merge (or auto)");
}

static void select_kernel(gs_plan *p, int family)
{
    const gsd::dev_arrays &a = p->lp.a;
    const gs_ctx *ctx = p->ctx;
    const int id = p->desc.sub_matrix_id;
    uint64_t ob = 0;
    auto add = [&](const char *pos, const char *name) { ob += arr_bytes(ctx, pos, name, id); };
    switch (family) {
    case GS_FAMILY_THREAD_TOTAL:
        require(a.t_first_row && a.t_first_nz && !a.t_nz_relative, "THREAD first_row_indices/first_nz_indices");
        require_iota(p, "THREAD_META", "thread_total");
        add("THREAD_META", "first_row_indices"), add("THREAD_META", "first_nz_indices");
        break;
    case GS_FAMILY_WARP_TOTAL:
        require(a.w_first_row && a.w_first_nz, "WARP first_row_indices/first_nz_indices");
        require_iota(p, "WARP_META", "warp_total");
        add("WARP_META", "first_row_indices"), add("WARP_META", "first_nz_indices");
        break;
    case GS_FAMILY_TBLOCK_TOTAL:
        require(a.tb_first_row && a.tb_first_nz, "TBLOCK first_row_indices/first_nz_indices");
        require_iota(p, "TBLOCK_META", "block_total");
        add("TBLOCK_META", "first_row_indices"), add("TBLOCK_META", "first_nz_indices");
        break;
    case GS_FAMILY_WARP_BITMAP:
        require(a.w_first_nz && a.w_bit_map && a.t_first_row_we && a.t_first_nz, "WARP first_nz_indices/bit_map_of_thread, THREAD first_row_indices_without_ending/first_nz_indices");
        GS_REQUIRE(p->desc.vector_width >= 1 && p->desc.vector_width <= 64, "VECTOR_WIDTH must be 1..64 for the bit map families");
        add("WARP_META", "first_nz_indices"), add("WARP_META", "bit_map_of_thread"), add("THREAD_META", "first_row_indices_without_ending");
        add("THREAD_META", a.t_nz_relative ? "first_nz_indices_relative_to_BMW" : "first_nz_indices");
        break;
    case GS_FAMILY_TBLOCK_BITMAP:
        require(a.tb_first_bmt && a.t_first_nz && !a.t_nz_relative && a.t_first_row_we && a.t_head && a.t_seg_offset,
                "TBLOCK first_BMT_indices, THREAD first_nz_indices/first_row_indices_without_ending/bit_map_of_thread/segment_offset");
        add("TBLOCK_META", "first_BMT_indices"), add("THREAD_META", "first_nz_indices"), add("THREAD_META", "first_row_indices_without_ending");
        add("THREAD_META", "bit_map_of_thread"), add("THREAD_META", "segment_offset");
        break;
    case GS_FAMILY_THREAD_BITMAP:
    case GS_FAMILY_WARP_SEGMENT:
        require(a.t_first_row && a.t_first_nz && !a.t_nz_relative && a.t_bit_map && a.t_seg_ptr && a.t_seg_empty_flag && a.t_seg_empty_rows,
                "THREAD first_row_indices/first_nz_indices/thread_bit_map/segment_ptr/segment_empty_flag/segment_empty_row_indices");
        GS_REQUIRE(p->desc.vector_width >= 1 && p->desc.vector_width <= 64, "VECTOR_WIDTH must be 1..64 for the bit map families");
        add("THREAD_META", "first_row_indices"), add("THREAD_META", "first_nz_indices"), add("THREAD_META", "thread_bit_map");
        add("THREAD_META", "segment_ptr"), add("THREAD_META", "segment_empty_flag"), add("THREAD_META", "segment_empty_row_indices");
        break;
    case GS_FAMILY_B200_ROW_SMEM: {
        if (p->row_plan) gsd::b200_row_free(p->row_plan), p->row_plan = nullptr;
        require_one_row_per_block(p, "b200_row_smem");
        gsd::launch_params lp = p->lp;
        // row alignment: vector loads need every row offset to be a multiple of 8 (sparse coarsen 8 padding)
        const char *lvl = (a.t_first_nz && a.t_first_row && !a.t_nz_relative) ? "THREAD_META" : (a.w_first_nz && a.w_first_row) ? "WARP_META" : "TBLOCK_META";
        const dev_array *rp = find_arr(ctx, lvl, "first_nz_indices", id);
        require(rp != nullptr, "a row-direction first_nz_indices array");
        lp.vector_width = (rp->align % 8 == 0) ? 8 : 1;
        cudaError_t e = gsd::b200_row_prepare(lp, ctx->sm_count, &p->row_plan, ctx->stream);
        if (e == cudaErrorInvalidValue) throw gs::Error("b200_row_smem: not built for this dense width or chain (needs N in {8,16,32} and a row-direction chain with row block 1)");
        CUDA_TRY(e);
        add(lvl, "first_row_indices"), add(lvl, "first_nz_indices");
        break;
    }
    case GS_FAMILY_B200_TCGEN05: {
        if (p->tile_plan) gsd::b200_tile_free(p->tile_plan), p->tile_plan = nullptr;
        GS_REQUIRE(p->half, "b200_tcgen05: the tensor-core path multiplies fp16 operands (HALF); fp32 values stay on the gather kernels");
        require_one_row_per_block(p, "b200_tcgen05");
        cudaError_t e = gsd::b200_tile_prepare(p->lp, ctx->sm_count, &p->tile_plan, ctx->stream);
        if (e == cudaErrorInvalidValue) throw gs::Error("b200_tcgen05: needs N in {16,32,64,128,256}, HALF values and a row-direction chain with row block 1");
        CUDA_TRY(e);
        break;
    }
    case GS_FAMILY_B200_SP24: {
        if (p->sp24_plan) gsd::b200_sp24_free(p->sp24_plan), p->sp24_plan = nullptr;
        GS_REQUIRE(p->half, "b200_sp24: the sparse tensor-core path multiplies fp16 operands (HALF); fp32 values stay on the gather kernels");
        require_one_row_per_block(p, "b200_sp24");
        cudaError_t e = gsd::b200_sp24_prepare(p->lp, ctx->sm_count, true, &p->sp24_plan, ctx->stream);
        if (e == cudaErrorInvalidValue) throw gs::Error("b200_sp24: needs N in {8,16,32,64,128,256}, HALF values and a row-direction chain with row block 1");
        CUDA_TRY(e);
        break;
    }
    case GS_FAMILY_B200_MERGE: {
        if (p->merge_plan) gsd::b200_merge_free(p->merge_plan), p->merge_plan = nullptr;
        // the finest level the chain built gives the ranges; the row pointer comes from nz_row_indices when the chain uploaded it (merge-path
        // and balanced chains register it), from the row of every chunk on the column-direction chains (every chunk lies in one row), or
        // from the ranges themselves when every block is one row
        const char *lvl = find_arr(ctx, "THREAD_META", "first_nz_indices", id) ? "THREAD_META" : find_arr(ctx, "WARP_META", "first_nz_indices", id) ? "WARP_META" : "TBLOCK_META";
        const dev_array *parts = find_arr(ctx, lvl, "first_nz_indices", id), *rows = find_arr(ctx, "GLOBAL_META", "nz_row_indices", id);
        const dev_array *fr = find_arr(ctx, lvl, "first_row_indices", id), *frwe = find_arr(ctx, lvl, "first_row_indices_without_ending", id);
        require(parts != nullptr && parts->len >= 2, "a first_nz_indices array at the thread, warp or tblock level");
        GS_REQUIRE(parts->dtype == GS_U32 && (!rows || rows->dtype == GS_U32) && (!frwe || frwe->dtype == GS_U32), "b200_merge: unexpected device type of an offset array");
        const bool col_dir = p->desc.family == GS_FAMILY_WARP_BITMAP || p->desc.family == GS_FAMILY_TBLOCK_BITMAP;
        const uint32_t *part_rows = !rows && col_dir && frwe && frwe->len + 1 == parts->len ? (const uint32_t *)frwe->ptr : nullptr;
        const bool parts_are_rows = !rows && !part_rows && fr && fr->iota;
        if (!rows && !part_rows && !parts_are_rows)
            throw gs::Error("b200_merge: needs GLOBAL nz_row_indices uploaded (merge-path and balanced chains register it), or a column-direction chain, or one row per block");
        gsd::launch_params lp = p->lp;
        lp.N = p->N;
        cudaError_t e = gsd::b200_merge_prepare(lp, (const uint32_t *)parts->ptr, (uint32_t)(parts->len - 1), rows ? (const uint32_t *)rows->ptr : nullptr, part_rows, parts_are_rows, &p->merge_plan,
                                                ctx->stream);
        if (e == cudaErrorInvalidValue) throw gs::Error("b200_merge: needs N in {1,2,4,8,16} or a multiple of 32 up to 512, and offsets that cover the stream in order");
        CUDA_TRY(e);
        add(lvl, "first_nz_indices");
        if (rows) add("GLOBAL_META", "nz_row_indices");
        break;
    }
    case GS_FAMILY_B200_NNZ: {
        require(a.t_first_row && a.t_first_nz && !a.t_nz_relative && a.t_bit_map && a.t_seg_ptr && a.t_seg_empty_flag && a.t_seg_empty_rows,
                "THREAD first_row_indices/first_nz_indices/thread_bit_map/segment_ptr/segment_empty_flag/segment_empty_row_indices");
        GS_REQUIRE(p->desc.vector_width >= 1 && p->desc.vector_width <= 64, "VECTOR_WIDTH must be 1..64 for the bit map families");
        gsd::launch_params lp = p->lp;
        lp.N = p->N;
        if (!gsd::b200_nnz_supported(lp)) throw gs::Error("b200_nnz: needs N in {32,64,128} and an nnz-direction bit-map chain (thread_bit_map / warp_segment)");
        if (p->nnz_plan) gsd::b200_nnz_free(p->nnz_plan), p->nnz_plan = nullptr;
        CUDA_TRY(gsd::b200_nnz_prepare(lp, &p->nnz_plan, ctx->stream));
        { // model-driven index compression: first_nz_indices of a padded nnz-direction level is coefficient * i: the kernel computes it
            const dev_array *fn = find_arr(ctx, "THREAD_META", "first_nz_indices", id);
            p->lp.first_nz_step = (fn && fn->model == (int)gs::LINEAR_COMPRESS && fn->model_intercept == 0 && !getenv("GS_NO_INDEX_MODEL")) ? (uint32_t)fn->model_coefficient : 0u;
        }
        add("THREAD_META", "first_row_indices"), add("THREAD_META", "first_nz_indices"), add("THREAD_META", "thread_bit_map");
        add("THREAD_META", "segment_ptr"), add("THREAD_META", "segment_empty_flag"), add("THREAD_META", "segment_empty_row_indices");
        break;
    }
    default: throw gs::Error("kernel family not built: " + std::string(gs_family_name(family)));
    }
    if (p->desc.has_row_permutation && family != GS_FAMILY_B200_TCGEN05 && family != GS_FAMILY_B200_SP24) add("GLOBAL_META", "original_nz_row_indices");
    p->offset_bytes = ob;
    p->kernel = family;
}

// ---- GS_KERNEL_AUTO: which kernel for this chain, matrix, N and value type (smallest useful slice) ----------------
// A cost model in microseconds per launch from the measured rates of this round's and round 1's profiles:
// every candidate that accepts the chain is priced, the cheapest is prepared, and a candidate whose prepare refuses the matrix
// (panel too large for shared memory, remainder blocks too large for a stage ...) falls to the next one.  The chain's own
// reference-shaped family is always the last candidate, so the call cannot fail on a valid chain.
struct auto_choice {
    int family;
    double us;
};

// 128 x 128 tiles of the row-level matrix that hold at least one entry (every row panel counts one at least): one pass over the
// pattern, at plan time
static uint64_t count_nonempty_tiles(const gs_plan *p, const char *lvl)
{
    const gs_ctx *ctx = p->ctx;
    const int id = p->desc.sub_matrix_id;
    const dev_array *rp = find_arr(ctx, lvl, "first_nz_indices", id), *col = find_arr(ctx, "GLOBAL_META", "nz_col_indices", id);
    if (!rp || !col || rp->len < 2) return 0;
    uint64_t n = 0; // counted on the device (sp24_build.cu): the arrays stay where they are
    CUDA_TRY(gsd::count_tiles128_dev(col->ptr, col->dtype == GS_U16, (const uint32_t *)rp->ptr, (uint32_t)(rp->len - 1), (uint32_t)p->desc.cols, col->len, ctx->stream, &n));
    return n;
}

static std::vector<auto_choice> auto_candidates(const gs_plan *p)
{
    const gsd::dev_arrays &a = p->lp.a;
    const gs_ctx *ctx = p->ctx;
    const int N = p->N, id = p->desc.sub_matrix_id;
    const double nnz = (double)a.nnz, rows = (double)p->desc.rows, cols = (double)p->desc.cols;
    const double bw = 6470.2e3; // bytes per microsecond: MEASURED_PEAKS.json hbm_gbs of this pool's B200s
    const double idxb = a.col_is_u16 ? 2 : 4, vb = p->esz();
    const double bc = (cols + rows) * N * vb;
    std::vector<auto_choice> c;
    const char *lvl = row_level(a);
    const dev_array *fr = lvl ? find_arr(ctx, lvl, "first_row_indices", id) : nullptr;
    const bool row_chain = fr && fr->iota;
    const bool bitmap_chain = a.t_first_row && a.t_first_nz && !a.t_nz_relative && a.t_bit_map && a.t_seg_ptr && a.t_seg_empty_flag && a.t_seg_empty_rows;
    if (row_chain) {
        const double mean_len = nnz / (rows > 0 ? rows : 1);
        if (N == 8 || N == 16 || N == 32) {
            // gather kernel: streams 4-8 B per non-zero at ~70 % of the HBM peak while the K x N panel fits shared memory, else its gathers
            // go to L2 (~2.2x); from N = 16 the shared-memory gathers and the FMAs bound it (13.5 M nnz x N = 32: 23 us); short rows
            // (one warp per row) waste lanes below ~64 entries
            const bool panel = cols * N * vb <= 200.0 * 1024;
            double t = (nnz * (idxb + vb) + bc) / (0.70 * bw) * (panel ? 1.0 : 2.2);
            t = std::max(t, nnz * N / 18.8e6);
            if (mean_len < 64) t *= 1.0 + (64 - mean_len) / 64.0;
            // the panel of B in global memory: the gathers decide.  Non-zeros per us by the bytes of one row of B, measured on C5's uniform 65 536^2
            // matrix (655 per row; profiles/r2_c5_sweep_narrow.md): 286 K at 16 B, 165 K at 32, 75 K at 64, 30 K at 128; shorter rows run at 0.6
            // (banded, 129 per row) and 0.3 - 0.4 (C1 stand-in, 37 per row) of that
            if (!panel) {
                const double rb = N * vb, rate = rb <= 16 ? 286e3 : rb <= 32 ? 165e3 : rb <= 64 ? 75e3 : 30e3;
                t = nnz / (rate * std::min(1.0, 0.25 + mean_len / 400.0));
            }
            c.push_back({GS_FAMILY_B200_ROW_SMEM, t + 2.5});
        }
        // the tensor-core layouts are built for weight-sized matrices: their host builders lay the whole rows x cols grid out before
        // empty tiles are dropped, so a 65 536^2 pattern (4.8 GB of image for 172 MB of non-zeros) never gets there
        const bool weight_sized = rows * cols * 1.125 <= 2147483648.0;
        double nonempty128 = std::ceil(rows / 128) * std::ceil(cols / 128);
        if (weight_sized && p->half) nonempty128 = (double)count_nonempty_tiles(p, lvl);
        if (weight_sized && p->half && (N == 16 || N == 32 || N == 64 || N == 128 || N == 256)) {
            // tile kernel: one 2-byte store per non-zero, but every 128 x 64 tile costs its zero fill + tensor-core read whatever it holds:
            // ~3.0 ns per tile over the whole GPU up to N = 64, 3.9 at 128, 7.8 at 256 (profiles/r1_tile_summary.md)
            const double tiles = std::ceil(rows / 128) * std::ceil(cols / 64);
            const double per_tile = N <= 64 ? 3.0e-3 : N == 128 ? 3.9e-3 : 7.8e-3;
            c.push_back({GS_FAMILY_B200_TCGEN05, std::max(tiles * per_tile, (nnz * 4 + bc) / (0.55 * bw)) + 4.0});
        }
        if (weight_sized && p->half && (N == 8 || N == 16 || N == 32 || N == 64 || N == 128 || N == 256)) {
            // 2:4 kernel: 18 432 B per non-empty 128 x 128 tile whatever it holds (counted on the host from the pattern; every row panel keeps one), ~90 % of the HBM peak up to N = 32, then the L2 -> SM intake of B's
            // slices bounds it (profiles/r1_sp24_summary.md); what 2:4 cannot hold (expected share from the density, exact for a
            // 2:4-pruned matrix only after prepare) is multiplied by remainder warps at ~5 M entry-columns per us
            const double d = nnz / (rows * cols);
            const double tiles = std::ceil(rows / 128) * std::ceil(cols / 128);
            const double p_empty = 1.0 - nonempty128 / tiles;
            const double g = N <= 16 ? 1.0 : N == 32 ? 1.1 : N == 64 ? 1.35 : N == 128 ? 2.0 : 4.3;
            // expected entries beyond two per group of four columns under independent keep-probability d
            const double q = std::min(d, 1.0), over = 4 * q * q * q * (1 - q) * 1 + q * q * q * q * 2;
            const double rem = d > 0.45 && d < 0.55 ? 0.0 : rows * cols / 4 * over; // a 2:4-pruned weight sits at exactly 0.5: no remainder
            double t = tiles * (1 - p_empty) * 18432.0 / (0.90 * bw) * g + bc / bw + rem * N / 5.1e6;
            if (N >= 128 && rem > 0) t += (rem * 6) / (0.5 * bw) + 6.0; // separate gather launch
            c.push_back({GS_FAMILY_B200_SP24, t + 1.0});
        }
    }
    if (gsd::b200_merge_supported(N) && (find_arr(ctx, "GLOBAL_META", "nz_row_indices", id) || row_chain || p->desc.family == GS_FAMILY_WARP_BITMAP || p->desc.family == GS_FAMILY_TBLOCK_BITMAP)) {
        // range kernel: A streamed once at ~60 % of the HBM peak, B's rows gathered through L1/L2 (N * value bytes per non-zero); the only
        // B200 kernel for merge-path / balanced / column-direction chains and for wide N on matrices too large for the tensor-core layouts
        const double gather = nnz * N * vb, l2_bw = 2.5 * bw;
        const bool b_in_l2 = cols * N * vb <= 96e6;
        // a segment is walked serially by its lanes; non-zeros per microsecond measured on C5's 65 536^2 uniform matrix (profiles/r2_c5_sweep.md,
        // r2_c5_sweep_narrow.md): the chunk kernel 110 K at N = 1, 83 K at 2, 137 K at 4 (the first width with a 16-byte row), 100 K at 8, 75 K at
        // 16, 68 K at 32, fp32 and fp16 alike (C1 stand-in: 100 / 89 / 77 K at N = 8 / 16 / 32); the group kernel 35.5 K at N = 64, 29 K at 128,
        // 5.5 K at 512, where fp16 values load 2 bytes per lane and run at about half of that
        static const double mn[] = {1, 2, 4, 8, 16, 32, 64, 128, 256, 512}, mr[] = {110e3, 83e3, 137e3, 100e3, 75e3, 68e3, 58e3, 28.6e3, 13.9e3, 5.5e3};
        double rate = mr[9];
        for (int i = 0; i < 10; i++)
            if (N <= mn[i]) {
                rate = mr[i];
                break;
            }
        // the chunk variant also takes rows of B up to 256 bytes (fp32 N = 64: 727 us on the uniform matrix, 58 K per us; fp16 N = 64 / 128: 796 / 1 000 us)
        if (p->half && N == 64) rate = 53e3;
        else if (p->half && N == 128) rate = 42e3;
        else if (p->half && N >= 256) rate *= 0.5;
        if (p->half && N <= 8) rate *= 0.75; // one lane per segment (C1 stand-in: 24.8 us against 17.8 in fp32)
        const double walk = nnz / rate;
        c.push_back({GS_FAMILY_B200_MERGE, std::max({(nnz * (idxb + vb) + bc) / (0.60 * bw), gather / (b_in_l2 ? l2_bw : 0.9 * bw), walk}) + 3.0});
    }
    if (bitmap_chain && (N == 32 || N == 64 || N == 128))
        c.push_back({GS_FAMILY_B200_NNZ, (nnz * (idxb + vb) + bc) / (0.25 * bw) + 3.0}); // bound by the gathers of B rows through L2
    // the chain's own reference-shaped family competes where it was measured to win: narrow N on the row-parallel total chains (same
    // sweep: thread_total 124 K non-zeros per us at N <= 2, 80 K at 4, 67 K at 8, halving per doubling of N; warp_total 170 K at N = 1,
    // 155 K at 2, 76 K at 4)
    if (row_chain && (p->desc.family == GS_FAMILY_THREAD_TOTAL || p->desc.family == GS_FAMILY_WARP_TOTAL)) {
        const bool tt = p->desc.family == GS_FAMILY_THREAD_TOTAL;
        const double r = tt ? (N <= 2 ? 124e3 : N <= 4 ? 80e3 : 67e3 * 8.0 / N) : (N <= 1 ? 170e3 : N <= 2 ? 155e3 : 76e3 * 4.0 / N);
        c.push_back({p->desc.family, nnz / r + 3.0});
    }
    std::sort(c.begin(), c.end(), [](const auto_choice &x, const auto_choice &y) { return x.us < y.us; });
    c.push_back({p->desc.family, 1e30});
    return c;
}

static void select_auto(gs_plan *p)
{
    std::string why;
    for (const auto_choice &c : auto_candidates(p)) {
        try {
            select_kernel(p, c.family);
            return;
        } catch (const gs::Error &e) {
            why += std::string(why.empty() ? "" : "; ") + gs_family_name(c.family) + ": " + e.what();
            cudaGetLastError();
        }
    }
    throw gs::Error("auto: no kernel accepts this chain (" + why + ")");
}

static bool uses_atomics(int family)
{
    return family == GS_FAMILY_WARP_BITMAP || family == GS_FAMILY_TBLOCK_BITMAP || family == GS_FAMILY_THREAD_BITMAP || family == GS_FAMILY_WARP_SEGMENT ||
           family == GS_FAMILY_B200_NNZ || family == GS_FAMILY_B200_MERGE;
}

// which: 0 = ctx->stream, else one of gs_spmm_host_async's other launch streams
static void enqueue(gs_plan *p, int which = 0)
{
    GS_REQUIRE(p->B_use != nullptr, "gs_set_dense has not been called");
    gs_ctx *ctx = p->ctx;
    gsd::launch_params lp = p->lp;
    lp.B = p->B_use, lp.C = p->C_out(), lp.N = p->N;
    if (which == 0) ctx->join_async();
    else ctx->dirty_extra[which] = true;
    cudaStream_t st = which ? ctx->launch_extra[which] : ctx->stream;
    bool &chain = which ? ctx->chain_extra[which] : ctx->row_chain;
    static const bool no_pdl = getenv("GS_NO_PDL") != nullptr;
    const bool b200_chain = p->kernel == GS_FAMILY_B200_ROW_SMEM || p->kernel == GS_FAMILY_B200_TCGEN05 || p->kernel == GS_FAMILY_B200_SP24;
    if (!b200_chain) chain = false;
    if (uses_atomics(p->kernel) && !(p->kernel == GS_FAMILY_B200_MERGE && !gsd::b200_merge_needs_zero(p->merge_plan))) CUDA_TRY(cudaMemsetAsync(p->C_out(), 0, (size_t)p->desc.rows * p->N * p->esz(), st));
    // the attribute is always set for the B200 kernels; what differs is whether the launch may read B at once (the chain is intact: the
    // operation before it on this stream was one of ours and B did not change hands) or must wait for the predecessor on the device
    lp.pdl = !no_pdl, lp.dep_wait = !chain;
    if (p->kernel == GS_FAMILY_B200_ROW_SMEM) CUDA_TRY(gsd::b200_row_launch(p->row_plan, lp, st));
    else if (p->kernel == GS_FAMILY_B200_TCGEN05) CUDA_TRY(gsd::b200_tile_launch(p->tile_plan, lp, st));
    else if (p->kernel == GS_FAMILY_B200_SP24) CUDA_TRY(gsd::b200_sp24_launch(p->sp24_plan, lp, st));
    else if (p->kernel == GS_FAMILY_B200_NNZ) CUDA_TRY(gsd::b200_nnz_launch(lp, p->nnz_plan, st));
    else if (p->kernel == GS_FAMILY_B200_MERGE) CUDA_TRY(gsd::b200_merge_launch(p->merge_plan, lp, st));
    else CUDA_TRY(gsd::launch_reference_family(p->kernel, lp, st));
    if (b200_chain) chain = true;
}

extern "C" {

gs_status gs_create(gs_ctx **out, int device_id)
{
    GS_API_BEGIN
    GS_REQUIRE(out, "null argument");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) return fail(GS_ERR_CUDA, std::string("CUDA: no usable device (") + cudaGetErrorString(e) + "); the SpMM path has no CPU fallback");
    GS_REQUIRE(device_id >= 0 && device_id < n, "device id out of range");
    CUDA_TRY(cudaSetDevice(device_id));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device_id));
    if (prop.major != 10) return fail(GS_ERR_CUDA, std::string("CUDA: device is sm_") + std::to_string(prop.major) + std::to_string(prop.minor) + ", this library is built for sm_100a only");
    gs_ctx *c = new gs_ctx;
    c->device = device_id, c->sm_count = prop.multiProcessorCount;
    e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete c;
        CUDA_TRY(e);
    }
    *out = c;
    GS_API_END
}

void gs_destroy(gs_ctx *ctx)
{
    if (!ctx) return;
    gs_comm_destroy(ctx);
    cudaSetDevice(ctx->device);
    for (auto &kv : ctx->arrays) cudaFree(kv.second.ptr);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
    if (ctx->copy_out) cudaStreamDestroy(ctx->copy_out);
    for (int i = 1; i < gs_ctx::MAX_LAUNCH; i++) {
        if (ctx->launch_extra[i]) cudaStreamDestroy(ctx->launch_extra[i]);
        if (ctx->ev_join[i]) cudaEventDestroy(ctx->ev_join[i]);
    }
    delete ctx;
}

void *gs_stream(gs_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

gs_status gs_upload_array(gs_ctx *ctx, const char *key, const void *host, uint64_t len, int dtype, int half)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && key, "null argument");
    upload(ctx, key, host, len, dtype, half);
    GS_API_END
}

gs_status gs_upload_session(gs_ctx *ctx, const gsh_session *s, int half) { return gs_upload_session_as(ctx, s, half, -1); }

gs_status gs_upload_session_as(gs_ctx *ctx, const gsh_session *s, int half, int as_id)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && s, "null argument");
    const kernel_plan &kp = s->cg->plan();
    // a session replaces the whole array set of its sub-matrix id: arrays of an earlier chain under the
    // same id must not leak into the next plan's bindings
    {
        CUDA_TRY(cudaSetDevice(ctx->device));
        const std::string suffix = "_" + std::to_string(as_id >= 0 ? as_id : s->cg->get_sub_matrix_id());
        for (auto it = ctx->arrays.begin(); it != ctx->arrays.end();) {
            const std::string &k = it->first;
            if (k.size() > suffix.size() && k.compare(k.size() - suffix.size(), suffix.size(), suffix) == 0) {
                cudaFree(it->second.ptr);
                it = ctx->arrays.erase(it);
            } else ++it;
        }
    }
    for (const auto &d : kp.deps) {
        const universal_array &a = s->ms->get(d.pos, d.name, d.sub_matrix_id);
        std::string key = meta_data_set::item_name(d.pos, d.name, as_id >= 0 ? as_id : d.sub_matrix_id);
        if (a.is_float()) upload(ctx, key, a.f.data(), a.f.size(), GS_F64, half);
        else upload(ctx, key, a.u.data(), a.u.size(), GS_U64, half);
    }
    // the row of every stored non-zero travels too, whatever the chain registered: it is what lets the range kernel (b200_merge) run on
    // any chain.  It is not one of the chain's kernel inputs, so SURVEY 8(d)'s algorithmic bytes do not count it.
    {
        const int sid = s->cg->get_sub_matrix_id();
        bool listed = false;
        for (const auto &d : kp.deps) listed = listed || d.name == "nz_row_indices";
        if (!listed && s->ms->is_exist(GLOBAL_META, "nz_row_indices", sid)) {
            const universal_array &a = s->ms->get(GLOBAL_META, "nz_row_indices", sid);
            upload(ctx, meta_data_set::item_name(GLOBAL_META, "nz_row_indices", as_id >= 0 ? as_id : sid), a.u.data(), a.u.size(), GS_U64, half);
        }
    }
    GS_API_END
}

// ---- (f2) the chain built on the device ----------------------------------------------------------------------------------------------------------
static void drop_arrays_of_id(gs_ctx *ctx, int id)
{
    const std::string suffix = "_" + std::to_string(id);
    for (auto it = ctx->arrays.begin(); it != ctx->arrays.end();) {
        const std::string &k = it->first;
        if (k.size() > suffix.size() && k.compare(k.size() - suffix.size(), suffix.size(), suffix) == 0) {
            cudaFree(it->second.ptr);
            it = ctx->arrays.erase(it);
        } else ++it;
    }
}

static void register_built(gs_ctx *ctx, gsd::transform_out &out, int as_id)
{
    for (gsd::built_array &a : out.arrays) {
        dev_array da;
        da.ptr = a.ptr, da.len = a.len, da.dtype = a.dtype, da.max_value = a.max_value, da.align = a.align, da.iota = a.iota;
        if (a.linear) da.model = (int)gs::LINEAR_COMPRESS, da.model_coefficient = a.linear_step, da.model_intercept = 0;
        ctx->arrays[a.pos + "_" + a.name + "_" + std::to_string(as_id)] = da;
        a.ptr = nullptr; // the context owns it now
    }
    out.arrays.clear();
}

gs_status gs_transform_device(gs_ctx *ctx, int family, uint64_t rows, uint64_t cols, uint64_t nnz, const uint32_t *row_dev, const uint32_t *col_dev, const float *val_dev, int N,
                              int sparse_coarsen, int coarsen, int half, int as_id, gs_plan_desc *desc_out)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && desc_out && row_dev && col_dev && val_dev, "null argument");
    GS_REQUIRE(N >= 1 && sparse_coarsen >= 1 && coarsen >= 1, "N, sparse_coarsen and coarsen must be >= 1");
    GS_REQUIRE(as_id >= 0, "as_id must be >= 0");
    if (family < GS_FAMILY_THREAD_TOTAL || family > GS_FAMILY_TBLOCK_BITMAP)
        return fail(GS_ERR_UNSUPPORTED, "gs_transform_device builds the seven driver chains (the reference families); other operator sequences are built on the host (gsh_session_add_and_run)");
    CUDA_TRY(cudaSetDevice(ctx->device));
    ctx->join_async();
    const double bound = (double)gs::get_config().as_integer("PADDING_RATE_UP_BOUND");
    gsd::coo_dev m;
    m.row = row_dev, m.col = col_dev, m.val = val_dev, m.rows = rows, m.cols = cols, m.nnz = nnz;
    gsd::transform_out out;
    // the launch shape the driver would have passed to grid_block_operator (token_test.cc:1042-1077, :1215-1232, :1342-1375)
    gs_plan_desc d;
    memset(&d, 0, sizeof d);
    d.family = family, d.sub_matrix_id = as_id, d.sparse_coarsen = sparse_coarsen, d.coarsen = coarsen;
    d.rows = rows, d.cols = cols, d.origin_nnz = nnz;
    unsigned x = 0, y = 0, grid_x = 0;
    cudaError_t e = cudaSuccess;
    if (family == GS_FAMILY_THREAD_TOTAL) {
        x = (unsigned)std::min(N / coarsen, 32);
        GS_REQUIRE(x >= 1, "N / coarsen must be >= 1");
        y = 128 / x, grid_x = (unsigned)(rows / y + 1);
        d.thread_for_row = 1;
        e = gsd::transform_row_chain(m, true, (uint32_t)sparse_coarsen, half != 0, "THREAD_META", bound, ctx->stream, out);
    } else if (family == GS_FAMILY_WARP_TOTAL) {
        y = (unsigned)std::min(N / coarsen, 32);
        GS_REQUIRE(y >= 1, "N / coarsen must be >= 1");
        x = std::max(256 / y, 32u), grid_x = (unsigned)rows;
        d.sparse_coarsen = 1; // (the warp / tblock total tokens carry no sparse coarsening: compile() leaves the default)
        e = gsd::transform_row_chain(m, false, 0, half != 0, "WARP_META", bound, ctx->stream, out);
    } else if (family == GS_FAMILY_TBLOCK_TOTAL) { // token_test.cc:1458-1514
        x = (unsigned)std::min(N, 32);
        y = 256 / x, grid_x = (unsigned)rows;
        d.sparse_coarsen = 1;
        e = gsd::transform_row_chain(m, false, 0, half != 0, "TBLOCK_META", bound, ctx->stream, out);
    } else if (family == GS_FAMILY_WARP_BITMAP) { // token_test.cc:1253-1315
        y = (unsigned)std::min(N / coarsen, 32);
        GS_REQUIRE(y >= 1, "N / coarsen must be >= 1");
        x = std::max(128 / y, 32u), grid_x = (unsigned)rows;
        d.need_warp_reduction = 1;
        e = gsd::transform_col_chain(m, 64, x, true, half != 0, bound, ctx->stream, out);
    } else if (family == GS_FAMILY_TBLOCK_BITMAP) { // token_test.cc:1518-1582
        x = (unsigned)std::min(N / coarsen, 32);
        GS_REQUIRE(x >= 1, "N / coarsen must be >= 1");
        y = 256 / x, grid_x = (unsigned)rows;
        e = gsd::transform_col_chain(m, 64, y, false, half != 0, bound, ctx->stream, out); // (the driver hands the operator block[1] before grid_block_operator rounds its own copy)
    } else if (family == GS_FAMILY_WARP_SEGMENT) { // token_test.cc:1393-1455
        x = (unsigned)std::min(N, 32);
        y = 256 / x, grid_x = (unsigned)(nnz / 128 + 1);
        d.need_warp_reduction = 1;
        e = gsd::transform_nnz_chain(m, x, true, half != 0, bound, ctx->stream, out);
    } else {
        x = (unsigned)std::min(N / coarsen, 32);
        GS_REQUIRE(x >= 1, "N / coarsen must be >= 1");
        y = 128 / x, grid_x = (unsigned)(nnz / y + 1);
        e = gsd::transform_nnz_chain(m, x, false, half != 0, bound, ctx->stream, out);
    }
    if (e != cudaSuccess) {
        out.release();
        cudaGetLastError();
        if (!out.error.empty()) return fail(out.refused ? GS_ERR_PADDING_RATE : GS_ERR_INVALID, out.error);
        return fail(GS_ERR_CUDA, std::string("device transform: ") + cudaGetErrorString(e));
    }
    d.vector_width = (int)x;
    d.block[0] = x, d.block[1] = y % 2 ? y + 1 : y; // grid_block_operator.cc:13-33
    const unsigned per = ((family == GS_FAMILY_WARP_TOTAL || family == GS_FAMILY_WARP_SEGMENT || family == GS_FAMILY_WARP_BITMAP) ? d.block[1] : d.block[0]) * (unsigned)coarsen; // a WARP first_nz_indices exists
    d.grid[0] = grid_x, d.grid[1] = (unsigned)((N + per - 1) / per);
    d.has_row_permutation = out.has_perm, d.stored_nnz = out.stored_nnz;
    drop_arrays_of_id(ctx, as_id);
    // what compile() registers for the family, in its order (operators.cc: code_generator::compile)
    std::vector<std::pair<const char *, const char *>> deps = {{"GLOBAL_META", "nz_col_indices"}, {"GLOBAL_META", "nz_vals"}};
    if (out.has_perm) deps.push_back({"GLOBAL_META", "original_nz_row_indices"});
    const char *lvl = family == GS_FAMILY_WARP_TOTAL ? "WARP_META" : family == GS_FAMILY_TBLOCK_TOTAL ? "TBLOCK_META" : "THREAD_META";
    if (family == GS_FAMILY_WARP_SEGMENT) deps.push_back({"WARP_META", "first_BMT_indices"});
    if (family == GS_FAMILY_WARP_BITMAP)
        for (const auto &k : {std::pair<const char *, const char *>{"WARP_META", "first_row_indices"}, {"WARP_META", "first_nz_indices"}, {"WARP_META", "first_BMT_indices"}, {"WARP_META", "bit_map_of_thread"},
                              {"THREAD_META", "first_row_indices_without_ending"}, {"THREAD_META", "first_nz_indices_relative_to_BMW"}})
            deps.push_back(k);
    else if (family == GS_FAMILY_TBLOCK_BITMAP)
        for (const auto &k : {std::pair<const char *, const char *>{"TBLOCK_META", "first_BMT_indices"}, {"THREAD_META", "first_nz_indices"}, {"THREAD_META", "first_row_indices_without_ending"},
                              {"THREAD_META", "bit_map_of_thread"}, {"THREAD_META", "segment_offset"}})
            deps.push_back(k);
    else deps.push_back({lvl, "first_row_indices"}), deps.push_back({lvl, "first_nz_indices"});
    if (family == GS_FAMILY_THREAD_BITMAP)
        for (const char *n : {"thread_bit_map", "segment_ptr", "segment_empty_flag", "segment_empty_row_indices"}) deps.push_back({"THREAD_META", n});
    if (family == GS_FAMILY_WARP_SEGMENT)
        for (const char *n : {"thread_bit_map", "segment_ptr", "segment_offset", "segment_empty_flag", "segment_empty_row_indices"}) deps.push_back({"THREAD_META", n});
    for (const auto &k : deps) {
        const std::string key = std::string(k.first) + "_" + k.second + "_" + std::to_string(as_id);
        GS_REQUIRE(key.size() < 64, "key too long");
        strcpy(d.deps[d.n_deps++], key.c_str());
    }
    register_built(ctx, out, as_id);
    *desc_out = d;
    GS_API_END
}

gs_status gs_transform_device_merge_path(gs_ctx *ctx, int level, uint32_t work_size, int sort, uint64_t rows, uint64_t cols, uint64_t nnz, const uint32_t *row_dev, const uint32_t *col_dev,
                                         const float *val_dev, int coo_on_host, int N, int coarsen, int vector_width, const uint32_t grid[2], const uint32_t block[2], int half, int as_id,
                                         gs_plan_desc *desc_out)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && desc_out && row_dev && col_dev && val_dev && grid && block && nnz > 0, "null argument");
    GS_REQUIRE(level >= 0 && level <= 2 && work_size >= 1 && N >= 1 && coarsen >= 1 && as_id >= 0, "bad argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    ctx->join_async();
    static const char *lv[3] = {"THREAD_META", "WARP_META", "TBLOCK_META"};
    struct staged { // the COO brought over when it lies on the host
        void *p[3] = {nullptr, nullptr, nullptr};
        ~staged()
        {
            for (void *q : p) cudaFree(q);
        }
    } st;
    if (coo_on_host) {
        const void *src[3] = {row_dev, col_dev, val_dev};
        for (int i = 0; i < 3; i++) {
            CUDA_TRY(cudaMalloc(&st.p[i], nnz * 4));
            CUDA_TRY(cudaMemcpyAsync(st.p[i], src[i], nnz * 4, cudaMemcpyHostToDevice, ctx->stream));
        }
        row_dev = (const uint32_t *)st.p[0], col_dev = (const uint32_t *)st.p[1], val_dev = (const float *)st.p[2];
    }
    gsd::coo_dev m;
    m.row = row_dev, m.col = col_dev, m.val = val_dev, m.rows = rows, m.cols = cols, m.nnz = nnz;
    gsd::transform_out out;
    const cudaError_t e = gsd::transform_merge_path_chain(m, sort != 0, work_size, lv[level], half != 0, (double)gs::get_config().as_integer("PADDING_RATE_UP_BOUND"), ctx->stream, out);
    if (e != cudaSuccess) {
        out.release();
        cudaGetLastError();
        if (!out.error.empty()) return fail(GS_ERR_INVALID, out.error);
        return fail(GS_ERR_CUDA, std::string("device transform: ") + cudaGetErrorString(e));
    }
    // what compile() gives for a total reduction over blocks that are not one row each (operators.cc): the range kernel's family and its three arrays
    gs_plan_desc d;
    memset(&d, 0, sizeof d);
    d.family = GS_FAMILY_B200_MERGE, d.sub_matrix_id = as_id, d.coarsen = coarsen, d.sparse_coarsen = 1, d.vector_width = vector_width;
    d.rows = rows, d.cols = cols, d.origin_nnz = nnz, d.stored_nnz = out.stored_nnz, d.has_row_permutation = out.has_perm;
    d.grid[0] = grid[0], d.grid[1] = grid[1], d.block[0] = block[0], d.block[1] = block[1];
    std::vector<std::pair<const char *, const char *>> deps = {{"GLOBAL_META", "nz_col_indices"}, {"GLOBAL_META", "nz_vals"}};
    if (out.has_perm) deps.push_back({"GLOBAL_META", "original_nz_row_indices"});
    deps.push_back({"GLOBAL_META", "nz_row_indices"}), deps.push_back({lv[level], "first_nz_indices"}), deps.push_back({lv[level], "first_row_indices_without_ending"});
    for (const auto &k : deps) {
        const std::string key = std::string(k.first) + "_" + k.second + "_" + std::to_string(as_id);
        GS_REQUIRE(key.size() < 64, "key too long");
        strcpy(d.deps[d.n_deps++], key.c_str());
    }
    drop_arrays_of_id(ctx, as_id);
    register_built(ctx, out, as_id);
    *desc_out = d;
    GS_API_END
}

gs_status gs_transform_host_coo(gs_ctx *ctx, int family, uint64_t rows, uint64_t cols, uint64_t nnz, const uint32_t *row, const uint32_t *col, const float *val, int N,
                                int sparse_coarsen, int coarsen, int half, int as_id, gs_plan_desc *desc_out, float *ms_device)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && row && col && val && nnz > 0, "null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    uint32_t *r = nullptr, *c = nullptr;
    float *v = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    gs_status st = GS_OK;
    auto cleanup = [&] {
        cudaFree(r), cudaFree(c), cudaFree(v);
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    };
    try {
        CUDA_TRY(cudaMalloc((void **)&r, nnz * 4));
        CUDA_TRY(cudaMalloc((void **)&c, nnz * 4));
        CUDA_TRY(cudaMalloc((void **)&v, nnz * 4));
        CUDA_TRY(cudaMemcpyAsync(r, row, nnz * 4, cudaMemcpyHostToDevice, ctx->stream));
        CUDA_TRY(cudaMemcpyAsync(c, col, nnz * 4, cudaMemcpyHostToDevice, ctx->stream));
        CUDA_TRY(cudaMemcpyAsync(v, val, nnz * 4, cudaMemcpyHostToDevice, ctx->stream));
        CUDA_TRY(cudaEventCreate(&e0));
        CUDA_TRY(cudaEventCreate(&e1));
        CUDA_TRY(cudaEventRecord(e0, ctx->stream));
        st = gs_transform_device(ctx, family, rows, cols, nnz, r, c, v, N, sparse_coarsen, coarsen, half, as_id, desc_out);
        CUDA_TRY(cudaEventRecord(e1, ctx->stream));
        CUDA_TRY(cudaEventSynchronize(e1));
        if (ms_device) CUDA_TRY(cudaEventElapsedTime(ms_device, e0, e1));
    } catch (...) {
        cleanup();
        throw;
    }
    cleanup();
    if (st != GS_OK) return st;
    GS_API_END
}

gs_status gs_download_array(gs_ctx *ctx, const char *key, void *host, uint64_t capacity_bytes, uint64_t *len, int *dtype)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && key, "null argument");
    auto it = ctx->arrays.find(key);
    if (it == ctx->arrays.end()) return fail(GS_ERR_NOT_FOUND, std::string("no device array ") + key);
    const dev_array &a = it->second;
    if (len) *len = a.len;
    if (dtype) *dtype = a.dtype;
    if (host) {
        GS_REQUIRE(capacity_bytes >= a.len * dsize(a.dtype), "buffer too small");
        CUDA_TRY(cudaSetDevice(ctx->device));
        ctx->join_async();
        CUDA_TRY(cudaMemcpyAsync(host, a.ptr, a.len * dsize(a.dtype), cudaMemcpyDeviceToHost, ctx->stream));
        CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    }
    GS_API_END
}

gs_status gs_plan_create(gs_ctx *ctx, const gs_plan_desc *desc, int N, int half, gs_plan **out)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && desc && out, "null argument");
    GS_REQUIRE(N >= 1, "N must be >= 1");
    GS_REQUIRE(desc->rows > 0 && desc->cols > 0, "empty shape");
    GS_REQUIRE(desc->family >= 0 && desc->family < GS_FAMILY_COUNT, "unknown family");
    CUDA_TRY(cudaSetDevice(ctx->device));
    gs_plan *p = new gs_plan;
    try {
        p->ctx = ctx, p->desc = *desc, p->N = N, p->half = half;
        p->lp.coarsen = desc->coarsen, p->lp.vector_width = desc->vector_width, p->lp.block_size = (int)desc->block[1];
        p->lp.N = N;
        bind_arrays(p);
        for (int i = 0; i < desc->n_deps && i < 16; i++) {
            const std::string key(desc->deps[i], strnlen(desc->deps[i], sizeof desc->deps[i]));
            if (key_has(key, "nz_col_indices") || key_has(key, "nz_vals")) continue;
            // the key carries the session's sub-matrix id; the arrays may live under another one (gs_upload_session_as)
            const size_t us = key.rfind('_');
            auto it = ctx->arrays.find(us == std::string::npos ? key : key.substr(0, us + 1) + std::to_string(desc->sub_matrix_id));
            if (it != ctx->arrays.end()) {
                p->chain_offset_bytes += it->second.len * dsize(it->second.dtype);
                if (it->second.model != 0) p->model_bytes += it->second.len * dsize(it->second.dtype);
            }
        }
        select_kernel(p, desc->family);
        size_t cbytes = (size_t)desc->rows * N * p->esz();
        CUDA_TRY(cudaMalloc(&p->C_dev, cbytes + 16));
        CUDA_TRY(cudaMemsetAsync(p->C_dev, 0, cbytes, ctx->mem_stream())); // rows no kernel writes (trimmed trailing empty rows) stay zero
        CUDA_TRY(cudaEventCreate(&p->ev0));
        CUDA_TRY(cudaEventCreate(&p->ev1));
        CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    } catch (...) {
        gs_plan_destroy(p);
        throw;
    }
    *out = p;
    GS_API_END
}

void gs_plan_destroy(gs_plan *p)
{
    if (!p) return;
    cudaSetDevice(p->ctx->device);
    if (p->row_plan) gsd::b200_row_free(p->row_plan);
    if (p->tile_plan) gsd::b200_tile_free(p->tile_plan);
    if (p->sp24_plan) gsd::b200_sp24_free(p->sp24_plan);
    if (p->merge_plan) gsd::b200_merge_free(p->merge_plan);
    if (p->nnz_plan) gsd::b200_nnz_free(p->nnz_plan);
    if (p->B_dev) cudaFree(p->B_dev);
    if (p->C_dev) cudaFree(p->C_dev);
    if (p->ev0) cudaEventDestroy(p->ev0);
    if (p->ev1) cudaEventDestroy(p->ev1);
    if (p->ev_b) cudaEventDestroy(p->ev_b);
    if (p->ev_k) cudaEventDestroy(p->ev_k);
    if (p->ev_c) cudaEventDestroy(p->ev_c);
    delete p;
}

gs_status gs_plan_set_kernel(gs_plan *p, int family)
{
    GS_API_BEGIN
    GS_REQUIRE(p, "null plan");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    p->ctx->row_chain = false; // prepare() puts copies and layout kernels on the stream
    if (family == GS_KERNEL_AUTO) select_auto(p);
    else select_kernel(p, family);
    GS_API_END
}

gs_status gs_set_dense(gs_plan *p, const void *B_host, int dtype)
{
    GS_API_BEGIN
    GS_REQUIRE(p && B_host, "null argument");
    GS_REQUIRE(dtype == GS_F32 || dtype == GS_F16, "dense B must be f32 or f16");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    size_t n = (size_t)p->desc.cols * p->N;
    std::vector<unsigned char> conv;
    const void *src = B_host;
    if ((dtype == GS_F16) != (p->half != 0)) {
        conv.resize(n * p->esz());
        for (size_t i = 0; i < n; i++) {
            if (p->half) ((__half *)conv.data())[i] = __float2half_rn(((const float *)B_host)[i]);
            else ((float *)conv.data())[i] = __half2float(((const __half *)B_host)[i]);
        }
        src = conv.data();
    }
    if (!p->B_dev) CUDA_TRY(cudaMalloc(&p->B_dev, n * p->esz() + 16));
    CUDA_TRY(cudaMemcpyAsync(p->B_dev, src, n * p->esz(), cudaMemcpyHostToDevice, p->ctx->mem_stream()));
    CUDA_TRY(cudaStreamSynchronize(p->ctx->stream));
    p->B_use = p->B_dev;
    GS_API_END
}

gs_status gs_set_dense_device(gs_plan *p, const void *B_dev)
{
    GS_API_BEGIN
    GS_REQUIRE(p && B_dev, "null argument");
    GS_REQUIRE(((uintptr_t)B_dev & 15) == 0, "device B must be 16-byte aligned");
    p->B_use = B_dev;
    // whoever produced this B on gs_stream() is not a launch of ours: the next launch must not start under programmatic dependent launch
    // (the B200 kernels read A and B before griddepcontrol.wait) -- break the chain when the pointer changes hands
    p->ctx->row_chain = false;
    for (int i = 1; i < gs_ctx::MAX_LAUNCH; i++) p->ctx->chain_extra[i] = false;
    GS_API_END
}

gs_status gs_launch(gs_plan *p)
{
    GS_API_BEGIN
    GS_REQUIRE(p, "null plan");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    enqueue(p);
    GS_API_END
}

gs_status gs_run(gs_plan *p, int repeat, float *ms_per_iter)
{
    GS_API_BEGIN
    GS_REQUIRE(p && repeat >= 1, "bad argument");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    CUDA_TRY(cudaEventRecord(p->ev0, p->ctx->stream));
    for (int i = 0; i < repeat; i++) enqueue(p);
    CUDA_TRY(cudaEventRecord(p->ev1, p->ctx->stream));
    CUDA_TRY(cudaEventSynchronize(p->ev1));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, p->ev0, p->ev1));
    if (ms_per_iter) *ms_per_iter = ms / repeat;
    GS_API_END
}

gs_status gs_get_output(gs_plan *p, void *C_host)
{
    GS_API_BEGIN
    GS_REQUIRE(p && C_host, "null argument");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    CUDA_TRY(cudaMemcpyAsync(C_host, p->C_out(), (size_t)p->desc.rows * p->N * p->esz(), cudaMemcpyDeviceToHost, p->ctx->mem_stream()));
    CUDA_TRY(cudaStreamSynchronize(p->ctx->stream));
    GS_API_END
}

void *gs_output_device(gs_plan *p) { return p ? p->C_out() : nullptr; }

gs_status gs_plan_set_output_device(gs_plan *p, void *C_dev)
{
    GS_API_BEGIN
    GS_REQUIRE(p, "null plan");
    GS_REQUIRE(((uintptr_t)C_dev & 15) == 0, "device C must be 16-byte aligned");
    p->C_ext = C_dev; // NULL: back to the plan's own buffer
    p->ctx->row_chain = false;
    GS_API_END
}

// ------------------------------------------------------------------------------------------- optional all-gather of C over NCCL
#define NCCL_TRY(expr)                                                                                                              \
    do {                                                                                                                            \
        int r__ = (expr);                                                                                                           \
        if (r__ != gsd::nccl::SUCCESS) throw gs::Error(std::string("NCCL: ") + (na.GetErrorString ? na.GetErrorString(r__) : "error") + " at " #expr); \
    } while (0)

gs_status gs_comm_unique_id(void *id128)
{
    GS_API_BEGIN
    GS_REQUIRE(id128, "null argument");
    const gsd::nccl::api &na = gsd::nccl::load();
    if (!na.ok) return fail(GS_ERR_UNSUPPORTED, na.why);
    gsd::nccl::unique_id id;
    NCCL_TRY(na.GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    GS_API_END
}

gs_status gs_comm_init(gs_ctx *ctx, int nranks, int rank, const void *id128)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && id128 && nranks >= 1 && rank >= 0 && rank < nranks, "bad argument");
    GS_REQUIRE(!ctx->comm, "the context already has a communicator");
    const gsd::nccl::api &na = gsd::nccl::load();
    if (!na.ok) return fail(GS_ERR_UNSUPPORTED, na.why);
    CUDA_TRY(cudaSetDevice(ctx->device));
    gsd::nccl::unique_id id;
    memcpy(&id, id128, sizeof id);
    NCCL_TRY(na.CommInitRank(&ctx->comm, nranks, id, rank));
    ctx->comm_rank = rank, ctx->comm_size = nranks;
    GS_API_END
}

gs_status gs_comm_allgather_rows(gs_ctx *ctx, void *C_full_dev, const uint64_t *row_cuts, uint64_t row_bytes)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx && C_full_dev && row_cuts && row_bytes > 0, "bad argument");
    GS_REQUIRE(ctx->comm, "gs_comm_init has not been called on this context");
    const gsd::nccl::api &na = gsd::nccl::load();
    CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->mem_stream(); // ordered behind the launches on gs_stream(); breaks the PDL chain
    NCCL_TRY(na.GroupStart());
    const int r = gsd::nccl::allgather_rows(na, ctx->comm, st, C_full_dev, row_cuts, ctx->comm_size, row_bytes);
    const int r2 = na.GroupEnd();
    NCCL_TRY(r);
    NCCL_TRY(r2);
    GS_API_END
}

void gs_comm_destroy(gs_ctx *ctx)
{
    if (!ctx || !ctx->comm) return;
    const gsd::nccl::api &na = gsd::nccl::load();
    cudaSetDevice(ctx->device);
    if (na.ok) na.CommDestroy(ctx->comm);
    ctx->comm = nullptr, ctx->comm_size = 0, ctx->comm_rank = -1;
}

gs_status gs_launch_no_alloc(gs_plan *p, const void *B_dev, void *C_dev, void *stream)
{
    GS_API_BEGIN
    GS_REQUIRE(p && B_dev && C_dev, "null argument");
    GS_REQUIRE(p->kernel == GS_FAMILY_B200_TCGEN05 && p->tile_plan, "gs_launch_no_alloc: the plan must run the b200_tcgen05 kernel (gs_plan_set_kernel)");
    GS_REQUIRE(((uintptr_t)B_dev & 15) == 0 && ((uintptr_t)C_dev & 15) == 0, "device B and C must be 16-byte aligned");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    const char *why = nullptr;
    cudaError_t e = gsd::b200_tile_launch_no_alloc(p->tile_plan, p->ctx->device, B_dev, C_dev, (cudaStream_t)stream, &why);
    if (e == cudaErrorInvalidValue && why) throw gs::Error(std::string("gs_launch_no_alloc: ") + why + " (the scratch array is fixed at " + std::to_string(gsd::b200_tile_no_alloc_scratch_bytes()) + " bytes)");
    CUDA_TRY(e);
    GS_API_END
}

gs_status gs_spmm_host(gs_plan *p, const void *B_host, void *C_host)
{
    GS_API_BEGIN
    GS_REQUIRE(p && B_host && C_host, "null argument");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    size_t nb = (size_t)p->desc.cols * p->N * p->esz(), nc = (size_t)p->desc.rows * p->N * p->esz();
    if (!p->B_dev) CUDA_TRY(cudaMalloc(&p->B_dev, nb + 16));
    CUDA_TRY(cudaMemcpyAsync(p->B_dev, B_host, nb, cudaMemcpyHostToDevice, p->ctx->mem_stream()));
    p->B_use = p->B_dev;
    enqueue(p);
    CUDA_TRY(cudaMemcpyAsync(C_host, p->C_out(), nc, cudaMemcpyDeviceToHost, p->ctx->mem_stream()));
    CUDA_TRY(cudaStreamSynchronize(p->ctx->stream));
    GS_API_END
}

gs_status gs_spmm_host_async(gs_plan *p, const void *B_host, void *C_host)
{
    GS_API_BEGIN
    GS_REQUIRE(p && B_host && C_host, "null argument");
    gs_ctx *ctx = p->ctx;
    CUDA_TRY(cudaSetDevice(ctx->device));
    if (!ctx->copy_in) CUDA_TRY(cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking));
    if (!ctx->copy_out) CUDA_TRY(cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking));
    // launch streams the steps rotate over (GS_ASYNC_STREAMS, 1 ... 4).  Measured on C2 N=8 (profiles/r1_sp24_summary.md): 22.6 / 16.5 /
    // 11.9 / 11.1 us per step with 1 / 2 / 3 / 4 streams.  Every B200 launch needs all its CTAs resident (split-panel hand-off); two
    // fit an SM side by side, and the CTAs of a younger grid are dispatched only after those of the older grids, so a queued third
    // or fourth launch never holds slots an older one is waiting for.
    static const int n_launch = [] {
        const char *v = getenv("GS_ASYNC_STREAMS");
        const int n = v && *v ? atoi(v) : 4;
        return n < 1 ? 1 : n > gs_ctx::MAX_LAUNCH ? gs_ctx::MAX_LAUNCH : n;
    }();
    const int which = (int)(ctx->async_turn++ % (unsigned)n_launch);
    if (which && !ctx->launch_extra[which]) {
        CUDA_TRY(cudaStreamCreateWithFlags(&ctx->launch_extra[which], cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&ctx->ev_join[which], cudaEventDisableTiming));
    }
    cudaStream_t ls = which ? ctx->launch_extra[which] : ctx->stream;
    if (which == 0) ctx->join_async(); // only matters when a caller mixes this call with launches of its own on the main stream
    if (!p->ev_b) {
        CUDA_TRY(cudaEventCreateWithFlags(&p->ev_b, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&p->ev_k, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&p->ev_c, cudaEventDisableTiming));
        // order behind whatever the launch stream already holds for this plan
        CUDA_TRY(cudaEventRecord(p->ev_k, ctx->stream));
        CUDA_TRY(cudaEventRecord(p->ev_c, ctx->stream));
    }
    size_t nb = (size_t)p->desc.cols * p->N * p->esz(), nc = (size_t)p->desc.rows * p->N * p->esz();
    if (!p->B_dev) CUDA_TRY(cudaMalloc(&p->B_dev, nb + 16));
    // B of this plan: not before the plan's previous kernel has finished reading it
    CUDA_TRY(cudaStreamWaitEvent(ctx->copy_in, p->ev_k, 0));
    CUDA_TRY(cudaMemcpyAsync(p->B_dev, B_host, nb, cudaMemcpyHostToDevice, ctx->copy_in));
    CUDA_TRY(cudaEventRecord(p->ev_b, ctx->copy_in));
    p->B_use = p->B_dev;
    // kernel: after B has landed, the plan's previous C has been read back and the plan's previous launch (it may sit on the other
    // launch stream; the split-panel flags count launches of one plan in order) has finished
    CUDA_TRY(cudaStreamWaitEvent(ls, p->ev_b, 0));
    CUDA_TRY(cudaStreamWaitEvent(ls, p->ev_c, 0));
    if (p->last_async != which) CUDA_TRY(cudaStreamWaitEvent(ls, p->ev_k, 0)); // (stream order says so when the previous launch used this very stream)
    p->last_async = which;
    enqueue(p, which);
    CUDA_TRY(cudaEventRecord(p->ev_k, ls));
    CUDA_TRY(cudaStreamWaitEvent(ctx->copy_out, p->ev_k, 0));
    CUDA_TRY(cudaMemcpyAsync(C_host, p->C_out(), nc, cudaMemcpyDeviceToHost, ctx->copy_out));
    CUDA_TRY(cudaEventRecord(p->ev_c, ctx->copy_out));
    GS_API_END
}

gs_status gs_synchronize(gs_ctx *ctx)
{
    GS_API_BEGIN
    GS_REQUIRE(ctx, "null argument");
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    for (int i = 1; i < gs_ctx::MAX_LAUNCH; i++) {
        if (ctx->launch_extra[i]) CUDA_TRY(cudaStreamSynchronize(ctx->launch_extra[i]));
        ctx->dirty_extra[i] = false;
    }
    if (ctx->copy_in) CUDA_TRY(cudaStreamSynchronize(ctx->copy_in));
    if (ctx->copy_out) CUDA_TRY(cudaStreamSynchronize(ctx->copy_out));
    GS_API_END
}

uint64_t gs_plan_algorithmic_bytes(const gs_plan *p)
{
    if (!p) return 0;
    // SURVEY.md 8(d), the one number builder and judge both compute: nnz_stored*(idx+val) + the chain's offset arrays + K*N*sizeof(B) +
    // M*N*sizeof(C) (+ the row permutation when sort_operator ran) -- the CHAIN's arrays, whatever layout the launched kernel streams
    // (gs_plan_format_bytes reports that)
    uint64_t idx = p->lp.a.col_is_u16 ? 2 : 4;
    return p->lp.a.nnz * (idx + p->esz()) + p->chain_offset_bytes + (uint64_t)p->desc.cols * p->N * p->esz() + (uint64_t)p->desc.rows * p->N * p->esz();
}

uint64_t gs_plan_model_bytes(const gs_plan *p) { return p ? p->model_bytes : 0; }

uint64_t gs_plan_format_bytes(const gs_plan *p)
{
    if (!p) return 0;
    if (p->kernel == GS_FAMILY_B200_TCGEN05) return gsd::b200_tile_bytes(p->tile_plan, p->desc.rows, p->desc.cols);
    if (p->kernel == GS_FAMILY_B200_SP24) return gsd::b200_sp24_bytes(p->sp24_plan, p->desc.rows, p->desc.cols);
    uint64_t idx = p->lp.a.col_is_u16 ? 2 : 4;
    if (p->kernel == GS_FAMILY_B200_MERGE) return gsd::b200_merge_bytes(p->merge_plan, p->lp.a.nnz, idx, p->esz(), p->desc.rows, p->desc.cols);
    return p->lp.a.nnz * (idx + p->esz()) + p->offset_bytes + (uint64_t)p->desc.cols * p->N * p->esz() + (uint64_t)p->desc.rows * p->N * p->esz();
}

int gs_plan_launch_count(const gs_plan *p) { return !p ? 0 : p->kernel == GS_FAMILY_B200_SP24 ? gsd::b200_sp24_launches(p->sp24_plan) : 1; }

gs_status gs_plan_sp24_probe(const gs_plan *p, uint32_t *out512)
{
    GS_API_BEGIN
    GS_REQUIRE(p && out512 && p->kernel == GS_FAMILY_B200_SP24 && p->sp24_plan, "not a b200_sp24 plan");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    CUDA_TRY(cudaStreamSynchronize(p->ctx->stream));
    if (gsd::b200_sp24_probe(p->sp24_plan, out512) != cudaSuccess) throw gs::Error("b200_sp24: the plan was not created with GS_SP24_PROBE=1");
    GS_API_END
}

const char *gs_plan_kernel_name(const gs_plan *p)
{
    if (!p) return "";
    if (p->kernel == GS_FAMILY_B200_ROW_SMEM) return gsd::b200_row_kernel_name(p->row_plan);
    if (p->kernel == GS_FAMILY_B200_TCGEN05) return gsd::b200_tile_kernel_name(p->tile_plan);
    if (p->kernel == GS_FAMILY_B200_SP24) return gsd::b200_sp24_kernel_name(p->sp24_plan);
    if (p->kernel == GS_FAMILY_B200_MERGE) return gsd::b200_merge_kernel_name(p->merge_plan);
    if (p->kernel == GS_FAMILY_B200_NNZ) {
        gsd::launch_params lp = p->lp;
        lp.N = p->N;
        return gsd::b200_nnz_kernel_name(lp, p->nnz_plan);
    }
    return gsd::reference_family_kernel_name(p->kernel);
}

gs_status gs_execute(gs_plan *p, int repeat, float *ms_total, float *gflops)
{
    GS_API_BEGIN
    GS_REQUIRE(p && repeat >= 1, "bad argument");
    CUDA_TRY(cudaSetDevice(p->ctx->device));
    enqueue(p); // one untimed warm-up launch, as the generated main() does (code_generator.cc:582)
    CUDA_TRY(cudaStreamSynchronize(p->ctx->stream));
    CUDA_TRY(cudaEventRecord(p->ev0, p->ctx->stream));
    for (int i = 0; i < repeat; i++) enqueue(p);
    CUDA_TRY(cudaEventRecord(p->ev1, p->ctx->stream));
    CUDA_TRY(cudaEventSynchronize(p->ev1));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, p->ev0, p->ev1));
    if (ms_total) *ms_total = ms;
    double rate = gs::get_config().as_float("FLOAT_RATE");
    if (gflops) *gflops = (float)(rate * (double)p->desc.origin_nnz * p->N * repeat / (ms * 1e-3) / 1e9);
    GS_API_END
}

} // extern "C"
