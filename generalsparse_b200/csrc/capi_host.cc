// gsh_* entry points of include/gsparse_b200.h: the host side (metadata store, operator graph).
// Pure C++; nothing here touches CUDA, so these calls work on a box without a GPU.
#include "capi_internal.hpp"
#include "host/arr_optimization.hpp"
#include "host/tile_format.hpp"
#include "host/sp24_format.hpp"

#include <algorithm>

#include <cstring>

using namespace gs;

namespace gs {
thread_local std::string g_last_error;

gs_status fail(gs_status st, const std::string &msg)
{
    g_last_error = msg;
    return st;
}

gs_status translate_exception()
{
    try {
        throw;
    } catch (const gs::Error &e) {
        std::string m = e.what();
        if (m.find("padding rate") != std::string::npos) return fail(GS_ERR_PADDING_RATE, m);
        if (m.find("no metadata item") != std::string::npos || m.find("did not produce") != std::string::npos) return fail(GS_ERR_NOT_FOUND, m);
        if (m.find("not built") != std::string::npos || m.find("unbuilt") != std::string::npos) return fail(GS_ERR_UNSUPPORTED, m);
        if (m.find("CUDA") != std::string::npos || m.find("cuda") != std::string::npos) return fail(GS_ERR_CUDA, m);
        return fail(GS_ERR_INVALID, m);
    } catch (const std::bad_alloc &) {
        return fail(GS_ERR_INTERNAL, "out of host memory");
    } catch (const std::exception &e) {
        return fail(GS_ERR_INTERNAL, e.what());
    } catch (...) {
        return fail(GS_ERR_INTERNAL, "unknown failure");
    }
}
} // namespace gs

#define GS_API_BEGIN try {
#define GS_API_END                                                                                                                 \
    return GS_OK;                                                                                                                  \
    }                                                                                                                              \
    catch (...) { return gs::translate_exception(); }

static gs_status copy_out(const std::string &s, char *buf, size_t n)
{
    if (!buf || n == 0) return fail(GS_ERR_INVALID, "null buffer");
    if (s.size() + 1 > n) return fail(GS_ERR_INVALID, "buffer too small");
    memcpy(buf, s.c_str(), s.size() + 1);
    return GS_OK;
}

static int to_gs_dtype(data_type t)
{
    switch (t) {
    case UNSIGNED_CHAR: return GS_U8;
    case UNSIGNED_SHORT: return GS_U16;
    case UNSIGNED_INT: return GS_U32;
    case UNSIGNED_LONG: return GS_U64;
    case FLOAT: return GS_F32;
    default: return GS_F64;
    }
}

extern "C" {

const char *gs_last_error(void) { return g_last_error.c_str(); }
const char *gs_version(void) { return "gsparse_b200 0.1 (sm_100a)"; }
const char *gs_family_name(int f)
{
    static const char *n[] = {"thread_total", "warp_total", "warp_bitmap", "thread_bit_map", "warp_segment",
                              "block_total", "block_bitmap", "b200_row_smem", "b200_tcgen05", "b200_nnz", "b200_sp24", "b200_merge"};
    return (f >= 0 && f < GS_FAMILY_COUNT) ? n[f] : "unknown";
}

gs_status gsh_metadata_from_mtx(const char *path, const char *float_precise, int reference_ones, gsh_metadata **out)
{
    GS_API_BEGIN
    GS_REQUIRE(path && out, "null argument");
    auto h = new gsh_metadata;
    try {
        h->ms = create_init_metadata_set_from_file(path, path, float_precise ? float_precise : "float", reference_ones != 0);
    } catch (...) {
        delete h;
        throw;
    }
    *out = h;
    GS_API_END
}

gs_status gsh_metadata_from_coo(uint64_t rows, uint64_t cols, uint64_t nnz, const int64_t *row, const int64_t *col, const float *vals,
                                gsh_metadata **out)
{
    GS_API_BEGIN
    GS_REQUIRE(row && col && out, "null argument");
    auto h = new gsh_metadata;
    try {
        h->ms = create_init_metadata_set_from_coo(rows, cols, nnz, row, col, vals, get_config().as_string("PRECISE_OF_FLOAT"));
    } catch (...) {
        delete h;
        throw;
    }
    *out = h;
    GS_API_END
}

gs_status gsh_metadata_row_block(const gsh_metadata *ms, uint64_t lo, uint64_t hi, gsh_metadata **out)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && out && lo < hi, "bad argument");
    const meta_data_set &src = *ms->ms;
    GS_REQUIRE(src.is_exist(GLOBAL_META, "nz_row_indices", 0) && !src.is_exist(GLOBAL_META, "original_nz_row_indices", 0) &&
                   src.count_of_metadata_of_diff_pos(TBLOCK_META, 0) + src.count_of_metadata_of_diff_pos(WARP_META, 0) + src.count_of_metadata_of_diff_pos(THREAD_META, 0) == 0,
               "row blocks are cut from the matrix as loaded, before any operator ran");
    const std::vector<uint64_t> &r = src.get(GLOBAL_META, "nz_row_indices", 0).u, &c = src.get(GLOBAL_META, "nz_col_indices", 0).u;
    const universal_array &v = src.get(GLOBAL_META, "nz_vals", 0);
    const uint64_t rows = src.scalar(GLOBAL_META, "origin_row_num", -1), cols = src.scalar(GLOBAL_META, "origin_col_num", -1);
    GS_REQUIRE(hi <= rows, "row block past the end of the matrix");
    // rows are sorted (struct.cc:125): the block is one contiguous run of the COO streams
    const uint64_t b = (uint64_t)(std::lower_bound(r.begin(), r.end(), lo) - r.begin()), e = (uint64_t)(std::lower_bound(r.begin(), r.end(), hi) - r.begin());
    std::vector<int64_t> br(e - b), bc(e - b);
    std::vector<float> bv(e - b);
    for (uint64_t i = b; i < e; i++) br[i - b] = (int64_t)(r[i] - lo), bc[i - b] = (int64_t)c[i], bv[i - b] = (float)v.f[i];
    auto h = new gsh_metadata;
    try {
        h->ms = create_init_metadata_set_from_coo(hi - lo, cols, e - b, br.data(), bc.data(), bv.data(), get_config().as_string("PRECISE_OF_FLOAT"));
    } catch (...) {
        delete h;
        throw;
    }
    *out = h;
    GS_API_END
}

gs_status gsh_metadata_row_nnz(const gsh_metadata *ms, uint64_t *out, uint64_t rows)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && out, "null argument");
    const meta_data_set &src = *ms->ms;
    GS_REQUIRE(rows == src.scalar(GLOBAL_META, "origin_row_num", -1), "rows must be origin_row_num");
    GS_REQUIRE(!src.is_exist(GLOBAL_META, "original_nz_row_indices", 0), "row lengths are taken from the matrix as loaded");
    std::fill(out, out + rows, (uint64_t)0);
    for (uint64_t x : src.get(GLOBAL_META, "nz_row_indices", 0).u) {
        GS_REQUIRE(x < rows, "row index out of range");
        out[x]++;
    }
    GS_API_END
}

void gsh_metadata_free(gsh_metadata *ms) { delete ms; }

uint64_t gsh_metadata_count(const gsh_metadata *ms) { return ms ? ms->ms->keys().size() : 0; }

gs_status gsh_metadata_key(const gsh_metadata *ms, uint64_t i, char *buf, size_t n)
{
    GS_API_BEGIN
    GS_REQUIRE(ms, "null metadata");
    auto k = ms->ms->keys();
    GS_REQUIRE(i < k.size(), "item index out of range");
    return copy_out(k[i], buf, n);
    GS_API_END
}

gs_status gsh_metadata_info(const gsh_metadata *ms, const char *key, uint64_t *len, int *mem_dtype, int *compressed_dtype)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && key, "null argument");
    const universal_array *a = ms->ms->find_key(key);
    if (!a) return fail(GS_ERR_NOT_FOUND, std::string("no metadata item ") + key);
    if (len) *len = a->get_len();
    if (mem_dtype) *mem_dtype = to_gs_dtype(a->type);
    if (compressed_dtype) *compressed_dtype = to_gs_dtype(a->get_compress_data_type(get_config().as_bool("DATA_TYPE_COMPRESS")));
    GS_API_END
}

gs_status gsh_metadata_read_u64(const gsh_metadata *ms, const char *key, uint64_t *dst, uint64_t cap)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && key && dst, "null argument");
    const universal_array *a = ms->ms->find_key(key);
    if (!a) return fail(GS_ERR_NOT_FOUND, std::string("no metadata item ") + key);
    GS_REQUIRE(!a->is_float(), "item holds values, read it as f64");
    GS_REQUIRE(cap >= a->get_len(), "destination too small");
    memcpy(dst, a->u.data(), a->u.size() * sizeof(uint64_t));
    GS_API_END
}

gs_status gsh_metadata_read_f64(const gsh_metadata *ms, const char *key, double *dst, uint64_t cap)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && key && dst, "null argument");
    const universal_array *a = ms->ms->find_key(key);
    if (!a) return fail(GS_ERR_NOT_FOUND, std::string("no metadata item ") + key);
    GS_REQUIRE(a->is_float(), "item holds integers, read it as u64");
    GS_REQUIRE(cap >= a->get_len(), "destination too small");
    memcpy(dst, a->f.data(), a->f.size() * sizeof(double));
    GS_API_END
}

gs_status gsh_metadata_set_vals(gsh_metadata *ms, int id, const float *vals, uint64_t n)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && vals, "null argument");
    universal_array &v = ms->ms->get(GLOBAL_META, "nz_vals", id);
    GS_REQUIRE(v.get_len() == n, "value count differs from the stored non-zero count");
    for (uint64_t i = 0; i < n; i++) v.f[i] = (double)vals[i];
    GS_API_END
}

gs_status gsh_metadata_dump(const gsh_metadata *ms, const char *dir)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && dir, "null argument");
    ms->ms->output_2_dir(dir);
    GS_API_END
}

gs_status gsh_metadata_check(const gsh_metadata *ms)
{
    GS_API_BEGIN
    GS_REQUIRE(ms, "null metadata");
    GS_REQUIRE(ms->ms->check(), "meta_data_set::check failed");
    GS_API_END
}

gs_status gsh_metadata_fit_model(const gsh_metadata *ms, const char *key, int *type, uint64_t *params)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && key && type && params, "null argument");
    const universal_array *a = ms->ms->find_key(key);
    if (!a) return fail(GS_ERR_NOT_FOUND, std::string("no metadata item ") + key);
    GS_REQUIRE(!a->is_float(), "models are fitted to integer items");
    const arr_model m = fit_arr_model(a->u, (uint64_t)get_config().as_integer("BRANCH_COMPRESS_MAX_SIZE"));
    *type = (int)m.type;
    params[0] = m.coefficient, params[1] = m.intercept, params[2] = m.type == BRANCH_COMPRESS ? m.low.size() : m.cycle, params[3] = 0;
    GS_API_END
}

gs_status gsh_config_load(const char *path)
{
    GS_API_BEGIN
    GS_REQUIRE(path, "null path");
    get_config().load(path);
    GS_API_END
}

gs_status gsh_config_set(const char *key, const char *json_scalar)
{
    GS_API_BEGIN
    GS_REQUIRE(key && json_scalar, "null argument");
    get_config().set(key, json_scalar);
    GS_API_END
}

gs_status gsh_config_get(const char *key, char *buf, size_t n)
{
    GS_API_BEGIN
    GS_REQUIRE(key, "null key");
    if (!get_config().has(key)) return fail(GS_ERR_NOT_FOUND, std::string("no config key ") + key);
    return copy_out(get_config().as_string(key), buf, n);
    GS_API_END
}

gs_status gsh_session_create(gsh_metadata *ms, int id, gsh_session **out)
{
    GS_API_BEGIN
    GS_REQUIRE(ms && out, "null argument");
    auto s = new gsh_session;
    try {
        s->ms = ms->ms;
        s->cg = std::make_shared<code_generator>(ms->ms, id);
    } catch (...) {
        delete s;
        throw;
    }
    *out = s;
    GS_API_END
}

void gsh_session_free(gsh_session *s) { delete s; }

static void harvest(gsh_session *s, const std::shared_ptr<basic_operator> &op)
{
    s->records.push_back(op->convert_to_string());
    for (const auto &r : op->get_data_transform_sequence()) s->records.push_back(r);
}

gs_status gsh_session_add_and_run(gsh_session *s, const char *op_name, const int64_t *args, int n_args)
{
    GS_API_BEGIN
    GS_REQUIRE(s && op_name && n_args >= 0 && (args || n_args == 0), "bad argument");
    std::vector<int64_t> a(args, args + n_args);
    auto op = make_operator(op_name, s->cg, s->ex.get_operator_context(), a);
    s->ex.add_and_run(op);
    harvest(s, op);
    GS_API_END
}

uint64_t gsh_session_record_count(const gsh_session *s) { return s ? s->records.size() : 0; }

gs_status gsh_session_record(const gsh_session *s, uint64_t i, char *buf, size_t n)
{
    GS_API_BEGIN
    GS_REQUIRE(s && i < s->records.size(), "record index out of range");
    return copy_out(s->records[i], buf, n);
    GS_API_END
}

gs_status gsh_session_run_family(gsh_session *s, int family, int N, int S, int C)
{
    GS_API_BEGIN
    GS_REQUIRE(s, "null session");
    run_family_chain(family, s->cg, s->ex, N, S, C);
    for (int st = 0; st < 3; st++)
        for (const auto &op : s->ex.get_operator_context()->read_operator_context_arr((operator_stage_type)st, s->cg->get_sub_matrix_id())) harvest(s, op);
    s->cg->compile();
    GS_API_END
}

gs_status gsh_session_compile(gsh_session *s)
{
    GS_API_BEGIN
    GS_REQUIRE(s, "null session");
    s->cg->compile();
    GS_API_END
}

gs_status gsh_session_generate_final_program(gsh_session *s, unsigned int repeat, uint64_t *id_out, char *path_buf, size_t path_len)
{
    GS_API_BEGIN
    GS_REQUIRE(s && id_out, "null argument");
    const unsigned long id = s->cg->generate_final_program(repeat);
    *id_out = id;
    if (path_buf) return copy_out(get_config().as_string("ROOT_PATH_STR") + "/data_source/" + std::to_string(id), path_buf, path_len);
    GS_API_END
}

gs_status gsh_session_plan_desc(const gsh_session *s, gs_plan_desc *out)
{
    GS_API_BEGIN
    GS_REQUIRE(s && out, "null argument");
    const kernel_plan &p = s->cg->plan();
    memset(out, 0, sizeof(*out));
    out->family = p.family;
    out->sub_matrix_id = s->cg->get_sub_matrix_id();
    out->sparse_coarsen = p.sparse_coarsen, out->coarsen = p.coarsen, out->vector_width = p.vector_width;
    out->need_warp_reduction = p.need_warp_reduction, out->thread_for_row = p.thread_for_row;
    out->has_row_permutation = p.has_row_permutation;
    for (int i = 0; i < 2; i++) out->grid[i] = p.grid[i], out->block[i] = p.block[i];
    out->rows = p.rows, out->cols = p.cols, out->origin_nnz = p.origin_nnz, out->stored_nnz = p.stored_nnz;
    GS_REQUIRE(p.deps.size() <= 16, "too many dependencies");
    out->n_deps = (int)p.deps.size();
    for (size_t i = 0; i < p.deps.size(); i++) {
        std::string k = meta_data_set::item_name(p.deps[i].pos, p.deps[i].name, p.deps[i].sub_matrix_id);
        GS_REQUIRE(k.size() < 64, "key too long");
        strcpy(out->deps[i], k.c_str());
    }
    GS_API_END
}

gs_status gsh_partition_rows_balanced(const uint64_t *row_nnz, uint64_t rows, int G, uint64_t *out)
{
    GS_API_BEGIN
    GS_REQUIRE(row_nnz && out && rows > 0 && G >= 1, "bad argument");
    std::vector<uint64_t> len(row_nnz, row_nnz + rows);
    uint64_t nnz = 0;
    for (uint64_t l : len) nnz += l;
    uint64_t per = (nnz + G - 1) / G;
    if (per == 0) per = 1;
    std::vector<uint64_t> b = steps::balanced_begin_rows(len, per);
    // the >= rule can leave fewer or more than G parts: merge a tail, or repeat the end for idle devices
    std::vector<uint64_t> cut(G + 1, rows);
    for (int g = 0; g <= G; g++) cut[g] = (size_t)g < b.size() ? b[g] : rows;
    cut[G] = rows;
    cut[0] = 0;
    for (int g = 0; g <= G; g++) out[g] = cut[g];
    GS_API_END
}

gs_status gsh_tile_stream_roundtrip(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val, uint32_t n_rows, uint32_t n_cols, uint32_t pass_warps, uint16_t *dense_out,
                                    uint64_t *stats)
{
    GS_API_BEGIN
    GS_REQUIRE(rowptr && col && val && dense_out && stats && n_rows > 0 && n_cols > 0, "bad argument");
    tile_stream ts;
    try {
        build_tile_stream(rowptr, col, val, n_rows, n_cols, pass_warps, ts);
    } catch (const std::exception &e) {
        throw gs::Error(e.what());
    }
    std::vector<uint16_t> dense;
    expand_tile_stream(ts, dense);
    std::copy(dense.begin(), dense.end(), dense_out);
    // the stores of the kernel's producers: pair slot j of the 32 lanes of warp w in sweep p
    const uint32_t pw = pass_warps ? pass_warps : 10, sweep = pw * 256u;
    uint64_t n_stores = 0, wavefronts = 0;
    for (size_t t = 0; t + 1 < ts.tile_ptr.size(); t++) {
        const uint32_t b = ts.tile_ptr[t], n = ts.tile_ptr[t + 1] - b;
        for (uint32_t base = 0; base < n; base += 256u) {
            (void)sweep;
            const uint32_t lanes = std::min<uint32_t>(32u, (n - base) / 8u);
            for (uint32_t j = 0; j < 8; j++) {
                uint32_t hist[32] = {0}, worst = 0;
                // two lanes that hit the same 32-bit word share a wavefront; distinct words of one bank do not
                std::vector<uint32_t> words;
                for (uint32_t l = 0; l < lanes; l++) words.push_back(ts.off[b + base + l * 8u + j] >> 2);
                std::sort(words.begin(), words.end());
                words.erase(std::unique(words.begin(), words.end()), words.end());
                for (uint32_t w : words) worst = std::max(worst, ++hist[w & 31u]);
                n_stores++, wavefronts += worst;
            }
        }
    }
    stats[0] = ts.off.size(), stats[1] = ts.nnz, stats[2] = ts.tile_ptr.size() - 1, stats[3] = n_stores, stats[4] = wavefronts;
    GS_API_END
}

gs_status gsh_sp24_roundtrip(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val, uint32_t n_rows, uint32_t n_cols, uint32_t group, uint16_t *dense_out, uint64_t *stats)
{
    GS_API_BEGIN
    GS_REQUIRE(rowptr && col && val && dense_out && stats && n_rows > 0 && n_cols > 0, "bad argument");
    sp24_image im;
    try {
        build_sp24_image(rowptr, col, val, n_rows, n_cols, group, im);
    } catch (const std::exception &e) {
        throw gs::Error(e.what());
    }
    std::vector<uint16_t> dense;
    expand_sp24_image(im, dense);
    std::copy(dense.begin(), dense.end(), dense_out);
    stats[0] = im.bytes.size(), stats[1] = im.nnz, stats[2] = im.rem_col.size(), stats[3] = (uint64_t)im.n_panels * im.k_tiles;
    GS_API_END
}

gs_status gsh_sp24_tiled_roundtrip(const uint32_t *rowptr, const uint32_t *col, const uint16_t *val, uint32_t n_rows, uint32_t n_cols, uint32_t row_bytes, uint16_t *dense_out, uint64_t *stats)
{
    GS_API_BEGIN
    GS_REQUIRE(rowptr && col && val && dense_out && stats && n_rows > 0 && n_cols > 0, "bad argument");
    sp24_image im;
    sp24_tiled t;
    std::vector<uint16_t> dense;
    uint64_t kept = 0;
    try {
        build_sp24_image(rowptr, col, val, n_rows, n_cols, 1, im);
        localize_sp24_remainder(im, row_bytes, t);
        // through the stream without its empty tiles and back: what the kept units hold is the whole matrix
        sp24_compact cp;
        GS_REQUIRE(compact_sp24_units(t.bytes.data(), t.unit_bytes, true, im.n_panels, im.k_tiles, 1.0, cp), "compaction refused");
        sp24_tiled back = t;
        std::fill(back.bytes.begin(), back.bytes.end(), (uint8_t)0);
        for (uint32_t p = 0; p < im.n_panels; p++)
            for (uint32_t u = cp.panel_ptr[p]; u < cp.panel_ptr[p + 1]; u++)
                std::copy_n(cp.bytes.begin() + (size_t)u * t.unit_bytes, t.unit_bytes, back.bytes.begin() + ((size_t)p * im.k_tiles + cp.unit_kt[u]) * t.unit_bytes);
        kept = cp.unit_kt.size();
        expand_sp24_tiled(im, back, dense);
    } catch (const std::exception &e) {
        throw gs::Error(e.what());
    }
    std::copy(dense.begin(), dense.end(), dense_out);
    stats[0] = t.bytes.size(), stats[1] = im.nnz, stats[2] = im.rem_col.size(), stats[3] = (uint64_t)im.n_panels * im.k_tiles, stats[4] = t.unit_bytes, stats[5] = t.max_entries, stats[6] = kept;
    GS_API_END
}

} // extern "C"
