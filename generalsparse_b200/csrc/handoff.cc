// gs_execute_binary: the reference's process boundary, kept as it is.
//
// In the reference, code_generator::generate_final_program(repeat) writes data_source/<id>/ -- one text file per array the kernel
// reads, named <POS>_<name>_<sub matrix id>, one decimal value per line (metadata_set.cc:517-571, struct.cc:1991) -- and
// execute_binary(path, matrix, &ms, &gflops) (executor.hpp:15, executor.cc:6-58) compiles and forks the generated program, which reads
// those files back (cuda_code/kernel_lib.hpp:704-781), times `repeat` launches, re-reads the matrix file for the host check
// (code_generator.cc:631-638) and leaves its numbers in perf_result.  This file is that second half without the compiler and the
// fork: the same directory, the same matrix file, the B200 kernels through the C ABI.  It serves both the library's own façade
// (gs::execute_binary in host/operators.hpp, after code_generator::generate_final_program) and the unmodified reference
// (integration/executor_b200.cc defines the reference's execute_binary with this call: INTEGRATION.md path A).
#include "capi_internal.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dirent.h>
#include <fstream>
#include <map>
#include <sstream>

using namespace gs;

#define GS_API_BEGIN try {
#define GS_API_END                                                                                             \
    return GS_OK;                                                                                              \
    }                                                                                                          \
    catch (...) { return gs::translate_exception(); }

#define ABI_TRY(expr)                                                                                          \
    do {                                                                                                       \
        if ((expr) != GS_OK) throw gs::Error(std::string(gs_last_error()));                                    \
    } while (0)

namespace {

struct handoff_array {
    std::string pos, name;
    int id = 0;
    std::vector<uint64_t> u;
    std::vector<double> f;
};

// "<POS>_META_<name>_<id>" -> parts; false for files that are not arrays (kernel_file.cpp, a.out, perf_result ...)
bool split_key(const std::string &file, handoff_array &a)
{
    static const char *const poses[] = {"GLOBAL_META", "TBLOCK_META", "WARP_META", "THREAD_META"};
    for (const char *p : poses) {
        const size_t n = strlen(p);
        if (file.compare(0, n, p) != 0 || file.size() <= n + 2 || file[n] != '_') continue;
        const size_t us = file.rfind('_');
        if (us == std::string::npos || us <= n) return false;
        const std::string tail = file.substr(us + 1);
        if (tail.empty() || tail.find_first_not_of("-0123456789") != std::string::npos) return false;
        a.pos = p, a.name = file.substr(n + 1, us - n - 1), a.id = atoi(tail.c_str());
        return true;
    }
    return false;
}

void read_values(const std::string &path, handoff_array &a)
{
    std::ifstream in(path);
    GS_REQUIRE(in.is_open(), "cannot open " + path);
    const bool is_float = a.name.find("nz_vals") != std::string::npos;
    std::string line;
    while (std::getline(in, line)) {
        if (line.empty()) continue;
        if (is_float) a.f.push_back(strtod(line.c_str(), nullptr));
        else a.u.push_back(strtoull(line.c_str(), nullptr, 10));
    }
    GS_REQUIRE(!(a.u.empty() && a.f.empty()), "empty array file " + path);
}

// the matrix file again, as the generated program reads it for its check: rows, cols and the non-zeros per row (values are 1 in the
// reference: struct.cc:188-195)
void read_matrix_rows(const std::string &path, uint64_t &rows, uint64_t &cols, std::vector<uint64_t> &row_nnz)
{
    std::ifstream in(path);
    GS_REQUIRE(in.is_open(), "cannot open matrix file " + path);
    std::string line;
    bool header = false;
    uint64_t nnz = 0;
    while (std::getline(in, line)) {
        if (line.empty() || line[0] == '%') continue;
        std::istringstream ss(line);
        if (!header) {
            ss >> rows >> cols;
            GS_REQUIRE(rows > 0 && cols > 0, "bad matrix header in " + path);
            row_nnz.assign(rows, 0);
            header = true;
            continue;
        }
        uint64_t r = 0, c = 0;
        ss >> r >> c;
        GS_REQUIRE(r >= 1 && r <= rows && c >= 1 && c <= cols, "matrix entry out of range in " + path);
        row_nnz[r - 1]++, nnz++;
    }
    GS_REQUIRE(header && nnz > 0, "matrix file holds no entries: " + path);
}

float half_bits_to_float(uint16_t h)
{
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {
            int e = -1;
            uint32_t m = man;
            do {
                e++;
                m <<= 1;
            } while (!(m & 0x400u));
            bits = sign | (uint32_t)(127 - 15 - e) << 23 | (m & 0x3ffu) << 13;
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | man << 13;
    else bits = sign | (exp - 15 + 127) << 23 | man << 13;
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

} // namespace

extern "C" gs_status gs_execute_binary(const char *execute_path, const char *matrix_file, int repeat, int kernel, float *ms_total, float *gflops)
{
    GS_API_BEGIN
    GS_REQUIRE(execute_path && matrix_file && repeat >= 1, "bad argument");
    // ---- the arrays of the hand-off directory ------------------------------------------------------------------------------
    std::vector<handoff_array> arrs;
    DIR *d = opendir(execute_path);
    GS_REQUIRE(d != nullptr, std::string("cannot open hand-off directory ") + execute_path);
    std::vector<std::string> files;
    while (dirent *e = readdir(d)) files.push_back(e->d_name);
    closedir(d);
    std::sort(files.begin(), files.end());
    int repeat_file = 0;
    for (const std::string &f : files) {
        handoff_array a;
        if (f == "gs_repeat") { // written by the façade's generate_final_program: the repeat count the reference bakes into its program text
            std::ifstream in(std::string(execute_path) + "/" + f);
            in >> repeat_file;
            continue;
        }
        if (!split_key(f, a)) continue;
        read_values(std::string(execute_path) + "/" + f, a);
        arrs.push_back(std::move(a));
    }
    GS_REQUIRE(!arrs.empty(), std::string("no metadata arrays in ") + execute_path);
    if (repeat_file > 0) repeat = repeat_file;
    const int id = arrs[0].id;
    auto has = [&](const char *pos, const char *name) {
        for (const handoff_array &a : arrs)
            if (a.pos == pos && a.name == name) return &a;
        return (const handoff_array *)nullptr;
    };
    GS_REQUIRE(has("GLOBAL_META", "nz_col_indices") && has("GLOBAL_META", "nz_vals"), "the hand-off holds no nz_col_indices / nz_vals");
    // ---- which family the registered arrays belong to (what compile() derives from the reduction tokens, operators.cc) --------
    int family = -1;
    if (has("THREAD_META", "thread_bit_map")) family = has("WARP_META", "first_BMT_indices") ? GS_FAMILY_WARP_SEGMENT : GS_FAMILY_THREAD_BITMAP;
    else if (has("WARP_META", "bit_map_of_thread")) family = GS_FAMILY_WARP_BITMAP;
    else if (has("TBLOCK_META", "first_BMT_indices") && has("THREAD_META", "bit_map_of_thread")) family = GS_FAMILY_TBLOCK_BITMAP;
    else if (has("THREAD_META", "first_nz_indices") && has("THREAD_META", "first_row_indices")) family = GS_FAMILY_THREAD_TOTAL;
    else if (has("WARP_META", "first_nz_indices") && has("WARP_META", "first_row_indices")) family = GS_FAMILY_WARP_TOTAL;
    else if (has("TBLOCK_META", "first_nz_indices") && has("TBLOCK_META", "first_row_indices")) family = GS_FAMILY_TBLOCK_TOTAL;
    GS_REQUIRE(family >= 0, "the arrays in the hand-off directory do not form one of the kernel families");

    // ---- shape and the expected result from the matrix file; N, HALF, VECTOR_WIDTH from the configuration (the reference driver writes
    // DENSE_MATRIX_SIZE and VECTOR_WIDTH back into ./global_config.json before it generates: token_test.cc:1630,1046) ------------
    uint64_t rows = 0, cols = 0;
    std::vector<uint64_t> row_nnz;
    read_matrix_rows(matrix_file, rows, cols, row_nnz);
    if (FILE *cfg = fopen("global_config.json", "r")) {
        fclose(cfg);
        ABI_TRY(gsh_config_load("global_config.json"));
    }
    const int N = (int)get_config().as_integer("DENSE_MATRIX_SIZE"), half = get_config().as_bool("HALF") ? 1 : 0;
    GS_REQUIRE(N >= 1, "DENSE_MATRIX_SIZE must be >= 1");
    gs_plan_desc desc;
    memset(&desc, 0, sizeof desc);
    desc.family = family, desc.sub_matrix_id = id, desc.sparse_coarsen = 1, desc.coarsen = 1;
    desc.vector_width = (int)get_config().as_integer("VECTOR_WIDTH");
    desc.thread_for_row = family == GS_FAMILY_THREAD_TOTAL, desc.has_row_permutation = has("GLOBAL_META", "original_nz_row_indices") != nullptr;
    desc.rows = rows, desc.cols = cols;
    for (uint64_t n : row_nnz) desc.origin_nnz += n;
    desc.stored_nnz = has("GLOBAL_META", "nz_col_indices")->u.size();
    if (family == GS_FAMILY_TBLOCK_BITMAP) { // BMTs per BMTB = the block's y dimension (token_test.cc:1543-1563)
        const handoff_array *fb = has("TBLOCK_META", "first_BMT_indices");
        desc.block[1] = fb->u.size() >= 2 ? (uint32_t)(fb->u[1] - fb->u[0]) : 1u;
    }
    for (const handoff_array &a : arrs) {
        if (desc.n_deps >= 16) break;
        snprintf(desc.deps[desc.n_deps++], sizeof desc.deps[0], "%s_%s_%d", a.pos.c_str(), a.name.c_str(), a.id);
    }

    // ---- upload, plan, run -------------------------------------------------------------------------------------------------------
    gs_ctx *ctx = nullptr;
    gs_plan *plan = nullptr;
    ABI_TRY(gs_create(&ctx, (int)get_config().as_integer("DEFAULT_DEVICE_ID")));
    try {
        for (const handoff_array &a : arrs) {
            const std::string key = a.pos + "_" + a.name + "_" + std::to_string(a.id);
            if (!a.f.empty()) ABI_TRY(gs_upload_array(ctx, key.c_str(), a.f.data(), a.f.size(), GS_F64, half));
            else ABI_TRY(gs_upload_array(ctx, key.c_str(), a.u.data(), a.u.size(), GS_U64, half));
        }
        ABI_TRY(gs_plan_create(ctx, &desc, N, half, &plan));
        if (kernel != GS_KERNEL_CHAIN) ABI_TRY(gs_plan_set_kernel(plan, kernel));
        std::vector<float> B((size_t)cols * N, 1.0f); // x_arr := 1 (code_generator.cc:464-467)
        ABI_TRY(gs_set_dense(plan, B.data(), GS_F32));
        float ms = 0, gf = 0;
        ABI_TRY(gs_execute(plan, repeat, &ms, &gf));
        // the generated program's check (code_generator.cc:631-638 -> check_result, cuda_code/kernel_lib.hpp:884-921): C[i][*] == nnz(row i), exact
        std::vector<unsigned char> out((size_t)rows * N * (half ? 2 : 4));
        ABI_TRY(gs_get_output(plan, out.data()));
        long wrong = 0;
        for (uint64_t i = 0; i < rows; i++)
            for (int j = 0; j < N; j++) {
                const double c = half ? (double)half_bits_to_float(((const uint16_t *)out.data())[i * N + j]) : (double)((const float *)out.data())[i * N + j];
                const double want = (double)row_nnz[i];
                if (want <= (half ? 2048.0 : 16777216.0) ? c != want : !(std::abs(c - want) <= 1e-2 * want)) wrong++;
            }
        printf("kernel: %s\n", gs_plan_kernel_name(plan));
        if (wrong) printf("wrong number:%ld\n", wrong);
        else printf("correct\n");
        // what the child process leaves in perf_result for the parent (executor.cc:37-49)
        if (FILE *pr = fopen((std::string(execute_path) + "/perf_result").c_str(), "w")) {
            fprintf(pr, "%g\n%g\n", ms, gf);
            fclose(pr);
        }
        if (ms_total) *ms_total = ms;
        if (gflops) *gflops = gf;
        gs_plan_destroy(plan), gs_destroy(ctx);
        if (wrong) return gs::fail(GS_ERR_INTERNAL, "gs_execute_binary: " + std::to_string(wrong) + " entries of C differ from the host check");
    } catch (...) {
        if (plan) gs_plan_destroy(plan);
        gs_destroy(ctx);
        throw;
    }
    GS_API_END
}

// the façade's free function, the reference's name and signature (executor.hpp:15): false on failure where the reference asserts
namespace gs {
bool execute_binary(const std::string &execute_path, const std::string &filename, float &total_execute_time, float &gflops)
{
    return gs_execute_binary(execute_path.c_str(), filename.c_str(), 10000, GS_KERNEL_AUTO, &total_execute_time, &gflops) == GS_OK;
}
} // namespace gs
