// token_test <matrix.mtx> <N> -- the reference driver's CLI (token_test.cc:1625-1847) on the B200 engine.
//
// Same positional arguments, same CWD convention (./global_config.json is honoured when present), same
// stdout lines ("<matrix> <gflops>" and "<matrix> : min time: <sec/iter>", token_test.cc:1086,1844).
// The reference's main runs test_spmm_thread_total(meta, name, 4, 1) (token_test.cc:1656-1657); that is the
// default here too.  Extra flags after the two positionals:
//   --family <thread_total|warp_total|warp_bitmap|thread_bit_map|warp_segment|block_total|block_bitmap>
//   --sparse-coarsen S  --coarsen C   the two integers every test_spmm_* takes
//   --ops "<operator>[,<arg>...];..." instead of a driver chain: any sequence of the reference's operators by class name (operator.hpp; arguments
//                                     = the constructor's, after code_generator_ptr, booleans as 0/1), run through operator_executer::add_and_run,
//                                     e.g. "sort_operator;merge_path_thread_operator,64;thread_total_reduce_operator,0,1,1;grid_block_operator,64,8,16,1"
//   --kernel <auto|reference|b200_row_smem|b200_tcgen05|b200_sp24|b200_nnz>
//                                     auto (default): the library picks the B200 kernel for the chain, the matrix and N
//                                     (gs_plan_set_kernel(GS_KERNEL_AUTO)); reference: the chain's own reference-shaped family
//   --half                            HALF storage (global_config.json "HALF": true does the same)
//   --values                          keep the file's values (the reference forces every value to 1)
//   --repeat R                        timed launches (reference: 10000)
//   --dump <dir>                      write every metadata array as text, one value per line (struct.cc:1991)
//   --gpus G                          the matrix is cut into G nnz-balanced row blocks, one per GPU 0..G-1 (gs_exec_*: B replicated,
//                                     C rows disjoint, no collective); --allgather: then replicate C on every GPU over NCCL and time it
//   --device-transform                the driver's chain (any --family) is built by device passes from the loader's COO
//                                     (gs_transform_host_coo) instead of the host transform steps; prints the device time
//   --no-parity                       skip the host check (the reference always checks, code_generator.cc:631-638)
//   --corrupt-check                   flip one element of the host copy of C before the check (shows that the check can fail)
// The host check is the reference's own: all-ones B, C[i][j] must equal the row's sum of values exactly in
// reference mode; it reads C back and compares on the host -- it is a check, not a compute fallback.  With HALF, C comes back
// in fp16: a row sum up to 2048 is an exact fp16 integer and must match exactly; above that (and with --values) the bound is
// the north star's 1e-2 (HALF) / 1e-5 (fp32) relative to the row's sum of |values|.
#include "../../include/gsparse_b200.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

// IEEE binary16 -> float on the host (the driver is plain C++: no cuda_fp16.h)
static float half_to_float(uint16_t h)
{
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else { // subnormal: normalise
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

static void die(const char *what)
{
    fprintf(stderr, "token_test: %s: %s\n", what, gs_last_error());
    exit(1);
}
#define OK(call)                 \
    do {                         \
        if ((call) != GS_OK) die(#call); \
    } while (0)

int main(int argc, char **argv)
{
    if (argc < 3) {
        fprintf(stderr, "usage: %s <matrix.mtx> <N> [--family F] [--sparse-coarsen S] [--coarsen C] [--kernel K] [--half] [--values] [--repeat R] [--dump DIR] [--no-parity]\n", argv[0]);
        return 2;
    }
    std::string mtx = argv[1];
    int N = atoi(argv[2]);
    std::string family = "thread_total", kernel, dump, ops;
    int S = 4, C = 1, repeat = 10000;
    bool half = false, values = false, parity = true, corrupt = false, allgather = false;
    int gpus = 1;
    bool device_transform = false;
    for (int i = 3; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char * { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--family") family = next();
        else if (a == "--sparse-coarsen") S = atoi(next());
        else if (a == "--coarsen") C = atoi(next());
        else if (a == "--kernel") kernel = next();
        else if (a == "--ops") ops = next();
        else if (a == "--half") half = true;
        else if (a == "--values") values = true;
        else if (a == "--repeat") repeat = atoi(next());
        else if (a == "--dump") dump = next();
        else if (a == "--no-parity") parity = false;
        else if (a == "--corrupt-check") corrupt = true;
        else if (a == "--gpus") gpus = atoi(next());
        else if (a == "--device-transform") device_transform = true;
        else if (a == "--allgather") allgather = true;
        else {
            fprintf(stderr, "token_test: unknown flag %s\n", a.c_str());
            return 2;
        }
    }
    FILE *cfg = fopen("global_config.json", "r");
    if (cfg) {
        fclose(cfg);
        OK(gsh_config_load("global_config.json"));
        char buf[64];
        if (gsh_config_get("HALF", buf, sizeof buf) == GS_OK && std::string(buf) == "true") half = true;
    }
    int fam = -1;
    for (int f = 0; f < 7; f++)
        if (family == gs_family_name(f)) fam = f;
    if (fam < 0) {
        fprintf(stderr, "token_test: unknown family %s\n", family.c_str());
        return 2;
    }

    gsh_metadata *ms = nullptr;
    OK(gsh_metadata_from_mtx(mtx.c_str(), "float", values ? 0 : 1, &ms));
    printf("finish read file\n");

    if (gpus > 1) {
        // ---- several GPUs, one process: the executor cuts the matrix as loaded, every block runs the chain on its own device ----
        std::vector<int> ids(gpus);
        for (int g = 0; g < gpus; g++) ids[g] = g;
        gs_exec *ex = nullptr;
        OK(gs_exec_create(&ex, ids.data(), gpus));
        const int k = kernel.empty() || kernel == "auto" ? GS_KERNEL_AUTO : kernel == "reference" ? GS_KERNEL_CHAIN : [&] {
            for (int f = 0; f < GS_FAMILY_COUNT; f++)
                if (kernel == gs_family_name(f)) return f;
            fprintf(stderr, "token_test: unknown kernel %s\n", kernel.c_str());
            exit(2);
            return 0;
        }();
        if (device_transform) OK(gs_exec_build_on_device(ex, ms, fam, N, S, C, half, k)); // every block's chain built by its GPU
        else OK(gs_exec_build(ex, ms, fam, N, S, C, half, k));
        uint64_t rows = 0, cols = 0, nnz = 0, one = 0;
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_origin_row_num_-1", &rows, 1));
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_origin_col_num_-1", &cols, 1));
        OK(gsh_metadata_info(ms, "GLOBAL_META_nz_row_indices_0", &nnz, nullptr, nullptr));
        (void)one;
        std::vector<uint64_t> cuts(gpus + 1), bnnz(gpus);
        OK(gs_exec_row_cuts(ex, cuts.data(), bnnz.data()));
        for (int g = 0; g < gpus; g++) printf("gpu %d: rows [%llu, %llu) nnz %llu kernel: %s\n", g, (unsigned long long)cuts[g], (unsigned long long)cuts[g + 1], (unsigned long long)bnnz[g], gs_exec_kernel_name(ex, g));
        std::vector<float> B((size_t)cols * N, 1.0f);
        OK(gs_exec_set_dense(ex, B.data(), GS_F32));
        if (allgather) OK(gs_exec_enable_allgather(ex));
        float ms_iter = 0;
        OK(gs_exec_run(ex, 1, &ms_iter, nullptr)); // warm-up
        OK(gs_exec_run(ex, repeat, &ms_iter, nullptr));
        char fr[64] = "2";
        gsh_config_get("FLOAT_RATE", fr, sizeof fr);
        const double gflops = atof(fr) * (double)nnz * N / (ms_iter * 1e-3) / 1e9;
        std::vector<unsigned char> out((size_t)rows * N * (half ? 2 : 4));
        OK(gs_exec_get_output(ex, out.data()));
        if (allgather) {
            float ag = 0;
            OK(gs_exec_allgather_output(ex, &ag));
            std::vector<unsigned char> rep(out.size());
            long diff = 0;
            for (int g = 0; g < gpus; g++) {
                OK(gs_exec_get_replica(ex, g, rep.data()));
                if (memcmp(rep.data(), out.data(), out.size()) != 0) diff++;
            }
            printf("allgather: %g ms, %ld of %d replicas differ from C\n", ag, diff, gpus);
        }
        if (parity) {
            std::vector<uint64_t> r(nnz);
            std::vector<double> v(nnz);
            OK(gsh_metadata_read_u64(ms, "GLOBAL_META_nz_row_indices_0", r.data(), nnz));
            OK(gsh_metadata_read_f64(ms, "GLOBAL_META_nz_vals_0", v.data(), nnz));
            std::vector<double> ref(rows, 0.0), absref(rows, 0.0);
            for (uint64_t i = 0; i < nnz; i++) ref[r[i]] += v[i], absref[r[i]] += v[i] < 0 ? -v[i] : v[i];
            if (corrupt && !out.empty()) out[out.size() / 2 + (half ? 1 : 3)] ^= 0x40u;
            long wrong = 0;
            const double tol = half ? 1e-2 : 1e-5;
            for (uint64_t i = 0; i < rows; i++)
                for (int j = 0; j < N; j++) {
                    const double c = half ? (double)half_to_float(((const uint16_t *)out.data())[i * N + j]) : (double)((const float *)out.data())[i * N + j];
                    const bool exact = !values && ref[i] <= (half ? 2048.0 : 16777216.0);
                    const double err = c > ref[i] ? c - ref[i] : ref[i] - c;
                    if (exact ? c != ref[i] : !(err <= tol * absref[i])) wrong++;
                }
            if (wrong) printf("wrong number:%ld\n", wrong);
            else printf("correct\n");
        }
        printf("%s %g\n", mtx.c_str(), gflops);
        printf("%s : min time: %g\n", mtx.c_str(), ms_iter * 1e-3);
        gs_exec_destroy(ex);
        gsh_metadata_free(ms);
        return 0;
    }
    gsh_session *s = nullptr;
    gs_ctx *ctx = nullptr;
    gs_plan_desc desc;
    if (device_transform) {
        // the driver's chain built by device passes from the loader's COO (gs_transform_host_coo) -- no host transform steps
        if (!ops.empty()) die("--device-transform builds the seven driver chains, not --ops");
        uint64_t nnz = 0, rows = 0, cols = 0;
        OK(gsh_metadata_info(ms, "GLOBAL_META_nz_row_indices_0", &nnz, nullptr, nullptr));
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_origin_row_num_-1", &rows, 1));
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_origin_col_num_-1", &cols, 1));
        std::vector<uint64_t> r64(nnz), c64(nnz);
        std::vector<double> v64(nnz);
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_nz_row_indices_0", r64.data(), nnz));
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_nz_col_indices_0", c64.data(), nnz));
        OK(gsh_metadata_read_f64(ms, "GLOBAL_META_nz_vals_0", v64.data(), nnz));
        std::vector<uint32_t> r32(nnz), c32(nnz);
        std::vector<float> v32(nnz);
        for (uint64_t i = 0; i < nnz; i++) r32[i] = (uint32_t)r64[i], c32[i] = (uint32_t)c64[i], v32[i] = (float)v64[i];
        OK(gs_create(&ctx, 0));
        float tms = 0;
        OK(gs_transform_host_coo(ctx, fam, rows, cols, nnz, r32.data(), c32.data(), v32.data(), N, S, C, half, 0, &desc, &tms));
        printf("device transform: %g ms\n", tms);
    } else {
    OK(gsh_session_create(ms, 0, &s));
    if (ops.empty()) OK(gsh_session_run_family(s, fam, N, S, C));
    else {
        char num[32];
        snprintf(num, sizeof num, "%d", N);
        OK(gsh_config_set("DENSE_MATRIX_SIZE", num)); // token_test.cc:1630
        snprintf(num, sizeof num, "%d", N < 32 ? N : 32);
        OK(gsh_config_set("VECTOR_WIDTH", num));
        size_t pos = 0;
        while (pos < ops.size()) {
            size_t end = ops.find(';', pos);
            if (end == std::string::npos) end = ops.size();
            std::string item = ops.substr(pos, end - pos);
            pos = end + 1;
            if (item.empty()) continue;
            std::vector<int64_t> args;
            size_t c0 = item.find(',');
            std::string name = item.substr(0, c0);
            while (c0 != std::string::npos) {
                size_t c1 = item.find(',', c0 + 1);
                args.push_back(atoll(item.substr(c0 + 1, c1 == std::string::npos ? std::string::npos : c1 - c0 - 1).c_str()));
                c0 = c1;
            }
            OK(gsh_session_add_and_run(s, name.c_str(), args.empty() ? nullptr : args.data(), (int)args.size()));
        }
        OK(gsh_session_compile(s));
    }
    for (uint64_t i = 0; i < gsh_session_record_count(s); i++) {
        char buf[1024];
        if (gsh_session_record(s, i, buf, sizeof buf) == GS_OK) printf("%s\n", buf);
    }
    if (!dump.empty()) OK(gsh_metadata_dump(ms, dump.c_str()));

    OK(gsh_session_plan_desc(s, &desc));
    OK(gs_create(&ctx, 0));
    OK(gs_upload_session(ctx, s, half));
    }
    gs_plan *plan = nullptr;
    OK(gs_plan_create(ctx, &desc, N, half, &plan));
    if (kernel.empty() || kernel == "auto") {
        if (gs_plan_set_kernel(plan, GS_KERNEL_AUTO) != GS_OK) die("--kernel auto");
    } else if (kernel != "reference") {
        int k = -1;
        for (int f = 0; f < GS_FAMILY_COUNT; f++)
            if (kernel == gs_family_name(f)) k = f;
        if (k < 0 || gs_plan_set_kernel(plan, k) != GS_OK) die("--kernel");
    }
    printf("kernel: %s\n", gs_plan_kernel_name(plan));
    std::vector<float> B((size_t)desc.cols * N, 1.0f); // x_arr := 1 (code_generator.cc:464-467)
    OK(gs_set_dense(plan, B.data(), GS_F32));
    float ms_total = 0, gflops = 0;
    OK(gs_execute(plan, repeat, &ms_total, &gflops));

    if (parity) {
        // the reference's check: CSR x ones on the host, exact compare (cuda_code/kernel_lib.hpp:859-921)
        uint64_t nnz = 0, rows = desc.rows;
        OK(gsh_metadata_info(ms, "GLOBAL_META_nz_row_indices_0", &nnz, nullptr, nullptr));
        std::vector<uint64_t> r(nnz), perm;
        std::vector<double> v(nnz);
        OK(gsh_metadata_read_u64(ms, "GLOBAL_META_nz_row_indices_0", r.data(), nnz));
        OK(gsh_metadata_read_f64(ms, "GLOBAL_META_nz_vals_0", v.data(), nnz));
        uint64_t plen = 0;
        if (gsh_metadata_info(ms, "GLOBAL_META_original_nz_row_indices_0", &plen, nullptr, nullptr) == GS_OK) {
            perm.resize(plen);
            OK(gsh_metadata_read_u64(ms, "GLOBAL_META_original_nz_row_indices_0", perm.data(), plen));
        }
        std::vector<double> ref(rows, 0.0), absref(rows, 0.0);
        for (uint64_t i = 0; i < nnz; i++) {
            const uint64_t row = perm.empty() ? r[i] : perm[r[i]];
            ref[row] += v[i], absref[row] += v[i] < 0 ? -v[i] : v[i];
        }
        std::vector<unsigned char> out((size_t)rows * N * (half ? 2 : 4));
        OK(gs_get_output(plan, out.data()));
        if (corrupt && !out.empty()) out[out.size() / 2 + (half ? 1 : 3)] ^= 0x40u; // an exponent bit of one element in the middle
        long wrong = 0;
        const double tol = half ? 1e-2 : 1e-5;
        for (uint64_t i = 0; i < rows; i++)
            for (int j = 0; j < N; j++) {
                const double c = half ? (double)half_to_float(((const uint16_t *)out.data())[i * N + j]) : (double)((const float *)out.data())[i * N + j];
                // reference mode: the exact compare of check_result (cuda_code/kernel_lib.hpp:884-921) wherever the sum is exact in C's type
                const bool exact = !values && ref[i] <= (half ? 2048.0 : 16777216.0);
                const double err = c > ref[i] ? c - ref[i] : ref[i] - c;
                if (exact ? c != ref[i] : !(err <= tol * absref[i])) wrong++;
            }
        if (wrong) printf("wrong number:%ld\n", wrong);
        else printf("correct\n");
    }
    printf("%s %g\n", mtx.c_str(), gflops);
    printf("%s : min time: %g\n", mtx.c_str(), ms_total * 1e-3 / repeat);
    gs_plan_destroy(plan);
    gs_destroy(ctx);
    if (s) gsh_session_free(s);
    gsh_metadata_free(ms);
    return 0;
}
