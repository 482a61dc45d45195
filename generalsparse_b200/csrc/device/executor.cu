// gs_exec_*: one SpMM over several GPUs of one box from ONE process -- the executor of north star (3).
//
// The reference's executor (executor.hpp:15, executor.cc:6-105) hands one generated program to one device; it has no multi-GPU
// path.  Here A is cut into nnz-balanced contiguous row blocks with the reference's balanced-interval rule
// (data_transform_common.cc:934-988 -> gsh_partition_rows_balanced: exactly a balanced_interval_row_direction_tblock_blocking with
// nnz_per_interval = ceil(nnz / G), one block per GPU), every block runs the chosen operator chain on its own meta_data_set and gets
// its own device context, stream and plan; B is replicated (one H2D per device), C rows are disjoint, and nothing crosses NVLink on
// the data path.  An all-gather of C over NCCL (ragged blocks: grouped broadcasts, in place in a full-size C per device) exists for
// callers that want C replicated; it is optional and timed separately.  Built on the public C ABI only.
#include "../capi_internal.hpp"
#include "nccl_dl.hpp"

#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

using namespace gs;

#define GS_API_BEGIN try {
#define GS_API_END                                                                                             \
    return GS_OK;                                                                                              \
    }                                                                                                          \
    catch (...) { return gs::translate_exception(); }

#define CUDA_TRY(expr)                                                                                         \
    do {                                                                                                       \
        cudaError_t e__ = (expr);                                                                              \
        if (e__ != cudaSuccess) throw gs::Error(std::string("CUDA: ") + cudaGetErrorString(e__) + " at " #expr); \
    } while (0)

// a failing C-ABI call inside the executor: keep its message
#define ABI_TRY(expr)                                                                                          \
    do {                                                                                                       \
        if ((expr) != GS_OK) throw gs::Error(std::string(gs_last_error()));                                    \
    } while (0)

struct gs_exec {
    int n = 0;
    std::vector<int> dev;
    std::vector<gs_ctx *> ctx;
    std::vector<gsh_metadata *> blk;
    std::vector<gsh_session *> sess;
    std::vector<gs_plan *> plan;
    std::vector<uint64_t> cuts, blk_nnz;
    uint64_t rows = 0, cols = 0, nnz = 0;
    int N = 0, half = 0;
    // all-gather state
    std::vector<void *> full_c;             // per device: rows x N values, the device's own block written in place
    std::vector<gsd::nccl::comm_t> comm;
    std::vector<cudaEvent_t> ev0, ev1;
    bool in_place = false;
    size_t esz() const { return half ? 2 : 4; }
    void clear_build()
    {
        for (gs_plan *p : plan) gs_plan_destroy(p);
        for (gsh_session *s : sess) gsh_session_free(s);
        for (gsh_metadata *m : blk) gsh_metadata_free(m);
        plan.clear(), sess.clear(), blk.clear();
        for (size_t g = 0; g < full_c.size(); g++)
            if (full_c[g]) cudaSetDevice(dev[g]), cudaFree(full_c[g]);
        full_c.clear();
    }
};

extern "C" {

gs_status gs_exec_create(gs_exec **out, const int *device_ids, int n_dev)
{
    GS_API_BEGIN
    GS_REQUIRE(out && device_ids && n_dev >= 1 && n_dev <= 64, "bad argument");
    for (int i = 0; i < n_dev; i++)
        for (int j = 0; j < i; j++) GS_REQUIRE(device_ids[i] != device_ids[j], "a device is listed twice");
    gs_exec *ex = new gs_exec;
    ex->n = n_dev;
    try {
        for (int g = 0; g < n_dev; g++) {
            gs_ctx *c = nullptr;
            ABI_TRY(gs_create(&c, device_ids[g]));
            ex->ctx.push_back(c), ex->dev.push_back(device_ids[g]);
        }
    } catch (...) {
        gs_exec_destroy(ex);
        throw;
    }
    *out = ex;
    GS_API_END
}

void gs_exec_destroy(gs_exec *ex)
{
    if (!ex) return;
    ex->clear_build();
    const gsd::nccl::api &na = gsd::nccl::load();
    for (size_t g = 0; g < ex->comm.size(); g++)
        if (ex->comm[g] && na.ok) cudaSetDevice(ex->dev[g]), na.CommDestroy(ex->comm[g]);
    for (size_t g = 0; g < ex->ev0.size(); g++) cudaSetDevice(ex->dev[g]), cudaEventDestroy(ex->ev0[g]), cudaEventDestroy(ex->ev1[g]);
    for (gs_ctx *c : ex->ctx) gs_destroy(c);
    delete ex;
}

int gs_exec_devices(const gs_exec *ex) { return ex ? ex->n : 0; }

gs_status gs_exec_build(gs_exec *ex, const gsh_metadata *ms, int family, int N, int sparse_coarsen, int coarsen, int half, int kernel)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && ms && N >= 1, "bad argument");
    ex->clear_build();
    const meta_data_set &src = *ms->ms;
    ex->rows = src.scalar(GLOBAL_META, "origin_row_num", -1), ex->cols = src.scalar(GLOBAL_META, "origin_col_num", -1);
    ex->nnz = src.scalar(GLOBAL_META, "origin_nnz_num", -1);
    ex->N = N, ex->half = half;
    std::vector<uint64_t> row_nnz(ex->rows);
    ABI_TRY(gsh_metadata_row_nnz(ms, row_nnz.data(), ex->rows));
    ex->cuts.assign(ex->n + 1, 0);
    ABI_TRY(gsh_partition_rows_balanced(row_nnz.data(), ex->rows, ex->n, ex->cuts.data()));
    ex->blk.assign(ex->n, nullptr), ex->sess.assign(ex->n, nullptr), ex->plan.assign(ex->n, nullptr), ex->blk_nnz.assign(ex->n, 0);
    for (int g = 0; g < ex->n; g++) {
        const uint64_t lo = ex->cuts[g], hi = ex->cuts[g + 1];
        for (uint64_t r = lo; r < hi; r++) ex->blk_nnz[g] += row_nnz[r];
        if (ex->blk_nnz[g] == 0) continue; // an idle device: the greedy rule ran out of rows (or the block holds empty rows only: C stays zero)
        ABI_TRY(gsh_metadata_row_block(ms, lo, hi, &ex->blk[g]));
        ABI_TRY(gsh_session_create(ex->blk[g], 0, &ex->sess[g]));
        ABI_TRY(gsh_session_run_family(ex->sess[g], family, N, sparse_coarsen, coarsen));
        gs_plan_desc d;
        ABI_TRY(gsh_session_plan_desc(ex->sess[g], &d));
        ABI_TRY(gs_upload_session(ex->ctx[g], ex->sess[g], half));
        ABI_TRY(gs_plan_create(ex->ctx[g], &d, N, half, &ex->plan[g]));
        if (kernel != GS_KERNEL_CHAIN) ABI_TRY(gs_plan_set_kernel(ex->plan[g], kernel));
    }
    GS_API_END
}

// The same partition, every block's chain built by its device (gs_transform_host_coo) instead of the host pipeline: the block's slice of
// the COO goes to the device as u32 / fp32 and comes out as the arrays gs_upload_session would have stored.  One host thread per device, so
// the G conversions and transfers overlap.  Families: those gs_transform_device builds.
gs_status gs_exec_build_on_device(gs_exec *ex, const gsh_metadata *ms, int family, int N, int sparse_coarsen, int coarsen, int half, int kernel)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && ms && N >= 1, "bad argument");
    ex->clear_build();
    const meta_data_set &src = *ms->ms;
    GS_REQUIRE(src.is_exist(GLOBAL_META, "nz_row_indices", 0) && !src.is_exist(GLOBAL_META, "original_nz_row_indices", 0), "row blocks are cut from the matrix as loaded, before any operator ran");
    ex->rows = src.scalar(GLOBAL_META, "origin_row_num", -1), ex->cols = src.scalar(GLOBAL_META, "origin_col_num", -1);
    ex->nnz = src.scalar(GLOBAL_META, "origin_nnz_num", -1);
    ex->N = N, ex->half = half;
    std::vector<uint64_t> row_nnz(ex->rows);
    ABI_TRY(gsh_metadata_row_nnz(ms, row_nnz.data(), ex->rows));
    ex->cuts.assign(ex->n + 1, 0);
    ABI_TRY(gsh_partition_rows_balanced(row_nnz.data(), ex->rows, ex->n, ex->cuts.data()));
    ex->blk.assign(ex->n, nullptr), ex->sess.assign(ex->n, nullptr), ex->plan.assign(ex->n, nullptr), ex->blk_nnz.assign(ex->n, 0);
    const std::vector<uint64_t> &r = src.get(GLOBAL_META, "nz_row_indices", 0).u, &c = src.get(GLOBAL_META, "nz_col_indices", 0).u;
    const std::vector<double> &v = src.get(GLOBAL_META, "nz_vals", 0).f;
    std::vector<std::string> err(ex->n);
    std::vector<std::thread> th;
    for (int g = 0; g < ex->n; g++) {
        const uint64_t lo = ex->cuts[g], hi = ex->cuts[g + 1];
        for (uint64_t q = lo; q < hi; q++) ex->blk_nnz[g] += row_nnz[q];
        if (ex->blk_nnz[g] == 0) continue;
        th.emplace_back([&, g, lo, hi] {
            // rows are sorted (struct.cc:125): the block is one contiguous run of the COO streams
            const uint64_t b = (uint64_t)(std::lower_bound(r.begin(), r.end(), lo) - r.begin()), e = (uint64_t)(std::lower_bound(r.begin(), r.end(), hi) - r.begin());
            std::vector<uint32_t> br(e - b), bc(e - b);
            std::vector<float> bv(e - b);
            for (uint64_t i = b; i < e; i++) br[i - b] = (uint32_t)(r[i] - lo), bc[i - b] = (uint32_t)c[i], bv[i - b] = (float)v[i];
            gs_plan_desc d;
            if (gs_transform_host_coo(ex->ctx[g], family, hi - lo, ex->cols, e - b, br.data(), bc.data(), bv.data(), N, sparse_coarsen, coarsen, half, 0, &d, nullptr) != GS_OK ||
                gs_plan_create(ex->ctx[g], &d, N, half, &ex->plan[g]) != GS_OK || (kernel != GS_KERNEL_CHAIN && gs_plan_set_kernel(ex->plan[g], kernel) != GS_OK))
                err[g] = gs_last_error();
        });
    }
    for (std::thread &t : th) t.join();
    for (int g = 0; g < ex->n; g++)
        if (!err[g].empty()) throw gs::Error("device " + std::to_string(ex->dev[g]) + ": " + err[g]);
    GS_API_END
}

gs_status gs_exec_row_cuts(const gs_exec *ex, uint64_t *cuts, uint64_t *block_nnz)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && cuts && !ex->cuts.empty(), "gs_exec_build has not run");
    for (int g = 0; g <= ex->n; g++) cuts[g] = ex->cuts[g];
    if (block_nnz)
        for (int g = 0; g < ex->n; g++) block_nnz[g] = ex->blk_nnz[g];
    GS_API_END
}

const char *gs_exec_kernel_name(const gs_exec *ex, int g) { return ex && g >= 0 && g < ex->n && ex->plan[g] ? gs_plan_kernel_name(ex->plan[g]) : ""; }

gs_status gs_exec_set_dense(gs_exec *ex, const void *B_host, int dtype)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && B_host && !ex->plan.empty(), "gs_exec_build has not run");
    for (int g = 0; g < ex->n; g++)
        if (ex->plan[g]) ABI_TRY(gs_set_dense(ex->plan[g], B_host, dtype)); // B replicated: one H2D per device
    GS_API_END
}

// every device launches `repeat` SpMMs of its block on its own stream, enqueued by a host thread of its own (a single thread would
// serialise 8 x repeat launch calls: ~4 us each against a 9 us kernel), started together; timed with CUDA events per device
gs_status gs_exec_run(gs_exec *ex, int repeat, float *ms_per_iter_max, float *ms_per_iter_dev)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && repeat >= 1 && !ex->plan.empty(), "bad argument");
    std::vector<float> ms(ex->n, 0.0f);
    std::vector<std::string> err(ex->n);
    std::atomic<int> ready{0};
    int active = 0;
    for (int g = 0; g < ex->n; g++) active += ex->plan[g] != nullptr;
    std::vector<std::thread> th;
    for (int g = 0; g < ex->n; g++) {
        if (!ex->plan[g]) continue;
        th.emplace_back([&, g] {
            cudaSetDevice(ex->dev[g]);
            ready.fetch_add(1);
            while (ready.load() < active) std::this_thread::yield();
            if (gs_run(ex->plan[g], repeat, &ms[g]) != GS_OK) err[g] = gs_last_error();
        });
    }
    for (std::thread &t : th) t.join();
    float mx = 0;
    for (int g = 0; g < ex->n; g++) {
        if (!err[g].empty()) throw gs::Error("device " + std::to_string(ex->dev[g]) + ": " + err[g]);
        mx = ms[g] > mx ? ms[g] : mx;
        if (ms_per_iter_dev) ms_per_iter_dev[g] = ms[g];
    }
    if (ms_per_iter_max) *ms_per_iter_max = mx;
    GS_API_END
}

gs_status gs_exec_get_output(gs_exec *ex, void *C_host)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && C_host && !ex->plan.empty(), "bad argument");
    const size_t rb = (size_t)ex->N * ex->esz();
    for (int g = 0; g < ex->n; g++) {
        unsigned char *dst = (unsigned char *)C_host + ex->cuts[g] * rb;
        if (ex->plan[g]) ABI_TRY(gs_get_output(ex->plan[g], dst));
        else memset(dst, 0, (ex->cuts[g + 1] - ex->cuts[g]) * rb);
    }
    GS_API_END
}

// ---- optional: C replicated on every device -------------------------------------------------------------------------------------
gs_status gs_exec_enable_allgather(gs_exec *ex)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && !ex->plan.empty(), "gs_exec_build has not run");
    const gsd::nccl::api &na = gsd::nccl::load();
    if (!na.ok) return gs::fail(GS_ERR_UNSUPPORTED, na.why);
    const size_t rb = (size_t)ex->N * ex->esz(), bytes = ex->rows * rb;
    if (ex->comm.empty()) {
        ex->comm.assign(ex->n, nullptr);
        const int r = na.CommInitAll(ex->comm.data(), ex->n, ex->dev.data());
        if (r != gsd::nccl::SUCCESS) {
            ex->comm.clear();
            throw gs::Error(std::string("NCCL: ncclCommInitAll: ") + na.GetErrorString(r));
        }
        ex->ev0.resize(ex->n), ex->ev1.resize(ex->n);
        for (int g = 0; g < ex->n; g++) {
            CUDA_TRY(cudaSetDevice(ex->dev[g]));
            CUDA_TRY(cudaEventCreate(&ex->ev0[g]));
            CUDA_TRY(cudaEventCreate(&ex->ev1[g]));
        }
    }
    if (ex->full_c.empty()) {
        ex->full_c.assign(ex->n, nullptr);
        // in place when every block starts on a 16-byte boundary of the full C (the kernels store 128-bit vectors): the plan writes its rows
        // straight into the full buffer; otherwise the block is copied there before the collective
        ex->in_place = true;
        for (int g = 0; g < ex->n; g++) ex->in_place = ex->in_place && (ex->cuts[g] * rb) % 16 == 0;
        for (int g = 0; g < ex->n; g++) {
            CUDA_TRY(cudaSetDevice(ex->dev[g]));
            CUDA_TRY(cudaMalloc(&ex->full_c[g], bytes + 16));
            CUDA_TRY(cudaMemset(ex->full_c[g], 0, bytes));
            if (ex->in_place && ex->plan[g]) ABI_TRY(gs_plan_set_output_device(ex->plan[g], (unsigned char *)ex->full_c[g] + ex->cuts[g] * rb));
        }
    }
    GS_API_END
}

gs_status gs_exec_allgather_output(gs_exec *ex, float *ms)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && !ex->full_c.empty() && !ex->comm.empty(), "gs_exec_enable_allgather has not run");
    const gsd::nccl::api &na = gsd::nccl::load();
    const size_t rb = (size_t)ex->N * ex->esz();
    for (int g = 0; g < ex->n; g++) {
        CUDA_TRY(cudaSetDevice(ex->dev[g]));
        cudaStream_t st = (cudaStream_t)gs_stream(ex->ctx[g]);
        if (!ex->in_place && ex->plan[g])
            CUDA_TRY(cudaMemcpyAsync((unsigned char *)ex->full_c[g] + ex->cuts[g] * rb, gs_output_device(ex->plan[g]), (ex->cuts[g + 1] - ex->cuts[g]) * rb, cudaMemcpyDeviceToDevice, st));
        CUDA_TRY(cudaEventRecord(ex->ev0[g], st));
    }
    int r = na.GroupStart();
    for (int g = 0; g < ex->n && r == gsd::nccl::SUCCESS; g++) r = gsd::nccl::allgather_rows(na, ex->comm[g], (cudaStream_t)gs_stream(ex->ctx[g]), ex->full_c[g], ex->cuts.data(), ex->n, rb);
    const int r2 = na.GroupEnd();
    if (r != gsd::nccl::SUCCESS || r2 != gsd::nccl::SUCCESS) throw gs::Error(std::string("NCCL: all-gather of C: ") + na.GetErrorString(r != gsd::nccl::SUCCESS ? r : r2));
    float mx = 0;
    for (int g = 0; g < ex->n; g++) {
        CUDA_TRY(cudaSetDevice(ex->dev[g]));
        CUDA_TRY(cudaEventRecord(ex->ev1[g], (cudaStream_t)gs_stream(ex->ctx[g])));
    }
    for (int g = 0; g < ex->n; g++) {
        CUDA_TRY(cudaSetDevice(ex->dev[g]));
        CUDA_TRY(cudaEventSynchronize(ex->ev1[g]));
        float t = 0;
        CUDA_TRY(cudaEventElapsedTime(&t, ex->ev0[g], ex->ev1[g]));
        mx = t > mx ? t : mx;
    }
    if (ms) *ms = mx;
    GS_API_END
}

gs_status gs_exec_get_replica(gs_exec *ex, int g, void *C_host)
{
    GS_API_BEGIN
    GS_REQUIRE(ex && C_host && g >= 0 && g < ex->n && !ex->full_c.empty(), "bad argument");
    CUDA_TRY(cudaSetDevice(ex->dev[g]));
    CUDA_TRY(cudaStreamSynchronize((cudaStream_t)gs_stream(ex->ctx[g]))); // gs streams do not synchronise with the legacy stream of cudaMemcpy
    CUDA_TRY(cudaMemcpy(C_host, ex->full_c[g], ex->rows * (size_t)ex->N * ex->esz(), cudaMemcpyDeviceToHost));
    GS_API_END
}

} // extern "C"
