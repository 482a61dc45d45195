// oracle/ref_spmm_shim.cc -- TEST INFRASTRUCTURE ONLY.
//
// extern "C" doors onto the reference's OWN host SpMM and result check, compiled from
// /root/reference/arm_code/kernel_lib.hpp where it lies (oracle/Makefile, -I$(REF)):
//   spmm_reference_host<Index,DType>   arm_code/kernel_lib.hpp:859-881 (== cuda_code/kernel_lib.hpp:859-881)
//   check_result<DType>                arm_code/kernel_lib.hpp:884-921
//   read_mtx_file                      arm_code/kernel_lib.hpp:923-1003
// Nothing of the header is copied here. check_result's printf names __half2float for its
// half branch (never taken for float); the overload below only lets the template instantiate.
static inline double __half2float(double x) { return x; }
#include "arm_code/kernel_lib.hpp"
#include <omp.h>

extern "C" {

// the reference loop exactly as shipped: serial, accumulation in float
void ref_spmm_host_f32(int M, int N, int K, const int *indptr, const int *indices, const float *vals,
                       const float *B, float *C)
{
    spmm_reference_host<int, float>(M, N, K, indptr, indices, vals, B, C);
}

// the same reference function applied to disjoint row ranges on all host threads
// (rows are independent; each call runs the reference loop on its own slice)
void ref_spmm_host_f32_mt(int M, int N, int K, const int *indptr, const int *indices, const float *vals,
                          const float *B, float *C, int threads)
{
    if (threads < 1) threads = omp_get_max_threads();
    int chunk = 256;
    int nchunk = (M + chunk - 1) / chunk;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int c = 0; c < nchunk; c++)
    {
        int r0 = c * chunk, r1 = r0 + chunk < M ? r0 + chunk : M;
        // indptr is not rebased: the reference indexes vals/indices with absolute offsets
        spmm_reference_host<int, float>(r1 - r0, N, K, indptr + r0, indices, vals, B, C + (long)r0 * N);
    }
}

// returns 1 when the reference's exact-equality check passes
int ref_check_result_f32(int M, int N, float *C, float *C_ref)
{
    return check_result<float>(M, N, C, C_ref) ? 1 : 0;
}

// the reference's .mtx -> CSR reader (pattern only); caller frees with ref_free
int ref_read_mtx_file(const char *filename, int *nrow, int *ncol, int *nnz, int **indptr, int **indices)
{
    vector<int> p, i;
    read_mtx_file(filename, *nrow, *ncol, *nnz, p, i);
    *indptr = (int *)malloc(sizeof(int) * p.size());
    *indices = (int *)malloc(sizeof(int) * (i.size() ? i.size() : 1));
    memcpy(*indptr, p.data(), sizeof(int) * p.size());
    memcpy(*indices, i.data(), sizeof(int) * i.size());
    return 0;
}
void ref_free(void *p) { free(p); }
int ref_max_threads() { return omp_get_max_threads(); }
}
