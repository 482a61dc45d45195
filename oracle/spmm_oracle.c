/* oracle/spmm_oracle.c -- TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * Plain-C restatement of the reference's CPU SpMM, the only result pin the reference has:
 *   spmm_reference_host   cuda_code/kernel_lib.hpp:859-881 (arm_code/kernel_lib.hpp: same lines)
 *       C zeroed; for i in rows ascending, for p in [indptr[i], indptr[i+1]) ascending,
 *       for j < N:  C[i*N+j] += val[p] * B[col[p]*N+j]     (B row-major K x N, C row-major M x N,
 *       accumulation in the element type, one rounding per multiply and per add)
 *   check_result          cuda_code/kernel_lib.hpp:884-921   exact compare (c - c_ref != c_ref - c)
 * Pinned by tests/test_oracle.py against oracle/_ref/libref_spmm.so (the reference's own
 * function compiled from /root/reference) and against the golden rule the reference program
 * checks (all-ones A and B => C[i][*] == nnz(row i), code_generator.cc:631-638).
 *
 * Extensions the reference does not have (it only ever multiplies ones): an fp64-accumulate
 * variant used as the tolerance anchor for random values, and a 64-bit-indptr variant.
 */
#include <stdint.h>
#include <string.h>
#include <stddef.h>

/* fp32, the reference's arithmetic. -ffp-contract is irrelevant here: written as two steps. */
void oracle_spmm_f32(int64_t M, int64_t N, const int64_t *indptr, const int32_t *indices,
                     const float *vals, const float *B, float *C)
{
    memset(C, 0, sizeof(float) * (size_t)(M * N));
    for (int64_t i = 0; i < M; i++)
        for (int64_t p = indptr[i]; p < indptr[i + 1]; p++) {
            const float v = vals[p];
            const float *b = B + (int64_t)indices[p] * N;
            float *c = C + i * N;
            for (int64_t j = 0; j < N; j++) {
                volatile float prod = v * b[j]; /* no fma: the reference is compiled without contraction on x86 */
                c[j] += prod;
            }
        }
}

/* fp64 accumulate of fp32 inputs: tolerance anchor (max rel err 1e-5 fp32 / 1e-2 HALF). */
void oracle_spmm_f64acc(int64_t M, int64_t N, const int64_t *indptr, const int32_t *indices,
                        const float *vals, const float *B, double *C)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < M; i++) {
        double *c = C + i * N;
        for (int64_t j = 0; j < N; j++) c[j] = 0.0;
        for (int64_t p = indptr[i]; p < indptr[i + 1]; p++) {
            const double v = vals[p];
            const float *b = B + (int64_t)indices[p] * N;
            for (int64_t j = 0; j < N; j++) c[j] += v * (double)b[j];
        }
    }
}

/* sum over each row of |val * B| in fp64: the abs floor of the tolerance (1e-6 * ||row||_1). */
void oracle_spmm_absrow(int64_t M, int64_t N, const int64_t *indptr, const int32_t *indices,
                        const float *vals, const float *B, double *A)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < M; i++) {
        double *c = A + i * N;
        for (int64_t j = 0; j < N; j++) c[j] = 0.0;
        for (int64_t p = indptr[i]; p < indptr[i + 1]; p++) {
            const double v = vals[p];
            const float *b = B + (int64_t)indices[p] * N;
            for (int64_t j = 0; j < N; j++) {
                double t = v * (double)b[j];
                c[j] += t < 0 ? -t : t;
            }
        }
    }
}

/* the reference loop on all host threads (rows are independent), fp32 arithmetic unchanged. */
void oracle_spmm_f32_mt(int64_t M, int64_t N, const int64_t *indptr, const int32_t *indices,
                        const float *vals, const float *B, float *C)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < M; i++) {
        float *c = C + i * N;
        for (int64_t j = 0; j < N; j++) c[j] = 0.0f;
        for (int64_t p = indptr[i]; p < indptr[i + 1]; p++) {
            const float v = vals[p];
            const float *b = B + (int64_t)indices[p] * N;
            for (int64_t j = 0; j < N; j++) {
                volatile float prod = v * b[j];
                c[j] += prod;
            }
        }
    }
}

/* check_result: count of entries failing the reference's exact compare. */
int64_t oracle_check_result_f32(int64_t M, int64_t N, const float *C, const float *C_ref)
{
    int64_t wrong = 0;
    for (int64_t i = 0; i < M * N; i++) {
        float c = C[i], r = C_ref[i];
        if (c - r != r - c) wrong++;
    }
    return wrong;
}
