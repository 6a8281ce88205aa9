// Layout ingestion: the t-innermost tensors FNO surrogates emit ([BS,F,Nx,Ny,Nt], handed on as the view
// pred.permute(0,1,4,2,3) -- SURVEY.md fact 6, Joint/Wave_Residuals_CP.py:216) brought to the
// [Nt][Nx*Ny] order the TMA star kernel tiles, one field at a time: per sample a [M = Nx*Ny][N = Nt]
// -> [N][M] transpose through shared memory.  A CTA copies 128 consecutive rows -- one contiguous
// 128*N*4-byte chunk, read fully coalesced -- into a tile padded to N+1 words per row (conflict-free
// column reads for any N), then writes N rows of 128 consecutive floats.
#include "common.cuh"

namespace cppre {
namespace {

constexpr int kTM = 128;
constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
transpose_inner_kernel(const float* __restrict__ src, int64_t src_batch_stride, float* __restrict__ dst, int32_t M,
                       int32_t N) {
    extern __shared__ float tile[];                  // [kTM][N + 1]
    const int b = blockIdx.y;
    const int m0 = blockIdx.x * kTM;
    const int rows = min(kTM, M - m0);
    const float* s = src + (int64_t)b * src_batch_stride + (int64_t)m0 * N;
    const int total = rows * N;
    for (int e = threadIdx.x; e < total; e += kThreads) {
        const int m = e / N, n = e - m * N;
        tile[m * (N + 1) + n] = __ldg(s + e);
    }
    __syncthreads();
    float* d = dst + (int64_t)b * N * M + m0;
    for (int idx = threadIdx.x; idx < N * kTM; idx += kThreads) {
        const int n = idx / kTM, m = idx - n * kTM;
        if (m < rows) d[(int64_t)n * M + m] = tile[m * (N + 1) + n];
    }
}

}  // namespace
}  // namespace cppre

using namespace cppre;

extern "C" int cppre_transpose_inner(const float* d_src, int64_t src_batch_stride, int64_t batch, int64_t M, int64_t N,
                                     float* d_dst, void* stream) {
    CPPRE_REQUIRE(batch >= 0 && M >= 0 && N >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (batch == 0 || M == 0 || N == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_src && d_dst, CPPRE_ERR_BAD_ARG, "null pointer");
    CPPRE_REQUIRE(N <= 256, CPPRE_ERR_UNSUPPORTED, "inner extent %lld > 256: use a plain copy", (long long)N);
    CPPRE_REQUIRE(M < (int64_t(1) << 31) && batch < 65536, CPPRE_ERR_UNSUPPORTED, "problem too large for one launch");
    const size_t smem = (size_t)kTM * (N + 1) * sizeof(float);
    static SmemOptIn memo;
    {
        const int rc = ensure_dynamic_smem(transpose_inner_kernel, memo, smem);
        if (rc != CPPRE_OK) return rc;
    }
    const dim3 grid((unsigned)((M + kTM - 1) / kTM), (unsigned)batch);
    transpose_inner_kernel<<<grid, kThreads, smem, (cudaStream_t)stream>>>(d_src, src_batch_stride, d_dst, (int32_t)M,
                                                                          (int32_t)N);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}
