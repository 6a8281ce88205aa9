// K4: pointwise spectrum arithmetic of the spectral path.  The FFTs themselves are cuFFT
// (through torch.fft) -- "an FFT only where it really is one"; this kernel replaces the chain of
// complex temporaries the reference builds between rfftn and irfftn
// [ref: Utils/fft_conv_pytorch/fft_conv.py:110-115; Utils/ConvOps_2d.py:207-210,263-268].
#include "common.cuh"

namespace cppre {
namespace {

constexpr int kThreads = 256;

__global__ void __launch_bounds__(kThreads)
spectral_kernel(const float2* __restrict__ sig, const float2* __restrict__ ker, int64_t batch, int64_t m,
                int flags, float eps, float2* __restrict__ out) {
    for (int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x; j < m; j += (int64_t)gridDim.x * kThreads) {
        float2 k = __ldg(ker + j);
        if (flags & CPPRE_SPEC_CONJ_FIRST) k.y = -k.y;
        if (flags & CPPRE_SPEC_INVERSE) {
            // 1 / (k + eps): eps is added to the real part (a real scalar plus a complex tensor)
            const float a = k.x + eps, b = k.y;
            const float d = a * a + b * b;
            k.x = a / d;
            k.y = -b / d;
        }
        if (flags & CPPRE_SPEC_CONJ_LAST) k.y = -k.y;
        for (int64_t i = 0; i < batch; ++i) {
            const float2 s = __ldg(sig + i * m + j);
            out[i * m + j] = make_float2(s.x * k.x - s.y * k.y, s.x * k.y + s.y * k.x);
        }
    }
}

}  // namespace
}  // namespace cppre

using namespace cppre;

extern "C" int cppre_spectral_pointwise(const float* d_sig, const float* d_ker, int64_t batch, int64_t m,
                                        int32_t flags, float eps, float* d_out, void* stream) {
    CPPRE_REQUIRE(batch >= 0 && m >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (batch * m == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_sig && d_ker && d_out, CPPRE_ERR_BAD_ARG, "null pointer");
    int64_t blocks = (m + kThreads - 1) / kThreads;
    if (blocks > 148 * 32) blocks = 148 * 32;
    spectral_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float2*>(d_sig), reinterpret_cast<const float2*>(d_ker), batch, m, flags, eps,
        reinterpret_cast<float2*>(d_out));
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}
