// K1, generic kernel: any residual program, any kernel extents <= 7 per axis, any padding,
// any input strides.  One thread per output cell; taps are read straight from global memory
// through L1/L2 with bounds checks (zero padding).  This is the functional-coverage path --
// shapes or tap patterns the TMA star kernel (stencil_tma.cu) does not take end up here.
#include "common.cuh"
#include "programs.cuh"

namespace cppre {

namespace {

constexpr int kThreads = 256;

struct GenericParams {
    Geometry g;
    FieldView f[CPPRE_MAX_FIELDS];
    EpiDev e;
    float scalars[CPPRE_MAX_SCALARS];
    int32_t chunks;  // blocks per sample
    int32_t dec_on;
    DecodeDev dec;
    DevTaps taps;
};

struct GenericCtx {
    const GenericParams& p;
    const float* base[CPPRE_MAX_FIELDS];
    int i2, i1, i0;

    template <int SLOT, int FIELD, int I>
    __device__ __forceinline__ float D() const {
        const FieldView& fv = p.f[FIELD];
        const float* b = base[FIELD];
        float acc = 0.0f;
        const int t1 = p.taps.begin[SLOT + 1];
        for (int t = p.taps.begin[SLOT]; t < t1; ++t) {
            const int j2 = i2 + p.taps.o2[t], j1 = i1 + p.taps.o1[t], j0 = i0 + p.taps.o0[t];
            if ((unsigned)j2 < (unsigned)p.g.N2 && (unsigned)j1 < (unsigned)p.g.N1 &&
                (unsigned)j0 < (unsigned)p.g.N0) {
                acc = fmaf(p.taps.c[t], __ldg(b + j2 * fv.s2 + j1 * fv.s1 + j0 * fv.s0), acc);
            }
        }
        return acc;
    }
    template <int FIELD, int I>
    __device__ __forceinline__ float V() const {
        const FieldView& fv = p.f[FIELD];
        return __ldg(base[FIELD] + i2 * fv.s2 + i1 * fv.s1 + i0 * fv.s0);
    }
};

template <int KIND>
__global__ void __launch_bounds__(kThreads)
generic_kernel(const __grid_constant__ GenericParams p) {
    const int b = blockIdx.x / p.chunks;
    const int chunk = blockIdx.x - b * p.chunks;
    const int64_t cells = (int64_t)p.g.O2 * p.g.O1 * p.g.O0;
    const int64_t cell = (int64_t)chunk * kThreads + threadIdx.x;
    const bool live = cell < cells;

    float v = 0.0f;
    bool counted = false;
    if (live) {
        const int o0 = (int)(cell % p.g.O0);
        const int64_t r = cell / p.g.O0;
        const int o1 = (int)(r % p.g.O1);
        const int o2 = (int)(r / p.g.O1);
        GenericCtx c{p, {}, o2, o1, o0};
#pragma unroll
        for (int f = 0; f < Prog<KIND>::NF; ++f) c.base[f] = p.f[f].p + (int64_t)b * p.f[f].sb;
        if (p.dec_on) v = Prog<KIND>::template eval<0>(DecodeCtx<GenericCtx>{c, p.dec}, p.scalars);
        else v = Prog<KIND>::template eval<0>(c, p.scalars);
        const float rf = p.e.ref ? __ldg(p.e.ref + b * p.e.rb + o2 * p.e.r2 + o1 * p.e.r1 + o0 * p.e.r0) : 0.0f;
        if (p.e.mode == CPPRE_EPI_COVER) {
            const float q = p.e.band_qp ? __ldg(p.e.band_qp) : p.e.band_q;
            const float h = p.e.weight ? __fmul_rn(q, __ldg(p.e.weight + cell)) : q;
            v = (v >= __fsub_rn(rf, h) && v <= __fadd_rn(rf, h)) ? 1.0f : 0.0f;
        } else {
            if (p.e.ref) v = __fsub_rn(v, rf);
            if (p.e.flags & CPPRE_FLAG_ABS) v = fabsf(v);
            if (p.e.weight) v = __fmul_rn(v, __ldg(p.e.weight + cell));
        }
        if (p.e.mode == CPPRE_EPI_STORE) {
            p.e.out[b * p.e.ob + o2 * p.e.o2 + o1 * p.e.o1 + o0 * p.e.o0] = v;
        } else {
            counted = true;
            if ((p.e.flags & CPPRE_FLAG_INTERIOR_0) && (o0 == 0 || o0 == p.g.O0 - 1)) counted = false;
            if ((p.e.flags & CPPRE_FLAG_INTERIOR_1) && (o1 == 0 || o1 == p.g.O1 - 1)) counted = false;
            if ((p.e.flags & CPPRE_FLAG_INTERIOR_2) && (o2 == 0 || o2 == p.g.O2 - 1)) counted = false;
        }
    }
    if (p.e.mode == CPPRE_EPI_COVER) {
        const uint32_t cnt = __reduce_add_sync(0xffffffffu, (counted && v != 0.0f) ? 1u : 0u);
        if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(reinterpret_cast<unsigned int*>(p.e.scores) + b, cnt);
    } else if (p.e.mode == CPPRE_EPI_SCORE) {
        if (p.e.reduce == CPPRE_REDUCE_MAX) {
            uint32_t m = counted ? score_bits(v) : 0u;
            m = __reduce_max_sync(0xffffffffu, m);
            if ((threadIdx.x & 31) == 0 && m) atomicMax(reinterpret_cast<unsigned int*>(p.e.scores) + b, m);
        } else {
            float s = counted ? v : 0.0f;
#pragma unroll
            for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(p.e.scores + b, s);
        }
    }
}

template <int KIND>
int launch_kind(const GenericParams& gp, int64_t blocks, cudaStream_t st) {
    generic_kernel<KIND><<<(unsigned)blocks, kThreads, 0, st>>>(gp);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

}  // namespace

int launch_generic(const Prepared& P, cudaStream_t st) {
    GenericParams gp;
    gp.g = P.g;
    for (int f = 0; f < CPPRE_MAX_FIELDS; ++f) gp.f[f] = P.fields[f];
    gp.e = P.epi;
    for (int i = 0; i < CPPRE_MAX_SCALARS; ++i) gp.scalars[i] = P.scalars[i];
    gp.taps = P.taps;
    gp.dec_on = P.dec_on;
    gp.dec = P.dec;
    const int64_t cells = (int64_t)P.g.O2 * P.g.O1 * P.g.O0;
    const int64_t chunks = (cells + kThreads - 1) / kThreads;
    const int64_t blocks = chunks * P.g.B;
    CPPRE_REQUIRE(chunks < (int64_t(1) << 31) && blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED,
                  "problem too large for one launch (%lld blocks)", (long long)blocks);
    gp.chunks = (int32_t)chunks;
    switch (P.kind) {
#define CPPRE_CASE(K) \
    case K:           \
        return launch_kind<K>(gp, blocks, st);
        CPPRE_CASE(CPPRE_PROG_LINEAR)
        CPPRE_CASE(CPPRE_PROG_BURGERS)
        CPPRE_CASE(CPPRE_PROG_NS_CONTINUITY)
        CPPRE_CASE(CPPRE_PROG_NS_MOMENTUM)
        CPPRE_CASE(CPPRE_PROG_MHD_CONTINUITY)
        CPPRE_CASE(CPPRE_PROG_MHD_MOMENTUM)
        CPPRE_CASE(CPPRE_PROG_MHD_ENERGY)
        CPPRE_CASE(CPPRE_PROG_MHD_INDUCTION)
        CPPRE_CASE(CPPRE_PROG_MHD_GAUSS)
        CPPRE_CASE(CPPRE_PROG_JOREK_CONTINUITY)
        CPPRE_CASE(CPPRE_PROG_JOREK_TEMPERATURE)
#undef CPPRE_CASE
    }
    set_error("unknown program kind %d", P.kind);
    return CPPRE_ERR_BAD_ARG;
}

}  // namespace cppre
