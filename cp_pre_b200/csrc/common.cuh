// Internal helpers shared by the libcppre translation units (not part of the C ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/cppre.h"

namespace cppre {

// ---- error plumbing ---------------------------------------------------------------------
void set_error(const char* fmt, ...);
void set_path(const char* name);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define CPPRE_CHECK_CUDA(expr)                                                             \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            ::cppre::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                               __FILE__, __LINE__);                                        \
            return CPPRE_ERR_CUDA;                                                         \
        }                                                                                  \
    } while (0)

#define CPPRE_REQUIRE(cond, code, ...)                                                     \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            ::cppre::set_error(__VA_ARGS__);                                               \
            return (code);                                                                 \
        }                                                                                  \
    } while (0)

// Checks the launch that was just enqueued (configuration errors surface here).
#define CPPRE_CHECK_LAUNCH()                                                               \
    do {                                                                                   \
        ::cppre::count_launch();                                                           \
        CPPRE_CHECK_CUDA(cudaGetLastError());                                              \
    } while (0)

// ---- dynamic shared memory opt-in ----------------------------------------------------------
// cudaFuncSetAttribute acts on the CURRENT device only, so the "already opted in" memo is kept per
// device: a process that launches on cuda:0 and later on cuda:1 configures the kernel on both.
constexpr int kMaxDevices = 64;
struct SmemOptIn {               // static storage: zero-initialised
    size_t bytes[kMaxDevices];   // largest dynamic shared memory size opted into, per device
    bool seen[kMaxDevices];
};
template <class Kernel>
int ensure_dynamic_smem(Kernel kern, SmemOptIn& memo, size_t smem, int carveout_percent = -1) {
    int dev = 0;
    CPPRE_CHECK_CUDA(cudaGetDevice(&dev));
    dev &= kMaxDevices - 1;
    if (!memo.seen[dev] || smem > memo.bytes[dev]) {
        if (smem > 48 * 1024)    // sizes up to 48 KB need no opt-in
            CPPRE_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (carveout_percent >= 0 && !memo.seen[dev])
            CPPRE_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carveout_percent));
        if (smem > memo.bytes[dev]) memo.bytes[dev] = smem;
        memo.seen[dev] = true;
    }
    return CPPRE_OK;
}

// ---- order-preserving float <-> uint32 key ------------------------------------------------
// Unsigned order of the key == total order np.sort uses: negatives (reversed), zeros,
// positives, +inf, then NaNs whose sign bit is clear.
__host__ __device__ __forceinline__ uint32_t float_to_key(uint32_t bits) {
    return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}
__host__ __device__ __forceinline__ uint32_t key_to_float_bits(uint32_t key) {
    return (key & 0x80000000u) ? (key & 0x7fffffffu) : ~key;
}

// ---- star (3-point-per-axis) coefficient set of one derivative slot ----------------------
enum Arm { ARM_C = 0, ARM_TM = 1, ARM_TP = 2, ARM_XM = 3, ARM_XP = 4, ARM_YM = 5, ARM_YP = 6 };
constexpr uint32_t ARMS_T = (1u << ARM_TM) | (1u << ARM_TP);
constexpr uint32_t ARMS_X = (1u << ARM_XM) | (1u << ARM_XP);
constexpr uint32_t ARMS_Y = (1u << ARM_YM) | (1u << ARM_YP);
constexpr uint32_t ARMS_C = (1u << ARM_C);
constexpr uint32_t ARMS_ALL = 0x7fu;

struct StarCoef {
    float c[CPPRE_MAX_SLOTS][8];  // [slot][arm], arm 7 unused
};

// ---- normaliser-decode fold (CPPRE_FLAG_DECODE) ----------------------------------------
struct DecodeDev {
    float a[CPPRE_MAX_FIELDS];                     // scale of field f
    float b[CPPRE_MAX_FIELDS];                     // offset of field f
    float bs[CPPRE_MAX_SLOTS][CPPRE_MAX_FIELDS];   // offset_f * (sum of the taps of slot s)
};

// ---- generic tap list ----------------------------------------------------------------------
struct DevTaps {
    int32_t begin[CPPRE_MAX_SLOTS + 1];
    int8_t o2[CPPRE_MAX_TAPS];
    int8_t o1[CPPRE_MAX_TAPS];
    int8_t o0[CPPRE_MAX_TAPS];
    float c[CPPRE_MAX_TAPS];
};

struct FieldView {
    const float* p;
    int64_t sb, s2, s1, s0;
};

struct EpiDev {
    int32_t mode, flags, reduce;
    const float* ref;
    int64_t rb, r2, r1, r0;
    const float* weight;
    float* out;
    int64_t ob, o2, o1, o0;
    float* scores;
    float band_q;
    const float* band_qp;   // device scalar overriding band_q (COVER), or null
};

struct Geometry {
    int32_t B, N2, N1, N0;  // input extents
    int32_t O2, O1, O0;     // output extents
};

// Host-side normalisation shared by the generic and the TMA path (stencil_common.cu).
struct Prepared {
    Geometry g;
    DevTaps taps;
    FieldView fields[CPPRE_MAX_FIELDS];
    EpiDev epi;
    float scalars[CPPRE_MAX_SCALARS];
    int kind, n_fields, n_slots;
    int in_bf16;   // CPPRE_FLAG_IN_BF16: fields[].p point to bfloat16 elements
    int weight_tin;      // the epilogue weight is laid out [N1][N0][N2] (t-innermost), not contiguous [N2][N1][N0]
    int weight_contig;   // ... contiguous (both can hold for degenerate extents)
    int dec_on;          // CPPRE_FLAG_DECODE
    DecodeDev dec;
};
int prepare(const cppre_program_t* prog, const int64_t shape[4], const cppre_field_t* fields,
            const cppre_epilogue_t* epi, Prepared* out);
int program_fields(int kind);
int program_slots(int kind);

int launch_generic(const Prepared& P, cudaStream_t stream);
// returns 1 when eligible (and fills nothing), 0 otherwise.  One translation unit per
// (epilogue mode, compiled-in operand set); "dyn" handles any operand combination.
#define CPPRE_TMA_TU(sfx)                                                   \
    int tma_eligible_##sfx(const cppre_program_t* prog, const Prepared& P); \
    int launch_tma_##sfx(const cppre_program_t* prog, const Prepared& P, cudaStream_t stream);
CPPRE_TMA_TU(store_dyn) CPPRE_TMA_TU(store_plain)
CPPRE_TMA_TU(score_dyn) CPPRE_TMA_TU(score_plain) CPPRE_TMA_TU(score_w)
CPPRE_TMA_TU(cover_dyn) CPPRE_TMA_TU(cover_w)
CPPRE_TMA_TU(store_dyn_bf16) CPPRE_TMA_TU(store_plain_bf16)
CPPRE_TMA_TU(score_dyn_bf16) CPPRE_TMA_TU(score_w_bf16)
CPPRE_TMA_TU(cover_dyn_bf16) CPPRE_TMA_TU(cover_w_bf16)
CPPRE_TMA_TU(store_dyn_dec) CPPRE_TMA_TU(score_dyn_dec) CPPRE_TMA_TU(cover_dyn_dec)
#undef CPPRE_TMA_TU
#define CPPRE_TIN_TU(sfx)                                                   \
    int tin_eligible_##sfx(const cppre_program_t* prog, const Prepared& P); \
    int launch_tin_##sfx(const cppre_program_t* prog, const Prepared& P, cudaStream_t stream);
CPPRE_TIN_TU(store) CPPRE_TIN_TU(score) CPPRE_TIN_TU(cover)
CPPRE_TIN_TU(store_dec) CPPRE_TIN_TU(score_dec) CPPRE_TIN_TU(cover_dec)
CPPRE_TIN_TU(store_x) CPPRE_TIN_TU(score_x) CPPRE_TIN_TU(cover_x)
#undef CPPRE_TIN_TU
inline int tin_eligible_plain(const cppre_program_t* prog, const Prepared& P) {
    return P.epi.mode == CPPRE_EPI_STORE ? tin_eligible_store(prog, P)
         : P.epi.mode == CPPRE_EPI_SCORE ? tin_eligible_score(prog, P) : tin_eligible_cover(prog, P);
}
inline int tin_eligible_flex(const cppre_program_t* prog, const Prepared& P) {     // odd Nt, unaligned views
    return P.epi.mode == CPPRE_EPI_STORE ? tin_eligible_store_x(prog, P)
         : P.epi.mode == CPPRE_EPI_SCORE ? tin_eligible_score_x(prog, P) : tin_eligible_cover_x(prog, P);
}
inline int tin_eligible(const cppre_program_t* prog, const Prepared& P) {
    return tin_eligible_plain(prog, P) || tin_eligible_flex(prog, P);
}
inline int launch_tin(const cppre_program_t* prog, const Prepared& P, cudaStream_t stream) {
    if (!tin_eligible_plain(prog, P))
        return P.epi.mode == CPPRE_EPI_STORE ? launch_tin_store_x(prog, P, stream)
             : P.epi.mode == CPPRE_EPI_SCORE ? launch_tin_score_x(prog, P, stream) : launch_tin_cover_x(prog, P, stream);
    if (P.dec_on)     // the decode fold lives in its own translation units: the plain kernels carry none of it
        return P.epi.mode == CPPRE_EPI_STORE ? launch_tin_store_dec(prog, P, stream)
             : P.epi.mode == CPPRE_EPI_SCORE ? launch_tin_score_dec(prog, P, stream) : launch_tin_cover_dec(prog, P, stream);
    return P.epi.mode == CPPRE_EPI_STORE ? launch_tin_store(prog, P, stream)
         : P.epi.mode == CPPRE_EPI_SCORE ? launch_tin_score(prog, P, stream) : launch_tin_cover(prog, P, stream);
}
#define CPPRE_FLAT_TU(sfx)                                                   \
    int flat_eligible_##sfx(const cppre_program_t* prog, const Prepared& P); \
    int launch_flat_##sfx(const cppre_program_t* prog, const Prepared& P, cudaStream_t stream);
CPPRE_FLAT_TU(store) CPPRE_FLAT_TU(score) CPPRE_FLAT_TU(cover)
#undef CPPRE_FLAT_TU
inline int flat_eligible(const cppre_program_t* prog, const Prepared& P) {
    return P.epi.mode == CPPRE_EPI_STORE ? flat_eligible_store(prog, P)
         : P.epi.mode == CPPRE_EPI_SCORE ? flat_eligible_score(prog, P) : flat_eligible_cover(prog, P);
}
inline int launch_flat(const cppre_program_t* prog, const Prepared& P, cudaStream_t stream) {
    return P.epi.mode == CPPRE_EPI_STORE ? launch_flat_store(prog, P, stream)
         : P.epi.mode == CPPRE_EPI_SCORE ? launch_flat_score(prog, P, stream) : launch_flat_cover(prog, P, stream);
}
inline int tma_eligible(const cppre_program_t* prog, const Prepared& P) {
    if (P.in_bf16)
        return P.epi.mode == CPPRE_EPI_STORE ? tma_eligible_store_dyn_bf16(prog, P)
             : P.epi.mode == CPPRE_EPI_SCORE ? tma_eligible_score_dyn_bf16(prog, P) : tma_eligible_cover_dyn_bf16(prog, P);
    return P.epi.mode == CPPRE_EPI_STORE ? tma_eligible_store_dyn(prog, P)
         : P.epi.mode == CPPRE_EPI_SCORE ? tma_eligible_score_dyn(prog, P) : tma_eligible_cover_dyn(prog, P);
}
inline int launch_tma(const cppre_program_t* prog, const Prepared& P, cudaStream_t stream) {
    const int oper = (P.epi.ref ? 1 : 0) | (P.epi.weight ? 2 : 0);
    if (P.dec_on)
        return P.epi.mode == CPPRE_EPI_STORE ? launch_tma_store_dyn_dec(prog, P, stream)
             : P.epi.mode == CPPRE_EPI_SCORE ? launch_tma_score_dyn_dec(prog, P, stream) : launch_tma_cover_dyn_dec(prog, P, stream);
    if (P.in_bf16) {
        switch (P.epi.mode) {
            case CPPRE_EPI_STORE:
                return oper == 0 ? launch_tma_store_plain_bf16(prog, P, stream) : launch_tma_store_dyn_bf16(prog, P, stream);
            case CPPRE_EPI_SCORE:
                return oper == 2 ? launch_tma_score_w_bf16(prog, P, stream) : launch_tma_score_dyn_bf16(prog, P, stream);
            default:
                return oper == 2 ? launch_tma_cover_w_bf16(prog, P, stream) : launch_tma_cover_dyn_bf16(prog, P, stream);
        }
    }
    switch (P.epi.mode) {
        case CPPRE_EPI_STORE:
            return oper == 0 ? launch_tma_store_plain(prog, P, stream) : launch_tma_store_dyn(prog, P, stream);
        case CPPRE_EPI_SCORE:
            return oper == 0 ? launch_tma_score_plain(prog, P, stream)
                 : oper == 2 ? launch_tma_score_w(prog, P, stream) : launch_tma_score_dyn(prog, P, stream);
        default:
            return oper == 2 ? launch_tma_cover_w(prog, P, stream) : launch_tma_cover_dyn(prog, P, stream);
    }
}

// ---- score reductions ------------------------------------------------------------------
// Non-negative floats (after |.|) are ordered like their bit patterns, and a NaN (sign
// cleared) compares above +inf, so an unsigned max both reduces and propagates NaN the way
// np.max does.
__device__ __forceinline__ uint32_t score_bits(float v) { return __float_as_uint(v); }

}  // namespace cppre
