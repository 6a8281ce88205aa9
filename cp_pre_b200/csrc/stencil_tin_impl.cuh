// K1t, the t-INNERMOST star kernel (sm_100a): the same fused residual programs as stencil_tma_impl.cuh,
// evaluated IN PLACE on the layout FNO surrogates emit -- physical [BS, F, Nx, Ny, Nt], handed on by the
// scripts as the view pred.permute(0,1,4,2,3) [ref: Joint/Wave_Residuals_CP.py:213-217, SURVEY fact 6] --
// instead of transposing every field first (one extra read + write per field, 34 % of the contiguous rate).
//
// Geometry.  Per sample and field the data is Nx rows of Q = Ny*Nt contiguous floats (y-lines of Nt values
// back to back).  A CTA owns one sample and a tile of TY consecutive y-lines, i.e. a contiguous run of
// TY*Nt floats of every x-row, and marches along x:
//   * x-1 / x / x+1 ride the 3-deep register pipeline (the role t plays in the contiguous kernel);
//   * t-1 / t+1 are the flat neighbours q-1 / q+1, switched off where q is the first / last instant of its
//     y-line (the reference's zero padding in t);
//   * y-1 / y+1 are the flat neighbours q-Nt / q+Nt, read from the shared-memory tile, which carries one
//     halo y-line on each side (zero-filled once where the line lies outside the field).
// Row tiles are contiguous in memory, so they are staged with plain bulk copies (cp.async.bulk, mbarrier
// complete_tx) -- no tensor map, hence no 16-byte rule on the y stride (Nt = 50 gives 200-byte lines, which
// cuTensorMapEncodeTiled rejects).  The accumulation order of the arms is the logical (t-, x-, y-, c, y+, x+,
// t+) of the contiguous kernel, so both kernels return bit-identical residuals.
// Output (STORE), reference residual and per-cell weight use the same physical order; reductions (SCORE /
// COVER) are per sample as before.
//
// This file is the body of one translation unit per epilogue mode (stencil_tin_<mode>.cu), each defining
// CPPRE_TIN_MODE and CPPRE_TIN_SUFFIX.  Optional operands (ref, weight) are decided at run time.
//
// Role 1 (stencil_flat_<mode>.cu, CPPRE_TIN_ROLE 1) applies the same flat-neighbour scheme to CONTIGUOUS
// [BS, Nt, Nx, Ny] data the TMA kernel cannot stage -- rows that are not a multiple of 16 bytes (N0 % 4 != 0,
// e.g. 254 x 254, 33 x 33, config 1's Nx = 197), unaligned bases or sample strides: march = t, lines = x, within
// a line = y.  Before it these requests fell to the one-thread-per-cell generic kernel at ~5 % of the star
// kernels' rate (ns momentum 20 vs 420 Gpoints/s).
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "programs.cuh"

namespace cppre {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxStages = 8;
// CPPRE_TIN_DECODE: the normaliser-decode fold (CPPRE_FLAG_DECODE, programs.cuh DecodeCtx) is compiled into
// this translation unit's kernels -- the layout FNO surrogates emit is also the one that arrives normalised.
#ifndef CPPRE_TIN_DECODE
#define CPPRE_TIN_DECODE 0
#endif
constexpr bool kDEC = CPPRE_TIN_DECODE != 0;
// CPPRE_TIN_ROLE: which logical axis plays which physical part.
//   0  t-innermost memory [B][Nx][Ny][Nt] (FNO layout): march = x, lines = y, within a line = t   (the default)
//   1  CONTIGUOUS memory [B][Nt][Nx][Ny] whose rows the TMA kernel cannot take (N0 % 4 != 0, unaligned base or
//      strides): march = t, lines = x, within a line = y.  Same flat-neighbour scheme; in addition this role copes
//      with rows that are not 16-byte aligned (the refilling warp stages them with 4-byte cp.async instead of one
//      bulk copy), odd line lengths, partially live quads and scalar epilogue accesses -- all decided at run time,
//      so aligned ragged shapes (e.g. 254 x 254) keep the bulk copies and the 128-bit epilogue.
#ifndef CPPRE_TIN_ROLE
#define CPPRE_TIN_ROLE 0
#endif
constexpr int kROLE = CPPRE_TIN_ROLE;
// CPPRE_TIN_FLEX: the unaligned / odd-line-length machinery (always on in role 1).  Role 0 has it in separate
// translation units (stencil_tinx_<mode>.cu) that take the FNO views the plain ones decline: odd Nt, bases or
// sample strides that are not 16-byte aligned.
#ifndef CPPRE_TIN_FLEX
#define CPPRE_TIN_FLEX (CPPRE_TIN_ROLE == 1)
#endif
constexpr bool kFLEX = CPPRE_TIN_FLEX != 0;
// logical arm served by each physical neighbour: I = within the line (flat -+1), M = march axis (register
// pipeline), L = neighbouring line (flat -+line length)
constexpr int A_IM = kROLE == 0 ? ARM_TM : ARM_YM, A_IP = kROLE == 0 ? ARM_TP : ARM_YP;
constexpr int A_MM = kROLE == 0 ? ARM_XM : ARM_TM, A_MP = kROLE == 0 ? ARM_XP : ARM_TP;
constexpr int A_LM = kROLE == 0 ? ARM_YM : ARM_XM, A_LP = kROLE == 0 ? ARM_YP : ARM_XP;

template <int NF>
struct alignas(16) TinParams {
    const float* f[NF];       // field base pointers (sample 0, x-row 0, flat index 0)
    int64_t fsb[NF];          // sample stride of each field, elements
    int64_t frow[NF];         // stride between consecutive rows of the march: lines * line length, or 0 for a field that
                              // is the same plane for every row and sample (role 1: JOREK's R(x, y) coefficient plane)
    StarCoef coef;
    float scalars[CPPRE_MAX_SCALARS];
    EpiDev e;                 // out / ref: sample strides ob / rb; x-row stride Q, flat stride 1 (checked by the host)
    int32_t Nt, Nx, Ny;       // logical extents
    int32_t TY;               // y-lines per tile
    int32_t tiles;            // ceil(Ny / TY)
    int32_t padl;             // leading pad elements: (padl + Nt) % 4 == 0, so centre element 0 is 16-byte aligned
    int32_t stages;
    uint32_t buf_bytes;       // one (stage, field) row tile, 128-aligned
    int32_t bulk;             // rows are 16-byte aligned: one bulk copy per (row, field); else 4-byte cp.async (kFLEX)
    int32_t vec;              // out / ref / weight quads are 16-byte aligned: 128-bit epilogue accesses (kFLEX)
    DecodeDev dec;            // read by the CPPRE_TIN_DECODE units only
};

template <int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (N > 0) {
        static_for<N - 1>(f);
        f(std::integral_constant<int, N - 1>{});
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bar_arrive(uint32_t id, uint32_t threads) {
    asm volatile("barrier.cta.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void bar_sync(uint32_t id, uint32_t threads) {
    asm volatile("barrier.cta.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// contiguous global -> shared bulk copy (16-byte aligned on both sides, size a multiple of 16)
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// 4-byte asynchronous copy (no alignment rule beyond the element), and "arrive on the mbarrier once all my prior
// cp.async have landed" (.noinc: the arrival is one of those the barrier was initialised to expect)
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_arrive(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ f2 lds64(uint32_t a) {
    float x, y;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x), "=f"(y) : "r"(a));
    return mk2(x, y);
}
__device__ __forceinline__ float lds32(uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}

// Neighbourhood of one thread: RX quads (4 consecutive flat elements each) per field.  Arm names are the
// LOGICAL ones of the coefficient table (T = time, X, Y), mapped to the physical neighbours listed above.
template <int NF, int RX, uint64_t MASKS, uint32_t QUAD_STRIDE>
struct TinCtx {
    const StarCoef& k;
    float (&prev)[NF][RX][4];      // x-1
    float (&cur)[NF][RX][4];
    float (&next)[NF][RX][4];      // x+1
    uint32_t cur_a;                // shared address of this thread's quad 0, field 0, in the slot of row x
    uint32_t buf;                  // bytes between the tiles of consecutive fields
    uint32_t nt_bytes;             // line length * 4
    uint32_t tmask[RX];            // bits 0..3: element i has a flat-1 neighbour in its line; bits 4..7: a flat+1 one
    bool odd;                      // kFLEX: odd line length (the line arms are then two 32-bit loads)

    // operand pair (elements i0, i1 of quad j, field FD) seen through the LOGICAL arm ARM
    template <int ARM, int FD, int IP>
    __device__ __forceinline__ f2 operand() const {
        constexpr int j = IP / 2, h = IP % 2, i0 = 2 * h, i1 = 2 * h + 1;
        const uint32_t fa = cur_a + FD * buf + j * QUAD_STRIDE + i0 * 4u;    // element i0 of quad j, field FD
        const uint32_t tm = tmask[j];
        if constexpr (ARM == ARM_C) {
            return mk2(cur[FD][j][i0], cur[FD][j][i1]);
        } else if constexpr (ARM == A_IM) {          // flat - 1, switched off at the first element of a line
            float a;
            if constexpr (h == 0) a = lds32(fa - 4u);
            else a = cur[FD][j][1];
            const float b = cur[FD][j][i0];
            return mk2((tm >> i0) & 1u ? a : 0.0f, (tm >> i1) & 1u ? b : 0.0f);
        } else if constexpr (ARM == A_IP) {          // flat + 1, switched off at the last element of a line
            const float a = cur[FD][j][i1];
            float b;
            if constexpr (h == 1) b = lds32(fa + 8u);
            else b = cur[FD][j][2];
            return mk2((tm >> (4 + i0)) & 1u ? a : 0.0f, (tm >> (4 + i1)) & 1u ? b : 0.0f);
        } else if constexpr (ARM == A_MM) {
            return mk2(prev[FD][j][i0], prev[FD][j][i1]);
        } else if constexpr (ARM == A_MP) {
            return mk2(next[FD][j][i0], next[FD][j][i1]);
        } else if constexpr (ARM == A_LM) {          // previous line: flat - line length (8-byte aligned when it is even)
            if (kFLEX && odd) return mk2(lds32(fa - nt_bytes), lds32(fa - nt_bytes + 4u));
            return lds64(fa - nt_bytes);
        } else {                                     // A_LP
            if (kFLEX && odd) return mk2(lds32(fa + nt_bytes), lds32(fa + nt_bytes + 4u));
            return lds64(fa + nt_bytes);
        }
    }
    template <int SLOT, int FD, int IP>
    __device__ __forceinline__ f2 D() const {
        constexpr uint32_t m = (uint32_t)((MASKS >> (7 * SLOT)) & 0x7f);
        const float* c = k.c[SLOT];
        // accumulation in the logical row-major tap order (t-, x-, y-, centre, y+, x+, t+), like every other kernel
        f2 acc = mk2(0.0f, 0.0f);
        if constexpr ((m >> ARM_TM) & 1) acc = pfma(operand<ARM_TM, FD, IP>(), c[ARM_TM], acc);
        if constexpr ((m >> ARM_XM) & 1) acc = pfma(operand<ARM_XM, FD, IP>(), c[ARM_XM], acc);
        if constexpr ((m >> ARM_YM) & 1) acc = pfma(operand<ARM_YM, FD, IP>(), c[ARM_YM], acc);
        if constexpr ((m >> ARM_C) & 1) acc = pfma(operand<ARM_C, FD, IP>(), c[ARM_C], acc);
        if constexpr ((m >> ARM_YP) & 1) acc = pfma(operand<ARM_YP, FD, IP>(), c[ARM_YP], acc);
        if constexpr ((m >> ARM_XP) & 1) acc = pfma(operand<ARM_XP, FD, IP>(), c[ARM_XP], acc);
        if constexpr ((m >> ARM_TP) & 1) acc = pfma(operand<ARM_TP, FD, IP>(), c[ARM_TP], acc);
        return acc;
    }
    template <int FD, int IP>
    __device__ __forceinline__ f2 V() const {
        return mk2(cur[FD][IP / 2][2 * (IP % 2)], cur[FD][IP / 2][2 * (IP % 2) + 1]);
    }
};

// Single-field programs run three CTAs per SM (80 registers, no spill: wave on 256 x 254 rows 443 -> 507 Gpoints/s);
// with more fields the cap costs a spill frame that outweighs the extra warps (NS momentum 342 -> 244), so they stay at two.
template <int KIND, int RX, uint64_t MASKS, int MODE>
__global__ void __launch_bounds__(kThreads, (Prog<KIND>::NF == 1 ? 3 : 2))
tin_kernel(const __grid_constant__ TinParams<Prog<KIND>::NF> p) {
    constexpr int NF = Prog<KIND>::NF;
    constexpr bool STORE = MODE == CPPRE_EPI_STORE;
    constexpr bool COVER = MODE == CPPRE_EPI_COVER;
    constexpr uint32_t QUAD_STRIDE = kThreads * 16u;      // bytes between a thread's consecutive quads

    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
    const uint32_t STAGE_BYTES = NF * p.buf_bytes;
    const uint32_t bars = base + (uint32_t)p.stages * STAGE_BYTES;

    const int tid = threadIdx.x;
    const int tr = tid >> 5;
    const int tile = blockIdx.x % p.tiles;
    const int b = blockIdx.x / p.tiles;
    const int Nt = p.Nt;
    const int64_t Q = (int64_t)p.Ny * Nt;                 // floats per x-row
    const int y0 = tile * p.TY;
    const int64_t q0 = (int64_t)y0 * Nt;                  // flat index of the tile's first centre element
    const int W = min(p.TY, p.Ny - y0) * Nt;              // live centre elements of this tile (a multiple of 4)

    // Shared-memory image of a row tile, element index s <-> flat index g = q0 - (Nt + padl) + s:
    //   [padl pad][halo line y0-1][TY centre lines][halo line y0+TY][pad to a multiple of 4]
    // The copy covers the part of it that lies inside [0, Q); the rest is zero-filled once (it is the same
    // for every x-row of the march) and is the reference's zero padding in y.
    const int lead = Nt + p.padl;
    const int L4 = (p.padl + (p.TY + 2) * Nt + 3) & ~3;
    const int64_t g_first = q0 - lead;                    // a multiple of 4 (may be negative)
    const int s_lo = g_first < 0 ? (int)(-g_first) : 0;
    const int s_hi = (int)min((int64_t)L4, Q - g_first);
    const uint32_t copy_bytes = (uint32_t)(s_hi - s_lo) * 4u;

    // rows that are not 16-byte aligned (kFLEX) are staged by the 32 lanes of the refilling warp with 4-byte
    // cp.async, each lane arriving on the slot's mbarrier when its copies have landed
    const bool coop = kFLEX && !p.bulk;
    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) mbar_init(bars + 8u * s, coop ? 32u : 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    for (int sf = 0; sf < p.stages * NF; ++sf) {
        const uint32_t t0 = base + (uint32_t)sf * p.buf_bytes;
        for (int s = tid; s < L4; s += kThreads)
            if (s < s_lo || s >= s_hi) asm volatile("st.shared.f32 [%0], %1;" ::"r"(t0 + 4u * s), "f"(0.0f) : "memory");
    }
    __syncthreads();

    const int NPL = p.Nx;
    auto issue = [&](int x, int stage) {
        const uint32_t bar = bars + 8u * stage;
        mbar_expect_tx(bar, copy_bytes * NF);
#pragma unroll
        for (int f = 0; f < NF; ++f)
            bulk_load(base + (uint32_t)stage * STAGE_BYTES + f * p.buf_bytes + 4u * s_lo,
                      p.f[f] + (int64_t)b * p.fsb[f] + (int64_t)x * p.frow[f] + g_first + s_lo, copy_bytes, bar);
    };
    auto issue_coop = [&](int x, int stage) {      // all 32 lanes of one warp
        const uint32_t bar = bars + 8u * stage;
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const float* src = p.f[f] + (int64_t)b * p.fsb[f] + (int64_t)x * p.frow[f] + g_first;
            const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES + f * p.buf_bytes;
            for (int s = s_lo + (tid & 31); s < s_hi; s += 32) cp_async4(dst + 4u * s, src + s);
        }
        cp_async_arrive(bar);
    };
    {
        const int pre = NPL < p.stages ? NPL : p.stages;
        if (coop) {
            if (tr == 0)
                for (int q = 0; q < pre; ++q) issue_coop(q, q);
        } else if (tid == 0) {
            for (int q = 0; q < pre; ++q) issue(q, q);
        }
    }

    float prev[NF][RX][4], cur[NF][RX][4], next[NF][RX][4];
#pragma unroll
    for (int f = 0; f < NF; ++f)
#pragma unroll
        for (int j = 0; j < RX; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) prev[f][j][i] = 0.0f;

    // per-quad constants: which elements exist, have t-neighbours, count in a reduction
    // shell flags by PHYSICAL axis: within the line / across lines / along the march (role 0: t / y / x; role 1: y / x / t)
    const bool in_i = (p.e.flags & (kROLE == 0 ? CPPRE_FLAG_INTERIOR_2 : CPPRE_FLAG_INTERIOR_0)) != 0;
    const bool in_l = (p.e.flags & (kROLE == 0 ? CPPRE_FLAG_INTERIOR_0 : CPPRE_FLAG_INTERIOR_1)) != 0;
    const bool in_m = (p.e.flags & (kROLE == 0 ? CPPRE_FLAG_INTERIOR_1 : CPPRE_FLAG_INTERIOR_2)) != 0;
    // live: quads with at least one element of the tile; emask (kFLEX): which of their elements exist (W need not be
    // a multiple of 4 when rows are unaligned); cmask: which of them count in a reduction
    uint32_t live = 0u, tmask[RX], cmask[RX], emask[RX];
#pragma unroll
    for (int j = 0; j < RX; ++j) {
        const int e = 4 * (j * kThreads + tid);
        tmask[j] = 0u;
        cmask[j] = 0u;
        emask[j] = 0u;
        if (e < W) {
            live |= 1u << j;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (kFLEX && e + i >= W) continue;
                emask[j] |= 1u << i;
                const int t = (e + i) % Nt, y = y0 + (e + i) / Nt;
                if (t != 0) tmask[j] |= 1u << i;
                if (t != Nt - 1) tmask[j] |= 1u << (4 + i);
                const bool shell = (in_i && (t == 0 || t == Nt - 1)) || (in_l && (y == 0 || y == p.Ny - 1));
                if (!shell) cmask[j] |= 1u << i;
            }
        }
    }
    (void)emask;
    const bool vec = !kFLEX || p.vec;             // 128-bit epilogue accesses
    const uint32_t cell0 = base + (uint32_t)(lead + 4 * tid) * 4u;   // quad 0, field 0, ring slot 0
    auto load_centres = [&](float (&dst)[NF][RX][4], uint32_t a) {
#pragma unroll
        for (int f = 0; f < NF; ++f)
#pragma unroll
            for (int j = 0; j < RX; ++j) {
                const float4 v = lds128(a + f * p.buf_bytes + j * QUAD_STRIDE);
                dst[f][j][0] = v.x; dst[f][j][1] = v.y; dst[f][j][2] = v.z; dst[f][j][3] = v.w;
            }
    };
    mbar_wait(bars, 0);
    load_centres(cur, cell0);

    const bool rmax = p.e.reduce == CPPRE_REDUCE_MAX;
    const uint32_t abs_mask = (p.e.flags & CPPRE_FLAG_ABS) ? 0x7fffffffu : 0xffffffffu;
    const bool has_ref = p.e.ref != nullptr, has_w = p.e.weight != nullptr;
    const float band_q = (COVER && p.e.band_qp != nullptr) ? __ldg(p.e.band_qp) : p.e.band_q;
    (void)band_q;
    const int64_t off = q0 + 4 * tid;                     // this thread's quad 0 within an x-row
    const float* refp = has_ref ? p.e.ref + (int64_t)b * p.e.rb + off : nullptr;
    const float* wp = has_w ? p.e.weight + off : nullptr;
    float* outp = STORE ? p.e.out + (int64_t)b * p.e.ob + off : nullptr;

    uint32_t acc_u = 0u;       // SCORE/max: running max of the bit images; COVER: count
    float acc_f = 0.0f;        // SCORE/sum

    int stage = 0, wturn = 0;
    int nstage = p.stages > 1 ? 1 : 0;
    uint32_t cur_a = cell0;
    uint32_t nxt_a = cell0 + (p.stages > 1 ? STAGE_BYTES : 0u);
    uint32_t nbar = bars + (p.stages > 1 ? 8u : 0u);
    uint32_t nphase = p.stages > 1 ? 0u : 1u;

    auto step = [&](const int x, float (&P)[NF][RX][4], float (&C)[NF][RX][4], float (&N)[NF][RX][4]) {
        float4 rq[RX], wq[RX];
#pragma unroll
        for (int j = 0; j < RX; ++j) {
            rq[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            wq[j] = make_float4(1.f, 1.f, 1.f, 1.f);
            if (live & (1u << j)) {
                if (vec) {
                    if (has_ref) rq[j] = __ldg(reinterpret_cast<const float4*>(refp + j * (kThreads * 4)));
                    if (has_w) wq[j] = __ldg(reinterpret_cast<const float4*>(wp + j * (kThreads * 4)));
                } else {
                    float* r4 = reinterpret_cast<float*>(&rq[j]);
                    float* w4 = reinterpret_cast<float*>(&wq[j]);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if ((emask[j] >> i) & 1u) {
                            if (has_ref) r4[i] = __ldg(refp + j * (kThreads * 4) + i);
                            if (has_w) w4[i] = __ldg(wp + j * (kThreads * 4) + i);
                        }
                }
            }
        }
        if (x + 1 < NPL) {
            mbar_wait(nbar, nphase);
            load_centres(N, nxt_a);
        } else {
#pragma unroll
            for (int f = 0; f < NF; ++f)
#pragma unroll
                for (int j = 0; j < RX; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) N[f][j][i] = 0.0f;
        }
        TinCtx<NF, RX, MASKS, QUAD_STRIDE> ctx{p.coef, P, C, N, cur_a, p.buf_bytes, (uint32_t)Nt * 4u, {}, (Nt & 1) != 0};
#pragma unroll
        for (int j = 0; j < RX; ++j) ctx.tmask[j] = tmask[j];
        const bool refiller = tr == wturn;
        const bool plane_live = STORE || !(in_m && (x == 0 || x == NPL - 1));
        if (plane_live) {
            static_for<RX>([&](auto J) {
                constexpr int j = decltype(J)::value;
                if (live & (1u << j)) {
                    f2 r01, r23;
                    if constexpr (kDEC) {
                        const DecodeCtx<std::remove_const_t<decltype(ctx)>> dctx{ctx, p.dec};
                        r01 = Prog<KIND>::template eval<2 * j + 0>(dctx, p.scalars);
                        r23 = Prog<KIND>::template eval<2 * j + 1>(dctx, p.scalars);
                    } else {
                        r01 = Prog<KIND>::template eval<2 * j + 0>(ctx, p.scalars);
                        r23 = Prog<KIND>::template eval<2 * j + 1>(ctx, p.scalars);
                    }
                    float r[4] = {r01.v.x, r01.v.y, r23.v.x, r23.v.y};
                    const float rr[4] = {rq[j].x, rq[j].y, rq[j].z, rq[j].w};
                    const float ww[4] = {wq[j].x, wq[j].y, wq[j].z, wq[j].w};
                    if constexpr (COVER) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float h = has_w ? __fmul_rn(band_q, ww[i]) : band_q;
                            bool in;
                            if (has_ref) in = r[i] >= __fsub_rn(rr[i], h) && r[i] <= __fadd_rn(rr[i], h);
                            else in = fabsf(r[i]) <= h;
                            acc_u += (in && ((cmask[j] >> i) & 1u)) ? 1u : 0u;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            float v = has_ref ? __fsub_rn(r[i], rr[i]) : r[i];
                            v = __uint_as_float(__float_as_uint(v) & abs_mask);
                            r[i] = has_w ? __fmul_rn(v, ww[i]) : v;
                        }
                        if constexpr (STORE) {
                            if (vec) {
                                *reinterpret_cast<float4*>(outp + j * (kThreads * 4)) = make_float4(r[0], r[1], r[2], r[3]);
                            } else {
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                    if ((emask[j] >> i) & 1u) outp[j * (kThreads * 4) + i] = r[i];
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if ((cmask[j] >> i) & 1u) {
                                    if (rmax) acc_u = max(acc_u, score_bits(r[i]));
                                    else acc_f += r[i];
                                }
                        }
                    }
                }
            });
        }
        // the tile of row x was read by the arms themselves: the slot is free only now
        if (!refiller) bar_arrive(stage + 1, kThreads);
        if constexpr (STORE) outp += Q;
        if (has_ref) refp += Q;
        if (has_w) wp += Q;
        if (refiller) {
            bar_sync(stage + 1, kThreads);
            if (x + p.stages < NPL) {
                if (coop) issue_coop(x + p.stages, stage);
                else if ((tid & 31) == 0) issue(x + p.stages, stage);
            }
        }
        wturn = (wturn + 1) & (kWarps - 1);
        stage = nstage;
        cur_a = nxt_a;
        nxt_a += STAGE_BYTES;
        nbar += 8u;
        if (++nstage == p.stages) { nstage = 0; nxt_a = cell0; nbar = bars; nphase ^= 1u; }
    };

    for (int x = 0; x < NPL; x += 3) {
        step(x, prev, cur, next);
        if (x + 1 < NPL) step(x + 1, cur, next, prev);
        if (x + 2 < NPL) step(x + 2, next, prev, cur);
    }

    if constexpr (!STORE) {
        if constexpr (COVER) {
            const uint32_t c = __reduce_add_sync(0xffffffffu, acc_u);
            if ((tid & 31) == 0 && c) atomicAdd(reinterpret_cast<unsigned int*>(p.e.scores) + b, c);
        } else if (rmax) {
            const uint32_t m = __reduce_max_sync(0xffffffffu, acc_u);
            if ((tid & 31) == 0 && m) atomicMax(reinterpret_cast<unsigned int*>(p.e.scores) + b, m);
        } else {
            float s = acc_f;
#pragma unroll
            for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if ((tid & 31) == 0) atomicAdd(p.e.scores + b, s);
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct TinHost {
    uint64_t masks;
    StarCoef coef;
};

// 3x3x3 "same"-padded star taps only (what every Taylor-2 operator of ConvOps_2d builds)
bool classify(const cppre_program_t* prog, const Prepared& P, TinHost* H) {
    const Geometry& g = P.g;
    if (g.O2 != g.N2 || g.O1 != g.N1 || g.O0 != g.N0) return false;
    H->masks = 0;
    memset(&H->coef, 0, sizeof(H->coef));
    for (int s = 0; s < P.n_slots; ++s) {
        const int* k = prog->ksize[s];
        const int* pd = prog->pad[s];
        const bool k333 = k[0] == 3 && k[1] == 3 && k[2] == 3 && pd[0] == 1 && pd[1] == 1 && pd[2] == 1;
        // 1-D + time data [B, Nt, 1, Nx] (ConvOps_1d): no x arm at all, one line per plane (role 1 only)
        const bool k313 = kFLEX && g.N1 == 1 && k[0] == 3 && k[1] == 1 && k[2] == 3 && pd[0] == 1 && pd[1] == 0 && pd[2] == 1;
        if (!(k333 || k313)) return false;
        for (int t = P.taps.begin[s]; t < P.taps.begin[s + 1]; ++t) {
            const int a = P.taps.o2[t], b = P.taps.o1[t], c = P.taps.o0[t];
            int arm = -1;
            if (a == 0 && b == 0 && c == 0) arm = ARM_C;
            else if (b == 0 && c == 0) arm = a < 0 ? ARM_TM : ARM_TP;
            else if (a == 0 && c == 0) arm = b < 0 ? ARM_XM : ARM_XP;
            else if (a == 0 && b == 0) arm = c < 0 ? ARM_YM : ARM_YP;
            if (arm < 0) return false;
            H->coef.c[s][arm] = P.taps.c[t];
            H->masks |= (uint64_t)1 << (7 * s + arm);
        }
    }
    return true;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Physical extents by role: (inner = line length, lines, march).
//   role 0: logical [B, N2 = Nt, N1 = Nx, N0 = Ny] stored as [B][Nx][Ny][Nt]: strides (sb, 1, Ny*Nt, Nt)
//   role 1: the same logical array stored contiguously [B][Nt][Nx][Ny]:    strides (sb, Nx*Ny, Ny, 1)
struct Phys {
    int inner, lines, march;
};
Phys phys(const Geometry& g) {
    return kROLE == 0 ? Phys{g.N2, g.N0, g.N1} : Phys{g.N0, g.N1, g.N2};
}
// a FIELD may also be stationary (role 1): one contiguous plane read for every row of the march and every sample
// (march and sample strides 0) -- a coefficient plane such as JOREK's R broadcast over time and batch
bool stationary(const Geometry& g, int64_t sb, int64_t s2, int64_t s1, int64_t s0) {
    const Phys ph = phys(g);
    const int64_t si = kROLE == 0 ? s2 : s0, sl = kROLE == 0 ? s0 : s1, sm = kROLE == 0 ? s1 : s2;
    return kFLEX && sm == 0 && sb == 0 && (ph.inner == 1 || si == 1) && (ph.lines == 1 || sl == ph.inner);
}
bool role_strides(const Geometry& g, int64_t sb, int64_t s2, int64_t s1, int64_t s0, bool field = false) {
    const Phys ph = phys(g);
    const int64_t plane = (int64_t)ph.lines * ph.inner;
    if (field && stationary(g, sb, s2, s1, s0)) return true;
    // strides on extent-1 axes are never used for addressing
    const int64_t si = kROLE == 0 ? s2 : s0, sl = kROLE == 0 ? s0 : s1, sm = kROLE == 0 ? s1 : s2;
    if ((ph.inner > 1 && si != 1) || (ph.lines > 1 && sl != ph.inner) || (ph.march > 1 && sm != plane)) return false;
    if (g.B > 1 && sb < plane * ph.march) return false;
    return kFLEX || g.B == 1 || sb % 4 == 0;
}

bool layout_ok(const Prepared& P) {
    const Geometry& g = P.g;
    const Phys ph = phys(g);
    if (P.in_bf16) return false;
    if (!kFLEX) {
        if (ph.inner < 4 || ph.inner % 2 != 0 || ph.inner > 512) return false;   // even lines: the line arms are 64-bit shared loads
        if (((int64_t)ph.lines * ph.inner) % 4 != 0 || ph.lines < 1 || ph.march < 1) return false;
    } else {
        // a tile is a whole number of lines covering a multiple of 4 elements: 1 / 2 / 4 lines at least
        const int q = ph.inner % 4 == 0 ? 1 : (ph.inner % 2 == 0 ? 2 : 4);
        if (ph.inner < 2 || (int64_t)q * ph.inner > kThreads * 4 || ph.lines < 1 || ph.march < 1) return false;
    }
    for (int f = 0; f < P.n_fields; ++f) {
        const FieldView& v = P.fields[f];
        if ((!kFLEX && !aligned16(v.p)) || (reinterpret_cast<uintptr_t>(v.p) & 3u) || !role_strides(g, v.sb, v.s2, v.s1, v.s0, true))
            return false;
    }
    const EpiDev& e = P.epi;
    if (CPPRE_TIN_MODE == CPPRE_EPI_STORE && ((!kFLEX && !aligned16(e.out)) || !role_strides(g, e.ob, e.o2, e.o1, e.o0))) return false;
    if (e.ref && ((!kFLEX && !aligned16(e.ref)) || !role_strides(g, e.rb, e.r2, e.r1, e.r0))) return false;
    if (e.weight && ((!kFLEX && !aligned16(e.weight)) || !(kROLE == 0 ? P.weight_tin : P.weight_contig))) return false;
    return true;
}

template <int KIND, int RX, uint64_t MASKS>
int launch_variant(const Prepared& P, const TinHost& H, cudaStream_t st) {
    constexpr int NF = Prog<KIND>::NF;
    TinParams<NF> tp;
    memset(&tp, 0, sizeof(tp));
    const Geometry& g = P.g;
    const Phys ph = phys(g);
    const int Nt = ph.inner, Nx = ph.march, Ny = ph.lines;      // physical roles, named after role 0
    for (int f = 0; f < NF; ++f) {
        const FieldView& v = P.fields[f];
        const bool still = stationary(g, v.sb, v.s2, v.s1, v.s0) && !(g.B == 1 && Nx == 1);
        tp.f[f] = v.p;
        tp.fsb[f] = (g.B > 1 && !still) ? v.sb : 0;
        tp.frow[f] = still ? 0 : (int64_t)Ny * Nt;
    }
    tp.coef = H.coef;
    memcpy(tp.scalars, P.scalars, sizeof(tp.scalars));
    tp.e = P.epi;
    tp.dec = P.dec;
    tp.Nt = Nt; tp.Nx = Nx; tp.Ny = Ny;
    // lines per tile: as many as RX quads per thread hold, TY*Nt a multiple of 4
    const int lq = Nt % 4 == 0 ? 1 : (Nt % 2 == 0 ? 2 : 4);     // lines per 4-element-aligned group
    int TY = (RX * kThreads * 4) / Nt / lq * lq;
    if (TY > Ny) TY = (Ny + lq - 1) / lq * lq;
    CPPRE_REQUIRE(TY >= 1, CPPRE_ERR_UNSUPPORTED, "lines of %d elements do not fit a row tile", Nt);
    tp.TY = TY;
    tp.tiles = (Ny + TY - 1) / TY;
    {
        // bulk copies need 16-byte aligned rows of every field; the 128-bit epilogue needs the same of out / ref / weight
        const int64_t Q = (int64_t)Ny * Nt;
        bool bulk = Q % 4 == 0, vec = Q % 4 == 0;
        for (int f = 0; f < NF; ++f) bulk = bulk && aligned16(P.fields[f].p) && (tp.fsb[f] % 4 == 0);
        if (CPPRE_TIN_MODE == CPPRE_EPI_STORE) vec = vec && aligned16(P.epi.out) && (g.B == 1 || P.epi.ob % 4 == 0);
        if (P.epi.ref) vec = vec && aligned16(P.epi.ref) && (g.B == 1 || P.epi.rb % 4 == 0);
        if (P.epi.weight) vec = vec && aligned16(P.epi.weight);
        tp.bulk = bulk ? 1 : 0;
        tp.vec = vec ? 1 : 0;
    }
    tp.padl = (4 - Nt % 4) % 4;
    const int lead = Nt + tp.padl;
    int elems = (tp.padl + (TY + 2) * Nt + 3) & ~3;
    if (elems < lead + RX * kThreads * 4) elems = lead + RX * kThreads * 4;     // idle lanes still load their quads
    tp.buf_bytes = ((uint32_t)elems * 4u + 127u) & ~127u;
    int stages = 4;
    const size_t per_stage = (size_t)NF * tp.buf_bytes;
    while (stages > 2 && stages * per_stage + 256 > 112 * 1024) --stages;
    tp.stages = stages;
    const size_t smem = stages * per_stage + 8 * kMaxStages + 128;
    CPPRE_REQUIRE(smem <= 227 * 1024, CPPRE_ERR_UNSUPPORTED, "row-tile ring of %zu bytes exceeds shared memory", smem);
    auto kern = tin_kernel<KIND, RX, MASKS, CPPRE_TIN_MODE>;
    static SmemOptIn memo;
    {
        const int rc = ensure_dynamic_smem(kern, memo, smem, cudaSharedmemCarveoutMaxShared);
        if (rc != CPPRE_OK) return rc;
    }
    const int64_t blocks = (int64_t)tp.tiles * g.B;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    kern<<<(unsigned)blocks, kThreads, smem, st>>>(tp);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

constexpr uint64_t pack(uint32_t a, uint32_t b = 0, uint32_t c = 0, uint32_t d = 0, uint32_t e = 0) {
    return (uint64_t)a | ((uint64_t)b << 7) | ((uint64_t)c << 14) | ((uint64_t)d << 21) | ((uint64_t)e << 28);
}
typedef int (*LaunchFn)(const Prepared&, const TinHost&, cudaStream_t);
struct Variant {
    int kind;
    uint64_t masks;
    LaunchFn fn;
};
// one quad per thread: two (tiles of 2048 elements) were measured and lose -- NS momentum 342 -> 260 Gpoints/s on
// the FNO layout, 180 -> 136 on ragged rows (spill frame at the 128-register cap), wave +4 %
#define V(K, M) Variant{K, M, &launch_variant<K, 1, M>}
// "shipped" = the kernels the reference constructors produce (D_y == D_t); then any star
constexpr uint64_t M_LIN = pack(ARMS_ALL);
constexpr uint64_t M_NSC = pack(ARMS_X, ARMS_T), M_NSC_FULL = pack(ARMS_ALL, ARMS_ALL);
constexpr uint64_t M_NSM = pack(ARMS_T, ARMS_X, ARMS_T, ARMS_C | ARMS_X | ARMS_Y);
constexpr uint64_t M_NSM_FULL = pack(ARMS_ALL, ARMS_ALL, ARMS_ALL, ARMS_ALL);
constexpr uint64_t M_MHD = pack(ARMS_T, ARMS_X, ARMS_T), M_MHD_FULL = pack(ARMS_ALL, ARMS_ALL, ARMS_ALL);
const Variant kVariants[] = {
    V(CPPRE_PROG_LINEAR, M_LIN),
#if CPPRE_TIN_ROLE == 1
    V(CPPRE_PROG_BURGERS, pack(ARMS_T, ARMS_Y, ARMS_C | ARMS_Y)), V(CPPRE_PROG_BURGERS, pack(ARMS_ALL, ARMS_ALL, ARMS_ALL)),
    // JOREK reduced MHD {D_t, D_R, D_Z, D_RR, D_ZZ} [ref: Marginal/JOREK_residuals_CP.py:203-209]: as the reference's
    // constructors build them (domain 'y' == 't'), then with D_Z / D_ZZ along y
    V(CPPRE_PROG_JOREK_CONTINUITY, pack(ARMS_T, ARMS_X, ARMS_T, ARMS_C | ARMS_X, ARMS_C | ARMS_T)),
    V(CPPRE_PROG_JOREK_CONTINUITY, pack(ARMS_T, ARMS_X, ARMS_Y, ARMS_C | ARMS_X, ARMS_C | ARMS_Y)),
    V(CPPRE_PROG_JOREK_TEMPERATURE, pack(ARMS_T, ARMS_X, ARMS_T, ARMS_C | ARMS_X, ARMS_C | ARMS_T)),
    V(CPPRE_PROG_JOREK_TEMPERATURE, pack(ARMS_T, ARMS_X, ARMS_Y, ARMS_C | ARMS_X, ARMS_C | ARMS_Y)),
#endif
    V(CPPRE_PROG_NS_CONTINUITY, M_NSC), V(CPPRE_PROG_NS_CONTINUITY, M_NSC_FULL),
    V(CPPRE_PROG_NS_MOMENTUM, M_NSM), V(CPPRE_PROG_NS_MOMENTUM, M_NSM_FULL),
    V(CPPRE_PROG_MHD_CONTINUITY, M_MHD), V(CPPRE_PROG_MHD_CONTINUITY, M_MHD_FULL),
    V(CPPRE_PROG_MHD_MOMENTUM, M_MHD), V(CPPRE_PROG_MHD_MOMENTUM, M_MHD_FULL),
    V(CPPRE_PROG_MHD_ENERGY, M_MHD), V(CPPRE_PROG_MHD_ENERGY, M_MHD_FULL),
    V(CPPRE_PROG_MHD_INDUCTION, M_MHD), V(CPPRE_PROG_MHD_INDUCTION, M_MHD_FULL),
    V(CPPRE_PROG_MHD_GAUSS, M_MHD), V(CPPRE_PROG_MHD_GAUSS, M_MHD_FULL),
};
#undef V

const Variant* pick(int kind, const TinHost& H) {
    for (const Variant& v : kVariants)
        if (v.kind == kind && (H.masks & ~v.masks) == 0) return &v;
    return nullptr;
}

}  // namespace

#define CPPRE_CAT_(a, b) a##b
#define CPPRE_CAT(a, b) CPPRE_CAT_(a, b)

#if CPPRE_TIN_ROLE == 1
#define CPPRE_TIN_ELIGIBLE flat_eligible_
#define CPPRE_TIN_LAUNCH launch_flat_
#else
#define CPPRE_TIN_ELIGIBLE tin_eligible_
#define CPPRE_TIN_LAUNCH launch_tin_
#endif

int CPPRE_CAT(CPPRE_TIN_ELIGIBLE, CPPRE_TIN_SUFFIX)(const cppre_program_t* prog, const Prepared& P) {
    TinHost H;
    if (kFLEX && P.dec_on) return 0;        // the decode fold of unaligned data is served by the generic kernel
    if (!classify(prog, P, &H) || !layout_ok(P)) return 0;
    return pick(P.kind, H) != nullptr;
}

int CPPRE_CAT(CPPRE_TIN_LAUNCH, CPPRE_TIN_SUFFIX)(const cppre_program_t* prog, const Prepared& P, cudaStream_t st) {
    TinHost H;
    CPPRE_REQUIRE(classify(prog, P, &H) && layout_ok(P), CPPRE_ERR_UNSUPPORTED,
                  "request does not qualify for the flat-line star kernel of this layout");
    const Variant* v = pick(P.kind, H);
    CPPRE_REQUIRE(v != nullptr, CPPRE_ERR_UNSUPPORTED, "no flat-line variant for program %d", P.kind);
    return v->fn(P, H, st);
}

}  // namespace cppre
