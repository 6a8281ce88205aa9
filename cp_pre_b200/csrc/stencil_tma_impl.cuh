// K1, TMA star kernel (sm_100a): the fast path for residual programs whose derivative slots
// are 3-point-per-axis "star" stencils (every operator ConvOps_1d/2d builds with Taylor order
// 2, and any additive composition of them).
//
// Data movement
//   * each CTA owns a tile of TX rows x 128 columns of one sample and marches along the slow
//     axis N2 (time for contiguous [BS,Nt,Nx,Ny] data);
//   * one elected thread stages plane t+k of every field into a ring of shared-memory tiles
//     with cp.async.bulk.tensor (TMA), completion signalled on an mbarrier.  Tiles carry a
//     +-1 row and +-16 byte (4 fp32 / 8 bf16 cells) column halo; out-of-range coordinates are
//     zero-filled by the TMA unit, which IS the reference's zero "same" padding (F.conv3d padding=k//2) -- no boundary
//     branches in the kernel;
//   * the t-1 / t / t+1 centre values live in registers (a 3-deep register pipeline), so a
//     plane is read from shared memory once for its centre values (128-bit LDS) plus the x+-1
//     rows (128-bit LDS) and two scalar y+-1 edge values per row;
//   * derivative coefficients sit in the kernel parameter (constant) bank and are used as
//     broadcast uniform-register operands of FFMA2 (two neighbouring cells per instruction,
//     programs.cuh); which arms of each slot exist is a template parameter, so zero taps cost
//     nothing;
//   * with a weight-only operand set the per-cell weight tile rides the same ring (WTMA);
//   * CPPRE_TMA_BF16: the field tiles are staged as bf16 and widened on the way into registers.
// Epilogue (fused, same kernel): minus reference residual, |.|, per-cell weight
// (1/modulation), then either a 128-bit coalesced store or a per-sample max/sum kept in
// registers, reduced with redux.sync / shuffles and one atomic per warp -- the Joint score
// never materialises the residual tensor.
//
// This file is the body of several translation units (stencil_tma_<mode>_<oper>.cu): each
// defines CPPRE_TMA_MODE (the epilogue mode), CPPRE_TMA_OPER (which optional epilogue operands
// exist: -1 = decided at run time, else bit 0 = ref, bit 1 = weight, compiled in) and
// CPPRE_TMA_SUFFIX (the suffix of the two functions exported to api.cu).
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "programs.cuh"

namespace cppre {

namespace {

constexpr int kTPR = 32;            // threads per tile row (one warp)
constexpr int kWarps = 8;           // thread rows per CTA
constexpr int kThreads = kTPR * kWarps;
constexpr int kTW = 4 * kTPR;       // tile width in cells (128)
constexpr int kMaxStages = 8;
constexpr int kDefaultL2Hints = 0;   // see StarParams::l2_hints; CPPRE_TMA_L2_HINTS overrides
// Field element type of this translation unit: fp32, or bf16 (CPPRE_TMA_BF16: the bf16-input
// mode -- tiles are staged as bf16, halving the field bytes read from HBM, and widened to fp32
// on the way into registers; all arithmetic and every output stay fp32).
#ifndef CPPRE_TMA_BF16
#define CPPRE_TMA_BF16 0
#endif
constexpr bool kBF16 = CPPRE_TMA_BF16 != 0;
// CPPRE_TMA_DECODE: the normaliser-decode fold (CPPRE_FLAG_DECODE, programs.cuh DecodeCtx) is compiled into
// this translation unit's kernels; the other units carry none of it.
#ifndef CPPRE_TMA_DECODE
#define CPPRE_TMA_DECODE 0
#endif
constexpr bool kDEC = CPPRE_TMA_DECODE != 0;
constexpr uint32_t kES = kBF16 ? 2u : 4u;      // bytes per staged field element
// smem row: 16 bytes of halo/pad each side (4 fp32 / 8 bf16 cells).  The TMA box must START on a
// 16-byte boundary of the global row (col0 - kHP), and the interior stays 16-byte aligned.
constexpr int kHP = 16 / (int)kES;
constexpr int kTWP = kTW + 2 * kHP;

template <int NF>
struct alignas(64) StarParams {
    CUtensorMap map[NF];
    CUtensorMap wmap;     // per-cell epilogue weight [N2][N1][N0] (used when it is staged by TMA, see WTMA)
    StarCoef coef;
    float scalars[CPPRE_MAX_SCALARS];
    EpiDev e;
    int32_t N2, N1, N0;   // logical extents (N1 = 1 in batch-rows mode)
    int32_t R;            // extent of the tile-row axis: N1, or B in batch-rows mode
    int32_t col_tiles, row_tiles;
    int32_t stages;
    uint32_t buf_bytes;   // bytes of one (stage, field) tile, 128-aligned
    int32_t l2_hints;     // bit 0: field tiles evict_first, bit 1: weight tile evict_last
    DecodeDev dec;        // read by the CPPRE_TMA_DECODE units only
};

template <int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (N > 0) {
        static_for<N - 1>(f);
        f(std::integral_constant<int, N - 1>{});
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
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
// Named barriers 1..stages (one per ring slot): consumers signal "done reading this slot" without
// blocking, the warp whose turn it is to refill the slot waits for all of them.
__device__ __forceinline__ void bar_arrive(uint32_t id, uint32_t threads) {
    asm volatile("barrier.cta.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void bar_sync(uint32_t id, uint32_t threads) {
    asm volatile("barrier.cta.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// L2 eviction-priority hints for the TMA loads: the field tiles are streamed once (evict_first), the
// per-cell weight tile (qhat / modulation, 13-26 MB) is re-read by every sample and should stay (evict_last)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void tma_load_3d_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                                 int c2, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(pol)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d_hint(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                                 int c2, int c3, uint64_t pol) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
        " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(pol)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

__device__ __forceinline__ float4 lds128(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ float lds32(uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
// four / one staged field cells -> fp32 registers (bf16: one 64-bit load, widening by shift / mask)
__device__ __forceinline__ float4 lds_cells4(uint32_t a) {
    if constexpr (kBF16) {
        uint32_t w0, w1;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(w0), "=r"(w1) : "r"(a));
        return make_float4(__uint_as_float(w0 << 16), __uint_as_float(w0 & 0xffff0000u),
                           __uint_as_float(w1 << 16), __uint_as_float(w1 & 0xffff0000u));
    } else {
        return lds128(a);
    }
}
__device__ __forceinline__ float lds_cell(uint32_t a) {
    if constexpr (kBF16) {
        uint16_t h;
        asm volatile("ld.shared.u16 %0, [%1];" : "=h"(h) : "r"(a));
        return __uint_as_float((uint32_t)h << 16);
    } else {
        return lds32(a);
    }
}

// two staged field cells (a 64-bit / 32-bit aligned pair) -> an fp32x2 operand
__device__ __forceinline__ f2 lds_cells2(uint32_t a) {
    if constexpr (kBF16) {
        uint32_t w;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(w) : "r"(a));
        return mk2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
    } else {
        float x, y;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x), "=f"(y) : "r"(a));
        return mk2(x, y);
    }
}

// Register-resident neighbourhood of one thread: RX rows x 4 columns per field.
// LAZY (programs with >= 4 fields): the x+-1 rows and the y+-1 edge cells are NOT preloaded into
// registers for all fields at once (6 fields x (2 x 4 + 2) values = 60 registers, which pushed the
// MHD momentum / energy kernels past the 128-register budget of 2 CTAs per SM and into local-memory
// spills); each derivative arm reads its two / one cells from the shared-memory tile at the point
// of use instead (64-bit / 32-bit ld.shared), so they live for one FFMA2 only.
template <int NF, int RX, uint64_t MASKS, bool LAZY, uint32_t BUF, uint32_t ROW>
struct StarCtx {
    const StarCoef& k;
    float (&prev)[NF][RX][4];
    float (&cur)[NF][RX][4];
    float (&next)[NF][RX][4];
    float (&xm)[NF][4];
    float (&xp)[NF][4];
    float (&ym)[NF][RX];
    float (&yp)[NF][RX];
    uint32_t cur_a;      // shared address of this thread's first centre cell of field 0 in the slot of plane t

    // Two neighbouring columns (2h, 2h+1) of row j per evaluation: IP = 2*j + h.  Every arm is
    // one FFMA2 whose coefficient is a broadcast uniform-register operand; pairs that are not a
    // register pair of the 128-bit shared-memory loads (the y+-1 arms) cost one extra MOV.
    template <int SLOT, int FD, int IP>
    __device__ __forceinline__ f2 D() const {
        constexpr int j = IP / 2, h = IP % 2, i0 = 2 * h, i1 = 2 * h + 1;
        constexpr uint32_t m = (uint32_t)((MASKS >> (7 * SLOT)) & 0x7f);
        constexpr int jm = j > 0 ? j - 1 : 0, jp = j < RX - 1 ? j + 1 : RX - 1;
        const float* c = k.c[SLOT];
        const uint32_t fa = cur_a + FD * BUF + j * ROW;     // this thread's cell 0 of row j, field FD
        // accumulation in row-major tap order (t-, x-, y-, centre, y+, x+, t+), like the oracle
        f2 acc = mk2(0.0f, 0.0f);
        if constexpr ((m >> ARM_TM) & 1) acc = pfma(mk2(prev[FD][j][i0], prev[FD][j][i1]), c[ARM_TM], acc);
        if constexpr ((m >> ARM_XM) & 1) {
            f2 v;
            if constexpr (j > 0) v = mk2(cur[FD][jm][i0], cur[FD][jm][i1]);
            else if constexpr (LAZY) v = lds_cells2(fa - ROW + i0 * kES);
            else v = mk2(xm[FD][i0], xm[FD][i1]);
            acc = pfma(v, c[ARM_XM], acc);
        }
        if constexpr ((m >> ARM_YM) & 1) {
            float e;
            if constexpr (h != 0) e = cur[FD][j][1];
            else if constexpr (LAZY) e = lds_cell(fa - kES);
            else e = ym[FD][j];
            acc = pfma(mk2(e, cur[FD][j][i0]), c[ARM_YM], acc);
        }
        if constexpr ((m >> ARM_C) & 1) acc = pfma(mk2(cur[FD][j][i0], cur[FD][j][i1]), c[ARM_C], acc);
        if constexpr ((m >> ARM_YP) & 1) {
            float e;
            if constexpr (h != 1) e = cur[FD][j][2];
            else if constexpr (LAZY) e = lds_cell(fa + 4u * kES);
            else e = yp[FD][j];
            acc = pfma(mk2(cur[FD][j][i1], e), c[ARM_YP], acc);
        }
        if constexpr ((m >> ARM_XP) & 1) {
            f2 v;
            if constexpr (j < RX - 1) v = mk2(cur[FD][jp][i0], cur[FD][jp][i1]);
            else if constexpr (LAZY) v = lds_cells2(fa + ROW + i0 * kES);
            else v = mk2(xp[FD][i0], xp[FD][i1]);
            acc = pfma(v, c[ARM_XP], acc);
        }
        if constexpr ((m >> ARM_TP) & 1) acc = pfma(mk2(next[FD][j][i0], next[FD][j][i1]), c[ARM_TP], acc);
        return acc;
    }
    template <int FD, int IP>
    __device__ __forceinline__ f2 V() const {
        return mk2(cur[FD][IP / 2][2 * (IP % 2)], cur[FD][IP / 2][2 * (IP % 2) + 1]);
    }
};

// OPER: which optional epilogue operands exist, known at compile time (bit 0 = ref, bit 1 =
// weight) or -1 = decided at run time (any combination; a few predicated instructions more).
template <int KIND, bool BATCH_ROWS, int RX, uint64_t MASKS, int MODE, int OPER>
__global__ void __launch_bounds__(kThreads, 2)
star_kernel(const __grid_constant__ StarParams<Prog<KIND>::NF> p) {
    constexpr int NF = Prog<KIND>::NF;
    constexpr bool STORE = MODE == CPPRE_EPI_STORE;
    constexpr bool SCORE = MODE == CPPRE_EPI_SCORE;
    constexpr bool COVER = MODE == CPPRE_EPI_COVER;
    constexpr int TX = kWarps * RX;               // rows per tile
    constexpr int HALO = BATCH_ROWS ? 0 : 1;      // row halo only when rows are a stencil axis
    constexpr int SR = TX + 2 * HALO;             // smem rows per tile
    constexpr uint32_t BOX_BYTES = (uint32_t)(SR * kTWP) * kES;
    constexpr uint32_t BUF_BYTES = (BOX_BYTES + 127u) & ~127u;   // one (stage, field) tile
    // WTMA: with a compiled-in weight-only operand set on 3-D problems in a reducing mode (the
    // marginal COVER / joint SCORE launches), the per-cell weight tile of plane t rides the same
    // TMA ring as the fields: the epilogue reads it with one 128-bit ld.shared instead of an
    // ld.global whose L2 round trip was 25 % of the COVER kernel's stall samples.
    constexpr bool WTMA = OPER == 2 && !BATCH_ROWS && !STORE;
    constexpr uint32_t WBUF_BYTES = WTMA ? (uint32_t)(TX * kTW * sizeof(float)) : 0u;
    constexpr uint32_t STAGE_BYTES = NF * BUF_BYTES + WBUF_BYTES;
    constexpr uint32_t ROW_BYTES = kTWP * kES;
    // see StarCtx: the halo operands of programs with many fields are read at the point of use
    constexpr bool LAZY = NF >= 4 && !BATCH_ROWS;
    (void)SCORE;

    // Shared memory is addressed by 32-bit shared-window addresses only (ld.shared / TMA /
    // mbarrier), so no generic-address arithmetic is left in the time loop.
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;   // TMA destinations: 128-byte aligned
    const uint32_t bars = base + (uint32_t)p.stages * STAGE_BYTES;  // mbarriers follow the ring

    const int tid = threadIdx.x;
    const int tc = tid & (kTPR - 1);
    const int tr = tid >> 5;

    // flattened grid: column tile fastest, then row tile, then outer (sample) index, so CTAs
    // that share halo rows are launched together and hit each other's lines in L2
    int bid = blockIdx.x;
    const int ct = bid % p.col_tiles;
    bid /= p.col_tiles;
    const int rt = bid % p.row_tiles;
    const int outer = bid / p.row_tiles;
    const int col0 = ct * kTW;
    const int row0 = rt * TX;

    if (tid == 0) {
        for (int s = 0; s < p.stages; ++s) mbar_init(bars + 8u * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    const int NPL = p.N2;
    const uint64_t pol_first = l2_policy_evict_first(), pol_last = l2_policy_evict_last();
    const bool hint_f = (p.l2_hints & 1) != 0, hint_w = (p.l2_hints & 2) != 0;
    auto issue = [&](int plane, int stage) {
        const uint32_t bar = bars + 8u * stage;
        mbar_expect_tx(bar, BOX_BYTES * NF + WBUF_BYTES);
        if constexpr (WTMA) {
            const uint32_t wdst = base + (uint32_t)stage * STAGE_BYTES + NF * BUF_BYTES;
            if (hint_w) tma_load_3d_hint(wdst, &p.wmap, bar, col0, row0, plane, pol_last);
            else tma_load_3d(wdst, &p.wmap, bar, col0, row0, plane);
        }
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES + f * BUF_BYTES;
            // map dims (N0, N2, B, 1) with samples as tile rows (batch-rows mode) or (N0, N1, N2, B)
            const int c1 = BATCH_ROWS ? plane : row0 - HALO, c2 = BATCH_ROWS ? row0 : plane, c3 = BATCH_ROWS ? 0 : outer;
            if (hint_f) tma_load_4d_hint(dst, &p.map[f], bar, col0 - kHP, c1, c2, c3, pol_first);
            else tma_load_4d(dst, &p.map[f], bar, col0 - kHP, c1, c2, c3);
        }
    };
    if (tid == 0) {
        const int pre = NPL < p.stages ? NPL : p.stages;
        for (int q = 0; q < pre; ++q) issue(q, q);
    }

    float prev[NF][RX][4], cur[NF][RX][4], next[NF][RX][4];
    float xm[NF][4], xp[NF][4], ym[NF][RX], yp[NF][RX];
#pragma unroll
    for (int f = 0; f < NF; ++f)
#pragma unroll
        for (int j = 0; j < RX; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) prev[f][j][i] = 0.0f;

    // shared address of this thread's first centre cell in ring slot 0
    const uint32_t cell0 = base + (uint32_t)((HALO + tr * RX) * kTWP + kHP + 4 * tc) * kES;
    // this thread's weight row 0 relative to its centre cell, within one ring slot
    const uint32_t woff = (NF * BUF_BYTES + (uint32_t)((tr * RX * kTW + 4 * tc) * sizeof(float))) - (cell0 - base);
    (void)woff;
    auto load_centres = [&](float (&dst)[NF][RX][4], uint32_t a) {
#pragma unroll
        for (int f = 0; f < NF; ++f)
#pragma unroll
            for (int j = 0; j < RX; ++j) {
                const float4 v = lds_cells4(a + f * BUF_BYTES + j * ROW_BYTES);
                dst[f][j][0] = v.x; dst[f][j][1] = v.y; dst[f][j][2] = v.z; dst[f][j][3] = v.w;
            }
    };

    mbar_wait(bars, 0);
    load_centres(cur, cell0);

    const int y = col0 + 4 * tc;                   // first column of this thread
    const bool col_live = y < p.N0;                // N0 % 4 == 0: a float4 is all in or all out
    const bool rmax = p.e.reduce == CPPRE_REDUCE_MAX;
    const bool in0 = (p.e.flags & CPPRE_FLAG_INTERIOR_0) != 0;
    const bool in1 = (p.e.flags & CPPRE_FLAG_INTERIOR_1) != 0;
    const bool in2 = (p.e.flags & CPPRE_FLAG_INTERIOR_2) != 0;
    // |.| as a bit mask: one LOP3 per value, no branch
    const uint32_t abs_mask = (p.e.flags & CPPRE_FLAG_ABS) ? 0x7fffffffu : 0xffffffffu;
    // reducing modes: which of the four columns count (y == 0 / y == N0-1 dropped by INTERIOR_0)
    const bool skip_first = !STORE && in0 && y == 0;
    const bool skip_last = !STORE && in0 && y + 3 == p.N0 - 1;
    const bool has_ref = OPER < 0 ? (p.e.ref != nullptr) : ((OPER & 1) != 0);
    const bool has_w = OPER < 0 ? (p.e.weight != nullptr) : ((OPER & 2) != 0);
    // COVER: q by value, or read once from the device scalar the quantile kernel wrote
    const float band_q = (COVER && p.e.band_qp != nullptr) ? __ldg(p.e.band_qp) : p.e.band_q;
    (void)band_q;

    // Per-thread row state, fixed over the time loop: which of the RX rows exist (live) and
    // which of them count in a reduction (cnt: not on the x-shell when INTERIOR_1 is set).
    // Epilogue operands are walked with one pointer each, advanced by the plane stride.
    const int row_first = row0 + tr * RX;
    uint32_t live = 0u, cnt_rows = 0u;
#pragma unroll
    for (int j = 0; j < RX; ++j) {
        const int row = row_first + j;
        if (col_live && row < p.R) {
            live |= 1u << j;
            if (!(in1 && !BATCH_ROWS && (row == 0 || row == p.N1 - 1))) cnt_rows |= 1u << j;
        }
    }
    const int b0 = BATCH_ROWS ? row_first : outer;
    const int x0 = BATCH_ROWS ? 0 : row_first;
    const int64_t ref_row = BATCH_ROWS ? p.e.rb : p.e.r1;
    const int64_t out_row = BATCH_ROWS ? p.e.ob : p.e.o1;
    const int64_t w_row = BATCH_ROWS ? 0 : p.N0;
    const int64_t w_plane = (int64_t)p.N1 * p.N0;
    const float* refp = has_ref ? p.e.ref + (int64_t)b0 * p.e.rb + (int64_t)x0 * p.e.r1 + y : nullptr;
    const float* wp = has_w ? p.e.weight + (int64_t)x0 * p.N0 + y : nullptr;
    float* outp = STORE ? p.e.out + (int64_t)b0 * p.e.ob + (int64_t)x0 * p.e.o1 + y : nullptr;

    uint32_t acc_u[RX];                            // SCORE/max: running max of the bit images; COVER: count
    float acc_f[RX];                               // SCORE/sum
#pragma unroll
    for (int j = 0; j < RX; ++j) { acc_u[j] = 0u; acc_f[j] = 0.0f; }

    // ring state: slot index (for the producer), shared addresses of this thread's cell in the
    // slots holding planes t and t+1, barrier address and parity guarding plane t+1
    int stage = 0;
    int wturn = 0;                                 // warp that refills the slot freed by this step
    int nstage = p.stages > 1 ? 1 : 0;
    uint32_t cur_a = cell0;
    uint32_t nxt_a = cell0 + (p.stages > 1 ? STAGE_BYTES : 0u);
    uint32_t nbar = bars + (p.stages > 1 ? 8u : 0u);
    uint32_t nphase = p.stages > 1 ? 0u : 1u;

    // One time step.  (P, C, N) are the register planes holding t-1, t, t+1; the caller rotates
    // the three arrays by NAME (the loop below is unrolled by three), so no register is moved.
    auto step = [&](const int t, float (&P)[NF][RX][4], float (&C)[NF][RX][4], float (&N)[NF][RX][4]) {
        // epilogue operands of row 0 first (their L2 latency overlaps the barrier wait and the
        // stencil math); row j+1's are requested just before row j is evaluated
        float4 rq[RX], wq[RX];
        auto fetch_operands = [&](auto J) {
            constexpr int j = decltype(J)::value;
            rq[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            wq[j] = make_float4(1.f, 1.f, 1.f, 1.f);
            if (live & (1u << j)) {
                if (has_ref) rq[j] = __ldg(reinterpret_cast<const float4*>(refp + j * ref_row));
                if constexpr (WTMA) wq[j] = lds128(cur_a + woff + j * (uint32_t)(kTW * sizeof(float)));
                else if (has_w) wq[j] = __ldg(reinterpret_cast<const float4*>(wp + j * w_row));
            }
        };
        fetch_operands(std::integral_constant<int, 0>{});
        if (t + 1 < NPL) {
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
        // x+-1 rows and y+-1 edge cells of plane t are fetched just before the row that needs
        // them (row_halo below), which keeps their registers live for one row only
        auto row_halo = [&](auto J) {
            constexpr int j = decltype(J)::value;
#pragma unroll
            for (int f = 0; f < NF; ++f) {
                const uint32_t a = cur_a + f * BUF_BYTES;
                if constexpr (!BATCH_ROWS) {
                    if constexpr (j == 0) {
                        const float4 lo = lds_cells4(a - ROW_BYTES);
                        xm[f][0] = lo.x; xm[f][1] = lo.y; xm[f][2] = lo.z; xm[f][3] = lo.w;
                    }
                    if constexpr (j == RX - 1) {
                        const float4 hi = lds_cells4(a + RX * ROW_BYTES);
                        xp[f][0] = hi.x; xp[f][1] = hi.y; xp[f][2] = hi.z; xp[f][3] = hi.w;
                    }
                } else if constexpr (j == 0) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) { xm[f][i] = 0.0f; xp[f][i] = 0.0f; }
                }
                ym[f][j] = lds_cell(a + j * ROW_BYTES - kES);
                yp[f][j] = lds_cell(a + j * ROW_BYTES + 4u * kES);
            }
        };
        const StarCtx<NF, RX, MASKS, LAZY, BUF_BYTES, ROW_BYTES> ctx{p.coef, P, C, N, xm, xp, ym, yp, cur_a};

        // Slot `stage` (plane t) is read for the last time by the halo loads of the last row.
        // Every warp but the refiller of this step signals that on the slot's named barrier and
        // carries on; nobody waits at the end of a step, so warps drift apart by up to the ring
        // depth and the scheduler always has math to issue while other warps wait for data.
        const bool refiller = tr == wturn;
        const bool plane_live = STORE || !(in2 && (t == 0 || t == NPL - 1));
        if (!plane_live && !refiller) bar_arrive(stage + 1, kThreads);
        if (plane_live) {
            static_for<RX>([&](auto J) {
                constexpr int j = decltype(J)::value;
                if constexpr (!LAZY) {
                    row_halo(J);
                    if constexpr (j == RX - 1)
                        if (!refiller) bar_arrive(stage + 1, kThreads);
                }
                if constexpr (j + 1 < RX) fetch_operands(std::integral_constant<int, j + 1>{});
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
                        // lo = c - q*h, hi = c + q*h with NumPy's separately rounded fp32 ops
                        uint32_t cnt = 0u;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float h = has_w ? __fmul_rn(band_q, ww[i]) : band_q;
                            bool in;
                            if (has_ref) {
                                const float lo = __fsub_rn(rr[i], h), hi = __fadd_rn(rr[i], h);
                                in = r[i] >= lo && r[i] <= hi;
                            } else {
                                // centre 0: -h <= r <= h is |r| <= h for every h (NaN and h < 0
                                // make both forms false), one compare instead of two
                                in = fabsf(r[i]) <= h;
                            }
                            const bool counted = !((i == 0 && skip_first) || (i == 3 && skip_last));
                            cnt += (in && counted) ? 1u : 0u;
                        }
                        acc_u[j] += (cnt_rows & (1u << j)) ? cnt : 0u;
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            float v = has_ref ? __fsub_rn(r[i], rr[i]) : r[i];
                            v = __uint_as_float(__float_as_uint(v) & abs_mask);
                            r[i] = has_w ? __fmul_rn(v, ww[i]) : v;
                        }
                        if constexpr (STORE) {
                            *reinterpret_cast<float4*>(outp + j * out_row) = make_float4(r[0], r[1], r[2], r[3]);
                        } else if (cnt_rows & (1u << j)) {
                            if (rmax) {
                                uint32_t m = max(score_bits(r[1]), score_bits(r[2]));
                                if (!skip_first) m = max(m, score_bits(r[0]));
                                if (!skip_last) m = max(m, score_bits(r[3]));
                                acc_u[j] = max(acc_u[j], m);
                            } else {
                                float s = r[1] + r[2];
                                if (!skip_first) s += r[0];
                                if (!skip_last) s += r[3];
                                acc_f[j] += s;
                            }
                        }
                    }
                }
            });
        }
        // LAZY: the tile of plane t is read by the arms themselves, so the slot is free only now
        if constexpr (LAZY)
            if (plane_live && !refiller) bar_arrive(stage + 1, kThreads);
        if constexpr (STORE) outp += p.e.o2;
        if (has_ref) refp += p.e.r2;
        if (has_w) wp += w_plane;
        if (refiller) {                             // wait until every warp is done with slot `stage`
            bar_sync(stage + 1, kThreads);
            if (tc == 0 && t + p.stages < NPL) issue(t + p.stages, stage);
        }
        wturn = (wturn + 1) & (kWarps - 1);
        stage = nstage;
        cur_a = nxt_a;
        nxt_a += STAGE_BYTES;
        nbar += 8u;
        if (++nstage == p.stages) { nstage = 0; nxt_a = cell0; nbar = bars; nphase ^= 1u; }
    };

    for (int t = 0; t < NPL; t += 3) {
        step(t, prev, cur, next);
        if (t + 1 < NPL) step(t + 1, cur, next, prev);
        if (t + 2 < NPL) step(t + 2, next, prev, cur);
    }

    if constexpr (!STORE) {
        static_for<RX>([&](auto J) {
            constexpr int j = decltype(J)::value;
            const int row = row_first + j;
            const int b = BATCH_ROWS ? row : outer;
            const bool ok = row < p.R;               // warp-uniform (one warp per tile row)
            if constexpr (COVER) {
                const uint32_t c = __reduce_add_sync(0xffffffffu, acc_u[j]);
                if (tc == 0 && ok && c) atomicAdd(reinterpret_cast<unsigned int*>(p.e.scores) + b, c);
            } else if (rmax) {
                const uint32_t m = __reduce_max_sync(0xffffffffu, acc_u[j]);
                if (tc == 0 && ok && m) atomicMax(reinterpret_cast<unsigned int*>(p.e.scores) + b, m);
            } else {
                float s = acc_f[j];
#pragma unroll
                for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                if (tc == 0 && ok) atomicAdd(p.e.scores + b, s);
            }
        });
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
                cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

struct StarHost {
    bool batch_rows;
    uint64_t masks;     // arms actually present, 7 bits per slot
    StarCoef coef;
};

// Classifies the request; returns false when some slot has a tap off the star.
bool classify(const cppre_program_t* prog, const Prepared& P, StarHost* H) {
    const Geometry& g = P.g;
    if (g.O2 != g.N2 || g.O1 != g.N1 || g.O0 != g.N0) return false;   // "same" padding only
    bool k333 = true, k313 = true;
    for (int s = 0; s < P.n_slots; ++s) {
        const int* k = prog->ksize[s];
        const int* pd = prog->pad[s];
        k333 = k333 && k[0] == 3 && k[1] == 3 && k[2] == 3 && pd[0] == 1 && pd[1] == 1 && pd[2] == 1;
        k313 = k313 && k[0] == 3 && k[1] == 1 && k[2] == 3 && pd[0] == 1 && pd[1] == 0 && pd[2] == 1;
    }
    if (!(k333 || (k313 && g.N1 == 1))) return false;
    H->batch_rows = !k333;
    H->masks = 0;
    memset(&H->coef, 0, sizeof(H->coef));
    for (int s = 0; s < P.n_slots; ++s) {
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

bool layout_ok(const Prepared& P, bool batch_rows) {
    const Geometry& g = P.g;
    if (g.N0 % 4 != 0 || g.N0 < 4) return false;
    const int64_t q = 16 / kES;                    // global strides must be multiples of 16 bytes
    for (int f = 0; f < P.n_fields; ++f) {
        const FieldView& v = P.fields[f];
        if (!aligned16(v.p) || v.s0 != 1) return false;
        if ((g.N1 > 1 && (v.s1 % q != 0 || v.s1 <= 0)) || (g.N2 > 1 && (v.s2 % q != 0 || v.s2 <= 0)) ||
            (g.B > 1 && (v.sb % q != 0 || v.sb <= 0)))
            return false;
    }
    const EpiDev& e = P.epi;
    if (CPPRE_TMA_MODE == CPPRE_EPI_STORE) {
        if (!aligned16(e.out) || e.o0 != 1) return false;
        if ((g.N1 > 1 && e.o1 % 4) || (g.N2 > 1 && e.o2 % 4) || (g.B > 1 && e.ob % 4)) return false;
    }
    if (e.ref) {
        if (!aligned16(e.ref) || e.r0 != 1) return false;
        if ((g.N1 > 1 && e.r1 % 4) || (g.N2 > 1 && e.r2 % 4) || (g.B > 1 && e.rb % 4)) return false;
    }
    if (e.weight && !aligned16(e.weight)) return false;
    (void)batch_rows;
    return true;
}

template <int KIND, bool BR, int RX, uint64_t MASKS>
int launch_variant(const Prepared& P, const StarHost& H, cudaStream_t st) {
    constexpr int NF = Prog<KIND>::NF;
    constexpr int TX = kWarps * RX;
    constexpr int HALO = BR ? 0 : 1;
    constexpr int SR = TX + 2 * HALO;
    EncodeTiledFn encode = get_encode();
    CPPRE_REQUIRE(encode != nullptr, CPPRE_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");

    StarParams<NF> sp;
    memset(&sp, 0, sizeof(sp));
    const Geometry& g = P.g;
    const int R = BR ? g.B : g.N1;
    for (int f = 0; f < NF; ++f) {
        const FieldView& v = P.fields[f];
        // Tensor-map dims, innermost first.  3-D problems: (N0, N1, N2, B), box kTWP x (TX+2) x 1 x 1
        // (kTWP = 136 fp32 / 144 bf16 cells).  Batch-rows (1-D+time): (N0, N2, B, 1), box kTWP x 1 x TX x 1 -- same smem image,
        // and the byte strides stay ascending for contiguous data.  A stride on an extent-1
        // axis is never used for addressing but must still be a positive multiple of 16 bytes.
        auto fix = [&](int64_t s, uint64_t extent) -> cuuint64_t {
            return (cuuint64_t)((extent > 1 && s > 0) ? s : (kBF16 ? (g.N0 + 7) / 8 * 8 : g.N0)) * kES;
        };
        cuuint64_t dims[4], strides[3];
        cuuint32_t box[4];
        if (BR) {
            dims[0] = g.N0; dims[1] = g.N2; dims[2] = g.B; dims[3] = 1;
            strides[0] = fix(v.s2, dims[1]); strides[1] = fix(v.sb, dims[2]); strides[2] = fix(0, 1);
            box[0] = kTWP; box[1] = 1; box[2] = SR; box[3] = 1;
        } else {
            dims[0] = g.N0; dims[1] = g.N1; dims[2] = g.N2; dims[3] = g.B;
            strides[0] = fix(v.s1, dims[1]); strides[1] = fix(v.s2, dims[2]); strides[2] = fix(v.sb, dims[3]);
            box[0] = kTWP; box[1] = SR; box[2] = 1; box[3] = 1;
        }
        const cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
        CUresult rc = encode(&sp.map[f], kBF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                             const_cast<float*>(v.p), dims,
                             strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CPPRE_REQUIRE(rc == CUDA_SUCCESS, CPPRE_ERR_UNSUPPORTED,
                      "cuTensorMapEncodeTiled rejected field %d (CUresult %d)", f, (int)rc);
    }
    constexpr bool WTMA = CPPRE_TMA_OPER == 2 && !BR && CPPRE_TMA_MODE != CPPRE_EPI_STORE;
    if (WTMA) {     // weight [N2][N1][N0], contiguous fp32: box 128 x TX x 1, no halo
        const cuuint64_t dims[3] = {(cuuint64_t)g.N0, (cuuint64_t)g.N1, (cuuint64_t)g.N2};
        const cuuint64_t strides[2] = {(cuuint64_t)g.N0 * sizeof(float), (cuuint64_t)g.N0 * g.N1 * sizeof(float)};
        const cuuint32_t box[3] = {(cuuint32_t)kTW, (cuuint32_t)TX, 1u};
        const cuuint32_t estr[3] = {1u, 1u, 1u};
        CUresult rc = encode(&sp.wmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(P.epi.weight), dims,
                             strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CPPRE_REQUIRE(rc == CUDA_SUCCESS, CPPRE_ERR_UNSUPPORTED,
                      "cuTensorMapEncodeTiled rejected the weight tensor (CUresult %d)", (int)rc);
    }
    sp.coef = H.coef;
    memcpy(sp.scalars, P.scalars, sizeof(sp.scalars));
    sp.e = P.epi;
    sp.dec = P.dec;
    sp.N2 = g.N2;
    sp.N1 = g.N1;
    sp.N0 = g.N0;
    sp.R = R;
    sp.col_tiles = (g.N0 + kTW - 1) / kTW;
    sp.row_tiles = (R + TX - 1) / TX;
    sp.buf_bytes = ((uint32_t)(SR * kTWP) * kES + 127u) & ~127u;

    // ring depth: as deep as fits while leaving room for >= 2 CTAs per SM
    static int env_stages = -1, env_hints = -1;
    if (env_stages < 0) {
        const char* s = getenv("CPPRE_TMA_STAGES");
        env_stages = s ? atoi(s) : 0;
        const char* h = getenv("CPPRE_TMA_L2_HINTS");       // bit 0: fields evict_first, bit 1: weight evict_last
        env_hints = h ? atoi(h) : kDefaultL2Hints;
    }
    sp.l2_hints = env_hints;
    int stages = env_stages > 0 ? env_stages : 4;
    const size_t per_stage = (size_t)NF * sp.buf_bytes + (WTMA ? (size_t)TX * kTW * sizeof(float) : 0);
    while (stages > 2 && stages * per_stage + 256 > 112 * 1024) --stages;      // 2 x (112 KB + 1 KB) <= 228 KB per SM
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) stages = 2;
    sp.stages = stages;
    const size_t smem = stages * per_stage + 8 * kMaxStages + 128;
    CPPRE_REQUIRE(smem <= 227 * 1024, CPPRE_ERR_UNSUPPORTED, "tile ring of %zu bytes exceeds shared memory", smem);

    auto kern = star_kernel<KIND, BR, RX, MASKS, CPPRE_TMA_MODE, CPPRE_TMA_OPER>;
    static SmemOptIn memo;          // per instantiation, per device
    {
        const int rc = ensure_dynamic_smem(kern, memo, smem, cudaSharedmemCarveoutMaxShared);
        if (rc != CPPRE_OK) return rc;
    }
    const int64_t blocks = (int64_t)sp.col_tiles * sp.row_tiles * (BR ? 1 : g.B);
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    kern<<<(unsigned)blocks, kThreads, smem, st>>>(sp);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

// Arm presets.  A request runs on the first preset whose arm sets contain the arms it uses.
constexpr uint64_t pack(uint32_t a, uint32_t b = 0, uint32_t c = 0, uint32_t d = 0) {
    return (uint64_t)a | ((uint64_t)b << 7) | ((uint64_t)c << 14) | ((uint64_t)d << 21);
}
constexpr uint32_t A1D = ARMS_C | ARMS_T | ARMS_Y;   // everything a 1-D+time star can hold

typedef int (*LaunchFn)(const Prepared&, const StarHost&, cudaStream_t);
struct Variant {
    int kind;
    bool batch_rows;
    int rx;
    uint64_t masks;
    LaunchFn fn;
};
#define V(K, BR, RX, M) Variant{K, BR, RX, M, &launch_variant<K, BR, RX, M>}

// "shipped" = the kernels the reference constructors produce on contiguous data (note
// D_y == D_t); "fixed" = D_y along N0; "full" = any star.
constexpr uint64_t M_LIN = pack(ARMS_ALL);
constexpr uint64_t M_LIN1D = pack(A1D);
constexpr uint64_t M_BURG = pack(ARMS_T, ARMS_Y, ARMS_C | ARMS_Y);
constexpr uint64_t M_BURG_FULL = pack(A1D, A1D, A1D);
constexpr uint64_t M_NSC = pack(ARMS_X, ARMS_T);
constexpr uint64_t M_NSC_FIX = pack(ARMS_X, ARMS_Y);
constexpr uint64_t M_NSC_FULL = pack(ARMS_ALL, ARMS_ALL);
constexpr uint64_t M_NSM = pack(ARMS_T, ARMS_X, ARMS_T, ARMS_C | ARMS_X | ARMS_Y);
constexpr uint64_t M_NSM_FIX = pack(ARMS_T, ARMS_X, ARMS_Y, ARMS_C | ARMS_X | ARMS_Y);
constexpr uint64_t M_NSM_FULL = pack(ARMS_ALL, ARMS_ALL, ARMS_ALL, ARMS_ALL);
constexpr uint64_t M_MHD = pack(ARMS_T, ARMS_X, ARMS_T);
constexpr uint64_t M_MHD_FIX = pack(ARMS_T, ARMS_X, ARMS_Y);
constexpr uint64_t M_MHD_FULL = pack(ARMS_ALL, ARMS_ALL, ARMS_ALL);

const Variant kVariants[] = {
    V(CPPRE_PROG_LINEAR, false, 1, M_LIN), V(CPPRE_PROG_LINEAR, false, 2, M_LIN),
    V(CPPRE_PROG_LINEAR, true, 1, M_LIN1D), V(CPPRE_PROG_LINEAR, true, 2, M_LIN1D),
    V(CPPRE_PROG_BURGERS, true, 1, M_BURG), V(CPPRE_PROG_BURGERS, true, 2, M_BURG),
    V(CPPRE_PROG_BURGERS, true, 1, M_BURG_FULL),
    V(CPPRE_PROG_NS_CONTINUITY, false, 1, M_NSC), V(CPPRE_PROG_NS_CONTINUITY, false, 2, M_NSC),
    V(CPPRE_PROG_NS_CONTINUITY, false, 1, M_NSC_FIX), V(CPPRE_PROG_NS_CONTINUITY, false, 1, M_NSC_FULL),
    V(CPPRE_PROG_NS_MOMENTUM, false, 1, M_NSM), V(CPPRE_PROG_NS_MOMENTUM, false, 2, M_NSM),
    V(CPPRE_PROG_NS_MOMENTUM, false, 1, M_NSM_FIX), V(CPPRE_PROG_NS_MOMENTUM, false, 1, M_NSM_FULL),
    V(CPPRE_PROG_MHD_CONTINUITY, false, 1, M_MHD), V(CPPRE_PROG_MHD_CONTINUITY, false, 2, M_MHD),
    V(CPPRE_PROG_MHD_CONTINUITY, false, 1, M_MHD_FIX), V(CPPRE_PROG_MHD_CONTINUITY, false, 1, M_MHD_FULL),
    V(CPPRE_PROG_MHD_MOMENTUM, false, 1, M_MHD), V(CPPRE_PROG_MHD_MOMENTUM, false, 1, M_MHD_FIX),
    V(CPPRE_PROG_MHD_MOMENTUM, false, 1, M_MHD_FULL),
    V(CPPRE_PROG_MHD_ENERGY, false, 1, M_MHD), V(CPPRE_PROG_MHD_ENERGY, false, 1, M_MHD_FIX),
    V(CPPRE_PROG_MHD_ENERGY, false, 1, M_MHD_FULL),
    V(CPPRE_PROG_MHD_INDUCTION, false, 1, M_MHD), V(CPPRE_PROG_MHD_INDUCTION, false, 1, M_MHD_FIX),
    V(CPPRE_PROG_MHD_INDUCTION, false, 1, M_MHD_FULL),
    V(CPPRE_PROG_MHD_GAUSS, false, 1, M_MHD), V(CPPRE_PROG_MHD_GAUSS, false, 2, M_MHD),
    V(CPPRE_PROG_MHD_GAUSS, false, 1, M_MHD_FIX), V(CPPRE_PROG_MHD_GAUSS, false, 1, M_MHD_FULL),
};
#undef V

int preferred_rx() {
    static int rx = -1;
    if (rx < 0) {
        const char* s = getenv("CPPRE_TMA_RX");
        rx = s ? atoi(s) : 0;
    }
    return rx;
}

const Variant* pick(int kind, const StarHost& H, const Geometry& g) {
    // first preset (table order = tightest first) that covers the arms in use; among equal
    // presets prefer the requested / heuristic rows-per-thread
    const Variant* best = nullptr;
    int want = preferred_rx();
    if (want <= 0) {
        const int R = H.batch_rows ? g.B : g.N1;
        want = R >= 16 ? 2 : 1;
    }
    for (const Variant& v : kVariants) {
        if (v.kind != kind || v.batch_rows != H.batch_rows) continue;
        if ((H.masks & ~v.masks) != 0) continue;
        if (!best) best = &v;
        else if (v.masks == best->masks && v.rx == want) best = &v;
        if (best->masks != v.masks) break;
    }
    return best;
}

}  // namespace

#define CPPRE_CAT_(a, b) a##b
#define CPPRE_CAT(a, b) CPPRE_CAT_(a, b)

int CPPRE_CAT(tma_eligible_, CPPRE_TMA_SUFFIX)(const cppre_program_t* prog, const Prepared& P) {
    StarHost H;
    if (!classify(prog, P, &H)) return 0;
    if (!layout_ok(P, H.batch_rows)) return 0;
    if (P.g.N2 < 1) return 0;
    return pick(P.kind, H, P.g) != nullptr;
}

int CPPRE_CAT(launch_tma_, CPPRE_TMA_SUFFIX)(const cppre_program_t* prog, const Prepared& P, cudaStream_t st) {
    StarHost H;
    CPPRE_REQUIRE(classify(prog, P, &H) && layout_ok(P, H.batch_rows), CPPRE_ERR_UNSUPPORTED,
                  "request does not qualify for the TMA star kernel");
    const Variant* v = pick(P.kind, H, P.g);
    CPPRE_REQUIRE(v != nullptr, CPPRE_ERR_UNSUPPORTED, "no TMA variant for program %d", P.kind);
    return v->fn(P, H, st);
}

}  // namespace cppre
