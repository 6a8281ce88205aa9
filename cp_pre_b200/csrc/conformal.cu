// K2 (per-cell statistics over the sample axis), K3 (order statistics along the sample axis)
// and K5 (coverage counts): the conformal-calibration half of the hot path.
//
// Layout: score / residual tensors are [n, cells] with cells contiguous, so "thread = cell,
// loop over samples" gives fully coalesced 128-byte warp accesses on every row.
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <mutex>
#include <type_traits>

#include "common.cuh"

namespace cppre {
namespace {

constexpr int kThreads = 256;

// V consecutive floats per thread: V = 4 (one 128-bit load) when every row is 16-byte aligned,
// V = 1 otherwise.
template <int V>
__device__ __forceinline__ void ldv(const float* p, float (&v)[V]) {
    if constexpr (V == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
        v[0] = __ldg(p);
    }
}

bool vec4_ok(int64_t cells, std::initializer_list<const void*> ptrs, std::initializer_list<int64_t> strides) {
    if (cells % 4) return false;
    for (const void* q : ptrs)
        if (reinterpret_cast<uintptr_t>(q) & 15u) return false;
    for (int64_t st : strides)
        if (st % 4) return false;
    return true;
}

// ------------------------------------------------------------------------------------------
// K2: shifted-sum moments in fp64, V cells per thread, 4 samples in flight per load stream
// ------------------------------------------------------------------------------------------
// FINAL: std (+ eps) and 1/std are written straight away (single-GPU modulation: one launch, no
// fp64 intermediates in HBM); otherwise (mean, M2) for the cross-rank merge.
// SPLIT: sample-lanes per cell.  1 = one thread streams the whole column.  8 = a block is 32 cell-lanes x
// 8 warps, warp w takes samples w, w+8, ... and the eight partial sums meet in shared memory: for grids
// with few cells (1-D Burgers: 102 400 cells = 100 blocks of the plain form on 148 SMs, 0.54 ms for
// 0.84 GB) this is what fills the machine.
template <int V, bool FINAL, int SPLIT>
__global__ void __launch_bounds__(kThreads)
cell_moments_kernel(const float* __restrict__ a, int64_t as, const float* __restrict__ b, int64_t bs,
                    int64_t n, int64_t cells, double* __restrict__ mean, double* __restrict__ m2,
                    float add_eps, float* __restrict__ sd, float* __restrict__ inv) {
    static_assert(SPLIT == 1 || SPLIT * 32 == kThreads, "SPLIT lanes are the warps of a block");
    __shared__ double part[SPLIT == 1 ? 1 : 2 * SPLIT * 32 * V];
    const int lane = threadIdx.x & 31, w = SPLIT == 1 ? 0 : threadIdx.x >> 5;
    const int64_t c = SPLIT == 1 ? ((int64_t)blockIdx.x * kThreads + threadIdx.x) * V
                                 : ((int64_t)blockIdx.x * 32 + lane) * V;
    const bool live = c < cells;
    if (SPLIT == 1 && !live) return;
    const float* pa = a + c;
    const float* pb = b ? b + c : nullptr;
    // shifted sums in fp64: d = x - x_0, so sum(d^2) - sum(d)^2/n does not cancel unless the
    // column mean is > 1e6 standard deviations away from its first sample
    auto sample = [&](int64_t i, float (&x)[V]) {
        ldv<V>(pa + i * as, x);
        if (pb) {
            float y[V];
            ldv<V>(pb + i * bs, y);
#pragma unroll
            for (int e = 0; e < V; ++e) x[e] = __fsub_rn(x[e], y[e]);
        }
    };
    float x0[V];
    double s1[V], s2[V];
#pragma unroll
    for (int e = 0; e < V; ++e) { x0[e] = 0.0f; s1[e] = 0.0; s2[e] = 0.0; }
    if (live && n > 0) {
        sample(0, x0);
        int64_t i = w;
        for (; i + 3 * SPLIT < n; i += 4 * SPLIT) {
            float x[4][V];
#pragma unroll
            for (int u = 0; u < 4; ++u) sample(i + u * SPLIT, x[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int e = 0; e < V; ++e) {
                    const double d = (double)x[u][e] - (double)x0[e];
                    s1[e] += d;
                    s2[e] = fma(d, d, s2[e]);
                }
        }
        for (; i < n; i += SPLIT) {
            float x[V];
            sample(i, x);
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const double d = (double)x[e] - (double)x0[e];
                s1[e] += d;
                s2[e] = fma(d, d, s2[e]);
            }
        }
    }
    if constexpr (SPLIT > 1) {
#pragma unroll
        for (int e = 0; e < V; ++e) {
            part[((0 * SPLIT + w) * 32 + lane) * V + e] = s1[e];
            part[((1 * SPLIT + w) * 32 + lane) * V + e] = s2[e];
        }
        __syncthreads();
        if (w != 0 || !live) return;
#pragma unroll
        for (int e = 0; e < V; ++e) {
            s1[e] = 0.0;
            s2[e] = 0.0;
            for (int q = 0; q < SPLIT; ++q) {
                s1[e] += part[((0 * SPLIT + q) * 32 + lane) * V + e];
                s2[e] += part[((1 * SPLIT + q) * 32 + lane) * V + e];
            }
        }
    }
#pragma unroll
    for (int e = 0; e < V; ++e) {
        if (n == 0) {
            if constexpr (!FINAL) { mean[c + e] = 0.0; m2[c + e] = 0.0; }
            continue;
        }
        double v = s2[e] - s1[e] * s1[e] / (double)n;
        v = v > 0.0 ? v : (v == v ? 0.0 : v);           // clamp round-off negatives, keep NaN
        if constexpr (FINAL) {                          // same arithmetic as cell_std_kernel
            const float s = __fadd_rn((float)sqrt(v / (double)n), add_eps);
            sd[c + e] = s;
            if (inv) inv[c + e] = __fdiv_rn(1.0f, s);
        } else {
            mean[c + e] = (double)x0[e] + s1[e] / (double)n;
            m2[c + e] = v;
        }
    }
}

template <bool FINAL>
int launch_moments(const float* d_a, int64_t as, const float* d_b, int64_t bs, int64_t n, int64_t cells, double* d_mean,
                   double* d_m2, float add_eps, float* d_std, float* d_inv, cudaStream_t st) {
    const bool v4 = vec4_ok(cells, {d_a, d_b}, {as, d_b ? bs : 0});
    const int64_t items = v4 ? cells / 4 : cells;
    // few cells and a long sample axis: spread each column over the 8 warps of a block
    const bool split = items < (int64_t)148 * 2048 / 2 && n >= 64;
    const int64_t blocks = split ? (items + 31) / 32 : (items + kThreads - 1) / kThreads;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
#define CPPRE_MOMENTS(V, SP) \
    cell_moments_kernel<V, FINAL, SP><<<(unsigned)blocks, kThreads, 0, st>>>(d_a, as, d_b, bs, n, cells, d_mean, d_m2, \
                                                                            add_eps, d_std, d_inv)
    if (v4 && split) CPPRE_MOMENTS(4, 8);
    else if (v4) CPPRE_MOMENTS(4, 1);
    else if (split) CPPRE_MOMENTS(1, 8);
    else CPPRE_MOMENTS(1, 1);
#undef CPPRE_MOMENTS
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

__global__ void __launch_bounds__(kThreads)
cell_std_kernel(const double* __restrict__ m2, double count, float add_eps, int64_t cells,
                float* __restrict__ sd, float* __restrict__ inv) {
    const int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (c >= cells) return;
    const float v = __fadd_rn((float)sqrt(m2[c] / count), add_eps);
    sd[c] = v;
    if (inv) inv[c] = __fdiv_rn(1.0f, v);
}

// scores[i] = max_c |a-b| / mod.  A block owns one cell chunk and S consecutive samples: the modulation of
// a cell is loaded (and its reciprocal approximated) once and used for S samples, which also puts S
// independent load streams in flight per thread.
// Exact IEEE quotients at a fraction of their cost: only the MAXIMUM quotient of a sample is wanted, so a
// cell is screened with q~ = |x| * rcp(mod) (one FMUL; relative error < 2^-21 incl. the MUFU.RCP error) and
// the correctly rounded division runs only where q~ reaches the running exact maximum less a 2^-19 guard --
// any cell that is skipped has an exact quotient strictly below the maximum already found.  Comparisons are
// written so that NaN (in x or mod) takes the exact path and wins the unsigned max, like np.max.
template <int V, int S>
__global__ void __launch_bounds__(kThreads)
joint_score_kernel(const float* __restrict__ a, int64_t as, const float* __restrict__ b, int64_t bs,
                   const float* __restrict__ mod, int64_t n, int64_t cells, int32_t chunks, float* scores) {
    const int64_t i0 = (int64_t)(blockIdx.x / chunks) * S;
    const int chunk = blockIdx.x % chunks;
    uint32_t m[S];
    float thr[S];
#pragma unroll
    for (int s = 0; s < S; ++s) { m[s] = 0u; thr[s] = 0.0f; }
    const int64_t step = (int64_t)chunks * kThreads * V;
    for (int64_t c = ((int64_t)chunk * kThreads + threadIdx.x) * V; c < cells; c += step) {
        float w[V], r[V], x[S][V];
        ldv<V>(mod + c, w);
#pragma unroll
        for (int s = 0; s < S; ++s) {
            // rows past the batch re-read the last sample (their maxima are never written)
            const int64_t i = i0 + s < n ? i0 + s : n - 1;
            ldv<V>(a + i * as + c, x[s]);
            if (b) {
                float y[V];
                ldv<V>(b + i * bs + c, y);
#pragma unroll
                for (int e = 0; e < V; ++e) x[s][e] = __fsub_rn(x[s][e], y[e]);
            }
        }
#pragma unroll
        for (int e = 0; e < V; ++e) asm("rcp.approx.f32 %0, %1;" : "=f"(r[e]) : "f"(w[e]));
#pragma unroll
        for (int s = 0; s < S; ++s)
#pragma unroll
            for (int e = 0; e < V; ++e) {
                const float ax = fabsf(x[s][e]);
                if (!(ax * r[e] < thr[s])) {           // near or above the running maximum, or NaN: exact quotient
                    // |x|/mod is >= 0 or NaN for mod > 0; fabsf clears a sign a negative mod would give
                    const uint32_t q = score_bits(fabsf(__fdiv_rn(ax, w[e])));
                    if (q > m[s]) {
                        m[s] = q;
                        // 1 - 2^-19; the relative-error argument needs normal numbers: a maximum below
                        // 1e-30 keeps the threshold at 0, i.e. every cell takes the exact path
                        thr[s] = __uint_as_float(q) >= 1e-30f ? __uint_as_float(q) * 0.999998093f : 0.0f;
                    }
                }
            }
    }
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const uint32_t v = __reduce_max_sync(0xffffffffu, m[s]);
        if ((threadIdx.x & 31) == 0 && v && i0 + s < n) atomicMax(reinterpret_cast<unsigned int*>(scores) + i0 + s, v);
    }
}

// ------------------------------------------------------------------------------------------
// K3: order statistics along the sample axis
// ------------------------------------------------------------------------------------------
constexpr int kMaxRanks = 32;

struct RankList {
    int32_t n;
    int32_t r[kMaxRanks];
};

// K3r: per-cell multi-rank select, register-resident bitonic network (the fast path, n <= 2048).
// A CTA stages a [N samples x TC cells] tile of order-preserving uint32 keys in shared memory
// with fully coalesced row loads, then every cell's column is spread over L lanes x R registers
// (N = L*R = next power of two >= n, padding keys sort last).  The network is the all-ascending
// bitonic form (each merge level starts with a "flip" comparator i <-> i^(k-1), then plain
// half-cleaners), so no comparator direction depends on the lane: strides below R are
// register-register min/max pairs with compile-time indices, strides >= R are one SHFL.BFLY plus
// one min-or-max.  After the sort the element of rank k sits in lane k / R, register k % R, so
// any number of ranks costs one sort.  Shared-memory rows are padded to TC+1 words: both the
// row-wise fill and the column-wise read-out are bank-conflict free.
template <int R>
__device__ __forceinline__ void ce_regs(uint32_t (&k)[R], int a, int b) {
    const uint32_t lo = min(k[a], k[b]), hi = max(k[a], k[b]);
    k[a] = lo;
    k[b] = hi;
}

// Levels and strides are template parameters (not unrolled loops) so that every register index
// is a compile-time constant even in the 64-register variant.
template <int L, int R, int STRIDE>
__device__ __forceinline__ void half_cleaners(uint32_t (&k)[R], const int j) {
    if constexpr (STRIDE > 0) {
        if constexpr (STRIDE < R) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if ((r & STRIDE) == 0) ce_regs<R>(k, r, r | STRIDE);
        } else {
            constexpr int lm = STRIDE / R;
            const bool lower = (j & lm) == 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint32_t o = __shfl_xor_sync(0xffffffffu, k[r], lm);
                k[r] = lower ? min(k[r], o) : max(k[r], o);
            }
        }
        half_cleaners<L, R, STRIDE / 2>(k, j);
    }
}

template <int L, int R, int SIZE>
__device__ __forceinline__ void bitonic_levels(uint32_t (&k)[R], const int j) {
    if constexpr (SIZE <= L * R) {
        // flip step: partner = s ^ (SIZE-1), s = j*R + r
        if constexpr (SIZE <= R) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if ((r & (SIZE >> 1)) == 0) ce_regs<R>(k, r, r ^ (SIZE - 1));
        } else {
            constexpr int lm = SIZE / R - 1;                 // lane xor mask (all lane bits of the block)
            const bool lower = (j & (SIZE / R / 2)) == 0;    // this lane holds the lower half
            uint32_t o[R];
#pragma unroll
            for (int r = 0; r < R; ++r) o[r] = __shfl_xor_sync(0xffffffffu, k[R - 1 - r], lm);
#pragma unroll
            for (int r = 0; r < R; ++r) k[r] = lower ? min(k[r], o[r]) : max(k[r], o[r]);
        }
        half_cleaners<L, R, SIZE / 4>(k, j);
        bitonic_levels<L, R, SIZE * 2>(k, j);
    }
}

template <int L, int R>
__device__ __forceinline__ void bitonic_sort_regs(uint32_t (&k)[R], const int j) {
    bitonic_levels<L, R, 2>(k, j);
}

struct PerCell {            // optional per-cell rank for the regsort kernel (cnt == nullptr: unused)
    const int32_t* cnt;     // [cells] #above (low 24 bits) + NaN flags (bits 24..)
    int32_t kt;
    uint32_t* overflow;     // counter, shared with K3w
    int32_t* overflow_cells;
    int32_t overflow_cap;
    int64_t cell_offset;    // global index of this call's first cell (for the overflow list)
};

// COMPACT (the owner's side of the multi-GPU marginal quantile): the rows are candidate blocks
// that are mostly -inf padding.  The fill drops the padding and packs each column's real
// candidates to the front of the tile (shared-memory atomics hand out the positions), so a
// network of N = L*R keys serves several times as many rows; a column with more than N real
// candidates is reported on the overflow list and redone by the caller.
template <int L, int R, int TC, bool COMPACT>
__global__ void __launch_bounds__(TC * L)
select_regsort_kernel(const float* __restrict__ x, int64_t xs, int32_t n, int64_t cells, RankList ranks,
                      float* __restrict__ out, const PerCell pc) {
    constexpr int N = L * R;
    constexpr int THREADS = TC * L;
    constexpr int WARPS = THREADS / 32;
    constexpr int CW = 32 / L;              // cells handled by one warp
    constexpr int LD = TC + 1;              // padded row length (words)
    constexpr int RPW = 32 / TC;            // tile rows covered by one warp-wide load (1 or 2)
    static_assert(L >= 1 && L <= 32 && (TC == 32 || TC == 16) && CW * WARPS == TC, "bad tiling");
    extern __shared__ uint32_t tile[];      // [N][LD] keys, then TC NaN flags
    uint32_t* nanflag = tile + N * LD;

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int64_t c0 = (int64_t)blockIdx.x * TC;
    if (tid < TC) nanflag[tid] = 0u;

    // fill: warp w takes rows (w*RPW + lane/TC) + WARPS*RPW*i, column lane % TC
    {
        const int col = lane % TC;
        const bool live = c0 + col < cells;
        const float* px = x + c0 + col;
        bool nan = false;
        if constexpr (COMPACT) {
            for (int e = tid; e < N * LD; e += THREADS) tile[e] = 0u;      // padding sorts first
            __syncthreads();
            for (int row = w * RPW + lane / TC; row < n; row += WARPS * RPW) {
                if (!live) continue;
                const uint32_t bits = __float_as_uint(__ldg(px + (int64_t)row * xs));
                if (bits == 0xff800000u) continue;                         // -inf: K3w padding
                const uint32_t pos = atomicAdd(&nanflag[col], 1u);         // nanflag doubles as the column count
                if (pos < (uint32_t)N) tile[pos * LD + col] = float_to_key(bits);
            }
        } else {
#pragma unroll 8
            for (int row = w * RPW + lane / TC; row < N; row += WARPS * RPW) {
                uint32_t key = 0xffffffffu;     // padding sorts last
                if (live && row < n) {
                    const uint32_t bits = __float_as_uint(__ldg(px + (int64_t)row * xs));
                    nan |= (bits & 0x7fffffffu) > 0x7f800000u;
                    key = float_to_key(bits);
                }
                tile[row * LD + col] = key;
            }
        }
        __syncthreads();                    // flags zeroed before anyone raises one
        if (!COMPACT && nan) nanflag[col] = 1u;
    }
    __syncthreads();

    // read-out: lane = i*L + j; cell-in-tile ct = w + WARPS*i; sample rows r*L + j
    const int j = lane % L, i = lane / L;
    const int ct = w + WARPS * i;
    uint32_t k[R];
#pragma unroll
    for (int r = 0; r < R; ++r) k[r] = tile[(r * L + j) * LD + ct];
    bitonic_sort_regs<L, R>(k, j);

    // write the sorted columns back (same conflict-free mapping as the read-out), then one
    // thread per cell picks every requested rank: rank k sits at lane k / R, register k % R
    __syncthreads();                        // every warp has its column in registers
#pragma unroll
    for (int r = 0; r < R; ++r) tile[(r * L + j) * LD + ct] = k[r];
    __syncthreads();
    if (tid < TC && c0 + tid < cells && pc.cnt != nullptr) {
        // per-cell form (multi-GPU marginal quantile): the (kt - #above)-th largest candidate; the
        // summed NaN flags of all ranks ride in the bits above 24
        const int32_t packed = pc.cnt[c0 + tid];
        const int want = pc.kt - (packed & 0xffffff);
        const int have = COMPACT ? (int)nanflag[tid] : n;      // real candidates in this column
        bool nan = (packed >> 24) != 0;
        if (!nan && (want < 1 || want > n || have > N)) {
            const uint32_t at = atomicAdd(pc.overflow, 1u);
            if (at < (uint32_t)pc.overflow_cap) pc.overflow_cells[at] = (int32_t)(pc.cell_offset + c0 + tid);
            nan = true;
        }
        // ascending position of the want-th largest; COMPACT: padding sorts first, and a rank
        // below the real candidates lands in the dropped -inf padding
        const int kq = nan ? 0 : (COMPACT ? N : n) - want;
        uint32_t v = tile[((kq % R) * L + kq / R) * LD + tid];
        if (COMPACT && want > have) v = 0x007fffffu;           // key(-inf)
        out[c0 + tid] = nan ? __uint_as_float(0x7fc00000u) : __uint_as_float(key_to_float_bits(v));
    } else if (tid < TC && c0 + tid < cells) {
        const bool has_nan = nanflag[tid] != 0u;
        for (int q = 0; q < ranks.n; ++q) {
            const int kq = ranks.r[q];
            const uint32_t v = tile[((kq % R) * L + kq / R) * LD + tid];
            out[(int64_t)q * cells + c0 + tid] = has_nan ? __uint_as_float(0x7fc00000u)
                                                         : __uint_as_float(key_to_float_bits(v));
        }
    }
}

template <int L, int R, int TC, bool COMPACT = false>
int launch_regsort(const float* x, int64_t xs, int64_t n, int64_t cells, const RankList& rl, float* out,
                   cudaStream_t st, const PerCell pc) {
    constexpr int N = L * R;
    const size_t smem = ((size_t)N * (TC + 1) + TC) * sizeof(uint32_t);
    auto kern = select_regsort_kernel<L, R, TC, COMPACT>;
    static SmemOptIn memo;
    {
        const int rc = ensure_dynamic_smem(kern, memo, smem);
        if (rc != CPPRE_OK) return rc;
    }
    const int64_t blocks = (cells + TC - 1) / TC;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    kern<<<(unsigned)blocks, TC * L, smem, st>>>(x, xs, (int32_t)n, cells, rl, out, pc);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

// n <= 2048: pick the smallest network that holds n keys.  Returns -1 when n is too large.
int dispatch_regsort(const float* x, int64_t xs, int64_t n, int64_t cells, const RankList& rl, float* out,
                     cudaStream_t st, const PerCell pc = PerCell{nullptr, 0, nullptr, nullptr, 0, 0}) {
    if (n <= 32) return launch_regsort<1, 32, 32>(x, xs, n, cells, rl, out, st, pc);
    if (n <= 64) return launch_regsort<2, 32, 32>(x, xs, n, cells, rl, out, st, pc);
    if (n <= 128) return launch_regsort<4, 32, 32>(x, xs, n, cells, rl, out, st, pc);
    if (n <= 256) return launch_regsort<8, 32, 32>(x, xs, n, cells, rl, out, st, pc);
    if (n <= 512) return launch_regsort<16, 32, 32>(x, xs, n, cells, rl, out, st, pc);
    if (n <= 1024) return launch_regsort<32, 32, 32>(x, xs, n, cells, rl, out, st, pc);
    if (n <= 2048) return launch_regsort<32, 64, 16>(x, xs, n, cells, rl, out, st, pc);
    return -1;
}

// Owner-side select over padded candidate blocks: network of `net` keys (a power of two chosen
// by the caller from the expected number of real candidates per cell), any number of rows.
int dispatch_compact(int net, const float* x, int64_t xs, int64_t n, int64_t cells, const RankList& rl, float* out,
                     cudaStream_t st, const PerCell pc) {
    switch (net) {
        case 32: return launch_regsort<1, 32, 32, true>(x, xs, n, cells, rl, out, st, pc);
        case 64: return launch_regsort<2, 32, 32, true>(x, xs, n, cells, rl, out, st, pc);
        case 128: return launch_regsort<4, 32, 32, true>(x, xs, n, cells, rl, out, st, pc);
        case 256: return launch_regsort<8, 32, 32, true>(x, xs, n, cells, rl, out, st, pc);
        case 512: return launch_regsort<16, 32, 32, true>(x, xs, n, cells, rl, out, st, pc);
        case 1024: return launch_regsort<32, 32, 32, true>(x, xs, n, cells, rl, out, st, pc);
        default: return -1;
    }
}

// Batcher's odd-even merge sort on R registers (ascending), every index a compile-time
// constant: 19 / 63 / 191 comparators for R = 8 / 16 / 32 against the bitonic network's
// 24 / 80 / 240.  Used where no lane crossing is needed (K3t).
//
// Compare-exchange flavours.  VIMNMX issues on the ALU pipe only (one warp instruction per two
// cycles per scheduler), which is what bounds a register sorting network.  The "B" flavour keeps
// the min on the ALU pipe and rebuilds the max on the FMA pipe as a + b - min(a, b) -- exact in
// modular uint32 arithmetic -- with two IMADs, so a mix of the two flavours keeps both pipes busy.
// `one` is the constant 1 passed as a kernel argument: a literal would let ptxas fold the two
// multiply-adds into one ALU-pipe IADD3, which is exactly what this flavour is there to avoid.
__device__ __forceinline__ void ce_mix(uint32_t& a, uint32_t& b, const bool fma_flavour, const uint32_t one) {
    const uint32_t lo = min(a, b);
    if (fma_flavour) {
        const uint32_t t = a * one + b;              // IMAD
        const uint32_t hi = lo * (0u - one) + t;     // IMAD: t - lo
        a = lo;
        b = hi;
    } else {
        const uint32_t hi = max(a, b);
        a = lo;
        b = hi;
    }
}

template <int R, int MIX>
struct OddEven {
    // comparator counter -> flavour: MIX of every 3 comparators take the FMA-pipe flavour
    template <int LO, int N, int STRIDE>
    __device__ __forceinline__ static void merge(uint32_t (&k)[R], int& n_ce, const uint32_t one) {
        constexpr int STEP = STRIDE * 2;
        if constexpr (STEP < N) {
            merge<LO, N, STEP>(k, n_ce, one);
            merge<LO + STRIDE, N, STEP>(k, n_ce, one);
#pragma unroll
            for (int i = LO + STRIDE; i + STRIDE < LO + N; i += STEP) {
                ce_mix(k[i], k[i + STRIDE], (n_ce % 3) < MIX, one);
                ++n_ce;
            }
        } else {
            ce_mix(k[LO], k[LO + STRIDE], (n_ce % 3) < MIX, one);
            ++n_ce;
        }
    }
    template <int LO, int N>
    __device__ __forceinline__ static void sort(uint32_t (&k)[R], int& n_ce, const uint32_t one) {
        if constexpr (N > 1) {
            sort<LO, N / 2>(k, n_ce, one);
            sort<LO + N / 2, N / 2>(k, n_ce, one);
            merge<LO, N, 1>(k, n_ce, one);
        }
    }
};

// bitonic merge (input bitonic, output ascending) with the same flavour mix
template <int R, int MIX, int STRIDE>
__device__ __forceinline__ void bitonic_merge_mix(uint32_t (&k)[R], int& n_ce, const uint32_t one) {
    if constexpr (STRIDE > 0) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            if ((r & STRIDE) == 0) {
                ce_mix(k[r], k[r | STRIDE], (n_ce % 3) < MIX, one);
                ++n_ce;
            }
        bitonic_merge_mix<R, MIX, STRIDE / 2>(k, n_ce, one);
    }
}

// K3t: per-cell order statistics near one END of the sorted column (the calibration case:
// alpha = 0.1 on n = 256 asks for the 24th largest).  One lane owns one cell -- a warp row load
// is 128 contiguous bytes, a CTA row 512 -- and streams its column in chunks of K samples:
// sort the chunk in registers (all-ascending bitonic network, compile-time indices, no
// shuffles), prune it against the running top-K (T[i] = max(T[i], C[K-1-i]) keeps the K largest
// of both and leaves a bitonic sequence) and re-sort T with one bitonic merge.  Per sample that
// is ~21 min/max operations for K = 32 against ~42 for the full sort, with no shared memory and
// no cross-lane traffic.  FROM_BOTTOM complements the keys, so "K largest" selects the K
// smallest.  Rank k of n lives at T[K - n + k] (top) / T[K - 1 - k] (bottom).
template <int K, bool FROM_BOTTOM, int MIX>
__global__ void __launch_bounds__(128)
select_topk_kernel(const float* __restrict__ x, int64_t xs, int32_t n, int64_t cells, RankList ranks,
                   float* __restrict__ out, const uint32_t one) {
    const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (c >= cells) return;
    const float* px = x + c;
    uint32_t T[K], C[K];
    // NaN detection costs one min per sample: a NaN with the sign bit clear has a key above
    // key(+inf) = 0xff800000 and surfaces at the top of T; one with the sign bit set has a key
    // below key(-inf) = 0x007fffff and is caught by the running minimum (and the other way
    // round for complemented keys).
    uint32_t kmin = 0xffffffffu;
    auto to_key = [](float v) -> uint32_t {
        const uint32_t bits = __float_as_uint(v);
        const uint32_t flip = (uint32_t)((int32_t)bits >> 31) | 0x80000000u;   // SHF + one LOP3
        return FROM_BOTTOM ? ~(bits ^ flip) : (bits ^ flip);
    };
    // `rows` real samples starting at px; full chunks take the branch-free path
    auto load_chunk = [&](uint32_t (&dst)[K], int rows) {
        if (rows >= K) {
            float v[K];
#pragma unroll
            for (int r = 0; r < K; ++r) { v[r] = __ldg(px); px += xs; }
#pragma unroll
            for (int r = 0; r < K; ++r) { dst[r] = to_key(v[r]); kmin = min(kmin, dst[r]); }
        } else {
#pragma unroll
            for (int r = 0; r < K; ++r) {
                dst[r] = 0u;                         // padding: below every real key
                if (r < rows) { dst[r] = to_key(__ldg(px)); px += xs; kmin = min(kmin, dst[r]); }
            }
        }
    };
    int n_ce = 0;                                    // compile-time after unrolling: picks the CE flavour
    load_chunk(T, n);
    OddEven<K, MIX>::template sort<0, K>(T, n_ce, one);
    for (int base = K; base < n; base += K) {
        load_chunk(C, n - base);
        n_ce = 0;
        OddEven<K, MIX>::template sort<0, K>(C, n_ce, one);
#pragma unroll
        for (int i = 0; i < K; ++i) T[i] = max(T[i], C[K - 1 - i]);
        bitonic_merge_mix<K, MIX, K / 2>(T, n_ce, one);
    }
    const bool nan = T[K - 1] > 0xff800000u || kmin < 0x007fffffu;
    for (int q = 0; q < ranks.n; ++q) {
        const int idx = FROM_BOTTOM ? K - 1 - ranks.r[q] : K - n + ranks.r[q];
        uint32_t v = 0u;
#pragma unroll
        for (int i = 0; i < K; ++i) v = (idx == i) ? T[i] : v;
        if (FROM_BOTTOM) v = ~v;
        out[(int64_t)q * cells + c] = nan ? __uint_as_float(0x7fc00000u) : __uint_as_float(key_to_float_bits(v));
    }
}

int topk_mix() {   // CPPRE_TOPK_MIX=0..3: comparators out of every 3 that use the FMA-pipe flavour
    static int mix = -1;
    if (mix < 0) {
        const char* e = getenv("CPPRE_TOPK_MIX");
        mix = e ? atoi(e) : 2;
        if (mix < 0 || mix > 3) mix = 2;
    }
    return mix;
}

template <int K, bool FROM_BOTTOM>
int launch_topk(const float* x, int64_t xs, int64_t n, int64_t cells, const RankList& rl, float* out,
                cudaStream_t st) {
    const int64_t blocks = (cells + 127) / 128;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    switch (topk_mix()) {
        case 0: select_topk_kernel<K, FROM_BOTTOM, 0><<<(unsigned)blocks, 128, 0, st>>>(x, xs, (int32_t)n, cells, rl, out, 1u); break;
        case 1: select_topk_kernel<K, FROM_BOTTOM, 1><<<(unsigned)blocks, 128, 0, st>>>(x, xs, (int32_t)n, cells, rl, out, 1u); break;
        case 3: select_topk_kernel<K, FROM_BOTTOM, 3><<<(unsigned)blocks, 128, 0, st>>>(x, xs, (int32_t)n, cells, rl, out, 1u); break;
        default: select_topk_kernel<K, FROM_BOTTOM, 2><<<(unsigned)blocks, 128, 0, st>>>(x, xs, (int32_t)n, cells, rl, out, 1u); break;
    }
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

// Every requested rank within K of one end of the column, K <= 64 and the column longer than
// the window (otherwise the full sort is as cheap): returns -1 when that does not hold.
int dispatch_topk(const float* x, int64_t xs, int64_t n, int64_t cells, const RankList& rl, float* out,
                  cudaStream_t st) {
    int lo = rl.r[0], hi = rl.r[0];
    for (int q = 1; q < rl.n; ++q) {
        lo = rl.r[q] < lo ? rl.r[q] : lo;
        hi = rl.r[q] > hi ? rl.r[q] : hi;
    }
    const int64_t need_top = n - lo, need_bot = (int64_t)hi + 1;   // window sizes that cover all ranks
    const bool top = need_top <= need_bot;
    const int64_t need = top ? need_top : need_bot;
    if (need > 64 || n <= 64 || (need > 32 && n <= 256)) return -1;
    if (need <= 8)
        return top ? launch_topk<8, false>(x, xs, n, cells, rl, out, st) : launch_topk<8, true>(x, xs, n, cells, rl, out, st);
    if (need <= 16)
        return top ? launch_topk<16, false>(x, xs, n, cells, rl, out, st) : launch_topk<16, true>(x, xs, n, cells, rl, out, st);
    if (need <= 32)
        return top ? launch_topk<32, false>(x, xs, n, cells, rl, out, st) : launch_topk<32, true>(x, xs, n, cells, rl, out, st);
    // 33..64 from one end of a long column (e.g. alpha = 0.1 on n = 512): still 3x cheaper than the full sort
    return top ? launch_topk<64, false>(x, xs, n, cells, rl, out, st) : launch_topk<64, true>(x, xs, n, cells, rl, out, st);
}

// K3w: window compaction, the second pass of the multi-GPU marginal quantile.  Every rank knows
// a per-cell interval [lo, hi] (ordered int32 keys, all-reduced) that contains the global
// quantile.  One lane owns one cell and streams its column once: counts the samples above hi and
// collects the samples inside the interval into W slots (staged in shared memory, slot-major, so
// the final write-out is coalesced).  Only those W candidates per cell and rank are exchanged
// instead of the whole score column.  Output blocks are laid out [cells/slab][W][slab] -- the
// send buffer of the all-to-all that moves slab r to rank r -- in sample order, -inf padded.
// cnt[c] = #{x > hi} + (hi is a NaN key, i.e. some rank holds a NaN in this cell ? 1 << 24 : 0).
__device__ __forceinline__ uint32_t ordered_key(float v) {
    const uint32_t bits = __float_as_uint(v);
    return bits ^ ((uint32_t)((int32_t)bits >> 31) | 0x80000000u);
}

// Fused exchange: with peers.n > 0 the candidate block of cell slab r is stored straight into
// rank r's receive buffer (peer memory over NVLink / NVSwitch, base pointers from a symmetric
// allocation) at [me][W][slab] -- the all-to-all happens inside this kernel, store by store,
// while the scan of the local scores keeps the SMs busy.
struct PeerBufs {
    float* base[16];
    int32_t n, me;
};

__global__ void __launch_bounds__(128)
window_compact_kernel(const float* __restrict__ x, int64_t xs, int32_t n, int64_t cells,
                      const int32_t* __restrict__ lo, const int32_t* __restrict__ hi, int32_t W, int64_t slab,
                      float* __restrict__ win, int32_t* __restrict__ cnt, uint32_t* overflow,
                      int32_t* __restrict__ overflow_cells, int32_t overflow_cap, const PeerBufs peers) {
    extern __shared__ float slots[];                 // [W][128]
    const int tid = threadIdx.x;
    const int64_t c = (int64_t)blockIdx.x * 128 + tid;
    if (c >= cells) return;
    const uint32_t klo = (uint32_t)lo[c] ^ 0x80000000u, khi = (uint32_t)hi[c] ^ 0x80000000u;
    const uint32_t span = khi - klo;                 // lo <= hi by construction
    const bool open = khi > 0xff800000u || klo > khi;   // hi is a NaN key (a rank saw a NaN): result is NaN anyway
    const float* px = x + c;
    // a NaN anywhere in this cell shows up as `open` on every rank: the local order statistic of
    // the rank that holds it is NaN, and NaN keys win the MAX all-reduce that produced hi
    uint32_t above = 0u, inside = 0u;
    auto take = [&](float v) {
        const uint32_t key = ordered_key(v);
        above += key > khi ? 1u : 0u;
        if (key - klo <= span && !open) {
            if (inside < (uint32_t)W) slots[inside * 128 + tid] = v;
            ++inside;
        }
    };
    int32_t i = 0;
    for (; i + 8 <= n; i += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { v[u] = __ldg(px); px += xs; }
#pragma unroll
        for (int u = 0; u < 8; ++u) take(v[u]);
    }
    for (; i < n; ++i) { take(__ldg(px)); px += xs; }
    if (inside > (uint32_t)W) {                      // reported, with the cell index: the caller redoes these exactly
        const uint32_t at = atomicAdd(overflow, 1u);
        if (at < (uint32_t)overflow_cap) overflow_cells[at] = (int32_t)c;
        inside = (uint32_t)W;
    }
    cnt[c] = (int32_t)(above | (open ? 1u << 24 : 0u));
    float* dst = peers.n > 0 ? peers.base[c / slab] + (int64_t)peers.me * W * slab + c % slab
                             : win + (c / slab) * (int64_t)W * slab + c % slab;
    for (int s = 0; s < W; ++s)
        dst[(int64_t)s * slab] = (uint32_t)s < inside ? slots[s * 128 + tid] : __uint_as_float(0xff800000u);
}

__global__ void __launch_bounds__(kThreads)
ordered_keys_kernel(const float* __restrict__ x, int64_t n, int32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i < n) out[i] = (int32_t)(ordered_key(x[i]) ^ 0x80000000u);
}

// ------------------------------------------------------------------------------------------
// K3x: the multi-GPU marginal quantile with every exchange fused into the kernels (peer memory
// over NVLink / NVSwitch, base pointers from CUDA symmetric allocations).  Three kernels, each
// followed by a barrier on the host side (a one-word all-reduce); no data-carrying collective:
//   peer_bcast_keys_kernel   key(a_r[c]) -> every peer's keys[me][c]
//   window_exchange_kernel   lo/hi = min/max of the w keys of the cell (read locally); scan the
//                            local column once; header (#above, #inside, NaN flag) + the
//                            candidates inside [lo, hi] -> the owner's recv[me][1+W][slab]; only
//                            as many candidate rows as the fullest cell of the warp are stored
//   owner_select_kernel      owner of a cell slab: pack the w candidate lists of each cell, sort
//                            one small network, pick the (kt - sum above)-th largest, store it
//                            into every peer's q[c]  (the all-gather of qhat, fused)
// ------------------------------------------------------------------------------------------
struct PeerPtrs {
    void* base[16];
    int32_t n, me;
};

template <int V>
__global__ void __launch_bounds__(kThreads)
peer_bcast_keys_kernel(const float* __restrict__ a, int64_t cells, int64_t cells_pad, const PeerPtrs peers) {
    const int64_t c = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * V;
    if (c >= cells) return;
    uint32_t k[V];
    if constexpr (V == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(a + c));
        k[0] = ordered_key(t.x); k[1] = ordered_key(t.y); k[2] = ordered_key(t.z); k[3] = ordered_key(t.w);
    } else {
        k[0] = ordered_key(__ldg(a + c));
    }
    for (int r = 0; r < peers.n; ++r) {
        uint32_t* dst = reinterpret_cast<uint32_t*>(peers.base[r]) + (int64_t)peers.me * cells_pad + c;
        if constexpr (V == 4) *reinterpret_cast<uint4*>(dst) = make_uint4(k[0], k[1], k[2], k[3]);
        else *dst = k[0];
    }
}

// src[0..n) (32-bit words) -> every listed peer at word offset `off`: the owner's slab of qhat pushed to
// the other ranks in one pass of wide stores (per-CTA 128-byte peer stores from the select kernel itself
// made that kernel wait on NVLink round trips: 0.45 -> 0.97 ms at 8 GPUs)
template <int V>
__global__ void __launch_bounds__(kThreads)
peer_copy_kernel(const uint32_t* __restrict__ src, int64_t n, int64_t off, const PeerPtrs peers) {
    const int64_t i = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * V;
    if (i >= n) return;
    if constexpr (V == 4) {
        const uint4 v = *reinterpret_cast<const uint4*>(src + i);
        for (int r = 0; r < peers.n; ++r) *reinterpret_cast<uint4*>(reinterpret_cast<uint32_t*>(peers.base[r]) + off + i) = v;
    } else {
        const uint32_t v = src[i];
        for (int r = 0; r < peers.n; ++r) (reinterpret_cast<uint32_t*>(peers.base[r]) + off)[i] = v;
    }
}

// header word of a (rank, cell) candidate list
__device__ __forceinline__ uint32_t pack_header(uint32_t above, uint32_t inside, bool open) {
    return (above & 0xffffffu) | (inside << 24) | (open ? 0x80000000u : 0u);
}

// T = lanes (cells) per CTA.  128 lanes x 60 candidate slots = 30.7 KB of shared memory per CTA; the
// carve-out is held at 164 KB (5 CTAs = 20 warps x 32 loads in flight) so that 64 KB of L1 remain for
// the lines in flight: with the carve-out at its maximum the same kernel measured 0.77 instead of
// 0.57 ms, and 64-lane CTAs (256-byte instead of 512-byte row pieces) were as slow.
template <int T>
__global__ void __launch_bounds__(T, 1024 / T)
window_exchange_kernel(const float* __restrict__ x, int64_t xs, int32_t n, int64_t cells,
                       const uint32_t* keys, int64_t cells_pad, int32_t key_mode, int32_t W, int64_t slab,
                       uint32_t* overflow, int32_t* __restrict__ overflow_cells, int32_t overflow_cap,
                       const PeerPtrs peers) {
    extern __shared__ float slots[];                 // [W][T]
    const int tid = threadIdx.x;
    const int64_t c = (int64_t)blockIdx.x * T + tid;
    const bool live = c < cells;
    uint32_t klo = 0xffffffffu, khi = 0u;
    if (live) {
        if (key_mode == 1) {                         // [~min, max] of signed ordered keys, MAX all-reduced
            klo = ~__ldcg(keys + c) ^ 0x80000000u;
            khi = __ldcg(keys + cells_pad + c) ^ 0x80000000u;
        } else {
            for (int r = 0; r < peers.n; ++r) {      // peer-written: read through L2, never the nc path
                const uint32_t k = __ldcg(keys + (int64_t)r * cells_pad + c);
                klo = min(klo, k);
                khi = max(khi, k);
            }
        }
    }
    const uint32_t span = khi - klo;
    const bool open = khi > 0xff800000u;             // some rank saw a NaN in this cell: the answer is NaN
    uint32_t above = 0u, inside = 0u;
    if (live && !open) {
        const float* px = x + c;
        // Branch-free per sample (10 instructions): key, `above` by predicate, and the candidate
        // store as a predicated st.shared at a running shared-window address -- a data-dependent
        // branch here costs a BSSY/BRA/BSYNC triple per sample and made the scan issue-bound.
        uint32_t sa = (uint32_t)__cvta_generic_to_shared(slots) + 4u * (uint32_t)tid;
        const uint32_t wmax = (uint32_t)W;
        auto take = [&](float v) {
            const uint32_t key = ordered_key(v);
            const uint32_t t = key - klo;
            asm volatile(
                "{\n\t.reg .pred a, q, p;\n\t"
                "setp.gt.u32 a, %3, %4;\n\t"             // above the bracket
                "@a add.u32 %0, %0, 1;\n\t"
                "setp.le.u32 q, %5, %6;\n\t"             // inside the bracket
                "setp.lt.and.u32 p, %1, %7, q;\n\t"      // and a slot is left
                "@p st.shared.f32 [%2], %8;\n\t"
                "@p add.u32 %2, %2, %9;\n\t"
                "@q add.u32 %1, %1, 1;\n\t}"
                : "+r"(above), "+r"(inside), "+r"(sa)
                : "r"(key), "r"(khi), "r"(t), "r"(span), "r"(wmax), "f"(v), "n"(4 * T)
                : "memory");
        };
        // two batches of 16 loads in flight per lane: the next batch is requested before the
        // current one is consumed (the asm statements above are ordering barriers for the
        // compiler, so the overlap has to be written out)
        float va[16], vb[16];
        auto load16 = [&](float (&v)[16]) {
#pragma unroll
            for (int u = 0; u < 16; ++u) { v[u] = __ldg(px); px += xs; }
        };
        auto take16 = [&](const float (&v)[16]) {
#pragma unroll
            for (int u = 0; u < 16; ++u) take(v[u]);
        };
        const int32_t nb = n / 16;
        if (nb > 0) load16(va);
        int32_t b = 0;
        for (; b + 2 <= nb; b += 2) {
            load16(vb);
            take16(va);
            if (b + 2 < nb) load16(va);
            take16(vb);
        }
        if (b < nb) take16(va);
        for (int32_t i = nb * 16; i < n; ++i) { take(__ldg(px)); px += xs; }
    }
    if (inside > (uint32_t)W) {                      // reported with the cell index: the caller redoes these exactly
        const uint32_t at = atomicAdd(overflow, 1u);
        if (at < (uint32_t)overflow_cap) overflow_cells[at] = (int32_t)c;
        inside = (uint32_t)W;
    }
    // Dense candidate rows.  The 32 candidate lists of a warp (= one 32-cell tile of the owner) are packed
    // into ONE stream, slot-major: for s = 0, 1, ... the lanes whose list is longer than s, in lane
    // order.  The owner rebuilds the same order from the 32 list lengths in the header row, so no index
    // travels.  Packing is done in place in the warp's own 32 columns of the slot array (a write never
    // lands beyond what the warp has already read: position <= 32*s + lane), and the tile's block
    //     [header row | ceil(sum(inside) / 32) dense rows]  of 128-byte lines
    // goes to the owner as whole lines only.  On |N(0,1)| scores at 8 GPUs that is 1 + 14/32 x 32 = ~15
    // words per cell instead of 1 + the longest list of the warp (~29): half the NVLink bytes.
    const int lane = tid & 31, wbase = tid & ~31;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t rows = __reduce_max_sync(0xffffffffu, inside);
    uint32_t total = 0u;
    for (uint32_t s = 0; s < rows; ++s) {
        const bool act = s < inside;
        const uint32_t bal = __ballot_sync(0xffffffffu, act);
        const float v = act ? slots[s * T + tid] : 0.0f;
        __syncwarp();
        if (act) {
            const uint32_t pos = total + __popc(bal & lt);
            slots[(pos >> 5) * T + wbase + (pos & 31u)] = v;
        }
        total += __popc(bal);
        __syncwarp();
    }
    const int64_t c0 = c - lane;                     // first cell of this warp's tile
    if (c0 >= cells) return;                         // warp-uniform
    const int64_t tiles_per_slab = slab >> 5;
    float* dst = reinterpret_cast<float*>(peers.base[c0 / slab]) +
                 ((int64_t)peers.me * tiles_per_slab + (c0 % slab) / 32) * (int64_t)(W + 1) * 32 + lane;
    dst[0] = __uint_as_float(live ? pack_header(above, inside, open) : 0u);
    const uint32_t dense_rows = (total + 31u) >> 5;
    for (uint32_t r = 0; r < dense_rows; ++r) dst[(r + 1) * 32] = slots[r * T + tid];
}

// Owner side.  recv = [w][1 + W][slab]: row 0 of block r holds rank r's headers, rows 1.. its
// candidates (only the first `inside` of a column are meaningful; the rest is stale memory and is
// never read).  One CTA handles TC cells with the register network of select_regsort_kernel:
// the fill packs each cell's candidates to the front of its tile column (shared-memory atomics
// hand out the positions), padding keys (0) sort first.
template <int L, int R, int TC>
__global__ void __launch_bounds__(TC * L)
owner_select_kernel(const float* recv, int32_t w, int32_t W, int64_t slab, int64_t cells, int32_t kt,
                    int64_t cell_offset, uint32_t* overflow, int32_t* __restrict__ overflow_cells,
                    int32_t overflow_cap, const PeerPtrs qpeers) {
    constexpr int N = L * R;
    constexpr int THREADS = TC * L;
    constexpr int WARPS = THREADS / 32;
    constexpr int CW = 32 / L;
    constexpr int LD = TC + 1;
    static_assert(TC == 32 && CW * WARPS == TC, "bad tiling");
    extern __shared__ uint32_t tile[];      // [N][LD] keys, then per column: count, above, flag; then the tile's longest list
    uint32_t* colcount = tile + N * LD;
    uint32_t* above = colcount + TC;
    uint32_t* flag = above + TC;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int64_t c0 = (int64_t)blockIdx.x * TC;
    if (tid < TC) { colcount[tid] = 0u; above[tid] = 0u; flag[tid] = 0u; }
    __syncthreads();
    // Warp `wid` takes the blocks of ranks wid, wid + WARPS, ...: lane = cell.  A block is the header row
    // (32 words) followed by the dense candidate stream of window_exchange_kernel; the slot-major order is
    // rebuilt from the 32 list lengths (ballot + popc), so lane `col` finds its s-th candidate at stream
    // position  sum_{s' < s} #{lists longer than s'} + #{lanes below col with a list longer than s}.
    // The tile is not cleared: the read-out below masks everything at or above the column's count.
    const int64_t tiles_per_slab = slab >> 5;
    {
        const int col = lane;
        const bool live = c0 + col < cells;
        const uint32_t lt = (1u << lane) - 1u;
        for (int r = wid; r < w; r += WARPS) {
            const float* blk = recv + ((int64_t)r * tiles_per_slab + blockIdx.x) * (int64_t)(W + 1) * 32;
            uint32_t l = 0u;
            if (live) {
                const uint32_t h = __float_as_uint(__ldcg(blk + col));
                atomicAdd(&above[col], h & 0xffffffu);
                if (h >> 31) flag[col] = 1u;
                l = (h >> 24) & 0x7fu;
            }
            const uint32_t rows = __reduce_max_sync(0xffffffffu, l);
            const float* stream = blk + 32;
            uint32_t total = 0u;
            // eight loads in flight per lane before any is consumed: one load per trip made the fill a
            // chain of L2 round trips (the whole kernel was latency-bound, 26 % issue-active)
            for (uint32_t sl0 = 0; sl0 < rows; sl0 += 8) {
                uint32_t bits[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const bool act = sl0 + u < l;
                    const uint32_t bal = __ballot_sync(0xffffffffu, act);
                    bits[u] = act ? __float_as_uint(__ldcg(stream + total + __popc(bal & lt))) : 0u;
                    total += __popc(bal);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (sl0 + u < l) {
                        const uint32_t pos = atomicAdd(&colcount[col], 1u);
                        if (pos < (uint32_t)N) tile[pos * LD + col] = float_to_key(bits[u]);
                    }
            }
        }
    }
    __syncthreads();
    // The network is sized by the fullest cell of THIS tile: N, N/2 or N/4 keys (the first RS
    // registers of every lane hold positions < RS*L, and the fill packs candidates from position
    // 0), so the capacity N only costs when a tile needs it.
    if (wid == 0) {
        const uint32_t m = __reduce_max_sync(0xffffffffu, colcount[lane]);
        if (lane == 0) flag[TC] = m;
    }
    __syncthreads();
    const uint32_t maxhave = flag[TC];
    const int j = lane % L, i = lane / L;
    const int ct = wid + WARPS * i;
    const uint32_t mine = colcount[ct];
    int RS = R;
    auto sort_sub = [&](auto RSC) {
        constexpr int rs = decltype(RSC)::value;
        uint32_t k[rs];
#pragma unroll
        for (int r = 0; r < rs; ++r) k[r] = (uint32_t)(r * L + j) < mine ? tile[(r * L + j) * LD + ct] : 0u;   // padding sorts first
        bitonic_sort_regs<L, rs>(k, j);
        __syncthreads();
#pragma unroll
        for (int r = 0; r < rs; ++r) tile[(r * L + j) * LD + ct] = k[r];
        RS = rs;
    };
    if constexpr (R >= 8) {
        if (maxhave * 4u <= (uint32_t)N) sort_sub(std::integral_constant<int, R / 4>{});
        else if (maxhave * 2u <= (uint32_t)N) sort_sub(std::integral_constant<int, R / 2>{});
        else sort_sub(std::integral_constant<int, R>{});
    } else {
        sort_sub(std::integral_constant<int, R>{});
    }
    __syncthreads();
    if (tid < TC && c0 + tid < cells) {
        const int NS = RS * L;
        const int want = kt - (int)above[tid];
        const int have = (int)colcount[tid];
        bool nan = flag[tid] != 0u;
        if (!nan && (want < 1 || want > have || have > N)) {   // bracket missed or a list was cut: redo exactly
            const uint32_t at = atomicAdd(overflow, 1u);
            if (at < (uint32_t)overflow_cap) overflow_cells[at] = (int32_t)(cell_offset + c0 + tid);
            nan = true;
        }
        const int kq = nan ? 0 : NS - want;
        const uint32_t v = tile[((kq % RS) * L + kq / RS) * LD + tid];
        const float q = nan ? __uint_as_float(0x7fc00000u) : __uint_as_float(key_to_float_bits(v));
        for (int r = 0; r < qpeers.n; ++r)
            reinterpret_cast<float*>(qpeers.base[r])[cell_offset + c0 + tid] = q;
    }
}

template <int L, int R>
int launch_owner(const float* recv, int32_t w, int32_t W, int64_t slab, int64_t cells, int32_t kt,
                 int64_t cell_offset, uint32_t* overflow, int32_t* overflow_cells, int32_t cap,
                 const PeerPtrs& qpeers, cudaStream_t st) {
    constexpr int N = L * R, TC = 32;
    const size_t smem = ((size_t)N * (TC + 1) + 3 * TC + 1) * sizeof(uint32_t);
    auto kern = owner_select_kernel<L, R, TC>;
    static SmemOptIn memo;
    {
        const int rc = ensure_dynamic_smem(kern, memo, smem);
        if (rc != CPPRE_OK) return rc;
    }
    const int64_t blocks = (cells + TC - 1) / TC;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    kern<<<(unsigned)blocks, TC * L, smem, st>>>(recv, w, W, slab, cells, kt, cell_offset, overflow, overflow_cells,
                                                 cap, qpeers);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int fill_peers(PeerPtrs* P, void* const* h_ptrs, int32_t n_peers, int32_t my_rank) {
    memset(P, 0, sizeof(*P));
    CPPRE_REQUIRE(h_ptrs != nullptr && n_peers >= 1 && n_peers <= 16 && my_rank >= 0 && my_rank < n_peers,
                  CPPRE_ERR_BAD_ARG, "bad peer list (n=%d me=%d)", n_peers, my_rank);
    for (int r = 0; r < n_peers; ++r) {
        CPPRE_REQUIRE(h_ptrs[r] != nullptr, CPPRE_ERR_BAD_ARG, "peer %d has no buffer", r);
        P->base[r] = h_ptrs[r];
    }
    P->n = n_peers;
    P->me = my_rank;
    return CPPRE_OK;
}

// K3b: per-cell single-rank radix select straight from global memory (any n).  8 passes of
// 4 bits; the 16 bucket counters are compared-and-added so they stay in registers.
__global__ void __launch_bounds__(kThreads)
select_radix_kernel(const float* __restrict__ x, int64_t xs, int64_t n, int64_t cells, RankList ranks,
                    float* __restrict__ out) {
    const int64_t c = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (c >= cells) return;
    const float* px = x + c;
    bool has_nan = false;
    for (int64_t i = 0; i < n; ++i)
        has_nan |= (__float_as_uint(__ldg(px + i * xs)) & 0x7fffffffu) > 0x7f800000u;
    for (int r = 0; r < ranks.n; ++r) {
        uint32_t prefix = 0u, mask = 0u;
        int64_t k = ranks.r[r];
        for (int shift = 28; shift >= 0; shift -= 4) {
            uint32_t cnt[16];
#pragma unroll
            for (int d = 0; d < 16; ++d) cnt[d] = 0u;
            for (int64_t i = 0; i < n; ++i) {
                const uint32_t key = float_to_key(__float_as_uint(__ldg(px + i * xs)));
                const bool m = (key & mask) == prefix;
                const uint32_t dig = (key >> shift) & 15u;
#pragma unroll
                for (int d = 0; d < 16; ++d) cnt[d] += (m && dig == (uint32_t)d) ? 1u : 0u;
            }
            uint32_t digit = 15u;
            bool found = false;
#pragma unroll
            for (int d = 0; d < 16; ++d) {
                if (!found) {
                    if (k < (int64_t)cnt[d]) { digit = d; found = true; }
                    else k -= cnt[d];
                }
            }
            prefix |= digit << shift;
            mask |= 15u << shift;
        }
        out[(int64_t)r * cells + c] = has_nan ? __uint_as_float(0x7fc00000u)
                                              : __uint_as_float(key_to_float_bits(prefix));
    }
}

// K3c: one 8-bit radix histogram pass over a flat array (Joint quantile across GPUs)
__global__ void __launch_bounds__(kThreads)
radix_hist_kernel(const float* __restrict__ x, int64_t n, uint32_t prefix, uint32_t mask, int shift,
                  uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0u;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
        const uint32_t key = float_to_key(__float_as_uint(__ldg(x + i)));
        if ((key & mask) == prefix) atomicAdd(&sh[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(hist + threadIdx.x, sh[threadIdx.x]);
}

// K3d: the same radix pass with its (prefix, mask) read from DEVICE state, and the digit choice as a
// kernel of its own, so the four passes of the Joint quantile run back to back on the stream with
// nothing but the all-reduce of the 257 counters between them (no host round trip).
//   state = {prefix, mask, k (remaining rank, 0-based), flags (bit 0: a NaN was seen, bit 1: rank out of range)}
//   hist  = 256 digit counters + 1 NaN counter (pass 0 only)
__global__ void __launch_bounds__(kThreads)
radix_hist_dev_kernel(const float* __restrict__ x, int64_t n, const uint32_t* __restrict__ state, int pass,
                      uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[257];
    sh[threadIdx.x] = 0u;
    if (threadIdx.x == 0) sh[256] = 0u;
    __syncthreads();
    const uint32_t prefix = state[0], mask = state[1];
    const int shift = 24 - 8 * pass;
    uint32_t nan = 0u;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
        const uint32_t bits = __float_as_uint(__ldg(x + i));
        const uint32_t key = float_to_key(bits);
        if (pass == 0) nan += (bits & 0x7fffffffu) > 0x7f800000u ? 1u : 0u;
        if ((key & mask) == prefix) atomicAdd(&sh[(key >> shift) & 255u], 1u);
    }
    if (pass == 0) {
        nan = __reduce_add_sync(0xffffffffu, nan);
        if ((threadIdx.x & 31) == 0 && nan) atomicAdd(&sh[256], nan);
    }
    __syncthreads();
    if (sh[threadIdx.x]) atomicAdd(hist + threadIdx.x, sh[threadIdx.x]);
    if (threadIdx.x == 0 && sh[256]) atomicAdd(hist + 256, sh[256]);
}

// The whole select in ONE launch when the scores live on one GPU: a single 1024-thread CTA runs the four
// passes back to back (the array is read from L2 after the first pass), histogram and state in shared memory.
__global__ void __launch_bounds__(1024)
radix_select_block_kernel(const float* __restrict__ x, int64_t n, uint32_t k, float* __restrict__ out) {
    __shared__ uint32_t hist[256];
    __shared__ uint32_t sel[3];          // prefix, remaining rank, flags (1: NaN seen, 2: rank out of range)
    const int t = threadIdx.x;
    if (t < 3) sel[t] = t == 1 ? k : 0u;
    uint32_t mask = 0u;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        if (t < 256) hist[t] = 0u;
        __syncthreads();
        const uint32_t prefix = sel[0];
        bool nan = false;
        for (int64_t i = t; i < n; i += 1024) {
            const uint32_t bits = __float_as_uint(__ldg(x + i));
            nan |= (bits & 0x7fffffffu) > 0x7f800000u;
            const uint32_t key = float_to_key(bits);
            if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
        }
        if (pass == 0 && nan) sel[2] = 1u;
        __syncthreads();
        if (t == 0) {                    // 256 counters: a serial scan costs less than another barrier round
            uint32_t kk = sel[1], acc = 0u;
            int d = 0;
            for (; d < 256; ++d) {
                if (kk < acc + hist[d]) break;
                acc += hist[d];
            }
            if (d == 256) { sel[2] |= 2u; d = 255; }
            sel[0] = prefix | ((uint32_t)d << shift);
            sel[1] = kk - acc;
        }
        mask |= 255u << shift;
        __syncthreads();
    }
    if (t == 0) *out = sel[2] ? __uint_as_float(0x7fc00000u) : __uint_as_float(key_to_float_bits(sel[0]));
}

__global__ void __launch_bounds__(kThreads)
radix_begin_kernel(uint32_t* __restrict__ state, uint32_t* __restrict__ hist, uint32_t k) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i < 257) hist[i] = 0u;
    if (i < 4) state[i] = i == 2 ? k : 0u;
}

// One CTA: picks the digit that holds rank k in the (all-reduced) histogram, narrows the state,
// clears the histogram for the next pass; after the last pass writes the selected value.
__global__ void __launch_bounds__(kThreads)
radix_pick_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ state, int pass, float* __restrict__ out) {
    __shared__ uint32_t warp_tot[8];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const uint32_t c = hist[t];
    // inclusive scan of the 256 counters: shuffles within a warp, then the 8 warp totals
    uint32_t s = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += v;
    }
    if (lane == 31) warp_tot[w] = s;
    __syncthreads();
    uint32_t before = 0u;
    for (int i = 0; i < w; ++i) before += warp_tot[i];
    const uint32_t incl = s + before, excl = incl - c;
    const uint32_t k = state[2];
    const uint32_t nan_seen = pass == 0 ? hist[256] : 0u;
    __syncthreads();                                 // everyone has read state / hist
    const int shift = 24 - 8 * pass;
    if (c != 0u && k >= excl && k < incl) {          // exactly one thread (or none: rank out of range)
        state[0] |= (uint32_t)t << shift;
        state[1] |= 255u << shift;
        state[2] = k - excl;
    }
    if (t == 255 && k >= incl) state[3] |= 2u;       // rank beyond the number of scores
    if (t == 0 && nan_seen) state[3] |= 1u;
    hist[t] = 0u;
    if (t == 0) hist[256] = 0u;
    if (pass == 3 && out != nullptr) {
        __syncthreads();
        if (t == 0) {
            __threadfence_block();
            const uint32_t flags = state[3];
            *out = flags ? __uint_as_float(0x7fc00000u) : __uint_as_float(key_to_float_bits(state[0]));
        }
    }
}

// ------------------------------------------------------------------------------------------
// K5: coverage
// ------------------------------------------------------------------------------------------
template <bool BAND, int V>
__global__ void __launch_bounds__(kThreads)
inside_kernel(const float* __restrict__ y, int64_t ys, const float* __restrict__ p0, int64_t s0,
              const float* __restrict__ p1, int64_t s1, float q, int64_t cells, int32_t chunks,
              uint32_t* __restrict__ inside) {
    const int64_t i = blockIdx.x / chunks;
    const int chunk = blockIdx.x % chunks;
    uint32_t cnt = 0u;
    const int64_t step = (int64_t)chunks * kThreads * V;
#pragma unroll 4
    for (int64_t c = ((int64_t)chunk * kThreads + threadIdx.x) * V; c < cells; c += step) {
        float v[V], lo[V], hi[V];
        ldv<V>(y + i * ys + c, v);
        if (BAND) {   // p0 = centre (or null), p1 = half width (or null); separately rounded ops
            float ctr[V], h[V];
#pragma unroll
            for (int e = 0; e < V; ++e) { ctr[e] = 0.0f; h[e] = q; }
            if (p0) ldv<V>(p0 + i * s0 + c, ctr);
            if (p1) {
                ldv<V>(p1 + c, h);
#pragma unroll
                for (int e = 0; e < V; ++e) h[e] = __fmul_rn(q, h[e]);
            }
#pragma unroll
            for (int e = 0; e < V; ++e) { lo[e] = __fsub_rn(ctr[e], h[e]); hi[e] = __fadd_rn(ctr[e], h[e]); }
        } else {
            ldv<V>(p0 + i * s0 + c, lo);
            ldv<V>(p1 + i * s1 + c, hi);
        }
#pragma unroll
        for (int e = 0; e < V; ++e) cnt += (v[e] >= lo[e] && v[e] <= hi[e]) ? 1u : 0u;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(inside + i, cnt);
}

int chunks_for(int64_t n, int64_t cells) {
    // enough blocks to fill the machine a few times, never less than 16 cells per thread
    int64_t want = (148 * 16 + n - 1) / (n > 0 ? n : 1);
    const int64_t cap = (cells + kThreads * 16 - 1) / (kThreads * 16);   // >= 16 cells per thread
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return (int)want;
}

}  // namespace
}  // namespace cppre

using namespace cppre;

extern "C" {

int cppre_cell_moments(const float* d_a, int64_t as, const float* d_b, int64_t bs, int64_t n,
                       int64_t cells, double* d_mean, double* d_m2, void* stream) {
    CPPRE_REQUIRE(n >= 0 && cells >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_mean && d_m2 && (d_a || n == 0), CPPRE_ERR_BAD_ARG, "null pointer");
    return launch_moments<false>(d_a, as, d_b, bs, n, cells, d_mean, d_m2, 0.0f, nullptr, nullptr, (cudaStream_t)stream);
}

int cppre_cell_modulation(const float* d_a, int64_t as, const float* d_b, int64_t bs, int64_t n, int64_t cells,
                          float add_eps, float* d_std, float* d_inv, void* stream) {
    CPPRE_REQUIRE(n >= 1 && cells >= 0, CPPRE_ERR_BAD_ARG, "bad size (n=%lld)", (long long)n);
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_a && d_std, CPPRE_ERR_BAD_ARG, "null pointer");
    return launch_moments<true>(d_a, as, d_b, bs, n, cells, nullptr, nullptr, add_eps, d_std, d_inv, (cudaStream_t)stream);
}

int cppre_cell_std(const double* d_m2, double count, float add_eps, int64_t cells, float* d_std,
                   float* d_inv, void* stream) {
    CPPRE_REQUIRE(cells >= 0 && count > 0, CPPRE_ERR_BAD_ARG, "bad size");
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_m2 && d_std, CPPRE_ERR_BAD_ARG, "null pointer");
    const int64_t blocks = (cells + kThreads - 1) / kThreads;
    cell_std_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(d_m2, count, add_eps, cells, d_std,
                                                                           d_inv);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_joint_score(const float* d_a, int64_t as, const float* d_b, int64_t bs, const float* d_mod,
                      int64_t n, int64_t cells, float* d_scores, void* stream) {
    CPPRE_REQUIRE(n >= 0 && cells >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (n == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_a && d_mod && d_scores, CPPRE_ERR_BAD_ARG, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    CPPRE_CHECK_CUDA(cudaMemsetAsync(d_scores, 0, sizeof(float) * (size_t)n, st));
    if (cells == 0) return CPPRE_OK;
    constexpr int S = 4;                                   // samples per block
    const int64_t groups = (n + S - 1) / S;
    const int chunks = chunks_for(groups, cells);
    CPPRE_REQUIRE(groups * chunks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    if (vec4_ok(cells, {d_a, d_b, d_mod}, {as, d_b ? bs : 0}))
        joint_score_kernel<4, S><<<(unsigned)(groups * chunks), kThreads, 0, st>>>(d_a, as, d_b, bs, d_mod, n, cells,
                                                                                  chunks, d_scores);
    else
        joint_score_kernel<1, S><<<(unsigned)(groups * chunks), kThreads, 0, st>>>(d_a, as, d_b, bs, d_mod, n, cells,
                                                                                  chunks, d_scores);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

size_t cppre_select_workspace_bytes(int64_t, int64_t) { return 0; }

int cppre_select_ranks(const float* d_scores, int64_t row_stride, int64_t n, int64_t cells,
                       const int32_t* h_ranks, int32_t n_ranks, float* d_out, void*, void* stream) {
    CPPRE_REQUIRE(n >= 1 && cells >= 0 && n_ranks >= 0, CPPRE_ERR_BAD_ARG, "bad size (n=%lld)", (long long)n);
    if (cells == 0 || n_ranks == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_scores && h_ranks && d_out, CPPRE_ERR_BAD_ARG, "null pointer");
    CPPRE_REQUIRE(n < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "n too large");
    cudaStream_t st = (cudaStream_t)stream;
    for (int r0 = 0; r0 < n_ranks; r0 += kMaxRanks) {
        RankList rl;
        rl.n = (n_ranks - r0 < kMaxRanks) ? n_ranks - r0 : kMaxRanks;
        for (int r = 0; r < rl.n; ++r) {
            CPPRE_REQUIRE(h_ranks[r0 + r] >= 0 && h_ranks[r0 + r] < n, CPPRE_ERR_BAD_ARG,
                          "rank %d out of range [0,%lld)", h_ranks[r0 + r], (long long)n);
            rl.r[r] = h_ranks[r0 + r];
        }
        float* out = d_out + (int64_t)r0 * cells;
        {
            const int rc = dispatch_topk(d_scores, row_stride, n, cells, rl, out, st);
            if (rc >= 0) {
                if (rc != CPPRE_OK) return rc;
                continue;
            }
        }
        if (n <= 2048) {
            const int rc = dispatch_regsort(d_scores, row_stride, n, cells, rl, out, st);
            if (rc != CPPRE_OK) return rc;
            continue;
        }
        const int64_t blocks = (cells + kThreads - 1) / kThreads;
        select_radix_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(d_scores, row_stride, n, cells, rl, out);
        CPPRE_CHECK_LAUNCH();
    }
    return CPPRE_OK;
}

int cppre_ordered_keys(const float* d_x, int64_t n, int32_t* d_out, void* stream) {
    CPPRE_REQUIRE(n >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (n == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_x && d_out, CPPRE_ERR_BAD_ARG, "null pointer");
    const int64_t blocks = (n + kThreads - 1) / kThreads;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    ordered_keys_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(d_x, n, d_out);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_window_compact(const float* d_scores, int64_t row_stride, int64_t n, int64_t cells,
                         const int32_t* d_lo, const int32_t* d_hi, int32_t slots, int64_t slab,
                         float* d_win, int32_t* d_cnt, uint32_t* d_overflow, int32_t* d_overflow_cells,
                         int32_t overflow_cap, float* const* h_peer_win, int32_t n_peers, int32_t my_rank,
                         void* stream) {
    PeerBufs peers;
    memset(&peers, 0, sizeof(peers));
    if (h_peer_win != nullptr) {
        CPPRE_REQUIRE(n_peers >= 1 && n_peers <= 16 && my_rank >= 0 && my_rank < n_peers, CPPRE_ERR_BAD_ARG,
                      "bad peer list (n=%d me=%d)", n_peers, my_rank);
        CPPRE_REQUIRE((cells + slab - 1) / slab <= n_peers, CPPRE_ERR_BAD_ARG, "more cell slabs than peers");
        for (int r = 0; r < n_peers; ++r) {
            CPPRE_REQUIRE(h_peer_win[r] != nullptr, CPPRE_ERR_BAD_ARG, "peer %d has no buffer", r);
            peers.base[r] = h_peer_win[r];
        }
        peers.n = n_peers;
        peers.me = my_rank;
    }
    CPPRE_REQUIRE(cells < (int64_t(1) << 31) && overflow_cap >= 0 && (overflow_cap == 0 || d_overflow_cells),
                  CPPRE_ERR_BAD_ARG, "bad overflow list");
    CPPRE_REQUIRE(n >= 0 && cells >= 0 && slots >= 1 && slots <= 96 && slab >= 1, CPPRE_ERR_BAD_ARG,
                  "bad size (n=%lld cells=%lld slots=%d slab=%lld)", (long long)n, (long long)cells, slots,
                  (long long)slab);
    CPPRE_REQUIRE(n < (int64_t(1) << 24), CPPRE_ERR_UNSUPPORTED, "n too large for the packed count");
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE((d_scores || n == 0) && d_lo && d_hi && (d_win || peers.n > 0) && d_cnt && d_overflow,
                  CPPRE_ERR_BAD_ARG, "null pointer");
    const int64_t blocks = (cells + 127) / 128;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    const size_t smem = (size_t)slots * 128 * sizeof(float);
    static SmemOptIn memo;
    {
        const int rc = ensure_dynamic_smem(window_compact_kernel, memo, smem);
        if (rc != CPPRE_OK) return rc;
    }
    window_compact_kernel<<<(unsigned)blocks, 128, smem, (cudaStream_t)stream>>>(
        d_scores, row_stride, (int32_t)n, cells, d_lo, d_hi, slots, slab, d_win, d_cnt, d_overflow,
        d_overflow_cells, overflow_cap, peers);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_select_from_top(const float* d_cand, int64_t row_stride, int64_t rows, int64_t cells,
                          const int32_t* d_cnt, int32_t kt, int32_t network, float* d_out,
                          uint32_t* d_overflow, int32_t* d_overflow_cells, int32_t overflow_cap,
                          int64_t cell_offset, void* stream) {
    CPPRE_REQUIRE(rows >= 1 && cells >= 0 && rows < (int64_t(1) << 31), CPPRE_ERR_BAD_ARG, "bad size");
    CPPRE_REQUIRE(overflow_cap >= 0 && (overflow_cap == 0 || d_overflow_cells), CPPRE_ERR_BAD_ARG, "bad overflow list");
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_cand && d_cnt && d_out && d_overflow, CPPRE_ERR_BAD_ARG, "null pointer");
    RankList rl;
    rl.n = 1;
    rl.r[0] = 0;
    const PerCell pc{d_cnt, kt, d_overflow, d_overflow_cells, overflow_cap, cell_offset};
    const int rc = network > 0
                       ? dispatch_compact(network, d_cand, row_stride, rows, cells, rl, d_out, (cudaStream_t)stream, pc)
                       : dispatch_regsort(d_cand, row_stride, rows, cells, rl, d_out, (cudaStream_t)stream, pc);
    CPPRE_REQUIRE(rc >= 0, CPPRE_ERR_UNSUPPORTED, "unsupported candidate select (rows=%lld network=%d)",
                  (long long)rows, network);
    return rc;
}

int cppre_peer_bcast_keys(const float* d_vals, int64_t cells, int64_t cells_pad, void* const* h_peer_keys,
                          int32_t n_peers, int32_t my_rank, void* stream) {
    PeerPtrs peers;
    const int rc = fill_peers(&peers, h_peer_keys, n_peers, my_rank);
    if (rc != CPPRE_OK) return rc;
    CPPRE_REQUIRE(cells >= 0 && cells_pad >= cells && cells_pad % 4 == 0, CPPRE_ERR_BAD_ARG, "bad size");
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_vals != nullptr, CPPRE_ERR_BAD_ARG, "null pointer");
    bool v4 = cells % 4 == 0 && (reinterpret_cast<uintptr_t>(d_vals) & 15u) == 0;
    for (int r = 0; r < n_peers; ++r) v4 = v4 && (reinterpret_cast<uintptr_t>(h_peer_keys[r]) & 15u) == 0;
    const int64_t items = v4 ? cells / 4 : cells;
    const int64_t blocks = (items + kThreads - 1) / kThreads;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    if (v4)
        peer_bcast_keys_kernel<4><<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(d_vals, cells, cells_pad, peers);
    else
        peer_bcast_keys_kernel<1><<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(d_vals, cells, cells_pad, peers);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_window_exchange(const float* d_scores, int64_t row_stride, int64_t n, int64_t cells,
                          const uint32_t* d_keys, int64_t cells_pad, int32_t key_mode, int32_t slots, int64_t slab,
                          void* const* h_peer_recv, int32_t n_peers, int32_t my_rank, uint32_t* d_overflow,
                          int32_t* d_overflow_cells, int32_t overflow_cap, void* stream) {
    CPPRE_REQUIRE(key_mode == 0 || key_mode == 1, CPPRE_ERR_BAD_ARG, "key_mode must be 0 or 1");
    PeerPtrs peers;
    const int rc = fill_peers(&peers, h_peer_recv, n_peers, my_rank);
    if (rc != CPPRE_OK) return rc;
    CPPRE_REQUIRE(n >= 0 && cells >= 0 && slots >= 1 && slots <= 96 && slab >= 1 && cells_pad >= cells,
                  CPPRE_ERR_BAD_ARG, "bad size (n=%lld cells=%lld slots=%d slab=%lld)", (long long)n,
                  (long long)cells, slots, (long long)slab);
    CPPRE_REQUIRE((cells + slab - 1) / slab <= n_peers, CPPRE_ERR_BAD_ARG, "more cell slabs than peers");
    CPPRE_REQUIRE(slab % 32 == 0, CPPRE_ERR_BAD_ARG, "slab must be a multiple of 32 cells (one owner tile per warp), got %lld",
                  (long long)slab);
    CPPRE_REQUIRE(cells < (int64_t(1) << 31) && overflow_cap >= 0 && (overflow_cap == 0 || d_overflow_cells),
                  CPPRE_ERR_BAD_ARG, "bad overflow list");
    CPPRE_REQUIRE(n < (int64_t(1) << 24), CPPRE_ERR_UNSUPPORTED, "n too large for the packed count");
    if (cells == 0) return CPPRE_OK;
    CPPRE_REQUIRE((d_scores || n == 0) && d_keys && d_overflow, CPPRE_ERR_BAD_ARG, "null pointer");
    constexpr int T = 128;
    const int64_t blocks = (cells + T - 1) / T;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    const size_t smem = (size_t)slots * T * sizeof(float);
    static SmemOptIn memo;
    {
        const int rc = ensure_dynamic_smem(window_exchange_kernel<T>, memo, smem, 72);
        if (rc != CPPRE_OK) return rc;
    }
    window_exchange_kernel<T><<<(unsigned)blocks, T, smem, (cudaStream_t)stream>>>(
        d_scores, row_stride, (int32_t)n, cells, d_keys, cells_pad, key_mode, slots, slab, d_overflow,
        d_overflow_cells, overflow_cap, peers);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_owner_select(const float* d_recv, int32_t n_ranks, int32_t slots, int64_t slab, int64_t cells,
                       int32_t kt, int32_t network, int64_t cell_offset, void* const* h_peer_q, int32_t n_peers,
                       uint32_t* d_overflow, int32_t* d_overflow_cells, int32_t overflow_cap, void* stream) {
    PeerPtrs qp;
    const int rc0 = fill_peers(&qp, h_peer_q, n_peers, 0);
    if (rc0 != CPPRE_OK) return rc0;
    CPPRE_REQUIRE(n_ranks >= 1 && n_ranks <= 16 && slots >= 1 && slots <= 96 && slab >= 1 && cells <= slab &&
                      slab % 32 == 0, CPPRE_ERR_BAD_ARG, "bad size (slab must be a multiple of 32)");
    CPPRE_REQUIRE(overflow_cap >= 0 && (overflow_cap == 0 || d_overflow_cells), CPPRE_ERR_BAD_ARG, "bad overflow list");
    if (cells <= 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_recv && d_overflow, CPPRE_ERR_BAD_ARG, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    switch (network) {
        case 32: return launch_owner<1, 32>(d_recv, n_ranks, slots, slab, cells, kt, cell_offset, d_overflow, d_overflow_cells, overflow_cap, qp, st);
        case 64: return launch_owner<2, 32>(d_recv, n_ranks, slots, slab, cells, kt, cell_offset, d_overflow, d_overflow_cells, overflow_cap, qp, st);
        case 128: return launch_owner<4, 32>(d_recv, n_ranks, slots, slab, cells, kt, cell_offset, d_overflow, d_overflow_cells, overflow_cap, qp, st);
        case 256: return launch_owner<8, 32>(d_recv, n_ranks, slots, slab, cells, kt, cell_offset, d_overflow, d_overflow_cells, overflow_cap, qp, st);
        case 512: return launch_owner<16, 32>(d_recv, n_ranks, slots, slab, cells, kt, cell_offset, d_overflow, d_overflow_cells, overflow_cap, qp, st);
        case 1024: return launch_owner<32, 32>(d_recv, n_ranks, slots, slab, cells, kt, cell_offset, d_overflow, d_overflow_cells, overflow_cap, qp, st);
        default: break;
    }
    CPPRE_REQUIRE(false, CPPRE_ERR_UNSUPPORTED, "network must be a power of two in [32, 1024] (got %d)", network);
    return CPPRE_ERR_UNSUPPORTED;
}

int cppre_peer_copy(const void* d_src, int64_t n_words, int64_t dst_offset_words, void* const* h_peer_dst,
                    int32_t n_peers, void* stream) {
    CPPRE_REQUIRE(n_words >= 0 && dst_offset_words >= 0 && n_peers >= 0, CPPRE_ERR_BAD_ARG, "bad size");
    if (n_words == 0 || n_peers == 0) return CPPRE_OK;
    PeerPtrs peers;
    const int rc = fill_peers(&peers, h_peer_dst, n_peers, 0);
    if (rc != CPPRE_OK) return rc;
    CPPRE_REQUIRE(d_src != nullptr, CPPRE_ERR_BAD_ARG, "null pointer");
    bool v4 = n_words % 4 == 0 && dst_offset_words % 4 == 0 && (reinterpret_cast<uintptr_t>(d_src) & 15u) == 0;
    for (int r = 0; r < n_peers; ++r) v4 = v4 && (reinterpret_cast<uintptr_t>(h_peer_dst[r]) & 15u) == 0;
    const int64_t items = v4 ? n_words / 4 : n_words;
    const int64_t blocks = (items + kThreads - 1) / kThreads;
    CPPRE_REQUIRE(blocks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    if (v4)
        peer_copy_kernel<4><<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(
            static_cast<const uint32_t*>(d_src), n_words, dst_offset_words, peers);
    else
        peer_copy_kernel<1><<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(
            static_cast<const uint32_t*>(d_src), n_words, dst_offset_words, peers);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_radix_hist(const float* d_x, int64_t n, uint32_t prefix, uint32_t mask, int32_t shift,
                     uint32_t* d_hist, void* stream) {
    CPPRE_REQUIRE(n >= 0 && shift >= 0 && shift <= 24, CPPRE_ERR_BAD_ARG, "bad argument");
    if (n == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_x && d_hist, CPPRE_ERR_BAD_ARG, "null pointer");
    int64_t blocks = (n + kThreads - 1) / kThreads;
    if (blocks > 148 * 8) blocks = 148 * 8;
    radix_hist_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(d_x, n, prefix, mask, shift, d_hist);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_radix_select(const float* d_x, int64_t n, int64_t k, float* d_out, void* stream) {
    CPPRE_REQUIRE(n >= 0 && k >= 0 && k < (int64_t(1) << 32), CPPRE_ERR_BAD_ARG, "bad argument");
    CPPRE_REQUIRE(d_out && (d_x || n == 0), CPPRE_ERR_BAD_ARG, "null pointer");
    radix_select_block_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(d_x, n, (uint32_t)k, d_out);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_radix_select_begin(uint32_t* d_state, uint32_t* d_hist, int64_t k, void* stream) {
    CPPRE_REQUIRE(d_state && d_hist, CPPRE_ERR_BAD_ARG, "null pointer");
    CPPRE_REQUIRE(k >= 0 && k < (int64_t(1) << 32), CPPRE_ERR_BAD_ARG, "rank out of range");
    radix_begin_kernel<<<2, kThreads, 0, (cudaStream_t)stream>>>(d_state, d_hist, (uint32_t)k);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_radix_select_hist(const float* d_x, int64_t n, const uint32_t* d_state, int32_t pass, uint32_t* d_hist,
                            void* stream) {
    CPPRE_REQUIRE(n >= 0 && pass >= 0 && pass <= 3, CPPRE_ERR_BAD_ARG, "bad argument");
    CPPRE_REQUIRE(d_state && d_hist && (d_x || n == 0), CPPRE_ERR_BAD_ARG, "null pointer");
    if (n == 0) return CPPRE_OK;
    int64_t blocks = (n + kThreads - 1) / kThreads;
    if (blocks > 148 * 8) blocks = 148 * 8;
    radix_hist_dev_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(d_x, n, d_state, pass, d_hist);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_radix_select_pick(uint32_t* d_hist, uint32_t* d_state, int32_t pass, float* d_out, void* stream) {
    CPPRE_REQUIRE(pass >= 0 && pass <= 3, CPPRE_ERR_BAD_ARG, "bad pass");
    CPPRE_REQUIRE(d_state && d_hist && (d_out || pass < 3), CPPRE_ERR_BAD_ARG, "null pointer");
    radix_pick_kernel<<<1, kThreads, 0, (cudaStream_t)stream>>>(d_hist, d_state, pass, d_out);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_inside_count(const float* d_y, int64_t ys, const float* d_lo, int64_t los, const float* d_hi,
                       int64_t his, int64_t n, int64_t cells, uint32_t* d_inside, void* stream) {
    CPPRE_REQUIRE(n >= 0 && cells >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (n == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_inside && (cells == 0 || (d_y && d_lo && d_hi)), CPPRE_ERR_BAD_ARG, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    CPPRE_CHECK_CUDA(cudaMemsetAsync(d_inside, 0, sizeof(uint32_t) * (size_t)n, st));
    if (cells == 0) return CPPRE_OK;
    const int chunks = chunks_for(n, cells);
    CPPRE_REQUIRE(n * chunks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    if (vec4_ok(cells, {d_y, d_lo, d_hi}, {ys, los, his}))
        inside_kernel<false, 4><<<(unsigned)(n * chunks), kThreads, 0, st>>>(d_y, ys, d_lo, los, d_hi, his, 0.0f,
                                                                            cells, chunks, d_inside);
    else
        inside_kernel<false, 1><<<(unsigned)(n * chunks), kThreads, 0, st>>>(d_y, ys, d_lo, los, d_hi, his, 0.0f,
                                                                            cells, chunks, d_inside);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_inside_count_band(const float* d_y, int64_t ys, const float* d_centre, int64_t cs,
                            const float* d_half, float q, int64_t n, int64_t cells, uint32_t* d_inside,
                            void* stream) {
    CPPRE_REQUIRE(n >= 0 && cells >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    if (n == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_inside && (cells == 0 || d_y), CPPRE_ERR_BAD_ARG, "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    CPPRE_CHECK_CUDA(cudaMemsetAsync(d_inside, 0, sizeof(uint32_t) * (size_t)n, st));
    if (cells == 0) return CPPRE_OK;
    const int chunks = chunks_for(n, cells);
    CPPRE_REQUIRE(n * chunks < (int64_t(1) << 31), CPPRE_ERR_UNSUPPORTED, "grid too large");
    if (vec4_ok(cells, {d_y, d_centre, d_half}, {ys, d_centre ? cs : 0}))
        inside_kernel<true, 4><<<(unsigned)(n * chunks), kThreads, 0, st>>>(d_y, ys, d_centre, cs, d_half, 0, q, cells,
                                                                           chunks, d_inside);
    else
        inside_kernel<true, 1><<<(unsigned)(n * chunks), kThreads, 0, st>>>(d_y, ys, d_centre, cs, d_half, 0, q, cells,
                                                                           chunks, d_inside);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// K3d without caller scratch (cppre_radix_select_ranks_noalloc)
//
// For hosts that cannot give the radix select its 261 words of state per rank because they have no
// device allocator to ask (all of HBM sits in somebody else's arena, or the binding only forwards
// field buffers, or cudaMalloc's implicit synchronisation is unwelcome where they call from).  The
// state lives in the array below: device memory that belongs to the library image, present on every
// GPU that loads it.
//
// `g_noalloc_scratch` as written in HOST code is only the host-side handle of the symbol, not a
// device address -- a kernel handed that value would fault, or quietly scribble over host memory on
// a system where the GPU can reach it.  The address kernels may use comes from cudaGetSymbolAddress,
// asked once per device with that device current (each device has its own copy of the array).
// ------------------------------------------------------------------------------------------
namespace {
constexpr int kNoallocSlotWords = 384;   // one rank: state at word 0 (4 words), counters at word 64 (257 words)
constexpr int kNoallocHistWord = 64;
static_assert(kNoallocHistWord + 257 <= kNoallocSlotWords, "slot too small for the counters");
static_assert(kNoallocSlotWords * 4 * CPPRE_NOALLOC_MAX_RANKS == CPPRE_NOALLOC_SCRATCH_BYTES,
              "header constant out of step with the array");
}  // namespace

static __device__ __align__(256) uint32_t g_noalloc_scratch[kNoallocSlotWords * CPPRE_NOALLOC_MAX_RANKS];

namespace {
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_addr[cppre::kMaxDevices];   // per device, filled on the first call there

// Device address of g_noalloc_scratch on the CURRENT device.
int noalloc_scratch(uint32_t** out) {
    int dev = -1;
    CPPRE_CHECK_CUDA(cudaGetDevice(&dev));
    CPPRE_REQUIRE(dev >= 0 && dev < cppre::kMaxDevices, CPPRE_ERR_UNSUPPORTED, "device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_addr[dev] == nullptr) {
        void* p = nullptr;
        CPPRE_CHECK_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        CPPRE_REQUIRE(p != nullptr, CPPRE_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_noalloc_addr[dev] = static_cast<uint32_t*>(p);
    }
    *out = g_noalloc_addr[dev];
    return CPPRE_OK;
}
}  // namespace

extern "C" int cppre_radix_select_ranks_noalloc(const float* d_x, int64_t n, const int64_t* h_ranks, int32_t n_ranks,
                                                float* d_out, void* stream) {
    CPPRE_REQUIRE(n >= 0 && n_ranks >= 0, CPPRE_ERR_BAD_ARG, "negative size");
    CPPRE_REQUIRE(n_ranks <= CPPRE_NOALLOC_MAX_RANKS, CPPRE_ERR_UNSUPPORTED,
                  "n_ranks = %d is more than CPPRE_NOALLOC_MAX_RANKS = %d: the built-in scratch holds one radix "
                  "state per rank; split the call or use cppre_radix_select_begin/hist/pick with your own scratch",
                  (int)n_ranks, CPPRE_NOALLOC_MAX_RANKS);
    CPPRE_REQUIRE(n < (int64_t(1) << 32), CPPRE_ERR_UNSUPPORTED,
                  "n = %lld does not fit the 32-bit counters of the built-in scratch (limit 2^32 - 1)", (long long)n);
    if (n_ranks == 0) return CPPRE_OK;
    CPPRE_REQUIRE(h_ranks && d_out && (d_x || n == 0), CPPRE_ERR_BAD_ARG, "null pointer");
    for (int r = 0; r < n_ranks; ++r)
        CPPRE_REQUIRE(h_ranks[r] >= 0 && h_ranks[r] < (int64_t(1) << 32), CPPRE_ERR_BAD_ARG, "rank out of range");
    uint32_t* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t blocks = (n + kThreads - 1) / kThreads;
    if (blocks > 148 * 8) blocks = 148 * 8;
    for (int r = 0; r < n_ranks; ++r) {
        uint32_t* state = scratch + (size_t)r * kNoallocSlotWords;
        uint32_t* hist = state + kNoallocHistWord;
        // the previous call's leftovers end here: state = {0, 0, k, 0}, all 257 counters = 0
        radix_begin_kernel<<<2, kThreads, 0, st>>>(state, hist, (uint32_t)h_ranks[r]);
        CPPRE_CHECK_LAUNCH();
        for (int pass = 0; pass < 4; ++pass) {
            if (n > 0) {
                radix_hist_dev_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(d_x, n, state, pass, hist);
                CPPRE_CHECK_LAUNCH();
            }
            radix_pick_kernel<<<1, kThreads, 0, st>>>(hist, state, pass, d_out + r);
            CPPRE_CHECK_LAUNCH();
        }
    }
    return CPPRE_OK;
}
