// Boundary-condition padding (BoundaryManager.pad_signal) as an index map, and the valid 2-D
// cross-correlation evaluated directly on the virtual padded image (no padded copy).
//
// The reference pads sequentially: left, right (last axis), then top, bottom (rows), each step
// acting on the tensor produced by the previous one [ref: Utils/boundary_conditions.py:103-176].
// That makes the map separable -- a row resolver and a column resolver -- and reproduces the
// reference's quirks by construction: a 'periodic' right/bottom ghost copies the first cells of
// the *already padded* tensor (i.e. the left/top ghosts), 'free_slip' pads nothing.
#include "common.cuh"

namespace cppre {
namespace {

constexpr int kThreads = 256;

struct Axis {
    int32_t n;         // original extent
    int32_t lo_w;      // effective ghost width on the low side (0 when the type pads nothing)
    int32_t hi_w;
    int32_t lo_t, hi_t;
    float lo_v, hi_v;
    int32_t padded;    // n + lo_w + hi_w
};

// Resolves index j of the low-padded axis (extent n + lo_w).  Returns the source index, or -1
// with *cv set when the cell is a Dirichlet constant.
__host__ __device__ __forceinline__ int resolve_low(const Axis& a, int j, float* cv) {
    if (j >= a.lo_w) return j - a.lo_w;
    switch (a.lo_t) {
        case CPPRE_BC_DIRICHLET: *cv = a.lo_v; return -1;
        case CPPRE_BC_NEUMANN: return 0;
        case CPPRE_BC_PERIODIC: return a.n - a.lo_w + j;
        default: return a.lo_w - j;   // symmetric (reflect, edge cell not repeated)
    }
}
__host__ __device__ __forceinline__ int resolve(const Axis& a, int j, float* cv) {
    const int n1 = a.n + a.lo_w;
    if (j < n1) return resolve_low(a, j, cv);
    const int k = j - n1;
    switch (a.hi_t) {
        case CPPRE_BC_DIRICHLET: *cv = a.hi_v; return -1;
        case CPPRE_BC_NEUMANN: return resolve_low(a, n1 - 1, cv);
        case CPPRE_BC_PERIODIC: return resolve_low(a, k, cv);   // first cells of the padded tensor
        default: return resolve_low(a, n1 - 2 - k, cv);
    }
}

struct PadParams {
    Axis rows, cols;
    int64_t n_img;
};

__global__ void __launch_bounds__(kThreads)
bc_pad_kernel(const float* __restrict__ in, float* __restrict__ out, PadParams p) {
    const int64_t per = (int64_t)p.rows.padded * p.cols.padded;
    const int64_t total = per * p.n_img;
    for (int64_t idx = (int64_t)blockIdx.x * kThreads + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * kThreads) {
        const int64_t img = idx / per;
        const int r = (int)((idx - img * per) / p.cols.padded);
        const int c = (int)(idx - img * per - (int64_t)r * p.cols.padded);
        float cv = 0.0f;
        // rows are padded last, so a Dirichlet row overrides whatever the columns hold
        const int sr = resolve(p.rows, r, &cv);
        float v = cv;
        if (sr >= 0) {
            const int sc = resolve(p.cols, c, &cv);
            v = sc >= 0 ? __ldg(in + (img * p.rows.n + sr) * p.cols.n + sc) : cv;
        }
        out[idx] = v;
    }
}

constexpr int kMaxK = 7;
struct ConvParams {
    Axis rows, cols;
    int64_t n_img;
    int32_t kh, kw, oh, ow;
    float taps[kMaxK * kMaxK];
};

__global__ void __launch_bounds__(kThreads)
bc_conv2d_kernel(const float* __restrict__ in, float* __restrict__ out, const __grid_constant__ ConvParams p) {
    const int64_t per = (int64_t)p.oh * p.ow;
    const int64_t total = per * p.n_img;
    for (int64_t idx = (int64_t)blockIdx.x * kThreads + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * kThreads) {
        const int64_t img = idx / per;
        const int r = (int)((idx - img * per) / p.ow);
        const int c = (int)(idx - img * per - (int64_t)r * p.ow);
        const float* src = in + img * p.rows.n * p.cols.n;
        float acc = 0.0f;
        for (int a = 0; a < p.kh; ++a) {
            float rcv = 0.0f;
            const int sr = resolve(p.rows, r + a, &rcv);
            for (int b = 0; b < p.kw; ++b) {
                const float w = p.taps[a * p.kw + b];
                if (w == 0.0f) continue;
                float v = rcv;
                if (sr >= 0) {
                    float ccv = 0.0f;
                    const int sc = resolve(p.cols, c + b, &ccv);
                    v = sc >= 0 ? __ldg(src + (int64_t)sr * p.cols.n + sc) : ccv;
                }
                acc = fmaf(w, v, acc);
            }
        }
        out[idx] = acc;
    }
}

int make_axis(const cppre_bc_t* bc, int lo, int hi, int64_t n, Axis* a) {
    a->n = (int32_t)n;
    a->lo_t = bc->type[lo];
    a->hi_t = bc->type[hi];
    a->lo_v = bc->value[lo];
    a->hi_v = bc->value[hi];
    for (int s : {lo, hi}) {
        CPPRE_REQUIRE(bc->type[s] >= CPPRE_BC_DIRICHLET && bc->type[s] <= CPPRE_BC_NONE, CPPRE_ERR_BAD_ARG,
                      "unknown boundary type %d", bc->type[s]);
        CPPRE_REQUIRE(bc->width[s] >= 0 && bc->width[s] <= 3, CPPRE_ERR_BAD_ARG, "ghost width %d out of range",
                      bc->width[s]);
    }
    a->lo_w = bc->type[lo] == CPPRE_BC_NONE ? 0 : bc->width[lo];
    a->hi_w = bc->type[hi] == CPPRE_BC_NONE ? 0 : bc->width[hi];
    const int64_t n1 = n + a->lo_w;
    // the same preconditions torch enforces: reflect needs pad < extent, periodic slices need
    // pad <= extent
    CPPRE_REQUIRE(!(a->lo_t == CPPRE_BC_SYMMETRIC && a->lo_w >= n && a->lo_w > 0), CPPRE_ERR_BAD_ARG,
                  "reflect padding %d must be smaller than the extent %lld", a->lo_w, (long long)n);
    CPPRE_REQUIRE(!(a->hi_t == CPPRE_BC_SYMMETRIC && a->hi_w >= n1 && a->hi_w > 0), CPPRE_ERR_BAD_ARG,
                  "reflect padding %d must be smaller than the extent %lld", a->hi_w, (long long)n1);
    CPPRE_REQUIRE(!(a->lo_t == CPPRE_BC_PERIODIC && a->lo_w > n), CPPRE_ERR_BAD_ARG,
                  "periodic padding %d exceeds the extent %lld", a->lo_w, (long long)n);
    CPPRE_REQUIRE(!(a->hi_t == CPPRE_BC_PERIODIC && a->hi_w > n1), CPPRE_ERR_BAD_ARG,
                  "periodic padding %d exceeds the extent %lld", a->hi_w, (long long)n1);
    CPPRE_REQUIRE(n >= 1 || (a->lo_t != CPPRE_BC_NEUMANN && a->hi_t != CPPRE_BC_NEUMANN) , CPPRE_ERR_BAD_ARG,
                  "replicate padding of an empty axis");
    a->padded = (int32_t)(n1 + a->hi_w);
    return CPPRE_OK;
}

int grid_for(int64_t total) {
    int64_t b = (total + kThreads - 1) / kThreads;
    if (b > 148 * 32) b = 148 * 32;
    return (int)(b < 1 ? 1 : b);
}

// Edge-value residual of a periodic boundary: out[img, j] = (u[img, wall line, j] - u[img, opposite line, j]) * scale,
// subtraction and scaling rounded separately like the reference's `(u[...,-1] - u[...,0]) * dx`
// [ref: Joint/NS_Residuals_CP.py:483-494].
__global__ void __launch_bounds__(kThreads)
edge_difference_kernel(const float* __restrict__ u, float* __restrict__ out, int64_t total, int64_t len, int64_t s_img,
                       int64_t s_along, int64_t off_a, int64_t off_b, float scale) {
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
        const int64_t img = i / len, j = i - img * len;
        const float* base = u + img * s_img + j * s_along;
        out[i] = __fmul_rn(__fsub_rn(__ldg(base + off_a), __ldg(base + off_b)), scale);
    }
}

}  // namespace
}  // namespace cppre

using namespace cppre;

extern "C" {

int cppre_bc_padded_shape(const cppre_bc_t* bc, int64_t h, int64_t w, int64_t* hp, int64_t* wp) {
    CPPRE_REQUIRE(bc && hp && wp && h >= 0 && w >= 0, CPPRE_ERR_BAD_ARG, "bad argument");
    Axis rows, cols;
    int rc = make_axis(bc, 0, 1, w, &cols);
    if (rc) return rc;
    rc = make_axis(bc, 2, 3, h, &rows);
    if (rc) return rc;
    *hp = rows.padded;
    *wp = cols.padded;
    return CPPRE_OK;
}

int cppre_bc_index_map(const cppre_bc_t* bc, int64_t h, int64_t w, int32_t* h_rows, float* h_row_vals,
                       int32_t* h_cols, float* h_col_vals) {
    CPPRE_REQUIRE(bc && h_rows && h_cols && h_row_vals && h_col_vals && h >= 0 && w >= 0, CPPRE_ERR_BAD_ARG,
                  "bad argument");
    Axis rows, cols;
    int rc = make_axis(bc, 0, 1, w, &cols);
    if (rc) return rc;
    rc = make_axis(bc, 2, 3, h, &rows);
    if (rc) return rc;
    for (int r = 0; r < rows.padded; ++r) {
        h_row_vals[r] = 0.0f;
        h_rows[r] = resolve(rows, r, &h_row_vals[r]);
    }
    for (int c = 0; c < cols.padded; ++c) {
        h_col_vals[c] = 0.0f;
        h_cols[c] = resolve(cols, c, &h_col_vals[c]);
    }
    return CPPRE_OK;
}

int cppre_bc_pad(const float* d_in, int64_t n_img, int64_t h, int64_t w, const cppre_bc_t* bc, float* d_out,
                 void* stream) {
    CPPRE_REQUIRE(bc && n_img >= 0 && h >= 0 && w >= 0, CPPRE_ERR_BAD_ARG, "bad argument");
    PadParams p;
    int rc = make_axis(bc, 0, 1, w, &p.cols);
    if (rc) return rc;
    rc = make_axis(bc, 2, 3, h, &p.rows);
    if (rc) return rc;
    p.n_img = n_img;
    const int64_t total = n_img * p.rows.padded * p.cols.padded;
    if (total == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_out && (d_in || h * w == 0), CPPRE_ERR_BAD_ARG, "null pointer");
    bc_pad_kernel<<<grid_for(total), kThreads, 0, (cudaStream_t)stream>>>(d_in, d_out, p);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_bc_conv2d(const float* d_in, int64_t n_img, int64_t h, int64_t w, const cppre_bc_t* bc,
                    const float* h_taps, int32_t kh, int32_t kw, float* d_out, void* stream) {
    CPPRE_REQUIRE(bc && h_taps && n_img >= 0 && h >= 0 && w >= 0, CPPRE_ERR_BAD_ARG, "bad argument");
    CPPRE_REQUIRE(kh >= 1 && kw >= 1, CPPRE_ERR_BAD_ARG, "bad kernel size");
    CPPRE_REQUIRE(kh <= kMaxK && kw <= kMaxK, CPPRE_ERR_UNSUPPORTED, "kernel larger than 7 is not supported");
    ConvParams p;
    int rc = make_axis(bc, 0, 1, w, &p.cols);
    if (rc) return rc;
    rc = make_axis(bc, 2, 3, h, &p.rows);
    if (rc) return rc;
    p.n_img = n_img;
    p.kh = kh;
    p.kw = kw;
    p.oh = p.rows.padded - kh + 1;
    p.ow = p.cols.padded - kw + 1;
    CPPRE_REQUIRE(p.oh >= 1 && p.ow >= 1, CPPRE_ERR_BAD_ARG, "kernel (%d,%d) larger than the padded image (%d,%d)",
                  kh, kw, p.rows.padded, p.cols.padded);
    for (int i = 0; i < kh * kw; ++i) p.taps[i] = h_taps[i];
    const int64_t total = n_img * p.oh * p.ow;
    if (total == 0) return CPPRE_OK;
    CPPRE_REQUIRE(d_in && d_out, CPPRE_ERR_BAD_ARG, "null pointer");
    bc_conv2d_kernel<<<grid_for(total), kThreads, 0, (cudaStream_t)stream>>>(d_in, d_out, p);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

int cppre_edge_difference(const float* d_u, int64_t n_img, int64_t h, int64_t w, int64_t s_img, int64_t s_row,
                          int64_t s_col, int32_t wall, float scale, float* d_out, void* stream) {
    CPPRE_REQUIRE(n_img >= 0 && h >= 0 && w >= 0, CPPRE_ERR_BAD_ARG, "bad argument");
    CPPRE_REQUIRE(wall >= CPPRE_WALL_TOP && wall <= CPPRE_WALL_RIGHT, CPPRE_ERR_BAD_ARG, "unknown wall %d", wall);
    const bool rows = wall == CPPRE_WALL_TOP || wall == CPPRE_WALL_BOTTOM;       // result runs along the columns
    const int64_t len = rows ? w : h;
    const int64_t total = n_img * len;
    if (total == 0) return CPPRE_OK;
    CPPRE_REQUIRE(h >= 1 && w >= 1, CPPRE_ERR_BAD_ARG, "an edge difference needs at least one row and one column");
    CPPRE_REQUIRE(d_u && d_out, CPPRE_ERR_BAD_ARG, "null pointer");
    // first / last line of the axis the wall closes; minuend = the named wall's line, as the reference writes it
    const int64_t first = 0, last = (rows ? h : w) - 1;
    const int64_t line = rows ? s_row : s_col, along = rows ? s_col : s_row;
    const bool named_is_first = wall == CPPRE_WALL_TOP || wall == CPPRE_WALL_LEFT;
    const int64_t off_a = (named_is_first ? first : last) * line, off_b = (named_is_first ? last : first) * line;
    edge_difference_kernel<<<grid_for(total), kThreads, 0, (cudaStream_t)stream>>>(d_u, d_out, total, len, s_img, along,
                                                                                   off_a, off_b, scale);
    CPPRE_CHECK_LAUNCH();
    return CPPRE_OK;
}

}  // extern "C"
