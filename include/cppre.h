/*
 * cppre.h -- C ABI of libcppre.so: the B200 (sm_100a) implementation of CP-PRE's hot path.
 *
 * The reference (gitvicky/CP-PRE @ 5b70067a) is pure Python: it has no FFI.  Each entry point
 * below therefore replaces a *library call site* of the reference; the citation after "replaces"
 * is the reference file:line a maintainer would re-point at this library (the ctypes stubs are
 * shown in INTEGRATION.md and implemented in cp_pre_b200/_lib.py).
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer named d_* is a DEVICE pointer (fp32 unless
 *    stated), every h_* a HOST pointer; strides are in ELEMENTS, int64;
 *  - work is enqueued on `stream` (a cudaStream_t passed as void*); no hidden synchronisation
 *    except where a host scalar is returned, which the doc comment states;
 *  - the library never allocates memory the caller does not own; scratch is passed in;
 *  - every function returns 0 on success or a CPPRE_ERR_* code; cppre_last_error() gives the
 *    thread-local message; no C++ exception crosses the ABI;
 *  - there is no CPU fallback: without a CUDA device every compute entry point fails with
 *    CPPRE_ERR_CUDA.
 */
#ifndef CPPRE_H_
#define CPPRE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CPPRE_VERSION 100 /* 0.1.0 */

#define CPPRE_MAX_FIELDS 6
#define CPPRE_MAX_SLOTS 6
#define CPPRE_MAX_SCALARS 12
#define CPPRE_MAX_TAPS 512 /* non-zero taps summed over all slots of one program */

enum {
    CPPRE_OK = 0,
    CPPRE_ERR_BAD_ARG = 1,     /* null pointer, negative size, inconsistent shapes */
    CPPRE_ERR_UNSUPPORTED = 2, /* valid request this build has no kernel for (e.g. kernel > 7) */
    CPPRE_ERR_CUDA = 3         /* a CUDA runtime/driver call failed; message has the string */
};

/* Residual programs: which combination of fields and derivative slots is evaluated per cell.
 * Slots are dense coefficient tensors taken from the caller's live ConvOperator.kernel
 * objects -- the library executes whatever tensor is there, it never
 * rebuilds them from operator names. */
enum {
    /* r = K0 (x) f0.                                   replaces F.conv3d  Utils/ConvOps_2d.py:149,
     * F.conv2d Utils/ConvOps_1d.py:150, F.conv1d Utils/ConvOps_0d.py:103, F.conv2d (valid)
     * Utils/ConvOps_Spatial.py:135 */
    CPPRE_PROG_LINEAR = 0,
    /* fields {u}; slots {D_t, D_x, D_xx}; scalars {dx, dt, nu*2*dt/dx}
     * r = s0*D_t(u) + s1*u*D_x(u) - s2*D_xx(u)          Joint/Burgers_Residuals_CP.py:182-187 */
    CPPRE_PROG_BURGERS = 1,
    /* fields {u,v}; slots {D_x, D_y}; scalars {dx/dy}
     * r = D_x(u) + s0*D_y(v)                            Joint/NS_Residuals_CP.py:222-228 */
    CPPRE_PROG_NS_CONTINUITY = 2,
    /* fields {u,v,p}; slots {D_t, D_x, D_y, D_xx_yy}; scalars {dx*dy, dt*dy, dt*dx, nu*dt}
     * r = res_x + res_y                                 Joint/NS_Residuals_CP.py:231-240,
     *                                                   Other_UQ/Evaluation/PRE_estimations.py:42-52 */
    CPPRE_PROG_NS_MOMENTUM = 3,
    /* fields {rho,u,v,p,Bx,By}; slots {D_t, D_x, D_y}; scalars {gamma}
     *                                                   Joint/MHD_Residuals_CP.py:225-278,
     *                                                   Other_UQ/Evaluation/PRE_estimations.py:70-80 */
    CPPRE_PROG_MHD_CONTINUITY = 4,
    CPPRE_PROG_MHD_MOMENTUM = 5,
    CPPRE_PROG_MHD_ENERGY = 6,
    CPPRE_PROG_MHD_INDUCTION = 7,
    CPPRE_PROG_MHD_GAUSS = 8,
    /* JOREK reduced MHD with a radius-dependent coefficient: fields {rho, phi, T, R}, where R is the
     * major-radius grid as a FIELD (pass the stride-0 broadcast view of the 1-D grid along N0, the
     * axis the reference's `R * (...)` broadcasts over); slots {D_t, D_R, D_Z, D_RR, D_ZZ};
     * scalars {D} (continuity) / {K, gamma} (temperature)   Marginal/JOREK_residuals_CP.py:211-242.
     * Served by the generic kernel (a stride-0 field cannot be a TMA tensor map). */
    CPPRE_PROG_JOREK_CONTINUITY = 9,
    CPPRE_PROG_JOREK_TEMPERATURE = 10,
    CPPRE_PROG_COUNT = 11
};

/* One input field: a strided fp32 view of logical shape [B, N2, N1, N0].
 * 2-D+time data [BS,Nt,Nx,Ny] maps to (B,N2,N1,N0) = (BS,Nt,Nx,Ny); 1-D+time [BS,Nt,Nx] to
 * (BS,Nt,1,Nx); ODE data [BS,Nt] to (BS,1,1,Nt); space-only [BS,1,Nx,Ny] to (BS,1,Nx,Ny).
 * Any strides are accepted (inputs are usually permuted views). */
typedef struct cppre_field {
    const float* d_data;
    int64_t stride[4]; /* elements, order [B, N2, N1, N0] */
    /* CPPRE_FLAG_DECODE: the program sees scale * x + offset instead of the stored x (the surrogate's output
     * normaliser, out_normalizer.decode, folded into the load; ignored without the flag). */
    float scale, offset;
} cppre_field_t;

typedef struct cppre_program {
    int32_t kind;     /* CPPRE_PROG_* */
    int32_t n_fields; /* must match the program kind */
    int32_t n_slots;
    int32_t ksize[CPPRE_MAX_SLOTS][3]; /* dense kernel dims (k2,k1,k0), each in 1..7 */
    int32_t pad[CPPRE_MAX_SLOTS][3];   /* zero padding per axis: k/2 = "same", 0 = valid */
    const float* h_taps[CPPRE_MAX_SLOTS]; /* HOST, dense row-major [k2][k1][k0]; cross-correlation */
    float scalars[CPPRE_MAX_SCALARS];
} cppre_program_t;

/* What happens to the residual r[b,cell] once computed -- fused into the stencil kernel so no
 * intermediate tensor makes an HBM round trip.
 *   v = r - ref (if d_ref);  v = |v| (if ABS);  v = v * weight[cell] (if d_weight)
 *   STORE: out[b,cell] = v                      replaces the residual tensors + np.abs
 *                                               Marginal/NS_Residuals_CP.py:289,327
 *   SCORE: scores[b] = reduce_cells(v)          replaces ncf_metric_joint (max of |a-b|/mod)
 *                                               Joint/NS_Residuals_CP.py:306 and the residual
 *                                               tensor it consumed
 *   COVER: counts[b] = #{cell : c - q*h <= r <= c + q*h}  (uint32, written through d_scores)
 *          with c = ref[b,cell] (0 if no d_ref), h = weight[cell] (1 if no d_weight), q = band_q;
 *          r is the raw residual (FLAG_ABS ignored); lo/hi are evaluated with separately
 *          rounded fp32 multiply / subtract / add, exactly as NumPy evaluates
 *          [pred_res - qhat*modulation, pred_res + qhat*modulation].  The prediction sets and the
 *          residual tensor of the validation samples are never materialised.
 *                                               replaces residual + sets + emp_cov / emp_cov_joint /
 *                                               filter_sims_joint  Joint/NS_Residuals_CP.py:312,336,
 *                                               Marginal/NS_Residuals_CP.py:311-312,336-337
 * INTERIOR_d excludes index 0 and N_d-1 of axis d from SCORE / COVER reductions (the reference's
 * [...,1:-1,1:-1,1:-1] crop, Joint/Wave_Residuals_CP.py:184).  */
enum { CPPRE_EPI_STORE = 0, CPPRE_EPI_SCORE = 1, CPPRE_EPI_COVER = 2 };
enum { CPPRE_REDUCE_MAX = 0, CPPRE_REDUCE_SUM = 1 };
enum {
    CPPRE_FLAG_ABS = 1,
    CPPRE_FLAG_INTERIOR_0 = 2, /* axis N0 */
    CPPRE_FLAG_INTERIOR_1 = 4, /* axis N1 */
    CPPRE_FLAG_INTERIOR_2 = 8, /* axis N2 */
    /* Normaliser-decode fold: every field is decoded as scale_f * x + offset_f before the
     * program reads it -- `pred = out_normalizer.decode(pred)` directly upstream of the residual call,
     * Joint/Wave_Residuals_CP.py:213-217, Tests/test_ns.py:134 -- without a pass over the tensor: a derivative
     * slot of the decoded field is scale_f * D(x) + offset_f * sum(taps) wherever the whole stencil lies inside
     * the field, and raw field factors are decoded at the output cell.  Results are those of the decoded field
     * on the INTERIOR (the cells the reference keeps, [...,1:-1,1:-1,1:-1]); cells whose stencil reaches into
     * the zero padding see the padding of the STORED field (zeros before the decode) and are unspecified:
     * crop them (STORE) or set the INTERIOR flags (SCORE / COVER).  fp32 fields only.
     * Because that interior is ONE cell deep, the fold is offered only while no non-zero tap lies more than one
     * cell from the centre (the 3-point stencils every reference script builds).  A wider stencil (Taylor order
     * 4 / 6) would reach the padding from the first kept layer, where the reference pads the DECODED field and
     * the fold would be off by offset_f * (taps outside the field): such a request is refused with
     * CPPRE_ERR_UNSUPPORTED -- decode the fields explicitly and call without the flag. */
    CPPRE_FLAG_DECODE = 32,
    /* bf16-input mode: every cppre_field_t::d_data points to bfloat16 elements (strides in
     * elements), staged as bf16 and widened to fp32 in the kernel; arithmetic, reference, weight
     * and every output stay fp32.  Halves the field bytes read (NS momentum STORE: 16 -> 10 B per
     * grid point).  Served by the TMA star kernel only: rows must be 16-byte aligned with
     * N0 % 8 == 0 and strides % 8 == 0, otherwise CPPRE_ERR_UNSUPPORTED.  Parity is stated
     * against the fp32 oracle evaluated on the bf16-rounded input. */
    CPPRE_FLAG_IN_BF16 = 16
};
/* Which kernel executes the program.  AUTO picks the TMA star kernel when the request
 * qualifies (3-point star taps, 16-byte aligned rows), the t-innermost star kernel when every field
 * (and the output / reference / weight) is a [B,Nt,Nx,Ny] view of [B,Nx,Ny,Nt] memory -- the layout FNO
 * surrogates emit, pred.permute(0,1,4,2,3)  Joint/Wave_Residuals_CP.py:216 --, the flat-line star kernel for
 * contiguous star requests whose rows the TMA unit cannot stage (N0 % 4 != 0 such as Nx = 197 of
 * Marginal/Advection_Residuals_CP.py:87,108, odd grids, unaligned bases; N0 <= 1024 / 512 / 256 for N0 % 4 == 0 /
 * even / odd) and the generic kernel otherwise (Taylor-4/6 taps, valid convolution, arbitrary strides). */
enum { CPPRE_PATH_AUTO = 0, CPPRE_PATH_GENERIC = 1, CPPRE_PATH_TMA = 2, CPPRE_PATH_TIN = 3, CPPRE_PATH_FLAT = 4 };

typedef struct cppre_epilogue {
    int32_t mode;   /* CPPRE_EPI_* */
    int32_t flags;  /* CPPRE_FLAG_* */
    int32_t reduce; /* CPPRE_REDUCE_* (SCORE only) */
    float band_q;   /* COVER: q */
    const float* d_ref;    /* optional [B, O2, O1, O0] with ref_stride, or NULL */
    int64_t ref_stride[4];
    const float* d_weight; /* optional contiguous [O2, O1, O0], or NULL */
    float* d_out;          /* STORE: [B, O2, O1, O0] with out_stride */
    int64_t out_stride[4];
    float* d_scores;       /* SCORE: [B] fp32 / COVER: [B] uint32; contiguous, overwritten */
    /* COVER: optional DEVICE scalar that replaces band_q.  The Joint flow selects qhat on the device
     * (cppre_radix_select / cppre_select_ranks) and hands the pointer straight to the coverage launch:
     * no host round trip between calibrate() and emp_cov_joint()  Joint/MHD_Residuals_CP.py:399-401 */
    const float* d_band_q;
    /* Element strides of d_weight, order [O2, O1, O0]; all zero = contiguous.  The only other order the
     * kernels take is the t-innermost one, (1, O0*O2, O2) -- the weight laid out like the [Nx][Ny][Nt]
     * fields it goes with (CPPRE_PATH_TIN); anything else is CPPRE_ERR_UNSUPPORTED. */
    int64_t weight_stride[3];
} cppre_epilogue_t;

/* ---- library state -------------------------------------------------------------------- */
int cppre_version(void);
const char* cppre_last_error(void);
/* Number of kernels this library has launched in the calling process (all threads). */
uint64_t cppre_launch_count(void);
/* Name of the kernel path the last cppre_residual call on this thread used: "tma" | "tin" | "flat" | "generic" | "empty". */
const char* cppre_last_path(void);

/* ---- K1: fused residual program ------------------------------------------------------- */
/* Output extent per axis: O_d = N_d + 2*pad_d - k_d + 1 (must agree across slots). */
int cppre_residual(const cppre_program_t* prog, const int64_t shape[4],
                   const cppre_field_t* fields, const cppre_epilogue_t* epi,
                   int path, void* stream);
/* Returns 1 if (prog, shape, fields, epi) qualifies for the t-innermost star kernel, else 0. */
int cppre_residual_tin_eligible(const cppre_program_t* prog, const int64_t shape[4],
                                const cppre_field_t* fields, const cppre_epilogue_t* epi);
/* Returns 1 if (prog, shape, fields, epi) qualifies for the TMA star kernel, else 0. */
int cppre_residual_tma_eligible(const cppre_program_t* prog, const int64_t shape[4],
                                const cppre_field_t* fields, const cppre_epilogue_t* epi);

/* ---- K2: per-cell statistics over the sample axis ------------------------------------- */
/* mean[c], m2[c] (sum of squared deviations) of x[i,c] = a[i,c] - b[i,c] over i < n.
 * a, b: fp32 [n, cells] with row strides; b may be NULL.  Accumulated in fp64.
 * replaces modulation_func = np.std(a-b, axis=0)   Tests/test_advection_inv_sampling_marginal.py:428,
 * Joint/NS_Residuals_CP.py:305 */
int cppre_cell_moments(const float* d_a, int64_t a_row_stride, const float* d_b,
                       int64_t b_row_stride, int64_t n, int64_t cells, double* d_mean,
                       double* d_m2, void* stream);
/* std[c] = sqrt(m2[c]/count) + add_eps (fp32); inv[c] = 1/std[c] if d_inv != NULL. */
int cppre_cell_std(const double* d_m2, double count, float add_eps, int64_t cells,
                   float* d_std, float* d_inv, void* stream);
/* Both in ONE launch (single-GPU modulation_func: nothing but std / 1/std is written):
 * std[c] = population std over i < n of a[i,c]-b[i,c], + add_eps; inv[c] = 1/std[c] if d_inv != NULL. */
int cppre_cell_modulation(const float* d_a, int64_t a_row_stride, const float* d_b, int64_t b_row_stride,
                          int64_t n, int64_t cells, float add_eps, float* d_std, float* d_inv, void* stream);
/* scores[i] = max_c |a[i,c]-b[i,c]| / mod[c]  (IEEE division, NaN-propagating max)
 * replaces ncf_metric_joint on stored residuals   Joint/NS_Residuals_CP.py:306 */
int cppre_joint_score(const float* d_a, int64_t a_row_stride, const float* d_b,
                      int64_t b_row_stride, const float* d_mod, int64_t n, int64_t cells,
                      float* d_scores, void* stream);

/* ---- K3: order statistics along the sample axis --------------------------------------- */
/* out[r, c] = the h_ranks[r]-th smallest (0-based) of scores[0..n-1, c]; NaN if the column
 * holds a NaN (np.quantile behaviour).  scores fp32 [n, cells] with row stride.
 * replaces np.quantile(..., axis=0, method='higher') inside calibrate()
 *          (Neural_PDE/UQ/inductive_cp.py, un-vendored; call sites Marginal/NS_Residuals_CP.py:311)
 * d_work: scratch of cppre_select_workspace_bytes(n, cells) bytes (may be NULL if that is 0). */
size_t cppre_select_workspace_bytes(int64_t n, int64_t cells);
int cppre_select_ranks(const float* d_scores, int64_t row_stride, int64_t n, int64_t cells,
                       const int32_t* h_ranks, int32_t n_ranks, float* d_out, void* d_work,
                       void* stream);
/* One 8-bit radix pass over a flat score array, for a quantile that spans several GPUs:
 * hist[d] += #{ i : (key(x_i) & mask) == prefix and ((key(x_i) >> shift) & 255) == d },
 * key = order-preserving uint32 image of the float.  d_hist: 256 x uint32, NOT cleared here.
 * The caller all-reduces d_hist over ranks (NCCL) between passes. */
int cppre_radix_hist(const float* d_x, int64_t n, uint32_t prefix, uint32_t mask, int32_t shift,
                     uint32_t* d_hist, void* stream);

/* The same radix select with its state on the DEVICE, so the four passes run back to back on the
 * stream and the Joint quantile never crosses to the host (the north_star's "single radix-select
 * over the per-sample scores" + "all-reduce radix histograms for the global quantile"):
 *     cppre_radix_select_begin(state, hist, k)                      state = {0, 0, k, 0}, hist = 0
 *     for pass in 0..3:  cppre_radix_select_hist(x, n, state, pass, hist)
 *                        [all-reduce SUM of the 257 words of hist over ranks, multi-GPU only]
 *                        cppre_radix_select_pick(hist, state, pass, out)
 * d_state: 4 x uint32 {key prefix, key mask, remaining rank, flags}; d_hist: 257 x uint32 (256 digit
 * bins + the NaN count of pass 0).  pick narrows the state, clears hist for the next pass and, on
 * pass 3, writes *d_out = the k-th smallest (0-based) of all scores, or NaN when some score is a
 * NaN (np.quantile) or k is out of range (flags bit 1).  d_out can be passed on as
 * cppre_epilogue_t::d_band_q.          replaces np.quantile inside calibrate() for 1-D scores
 *                                      Joint/MHD_Residuals_CP.py:399, Joint/NS_Residuals_CP.py:311 */
int cppre_radix_select_begin(uint32_t* d_state, uint32_t* d_hist, int64_t k, void* stream);
/* All four passes in ONE launch (one CTA; scores on one GPU, a few hundred thousand at most):
 * *d_out = the k-th smallest (0-based) of d_x[0..n), NaN rules as above. */
int cppre_radix_select(const float* d_x, int64_t n, int64_t k, float* d_out, void* stream);
int cppre_radix_select_hist(const float* d_x, int64_t n, const uint32_t* d_state, int32_t pass,
                            uint32_t* d_hist, void* stream);
int cppre_radix_select_pick(uint32_t* d_hist, uint32_t* d_state, int32_t pass, float* d_out, void* stream);

/* ---- K3d for callers that have no device allocator ---------------------------------------
 * No counterpart in the reference (it never manages device memory itself).
 *
 * The begin/hist/pick calls above need 261 words of device scratch (state + counters) per rank,
 * and somebody has to own them.  Some hosts of this library cannot provide that: a solver or
 * serving worker whose framework handed every byte of HBM to an arena the plugin call site has no
 * handle to, a coupling layer written in a language that only passes the field buffers through,
 * or a section of a time loop where the device-wide synchronisation hidden inside
 * cudaMalloc/cudaFree is not acceptable.  For those callers this entry point does the whole
 * device-state radix select -- begin, then four (histogram, pick) passes, for each requested rank
 * -- with the state and counters in a fixed array that is part of the library image itself.
 *
 *   d_out[r] = the h_ranks[r]-th smallest (0-based) of d_x[0..n),  r < n_ranks;
 *              NaN when some score is a NaN or the rank is >= n (the rules of
 *              cppre_radix_select, and bit-equal to it).
 *
 * d_x (n floats) and d_out (n_ranks floats) are the caller's; h_ranks is a host array (the alpha
 * grid of Marginal/NS_Residuals_CP.py:331-337 applied to Joint scores is the typical use).
 *
 * Limits and rules
 *  - n_ranks <= CPPRE_NOALLOC_MAX_RANKS and n < 2^32 (the counters are 32-bit); anything larger
 *    is refused with CPPRE_ERR_UNSUPPORTED and a message, nothing is launched.
 *  - The array is CPPRE_NOALLOC_SCRATCH_BYTES of device memory on every GPU that loads the
 *    library, held until the process exits whether or not this entry point is ever called.
 *  - The FIRST call on each device looks the array's device address up (cudaGetSymbolAddress,
 *    which may load the module); make that first call with the device current and before any
 *    stream capture begins.  Later calls only enqueue kernels on `stream`.
 *  - It calls no cudaMalloc / cudaFree / allocating library routine, on any call.
 *  - All calls on one device share the one array.  Each call re-initialises the part it uses,
 *    so it does not matter what the previous call left behind, but two calls that can be in
 *    flight on the GPU at the same time (different streams, different host threads) race on it:
 *    order them on one stream or with events. */
#define CPPRE_NOALLOC_MAX_RANKS 16
#define CPPRE_NOALLOC_SCRATCH_BYTES (CPPRE_NOALLOC_MAX_RANKS * 1536)
int cppre_radix_select_ranks_noalloc(const float* d_x, int64_t n, const int64_t* h_ranks, int32_t n_ranks,
                                     float* d_out, void* stream);

/* ---- K3w: multi-GPU marginal quantile (batch axis sharded over ranks) --------------------
 * No counterpart in the reference (it is single-process); these three calls replace the naive
 * all-to-all of the whole score tensor in front of np.quantile (Marginal/NS_Residuals_CP.py:311):
 *   1. every rank selects two local order statistics per cell (cppre_select_ranks), converts
 *      them to order-preserving int32 keys (cppre_ordered_keys) and all-reduces MIN / MAX: the
 *      global quantile of every cell lies in the resulting [lo, hi];
 *   2. cppre_window_compact streams the local scores once: cnt[c] = #{x > hi} (+ 1<<24 if the
 *      cell holds a NaN on some rank, i.e. hi is a NaN key) and the local samples inside [lo, hi]
 *      go into `slots` candidate rows (sample order, -inf padded), laid out
 *      [cells/slab][slots][slab] = the send buffer of an all-to-all that
 *      gives rank r the candidates of cell slab r; *d_overflow += #cells with more than `slots`
 *      samples inside, whose indices go to d_overflow_cells[old counter value] while that is
 *      below overflow_cap (the caller redoes those few cells from their full columns).
 *      FUSED EXCHANGE: with h_peer_win (HOST array of n_peers device pointers, each rank's
 *      [n_peers][slots][slab] receive buffer mapped into this process, e.g. from a CUDA
 *      symmetric-memory allocation) the block of slab r is stored directly into
 *      h_peer_win[r][my_rank] over NVLink and d_win is not used: the all-to-all is part of the
 *      kernel.  The all-reduce of cnt that follows orders those stores before the owner reads;
 *   3. after all-reducing cnt (SUM), the owner of a slab picks, per cell, the
 *      (kt - (cnt & 0xffffff))-th largest (1-based) of the gathered candidate rows
 *      (cppre_select_from_top); NaN where cnt >> 24 != 0.  network = 0 sorts all rows (<= 2048);
 *      network = 32..1024 (a power of two) first drops the -inf padding and packs the real
 *      candidates of each cell, then sorts `network` keys -- far fewer than rows, since most of
 *      a block is padding.  A cell whose rank falls outside [1, rows], or that holds more than
 *      `network` real candidates, is counted in *d_overflow and listed (cell_offset + index) in
 *      d_overflow_cells like in step 2.
 * A non-zero overflow counter means the caller must redo the listed cells (or, past the list
 * capacity, everything) on the exact slow path. */
int cppre_ordered_keys(const float* d_x, int64_t n, int32_t* d_out, void* stream);
int cppre_window_compact(const float* d_scores, int64_t row_stride, int64_t n, int64_t cells,
                         const int32_t* d_lo, const int32_t* d_hi, int32_t slots, int64_t slab,
                         float* d_win, int32_t* d_cnt, uint32_t* d_overflow, int32_t* d_overflow_cells,
                         int32_t overflow_cap, float* const* h_peer_win, int32_t n_peers, int32_t my_rank,
                         void* stream);
int cppre_select_from_top(const float* d_cand, int64_t row_stride, int64_t rows, int64_t cells,
                          const int32_t* d_cnt, int32_t kt, int32_t network, float* d_out,
                          uint32_t* d_overflow, int32_t* d_overflow_cells, int32_t overflow_cap,
                          int64_t cell_offset, void* stream);

/* K3x: the same protocol with EVERY exchange fused into the kernels (peer memory over NVLink /
 * NVSwitch; each h_peer_* is a HOST array of n_peers device pointers, rank r's buffer mapped into
 * this process, e.g. from a CUDA symmetric-memory allocation).  Between the three calls the
 * caller only needs a barrier (any one-word collective); no data-carrying collective is left:
 *   1. cppre_peer_bcast_keys: order-preserving uint32 key of this rank's local order statistic
 *      d_vals[c] -> h_peer_keys[r][my_rank * cells_pad + c] for every r (cells_pad % 4 == 0).
 *      Replaces cppre_ordered_keys + the MIN/MAX all-reduce: every rank then holds all n_peers
 *      keys of every cell and takes min / max itself.
 *   2. cppre_window_exchange: reads d_keys = this rank's [n_peers][cells_pad] key buffer (key_mode 0)
 *      or, key_mode 1, a [2][cells_pad] buffer holding [~min, max] of the signed keys of
 *      cppre_ordered_keys after a MAX all-reduce (26 MB through the switch instead of world x 13 MB
 *      of peer stores: the cheaper bracket exchange from 4 GPUs up), scans the
 *      local scores once and stores, for the owner r = c / slab of cell c, into
 *      h_peer_recv[r][my_rank][1 + slots][slab]: row 0 = header word
 *      (#{x > hi} in bits 0..23, #candidates in bits 24..30, bit 31 = the cell holds a NaN on
 *      some rank), rows 1.. = the candidates inside [lo, hi] in sample order.  Rows beyond a
 *      column's candidate count are NOT written (stale memory the owner never reads).
 *      Overflow list as in cppre_window_compact.
 *   3. cppre_owner_select: the owner of slab `cell_offset / slab` reads its own receive buffer
 *      d_recv [n_ranks][1 + slots][slab] (`cells` <= slab live cells), packs the candidate lists of
 *      each cell, sorts `network` keys (power of two, 32..1024) and stores the
 *      (kt - sum of #above)-th largest (1-based) into h_peer_q[r][cell_offset + c] for every r --
 *      the all-gather of qhat, fused.  NaN where a header carries the NaN bit; cells that cannot
 *      be answered (bracket missed, a list was cut, more than `network` candidates) are counted
 *      in *d_overflow, listed in d_overflow_cells and written as NaN for the caller to redo. */
/* n_words 32-bit words from d_src -> h_peer_dst[r] + dst_offset_words for every listed peer (wide stores
 * over NVLink): used to push the owner's slab of qhat to the other ranks after cppre_owner_select wrote
 * it locally (h_peer_q = {own buffer}); one pass of 128-bit stores instead of per-CTA peer stores. */
int cppre_peer_copy(const void* d_src, int64_t n_words, int64_t dst_offset_words, void* const* h_peer_dst,
                    int32_t n_peers, void* stream);
int cppre_peer_bcast_keys(const float* d_vals, int64_t cells, int64_t cells_pad, void* const* h_peer_keys,
                          int32_t n_peers, int32_t my_rank, void* stream);
int cppre_window_exchange(const float* d_scores, int64_t row_stride, int64_t n, int64_t cells,
                          const uint32_t* d_keys, int64_t cells_pad, int32_t key_mode, int32_t slots, int64_t slab,
                          void* const* h_peer_recv, int32_t n_peers, int32_t my_rank, uint32_t* d_overflow,
                          int32_t* d_overflow_cells, int32_t overflow_cap, void* stream);
int cppre_owner_select(const float* d_recv, int32_t n_ranks, int32_t slots, int64_t slab, int64_t cells,
                       int32_t kt, int32_t network, int64_t cell_offset, void* const* h_peer_q, int32_t n_peers,
                       uint32_t* d_overflow, int32_t* d_overflow_cells, int32_t overflow_cap, void* stream);

/* ---- K5: coverage --------------------------------------------------------------------- */
/* inside[i] = #{ c : lo[i,c] <= y[i,c] <= hi[i,c] }, uint32 per sample, overwritten.
 * lo / hi: row stride 0 broadcasts one [cells] row over all samples.
 * replaces emp_cov / emp_cov_joint / filter_sims_joint / filter_sims_within_bounds
 *          (Marginal/NS_Residuals_CP.py:312, Joint/Advection_Residuals_CP.py:327-329,
 *           Active_Learning/Advection_AL_Marginal.py:169-197) */
int cppre_inside_count(const float* d_y, int64_t y_row_stride, const float* d_lo,
                       int64_t lo_row_stride, const float* d_hi, int64_t hi_row_stride, int64_t n,
                       int64_t cells, uint32_t* d_inside, void* stream);
/* Same with the prediction set never materialised: lo = c - q*h, hi = c + q*h evaluated in fp32
 * exactly as NumPy would (separately rounded multiply, subtract, add); c = centre[i,c] or 0 when
 * d_centre == NULL; h = half[c] or 1 when d_half == NULL.
 * replaces [pred -/+ qhat*modulation] + emp_cov_joint   Joint/NS_Residuals_CP.py:312,336 */
int cppre_inside_count_band(const float* d_y, int64_t y_row_stride, const float* d_centre,
                            int64_t centre_row_stride, const float* d_half, float q, int64_t n,
                            int64_t cells, uint32_t* d_inside, void* stream);

/* ---- layout ingestion ------------------------------------------------------------------ */
/* d_dst[b][n][m] = d_src[b * src_batch_stride + m * N + n]: per sample a [M][N] -> [N][M] transpose
 * (N <= 256).  Brings the t-innermost tensors FNO surrogates emit ([BS,F,Nx,Ny,Nt], passed on as
 * pred.permute(0,1,4,2,3), Joint/Wave_Residuals_CP.py:216) to the [Nt][Nx*Ny] order K1 tiles; it
 * replaces the implicit .contiguous() copy F.conv3d makes of such a view (Utils/ConvOps_2d.py:149). */
int cppre_transpose_inner(const float* d_src, int64_t src_batch_stride, int64_t batch, int64_t M, int64_t N,
                          float* d_dst, void* stream);

/* ---- boundary conditions -------------------------------------------------------------- */
/* Per-side boundary codes, order {left, right, top, bottom}; left/right act on the last axis. */
enum {
    CPPRE_BC_DIRICHLET = 0,
    CPPRE_BC_NEUMANN = 1, /* also 'outflow' */
    CPPRE_BC_PERIODIC = 2,
    CPPRE_BC_SYMMETRIC = 3,
    CPPRE_BC_NONE = 4 /* 'free_slip': the reference pads nothing */
};
typedef struct cppre_bc {
    int32_t type[4];  /* CPPRE_BC_* for left, right, top, bottom */
    int32_t width[4]; /* ghost widths: kw/2, kw/2, kh/2, kh/2 */
    float value[4];   /* Dirichlet values */
} cppre_bc_t;
/* Padded extents produced by the reference's sequential left,right,top,bottom padding. */
int cppre_bc_padded_shape(const cppre_bc_t* bc, int64_t h, int64_t w, int64_t* hp, int64_t* wp);
/* Host-side view of the same index map the kernels use: h_rows[r] / h_cols[c] = source row /
 * column of padded row r / column c, or -1 where the cell is a Dirichlet constant whose value
 * is then in h_row_vals[r] / h_col_vals[c] (a row constant wins over a column constant: rows
 * are padded last).  Arrays of hp and wp entries (cppre_bc_padded_shape).  No GPU needed. */
int cppre_bc_index_map(const cppre_bc_t* bc, int64_t h, int64_t w, int32_t* h_rows, float* h_row_vals,
                       int32_t* h_cols, float* h_col_vals);
/* out[img, :, :] = pad(in[img, :, :]) for img < n_img; in [n_img,H,W], out [n_img,Hp,Wp], both
 * contiguous.   replaces BoundaryManager.pad_signal   Utils/boundary_conditions.py:80-182 */
int cppre_bc_pad(const float* d_in, int64_t n_img, int64_t h, int64_t w, const cppre_bc_t* bc,
                 float* d_out, void* stream);
/* Valid cross-correlation of the (virtual) padded image with a kh x kw kernel; no padded copy
 * is made.  out [n_img, Hp-kh+1, Wp-kw+1] contiguous.
 * replaces pad_signal + F.conv2d   Utils/boundary_conditions.py:184-211,
 *          Utils/VectorConvOps_Spatial.py:41-55 */
int cppre_bc_conv2d(const float* d_in, int64_t n_img, int64_t h, int64_t w, const cppre_bc_t* bc,
                    const float* h_taps, int32_t kh, int32_t kw, float* d_out, void* stream);

/* Edge-value residual of a periodic boundary: for every image u[img] (h rows x w columns, element strides s_img /
 * s_row / s_col -- any view), out[img, :] = (u[named wall's line] - u[opposite line]) * scale, contiguous
 * [n_img, w] for CPPRE_WALL_TOP / BOTTOM (rows 0 / h-1) and [n_img, h] for LEFT / RIGHT (columns 0 / w-1); the
 * subtraction and the scaling are rounded separately, like the reference's tensor expression.
 * replaces periodic_bc_residual(u, wall)   Joint/NS_Residuals_CP.py:483-494 (result fed to the Joint CP helpers) */
enum { CPPRE_WALL_TOP = 0, CPPRE_WALL_BOTTOM = 1, CPPRE_WALL_LEFT = 2, CPPRE_WALL_RIGHT = 3 };
int cppre_edge_difference(const float* d_u, int64_t n_img, int64_t h, int64_t w, int64_t s_img, int64_t s_row,
                          int64_t s_col, int32_t wall, float scale, float* d_out, void* stream);

/* ---- spectral path -------------------------------------------------------------------- */
enum {
    CPPRE_SPEC_CONJ_FIRST = 1, /* K <- conj(K) before anything else (correlation) */
    CPPRE_SPEC_INVERSE = 2,    /* K <- 1/(K + eps) */
    CPPRE_SPEC_CONJ_LAST = 4   /* K <- conj(K) after the inverse */
};
/* out[b, m] = sig[b, m] * g(ker[m]) on interleaved complex64; g per `flags`.
 * replaces the pointwise spectrum arithmetic of fft_conv / differentiate / integrate
 *          Utils/fft_conv_pytorch/fft_conv.py:110-115, Utils/ConvOps_2d.py:207-210,263-268 */
int cppre_spectral_pointwise(const float* d_sig, const float* d_ker, int64_t batch, int64_t m,
                             int32_t flags, float eps, float* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CPPRE_H_ */
