// Library state, argument normalisation and the cppre_residual dispatcher.
#include <cstring>

#include "common.cuh"

namespace cppre {

static thread_local char t_error[512] = "";
static thread_local const char* t_path = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_error, sizeof(t_error), fmt, ap);
    va_end(ap);
}
void set_path(const char* name) { t_path = name; }

static const int kFields[CPPRE_PROG_COUNT] = {1, 1, 2, 3, 3, 6, 6, 4, 2, 4, 4};
static const int kSlots[CPPRE_PROG_COUNT] = {1, 3, 2, 4, 3, 3, 3, 3, 3, 5, 5};
int program_fields(int kind) { return kFields[kind]; }
int program_slots(int kind) { return kSlots[kind]; }

int prepare(const cppre_program_t* prog, const int64_t shape[4], const cppre_field_t* fields,
            const cppre_epilogue_t* epi, Prepared* out) {
    CPPRE_REQUIRE(prog && shape && fields && epi && out, CPPRE_ERR_BAD_ARG, "null argument");
    CPPRE_REQUIRE(prog->kind >= 0 && prog->kind < CPPRE_PROG_COUNT, CPPRE_ERR_BAD_ARG,
                  "unknown program kind %d", prog->kind);
    CPPRE_REQUIRE(prog->n_fields == kFields[prog->kind], CPPRE_ERR_BAD_ARG,
                  "program %d takes %d fields, got %d", prog->kind, kFields[prog->kind],
                  prog->n_fields);
    CPPRE_REQUIRE(prog->n_slots == kSlots[prog->kind], CPPRE_ERR_BAD_ARG,
                  "program %d takes %d slots, got %d", prog->kind, kSlots[prog->kind],
                  prog->n_slots);
    for (int d = 0; d < 4; ++d)
        CPPRE_REQUIRE(shape[d] >= 0 && shape[d] < (int64_t(1) << 31), CPPRE_ERR_BAD_ARG,
                      "shape[%d]=%lld out of range", d, (long long)shape[d]);
    Prepared& P = *out;
    memset(&P, 0, sizeof(P));
    P.kind = prog->kind;
    P.n_fields = prog->n_fields;
    P.n_slots = prog->n_slots;
    P.g.B = (int)shape[0];
    P.g.N2 = (int)shape[1];
    P.g.N1 = (int)shape[2];
    P.g.N0 = (int)shape[3];
    const int N[3] = {P.g.N2, P.g.N1, P.g.N0};
    int O[3] = {0, 0, 0};
    int ntap = 0;
    for (int s = 0; s < prog->n_slots; ++s) {
        CPPRE_REQUIRE(prog->h_taps[s] != nullptr, CPPRE_ERR_BAD_ARG, "slot %d has no taps", s);
        P.taps.begin[s] = ntap;
        const int* k = prog->ksize[s];
        const int* p = prog->pad[s];
        for (int d = 0; d < 3; ++d) {
            CPPRE_REQUIRE(k[d] >= 1, CPPRE_ERR_BAD_ARG, "slot %d: kernel extent %d < 1", s, k[d]);
            CPPRE_REQUIRE(k[d] <= 7, CPPRE_ERR_UNSUPPORTED,
                          "slot %d: kernel extent %d > 7 is not supported", s, k[d]);
            CPPRE_REQUIRE(p[d] >= 0 && p[d] <= 6, CPPRE_ERR_BAD_ARG, "slot %d: bad padding %d", s,
                          p[d]);
            const int o = N[d] + 2 * p[d] - k[d] + 1;
            if (s == 0) O[d] = o;
            CPPRE_REQUIRE(o == O[d], CPPRE_ERR_BAD_ARG,
                          "slot %d: output extent %d on axis %d differs from slot 0 (%d)", s, o, d,
                          O[d]);
        }
        for (int a = 0; a < k[0]; ++a)
            for (int b = 0; b < k[1]; ++b)
                for (int c = 0; c < k[2]; ++c) {
                    const float v = prog->h_taps[s][(a * k[1] + b) * k[2] + c];
                    if (v == 0.0f) continue;  // NaN != 0: NaN taps are kept and propagate
                    CPPRE_REQUIRE(ntap < CPPRE_MAX_TAPS, CPPRE_ERR_UNSUPPORTED,
                                  "more than %d non-zero taps in one program", CPPRE_MAX_TAPS);
                    P.taps.o2[ntap] = (int8_t)(a - p[0]);
                    P.taps.o1[ntap] = (int8_t)(b - p[1]);
                    P.taps.o0[ntap] = (int8_t)(c - p[2]);
                    P.taps.c[ntap] = v;
                    ++ntap;
                }
    }
    for (int s = prog->n_slots; s <= CPPRE_MAX_SLOTS; ++s) P.taps.begin[s] = ntap;
    P.g.O2 = O[0] > 0 ? O[0] : 0;
    P.g.O1 = O[1] > 0 ? O[1] : 0;
    P.g.O0 = O[2] > 0 ? O[2] : 0;
    const bool same = (P.g.O2 == P.g.N2 && P.g.O1 == P.g.N1 && P.g.O0 == P.g.N0);
    CPPRE_REQUIRE(same || prog->kind == CPPRE_PROG_LINEAR, CPPRE_ERR_BAD_ARG,
                  "non-linear programs need same-size ('same' padded) derivative slots");
    for (int f = 0; f < prog->n_fields; ++f) {
        const bool empty = (shape[0] * shape[1] * shape[2] * shape[3] == 0);
        CPPRE_REQUIRE(fields[f].d_data != nullptr || empty, CPPRE_ERR_BAD_ARG, "field %d is null", f);
        P.fields[f] = FieldView{fields[f].d_data, fields[f].stride[0], fields[f].stride[1],
                                fields[f].stride[2], fields[f].stride[3]};
    }
    memcpy(P.scalars, prog->scalars, sizeof(P.scalars));
    P.dec_on = (epi->flags & CPPRE_FLAG_DECODE) ? 1 : 0;
    if (P.dec_on) {
        CPPRE_REQUIRE(!(epi->flags & CPPRE_FLAG_IN_BF16), CPPRE_ERR_UNSUPPORTED,
                      "the decode fold takes fp32 fields only");
        CPPRE_REQUIRE(same, CPPRE_ERR_UNSUPPORTED, "the decode fold needs 'same'-padded derivative slots");
        // The fold is only claimed on the one-cell-deep interior; a tap two or more cells out reaches the padding
        // from there, and the reference pads the DECODED field (offset, not zero, would be missing).
        for (int t = 0; t < ntap; ++t)
            CPPRE_REQUIRE(P.taps.o2[t] >= -1 && P.taps.o2[t] <= 1 && P.taps.o1[t] >= -1 && P.taps.o1[t] <= 1 &&
                              P.taps.o0[t] >= -1 && P.taps.o0[t] <= 1,
                          CPPRE_ERR_UNSUPPORTED,
                          "the decode fold needs every non-zero tap within one cell of the centre (3-point stencils); "
                          "decode the fields explicitly for wider stencils");
        for (int s = 0; s < prog->n_slots; ++s) {
            double sum = 0.0;       // sum of the slot's taps: what the slot returns on a constant field of ones
            for (int t = P.taps.begin[s]; t < P.taps.begin[s + 1]; ++t) sum += (double)P.taps.c[t];
            for (int f = 0; f < prog->n_fields; ++f) P.dec.bs[s][f] = (float)((double)fields[f].offset * sum);
        }
        for (int f = 0; f < prog->n_fields; ++f) {
            P.dec.a[f] = fields[f].scale;
            P.dec.b[f] = fields[f].offset;
        }
    }
    CPPRE_REQUIRE(epi->mode == CPPRE_EPI_STORE || epi->mode == CPPRE_EPI_SCORE || epi->mode == CPPRE_EPI_COVER,
                  CPPRE_ERR_BAD_ARG, "unknown epilogue mode %d", epi->mode);
    P.epi.mode = epi->mode;
    P.in_bf16 = (epi->flags & CPPRE_FLAG_IN_BF16) ? 1 : 0;
    P.epi.flags = epi->flags;
    P.epi.reduce = epi->reduce;
    P.epi.ref = epi->d_ref;
    P.epi.rb = epi->ref_stride[0];
    P.epi.r2 = epi->ref_stride[1];
    P.epi.r1 = epi->ref_stride[2];
    P.epi.r0 = epi->ref_stride[3];
    P.epi.weight = epi->d_weight;
    {
        const int64_t* ws = epi->weight_stride;
        const int64_t o2 = P.g.O2, o1 = P.g.O1, o0 = P.g.O0;
        const bool unset = ws[0] == 0 && ws[1] == 0 && ws[2] == 0;
        auto fits = [&](int64_t a2, int64_t a1, int64_t a0) {   // strides on extent-1 axes are free
            return (o2 <= 1 || ws[0] == a2) && (o1 <= 1 || ws[1] == a1) && (o0 <= 1 || ws[2] == a0);
        };
        P.weight_contig = unset || fits(o1 * o0, o0, 1);
        P.weight_tin = !unset && fits(1, o0 * o2, o2);
        CPPRE_REQUIRE(!epi->d_weight || P.weight_contig || P.weight_tin, CPPRE_ERR_UNSUPPORTED,
                      "d_weight must be contiguous [O2,O1,O0] or t-innermost (strides 1, O0*O2, O2)");
    }
    P.epi.out = epi->d_out;
    P.epi.ob = epi->out_stride[0];
    P.epi.o2 = epi->out_stride[1];
    P.epi.o1 = epi->out_stride[2];
    P.epi.o0 = epi->out_stride[3];
    P.epi.scores = epi->d_scores;
    P.epi.band_q = epi->band_q;
    P.epi.band_qp = epi->d_band_q;
    const bool empty_out = (int64_t)P.g.B * P.g.O2 * P.g.O1 * P.g.O0 == 0;
    if (epi->mode == CPPRE_EPI_STORE) {
        CPPRE_REQUIRE(epi->d_out != nullptr || empty_out, CPPRE_ERR_BAD_ARG, "STORE epilogue without d_out");
    } else if (epi->mode == CPPRE_EPI_COVER) {
        CPPRE_REQUIRE(epi->d_scores != nullptr || P.g.B == 0, CPPRE_ERR_BAD_ARG,
                      "COVER epilogue without d_scores");
    } else {
        CPPRE_REQUIRE(epi->d_scores != nullptr || P.g.B == 0, CPPRE_ERR_BAD_ARG,
                      "SCORE epilogue without d_scores");
        CPPRE_REQUIRE(epi->reduce == CPPRE_REDUCE_MAX || epi->reduce == CPPRE_REDUCE_SUM,
                      CPPRE_ERR_BAD_ARG, "unknown reduction %d", epi->reduce);
        CPPRE_REQUIRE(epi->reduce != CPPRE_REDUCE_MAX || (epi->flags & CPPRE_FLAG_ABS),
                      CPPRE_ERR_UNSUPPORTED, "max-reduction is defined on |residual| (set CPPRE_FLAG_ABS)");
    }
    return CPPRE_OK;
}

}  // namespace cppre

using namespace cppre;

extern "C" {

int cppre_version(void) { return CPPRE_VERSION; }
const char* cppre_last_error(void) { return t_error; }
uint64_t cppre_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
const char* cppre_last_path(void) { return t_path; }

int cppre_residual_tin_eligible(const cppre_program_t* prog, const int64_t shape[4],
                                const cppre_field_t* fields, const cppre_epilogue_t* epi) {
    Prepared P;
    if (prepare(prog, shape, fields, epi, &P) != CPPRE_OK) return 0;
    return tin_eligible(prog, P);
}

int cppre_residual_tma_eligible(const cppre_program_t* prog, const int64_t shape[4],
                                const cppre_field_t* fields, const cppre_epilogue_t* epi) {
    Prepared P;
    if (prepare(prog, shape, fields, epi, &P) != CPPRE_OK) return 0;
    return tma_eligible(prog, P);
}

int cppre_residual(const cppre_program_t* prog, const int64_t shape[4],
                   const cppre_field_t* fields, const cppre_epilogue_t* epi, int path,
                   void* stream) {
    Prepared P;
    int rc = prepare(prog, shape, fields, epi, &P);
    if (rc != CPPRE_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (P.epi.mode != CPPRE_EPI_STORE && P.g.B > 0) {
        CPPRE_CHECK_CUDA(cudaMemsetAsync(P.epi.scores, 0, sizeof(float) * (size_t)P.g.B, st));
    }
    if ((int64_t)P.g.B * P.g.O2 * P.g.O1 * P.g.O0 == 0) {  // empty input: nothing to launch
        set_path("empty");
        return CPPRE_OK;
    }
    CPPRE_REQUIRE(path == CPPRE_PATH_AUTO || path == CPPRE_PATH_GENERIC || path == CPPRE_PATH_TMA ||
                      path == CPPRE_PATH_TIN || path == CPPRE_PATH_FLAT, CPPRE_ERR_BAD_ARG, "unknown path %d", path);
    if (path == CPPRE_PATH_TIN || (path == CPPRE_PATH_AUTO && tin_eligible(prog, P))) {
        CPPRE_REQUIRE(tin_eligible(prog, P), CPPRE_ERR_UNSUPPORTED,
                      "request does not qualify for the t-innermost star kernel");
        set_path("tin");
        return launch_tin(prog, P, st);
    }
    CPPRE_REQUIRE(!P.epi.weight || P.weight_contig, CPPRE_ERR_UNSUPPORTED,
                  "a t-innermost weight is served by the t-innermost star kernel only");
    const int ok = tma_eligible(prog, P);
    if (path == CPPRE_PATH_TMA)
        CPPRE_REQUIRE(ok, CPPRE_ERR_UNSUPPORTED, "request does not qualify for the TMA star kernel");
    if (ok && path != CPPRE_PATH_GENERIC && path != CPPRE_PATH_FLAT) {
        set_path("tma");
        rc = launch_tma(prog, P, st);
        // the driver may still refuse an exotic stride pattern when encoding the tensor map;
        // nothing has been launched in that case, so AUTO takes the generic kernel instead
        if (rc != CPPRE_ERR_UNSUPPORTED || path == CPPRE_PATH_TMA) return rc;
    }
    // contiguous star requests the TMA unit cannot stage (rows that are not a multiple of 16 bytes, unaligned
    // bases): the flat-line star kernel, before the one-thread-per-cell generic kernel
    if (path == CPPRE_PATH_FLAT || (path == CPPRE_PATH_AUTO && flat_eligible(prog, P))) {
        CPPRE_REQUIRE(flat_eligible(prog, P), CPPRE_ERR_UNSUPPORTED,
                      "request does not qualify for the flat-line star kernel (contiguous 3x3x3 star programs)");
        set_path("flat");
        return launch_flat(prog, P, st);
    }
    CPPRE_REQUIRE(!P.in_bf16, CPPRE_ERR_UNSUPPORTED,
                  "bf16 fields are served by the TMA star kernel only (star stencils, 'same' padding, N0 %% 8 == 0, "
                  "16-byte aligned rows and strides)");
    set_path("generic");
    return launch_generic(P, st);
}

}  // extern "C"
