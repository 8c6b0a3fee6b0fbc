// extern "C" entry points of libgpk (see include/gpk.h) and the host-side drivers that sequence the
// kernels of the blocked factorisation.  No torch types, no allocation, no CPU fallback.
#include <atomic>
#include <cstdarg>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/gpk.h"
#include "gpk_common.cuh"

#include <cstdlib>

namespace gpk {

static std::mutex g_err_mu;
static char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((unsigned long long)n); }

// ---- optional per-kernel-class device timing (bench.py's roofline) ---------------------------------
// While enabled, every launch is followed by an event on the launching stream; with an in-order stream
// kept busy, the gap between consecutive events is that kernel's duration.
struct TimingRec { cudaEvent_t ev; int cls; };
static bool g_timing = false;
static cudaStream_t g_cur_stream = nullptr;
static std::vector<TimingRec> g_recs;
static std::vector<cudaEvent_t> g_pool;
static std::vector<std::string> g_cls_names;
static std::vector<double> g_cls_ms;
static std::vector<long long> g_cls_n;

void set_stream(cudaStream_t st) { g_cur_stream = st; }

static cudaEvent_t take_event() {
    if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}

static int class_id(const char* what) {
    for (size_t i = 0; i < g_cls_names.size(); ++i) if (g_cls_names[i] == what) return (int)i;
    g_cls_names.push_back(what); g_cls_ms.push_back(0.0); g_cls_n.push_back(0);
    return (int)g_cls_names.size() - 1;
}

static void timing_mark(const char* what) {
    TimingRec r{take_event(), class_id(what)};
    cudaEventRecord(r.ev, g_cur_stream);
    g_recs.push_back(r);
}

int check_launch(const char* what) {
    count_launch(1);
    if (g_timing) timing_mark(what);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return 1;
    }
    return 0;
}

constexpr int MAX_GROUPS = 8;
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// `precision` argument of the C ABI (include/gpk.h GPK_PREC_*) -> digit groups kept by the digit GEMM (0 = native FP64)
static int groups_of(int precision) {
    switch (precision) {
        case GPK_PREC_EMU34: return 8;
        case GPK_PREC_EMU28: return 7;
        case GPK_PREC_EMU21: return 6;
        case GPK_PREC_EMU15: return 5;
        case GPK_PREC_EMU28_34: return 7;       // Cholesky phase (the inverse phase and the prediction GEMM use inverse_groups_of)
        case GPK_PREC_EMU34_28: return 8;
        default: return 0;
    }
}
// digit groups of the inverse phase (and of the prediction GEMM): the mixed modes differ from groups_of here
static int inverse_groups_of(int precision) {
    if (precision == GPK_PREC_EMU28_34) return 8;
    if (precision == GPK_PREC_EMU34_28) return 7;
    return groups_of(precision);
}
static bool bad_precision(const char* who, int precision, int N) {
    if (precision < GPK_PREC_FP64 || precision > GPK_PREC_EMU34_28) {
        set_error("%s: precision %d (0 native FP64, 1..4 emulated FP64 with 34 / 28 / 21 / 15 digit products, 5 / 6 mixed 28+34 / 34+28)", who, precision);
        return true;
    }
    if (precision != GPK_PREC_FP64 && round_up(N, NB) >= 16384) {
        // a group accumulator receives up to 8 digit products of magnitude <= 2^14 per k: exact in int32 for K < 2^14 only
        set_error("%s: the emulated-FP64 modes support N <= 16256 (int32 accumulators); use precision 0", who);
        return true;
    }
    return false;
}

BlockedLayout blocked_layout(int P, int N, int d, int emulated) {
    BlockedLayout L{};
    L.P = P; L.N = N; L.d = d;
    L.Np = round_up(N, NB);
    L.T = L.Np / NB;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
    const size_t Np = L.Np, T = L.T;
    L.mat = take((size_t)P * Np * Np * 8);
    L.dinv = take((size_t)P * T * NB * NB * 8);
    L.dinvT = take((size_t)P * T * NB * NB * 8);
    L.mt = take((size_t)P * NB * Np * 8);
    L.xs = take((size_t)P * d * Np * 8);
    L.sq = take((size_t)P * Np * 8);
    L.cn = take((size_t)P * Np * 8);
    L.z = take((size_t)P * Np * 8);
    L.alpha = take((size_t)P * Np * 8);
    L.pkv = take((size_t)P * T * Np * 8);
    L.pu = take((size_t)P * Np * 8);
    L.pt = take((size_t)P * Np * 8);
    L.pw = take((size_t)P * Np * 8);
    L.probe = take((size_t)P * 8);
    L.fail = take((size_t)P);
    L.oz_slices = L.oz_mt = L.oz_rows = L.oz_mtsc = L.oz_sdig = L.oz_ddig = L.oz_dsc = L.oz_cnt = 0;
    if (emulated) {
        L.oz_slices = take((size_t)P * oz_slice_bytes(L.Np));
        L.oz_mt = take((size_t)P * oz_mt_bytes(L.Np));
        L.oz_rows = take((size_t)4 * Np * 8);
        L.oz_mtsc = take((size_t)3 * P * NB * 8);
        L.oz_sdig = take((size_t)P * oz_panel_bytes(L.Np));
        L.oz_ddig = take((size_t)P * oz_ddig_bytes(L.Np));
        L.oz_dsc = take((size_t)P * T * 2 * NB * 8);
        L.oz_cnt = take((4 * T + 8) * sizeof(int));
    }
    L.total = off;
    return L;
}

// Buffers of the emulated-FP64 mode (null slices = native FP64 DMMA contractions).
struct OzBuffers {
    uint8_t* slices = nullptr;
    uint8_t* mtslices = nullptr;
    uint8_t* sdig = nullptr;
    uint8_t* ddig = nullptr;
    double* dsc = nullptr;
    int* cnt = nullptr;          // tile counters; next() hands out one per digit-GEMM launch
    mutable int cnt_used = 0;
    int* next() const { return cnt + cnt_used++; }
    double *rowsig = nullptr, *rowinv = nullptr, *gam = nullptr, *gaminv = nullptr, *mtsig = nullptr;
    unsigned long long* mtmax = nullptr;      // row maxima of the phase-2 panel (bit patterns of non-negative doubles),
    size_t mtmax_stride = 0;                  // two buffers (parity of the block column) of mtmax_stride entries each
    size_t mtmax_bytes = 0;
};

static OzBuffers oz_buffers(char* ws, const BlockedLayout& L) {
    OzBuffers o;
    if (!L.oz_slices && !L.oz_rows) return o;
    o.slices = (uint8_t*)(ws + L.oz_slices);
    o.mtslices = (uint8_t*)(ws + L.oz_mt);
    double* r = (double*)(ws + L.oz_rows);
    o.rowsig = r; o.rowinv = r + L.Np; o.gam = r + 2 * (size_t)L.Np; o.gaminv = r + 3 * (size_t)L.Np;
    double* m = (double*)(ws + L.oz_mtsc);
    o.mtsig = m;
    o.mtmax = (unsigned long long*)(m + (size_t)L.P * NB);
    o.mtmax_stride = (size_t)L.P * NB;
    o.mtmax_bytes = 2 * (size_t)L.P * NB * 8;
    o.sdig = (uint8_t*)(ws + L.oz_sdig);
    o.ddig = (uint8_t*)(ws + L.oz_ddig);
    o.dsc = (double*)(ws + L.oz_dsc);
    o.cnt = (int*)(ws + L.oz_cnt);
    return o;
}

// Blocked left-looking Cholesky (phase 1).  mode: what the diagonal kernel leaves in the diagonal blocks.
static int factor_phase(double* mat, int Np, int T, int P, double* dinv, double* dinvT, int* info,
                        const unsigned char* active, int mode, const OzBuffers& oz, cudaStream_t st) {
    for (int J = 0; J < T; ++J) {
        if (!oz.slices) {
            if (launch_syrk_update(mat, Np, T, J, P, active, st)) return 1;
            if (launch_diag_block(mat, Np, J, P, dinv, dinvT, T, info, active, mode, st)) return 1;
            if (launch_trsm_panel(mat, Np, T, J, P, dinv, active, st)) return 1;
            continue;
        }
        // emulated FP64: every contraction on the digit GEMM.  The SYRK epilogue leaves the digits of S_IJ, the
        // TRSM reads them with the digits of D_J and leaves L_IJ and its digits for the later SYRKs.
        if (J > 0) {
            if (launch_oz_syrk(mat, Np, T, J, P, oz.slices, oz.rowsig, oz.rowinv, oz.sdig, oz.next(), active, st)) return 1;
        }
        if (launch_diag_block(mat, Np, J, P, dinv, dinvT, T, info, active, mode, st)) return 1;
        // (fusing this slicer into the diagonal kernel was measured in round 2: one CTA per particle slices the two operand
        // sets in 14 us where the separate launch spreads them over 2 P CTAs; C2 2.65 -> 2.70 ms per swarm: kept separate)
        if (launch_oz_slice_d(dinv, dinvT, T, J, P, oz.rowsig, oz.ddig, oz.dsc, active, st)) return 1;
        if (J == 0) {        // no SYRK ran: the first block column is still the raw covariance, no digits of it exist
            if (launch_trsm_panel(mat, Np, T, J, P, dinv, active, st)) return 1;
            if (launch_oz_slice_l(mat, Np, T, J, P, oz.rowinv, oz.slices, active, st)) return 1;
        } else {
            if (launch_oz_trsm(mat, Np, T, J, P, oz.sdig, oz.ddig, oz.dsc, oz.rowsig, oz.rowinv, oz.slices, oz.next(), active, st)) return 1;
        }
    }
    return 0;
}

// Blocked triangular inverse (phase 2), in place: L (with D_I already in the diagonal blocks) -> W = L^-1.
static int inverse_phase(double* mat, int Np, int T, int P, const double* dinvT, double* mt,
                         const unsigned char* active, const OzBuffers& oz, cudaStream_t st) {
    if (oz.slices) {
        if (cudaMemsetAsync(oz.mtmax, 0, oz.mtmax_bytes, st) != cudaSuccess) { set_error("memset of the panel maxima failed"); return 1; }
        if (launch_oz_slice_w(mat, Np, T, 0, 1, P, oz.gam, oz.slices, active, st)) return 1;
    }
    for (int J = T - 2; J >= 0; --J) {
        if (oz.slices) {
            if (launch_oz_invp(mt, Np, T, J, P, oz.ddig, oz.dsc, oz.slices, oz.rowsig, oz.mtmax + (size_t)(J & 1) * oz.mtmax_stride,
                               oz.gaminv, oz.next(), active, st)) return 1;
            if (launch_oz_mt(mt, Np, T, J, P, oz.gaminv, oz.mtmax, oz.mtmax_stride, oz.mtsig, oz.mtslices, active, st)) return 1;
            if (launch_oz_invu(mat, Np, T, J, P, oz.slices, oz.mtslices, oz.mtsig, oz.gam, 1, oz.next(), active, st)) return 1;
        } else {
            if (launch_inv_panel(mat, Np, T, J, P, dinvT, mt, active, st)) return 1;
            if (launch_inv_update(mat, Np, T, J, P, mt, active, st)) return 1;
        }
    }
    return 0;
}

// ---- concurrent particle groups ------------------------------------------------------------------
static int g_groups = 2;
static int g_min_group = 16;       // particles per group below which a swarm is not split further (gpk_set_groups)
static cudaStream_t g_group_streams[MAX_GROUPS] = {};
static cudaEvent_t g_join[MAX_GROUPS] = {};
static cudaEvent_t g_fork = nullptr;
static int g_groups_ready = 0;

static int ensure_group_streams(int G) {
    if (g_fork == nullptr && cudaEventCreateWithFlags(&g_fork, cudaEventDisableTiming) != cudaSuccess) { set_error("event create failed"); return 1; }
    for (; g_groups_ready < G; ++g_groups_ready) {
        if (cudaStreamCreateWithFlags(&g_group_streams[g_groups_ready], cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&g_join[g_groups_ready], cudaEventDisableTiming) != cudaSuccess) {
            set_error("group stream create failed");
            return 1;
        }
    }
    return 0;
}

static bool bad_dims(const char* who, int N, int d, int S) {
    if (N < 1 || d < 1 || d > MAX_D || S < 0 || S > MAX_S) {
        set_error("%s: unsupported sizes N=%d d=%d S=%d (d<=%d, S<=%d)", who, N, d, S, MAX_D, MAX_S);
        return true;
    }
    return false;
}

}  // namespace gpk

using namespace gpk;

extern "C" {

int gpk_version(void) { return GPK_VERSION; }
const char* gpk_last_error_string(void) { return g_err; }
unsigned long long gpk_launch_count(void) { return g_launches.load(); }
int gpk_padded_n(int N) { return round_up(N, NB); }

int gpk_cov_batched(const double* X1, const double* X2, const double* e, const double* ls, int P, int N1, int N2,
                    int d, double* K_out, long long ldk, long long strideK, void* stream) {
    if (bad_dims("gpk_cov_batched", N1, d, 0) || N2 < 1 || P < 1) return 2;
    set_stream((cudaStream_t)stream);
    return launch_cov_full(X1, X2, e, ls, P, N1, N2, d, K_out, ldk, strideK, (cudaStream_t)stream);
}

size_t gpk_potrf_workspace_bytes(int P, int Np) {
    const size_t T = Np / NB;
    return 2 * align256((size_t)P * T * NB * NB * 8);
}

int gpk_potrf_batched(double* K_inout, int P, int Np, int* info, void* workspace, size_t workspace_bytes,
                      void* stream) {
    if (Np < NB || Np % NB != 0 || P < 1) { set_error("gpk_potrf_batched: Np=%d must be a positive multiple of %d", Np, NB); return 2; }
    if (workspace_bytes < gpk_potrf_workspace_bytes(P, Np)) { set_error("gpk_potrf_batched: workspace too small"); return 3; }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    const int T = Np / NB;
    double* dinv = (double*)workspace;
    double* dinvT = (double*)((char*)workspace + align256((size_t)P * T * NB * NB * 8));
    if (cudaMemsetAsync(info, 0, (size_t)P * sizeof(int), st) != cudaSuccess) { set_error("memset info failed"); return 1; }
    return factor_phase(K_inout, Np, T, P, dinv, dinvT, info, nullptr, DIAG_WRITE_FACTOR, OzBuffers{}, st);
}

// ---- staged interface (gpk_fit_alpha_batched, gpk_selfpredict_counts): the stages gpk_loss_batched fuses,
// callable one by one on factors produced by gpk_potrf_batched.  Native FP64 arithmetic (the digit state of the
// emulated mode is not part of this interface).
struct StagedLayout { size_t mat, dinv, dinvT, mt, z, cn, alpha, info, total; };
static StagedLayout staged_layout(int P, int Np) {
    StagedLayout L{};
    const size_t T = Np / NB;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
    L.mat = take((size_t)P * Np * Np * 8);
    L.dinv = take((size_t)P * T * NB * NB * 8);
    L.dinvT = take((size_t)P * T * NB * NB * 8);
    L.mt = take((size_t)P * NB * Np * 8);
    L.z = take((size_t)P * Np * 8);
    L.cn = take((size_t)P * Np * 8);
    L.alpha = take((size_t)P * Np * 8);
    L.info = take((size_t)P * sizeof(int));
    L.total = off;
    return L;
}

// W = L^-1 into the workspace copy (lower block triangle): diagonal blocks by substitution, the rest by the blocked
// inverse of the fused path (native DMMA contractions)
static int staged_inverse(const double* L, int P, int Np, char* ws, const StagedLayout& S, const unsigned char* active,
                          cudaStream_t st) {
    const int T = Np / NB;
    double* mat = (double*)(ws + S.mat);
    if (cudaMemcpyAsync(mat, L, (size_t)P * Np * Np * 8, cudaMemcpyDeviceToDevice, st) != cudaSuccess) { set_error("copy of the factors failed"); return 1; }
    if (cudaMemsetAsync(ws + S.info, 0, (size_t)P * sizeof(int), st) != cudaSuccess) { set_error("memset info failed"); return 1; }
    if (launch_tri_inv_diag(L, mat, Np, T, P, (double*)(ws + S.dinv), (double*)(ws + S.dinvT), (int*)(ws + S.info), active, st)) return 1;
    return inverse_phase(mat, Np, T, P, (const double*)(ws + S.dinvT), (double*)(ws + S.mt), active, OzBuffers{}, st);
}

size_t gpk_staged_workspace_bytes(int P, int Np) {
    if (P < 1 || Np < NB || Np % NB != 0) return 0;
    return staged_layout(P, Np).total;
}

int gpk_fit_alpha_batched(const double* L, const double* y, int P, int N, int Np, double* alpha_out, void* workspace,
                          size_t workspace_bytes, void* stream) {
    if (P < 1 || N < 1 || Np < NB || Np % NB != 0 || N > Np) { set_error("gpk_fit_alpha_batched: P=%d N=%d Np=%d", P, N, Np); return 2; }
    const StagedLayout S = staged_layout(P, Np);
    if (workspace == nullptr || workspace_bytes < S.total) { set_error("gpk_fit_alpha_batched: workspace too small"); return 3; }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    char* ws = (char*)workspace;
    if (staged_inverse(L, P, Np, ws, S, nullptr, st)) return 1;
    const double* mat = (const double*)(ws + S.mat);
    if (launch_wy(mat, Np, N, P, y, (double*)(ws + S.z), nullptr, nullptr, nullptr, st)) return 1;
    return launch_colnorm_alpha(mat, Np, P, (const double*)(ws + S.z), (double*)(ws + S.cn), alpha_out, nullptr, nullptr, nullptr, st);
}

int gpk_selfpredict_counts(const double* L, const double* alpha, const double* y, const double* e,
                           const double* sigma_vals, const unsigned char* active, int P, int N, int Np, int S,
                           int* counts_out, double* resid_out, double* sig_out, void* workspace, size_t workspace_bytes,
                           void* stream) {
    (void)y;        // mu - y = -e^2 alpha: the targets enter through alpha only
    if (P < 1 || N < 1 || Np < NB || Np % NB != 0 || N > Np || S < 0 || S > MAX_S) { set_error("gpk_selfpredict_counts: P=%d N=%d Np=%d S=%d", P, N, Np, S); return 2; }
    const StagedLayout SL = staged_layout(P, Np);
    if (workspace == nullptr || workspace_bytes < SL.total) { set_error("gpk_selfpredict_counts: workspace too small"); return 3; }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    char* ws = (char*)workspace;
    if (staged_inverse(L, P, Np, ws, SL, active, st)) return 1;
    const double* mat = (const double*)(ws + SL.mat);
    // column norms of W (diag of K^-1); the alpha this kernel also forms from z goes to scratch
    if (cudaMemsetAsync(ws + SL.z, 0, (size_t)P * Np * 8, st) != cudaSuccess) { set_error("memset z failed"); return 1; }
    if (launch_colnorm_alpha(mat, Np, P, (const double*)(ws + SL.z), (double*)(ws + SL.cn), (double*)(ws + SL.alpha), nullptr, nullptr, active, st)) return 1;
    return launch_counts(alpha, (const double*)(ws + SL.cn), e, sigma_vals, P, N, Np, S, counts_out, resid_out, sig_out,
                         nullptr, 0.0, nullptr, nullptr, 1, active, st);
}

size_t gpk_loss_workspace_bytes(int P, int N, int d, int precision) {
    if (N <= SMALL_N_MAX) return 256;
    size_t best = 0;                      // enough for any particle-group split (gpk_set_groups)
    for (int G = 1; G <= MAX_GROUPS; ++G) {
        const size_t b = (size_t)G * blocked_layout((P + G - 1) / G, N, d, precision != GPK_PREC_FP64).total;
        if (b > best) best = b;
    }
    return best;
}

// One pass of the blocked path over the particles of one group: covariance -> Cholesky -> inverse -> alpha, column norms ->
// counts.  emu: the contractions run on the digit GEMM (oz_set_groups done by the caller).  probe: the residual probe is
// formed on the way (rho -> ws.probe, rho > tol -> ws.fail).  zero_inactive = 0: particles outside `act` are left untouched.
static int blocked_pass(const double* X, const double* y, const double* e, const double* lsp, const unsigned char* act,
                        const double* sigma_vals, int Pc, int N, int d, int S, int* counts, int* info, double* resid_out,
                        double* sig_out, char* ws, const BlockedLayout& L, int emu, bool probe, double tol,
                        int zero_inactive, cudaStream_t gs, int inv_groups = 0) {
    OzBuffers oz;
    if (emu) {
        oz = oz_buffers(ws, L);
        if (cudaMemsetAsync(oz.cnt, 0, (4 * L.T + 8) * sizeof(int), gs) != cudaSuccess) { set_error("memset counters failed"); return 1; }
        if (launch_oz_prep(e, N, L.Np, oz.rowsig, oz.rowinv, oz.gam, oz.gaminv, gs)) return 1;
    }
    double* mat = (double*)(ws + L.mat);
    double* dinv = (double*)(ws + L.dinv);
    double* dinvT = (double*)(ws + L.dinvT);
    double* mt = (double*)(ws + L.mt);
    double* xs = (double*)(ws + L.xs);
    double* sq = (double*)(ws + L.sq);
    double* cn = (double*)(ws + L.cn);
    double* z = (double*)(ws + L.z);
    double* alpha = (double*)(ws + L.alpha);
    double* pkv = probe ? (double*)(ws + L.pkv) : nullptr;
    double* pu = probe ? (double*)(ws + L.pu) : nullptr;
    double* pt = probe ? (double*)(ws + L.pt) : nullptr;
    double* pw = probe ? (double*)(ws + L.pw) : nullptr;
    if (launch_scale_points(X, lsp, Pc, N, L.Np, d, xs, sq, act, gs)) return 1;
    if (launch_cov_lower(xs, sq, e, Pc, N, L.Np, d, mat, pkv, act, gs)) return 1;
    if (probe && launch_probe_sum(pkv, L.Np, L.T, Pc, pu, act, gs)) return 1;
    if (factor_phase(mat, L.Np, L.T, Pc, dinv, dinvT, info, act, DIAG_WRITE_INVERSE, oz, gs)) return 1;
    const int fac_groups = oz_get_groups();
    if (emu && inv_groups) oz_set_groups(inv_groups);           // mixed modes: another digit-product count for the inverse
    const int rc_inv = inverse_phase(mat, L.Np, L.T, Pc, dinvT, mt, act, oz, gs);
    if (emu && inv_groups) oz_set_groups(fac_groups);
    if (rc_inv) return 1;
    if (launch_wy(mat, L.Np, N, Pc, y, z, pu, pt, act, gs)) return 1;
    if (launch_colnorm_alpha(mat, L.Np, Pc, z, cn, alpha, pt, pw, act, gs)) return 1;
    return launch_counts(alpha, cn, e, sigma_vals, Pc, N, L.Np, S, counts, resid_out, sig_out, pw, tol,
                         (double*)(ws + L.probe), (unsigned char*)(ws + L.fail), zero_inactive, act, gs);
}

int gpk_loss_batched(const double* X, const double* y, const double* e, const double* ls,
                     const unsigned char* active, const double* sigma_vals, int P, int N, int d, int S,
                     int precision, int flags, double check_tol, int* counts_out, int* info, double* resid_out,
                     double* sig_out, double* probe_out, unsigned char* fallback_out, void* workspace,
                     size_t workspace_bytes, void* stream) {
    if (bad_dims("gpk_loss_batched", N, d, S) || P < 1) return 2;
    if (bad_precision("gpk_loss_batched", precision, N)) return 2;
    if (flags & ~(GPK_FLAG_REFERENCE_FORM | GPK_FLAG_KEEP_INACTIVE)) { set_error("gpk_loss_batched: flags 0x%x", flags); return 2; }
    const int form = (flags & GPK_FLAG_REFERENCE_FORM) ? 1 : 0;
    const int zero_inactive = (flags & GPK_FLAG_KEEP_INACTIVE) ? 0 : 1;
    if (form && N > SMALL_N_MAX) {
        set_error("gpk_loss_batched: the reference-formula form (N^3 solve) exists on the one-CTA path only, N <= %d", SMALL_N_MAX);
        return 2;
    }
    if (!zero_inactive && (active == nullptr || N <= SMALL_N_MAX)) {
        set_error("gpk_loss_batched: GPK_FLAG_KEEP_INACTIVE needs an active mask and the blocked path (N > %d)", SMALL_N_MAX);
        return 2;
    }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    if (zero_inactive) {
        if (cudaMemsetAsync(info, 0, (size_t)P * sizeof(int), st) != cudaSuccess) { set_error("memset info failed"); return 1; }
    } else if (launch_clear_info(info, active, P, st)) {
        return 1;
    }
    if (N <= SMALL_N_MAX) {
        // native FP64 in one CTA per particle: nothing is emulated, nothing to check
        if (probe_out && cudaMemsetAsync(probe_out, 0, (size_t)P * 8, st) != cudaSuccess) { set_error("memset probe failed"); return 1; }
        if (fallback_out && cudaMemsetAsync(fallback_out, 0, (size_t)P, st) != cudaSuccess) { set_error("memset fallback failed"); return 1; }
        return launch_small_loss(X, y, e, ls, active, sigma_vals, P, N, d, S, counts_out, info, resid_out, sig_out, form, st);
    }

    // blocked path: waves of as many particles as the workspace holds, each wave split into independent
    // particle groups on concurrent streams (the chain SYRK -> diag -> TRSM of one group is serial; kernels of
    // other groups fill its wave tails and hide the latency-bound diagonal kernel)
    const int emu = precision != GPK_PREC_FP64;
    if (emu) oz_set_groups(groups_of(precision));
    const bool probe = check_tol != 0.0 || probe_out != nullptr;
    const bool fallback = emu && check_tol > 0.0;              // check_tol < 0: flags only, the caller re-evaluates
    const double tol = check_tol != 0.0 ? (check_tol > 0.0 ? check_tol : -check_tol) : 1e300;
    const size_t per1 = blocked_layout(1, N, d, emu).total;
    if (workspace == nullptr || workspace_bytes < per1) { set_error("gpk_loss_batched: workspace too small (%zu < %zu)", workspace_bytes, per1); return 3; }
    // particle groups: measured on one B200 at N = 4096 (profiles/r02_bench_v4.json / v5, strong_scaling_slices_1gpu):
    // two groups are within noise of the best setting at every swarm size (256 particles: 1 / 2 / 4 groups = 1311 / 1307 /
    // 1286 evals/s; 32 particles: 1113-1127 / 1167-1199 / 1186-1194): one group's latency-bound diagonal-block kernel
    // overlaps the other's digit GEMMs, while each persistent GEMM launch keeps enough tiles for the 148 SMs
    int G = g_timing ? 1 : g_groups;
    if (G > P / g_min_group) G = P / g_min_group;
    if (G < 1) G = 1;
    if (G > 1 && ensure_group_streams(G)) return 1;
    auto fits = [&](int pw) { const int pg = (pw + G - 1) / G; return (size_t)G * blocked_layout(pg, N, d, emu).total <= workspace_bytes; };
    int Pw = P;
    while (Pw > 1 && !fits(Pw)) { --Pw; }
    if (!fits(Pw)) { G = 1; Pw = 1; }
    for (int p0 = 0; p0 < P; p0 += Pw) {
        const int Pwave = (P - p0 < Pw) ? (P - p0) : Pw;
        const int pg = (Pwave + G - 1) / G;
        const size_t gbytes = blocked_layout(pg, N, d, emu).total;
        if (G > 1 && cudaEventRecord(g_fork, st) != cudaSuccess) { set_error("fork event failed"); return 1; }
        for (int g = 0; g < G; ++g) {
            const int q0 = p0 + g * pg;
            const int Pc = (p0 + Pwave - q0 < pg) ? (p0 + Pwave - q0) : pg;
            if (Pc <= 0) break;
            cudaStream_t gs = (G > 1) ? g_group_streams[g] : st;
            if (G > 1) cudaStreamWaitEvent(gs, g_fork, 0);
            const BlockedLayout L = blocked_layout(Pc, N, d, emu);
            char* ws = (char*)workspace + (size_t)g * gbytes;
            const unsigned char* act = active ? active + q0 : nullptr;
            const double* lsp = ls + (size_t)q0 * d;
            int* cnt = counts_out + (size_t)q0 * S;
            double* rs = resid_out ? resid_out + (size_t)q0 * N : nullptr;
            double* sg = sig_out ? sig_out + (size_t)q0 * N : nullptr;
            if (blocked_pass(X, y, e, lsp, act, sigma_vals, Pc, N, d, S, cnt, info + q0, rs, sg, ws, L, emu, probe, tol, zero_inactive, gs,
                             emu ? inverse_groups_of(precision) : 0)) return 1;
            if (probe_out && cudaMemcpyAsync(probe_out + q0, ws + L.probe, (size_t)Pc * 8, cudaMemcpyDeviceToDevice, gs) != cudaSuccess) { set_error("copy of the probe values failed"); return 1; }
            if (fallback_out) {
                cudaError_t ce = probe ? cudaMemcpyAsync(fallback_out + q0, ws + L.fail, (size_t)Pc, cudaMemcpyDeviceToDevice, gs)
                                       : cudaMemsetAsync(fallback_out + q0, 0, (size_t)Pc, gs);
                if (ce != cudaSuccess) { set_error("copy of the fallback flags failed"); return 1; }
            }
            if (fallback) {
                // checked mixed precision: the particles whose probe failed are evaluated again by the native FP64 (DMMA)
                // kernels, in the same call and on the same stream, with the fail flags as the active mask.  No host
                // synchronisation (the call stays asynchronous and capturable): when nothing failed, every kernel of this
                // pass exits at its first instruction.  The flags are copied first: this pass reuses the workspace.
                unsigned char* mask = (unsigned char*)(ws + L.fail);
                if (launch_clear_info(info + q0, mask, Pc, gs)) return 1;
                if (blocked_pass(X, y, e, lsp, mask, sigma_vals, Pc, N, d, S, cnt, info + q0, rs, sg, ws, L, 0, false, 0.0, 0, gs)) return 1;
            }
            if (G > 1) {
                cudaEventRecord(g_join[g], gs);
                cudaStreamWaitEvent(st, g_join[g], 0);
            }
        }
    }
    return 0;
}

size_t gpk_fit_workspace_bytes(int N, int d, int precision) {
    (void)d;
    const size_t Np = round_up(N, NB), T = Np / NB;
    size_t b = 2 * align256(T * NB * NB * 8) + align256(NB * Np * 8) + 2 * align256(Np * 8);
    b += align256(T * Np * 8) + 3 * align256(Np * 8) + 2 * align256(8);                 // residual probe
    if (precision != GPK_PREC_FP64)
        b += align256(oz_slice_bytes((int)Np)) + align256(oz_mt_bytes((int)Np)) + align256(4 * Np * 8) + align256(3 * NB * 8) +
             align256(oz_panel_bytes((int)Np)) + align256(oz_ddig_bytes((int)Np)) + align256(T * 2 * NB * 8) +
             align256((4 * T + 8) * sizeof(int));
    return b;
}

int gpk_fit(const double* X, const double* y, const double* e, const double* ls, int N, int d, int precision,
            double* W_out, double* alpha_out, double* xs_out, double* sq_out, int* info, double* probe_out,
            void* workspace, size_t workspace_bytes, void* stream) {
    if (bad_dims("gpk_fit", N, d, 0)) return 2;
    if (bad_precision("gpk_fit", precision, N)) return 2;
    if (workspace_bytes < gpk_fit_workspace_bytes(N, d, precision)) { set_error("gpk_fit: workspace too small"); return 3; }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    const int emu = precision != GPK_PREC_FP64;
    if (emu) oz_set_groups(groups_of(precision));
    const int Np = round_up(N, NB), T = Np / NB;
    char* ws = (char*)workspace;
    size_t off = 0;
    double* dinv = (double*)(ws + off); off += align256((size_t)T * NB * NB * 8);
    double* dinvT = (double*)(ws + off); off += align256((size_t)T * NB * NB * 8);
    double* mt = (double*)(ws + off); off += align256((size_t)NB * Np * 8);
    double* z = (double*)(ws + off); off += align256((size_t)Np * 8);
    double* cn = (double*)(ws + off); off += align256((size_t)Np * 8);
    double* pkv = (double*)(ws + off); off += align256((size_t)T * Np * 8);
    double* pu = (double*)(ws + off); off += align256((size_t)Np * 8);
    double* pt = (double*)(ws + off); off += align256((size_t)Np * 8);
    double* pw = (double*)(ws + off); off += align256((size_t)Np * 8);
    double* rho = (double*)(ws + off); off += align256(8);
    unsigned char* fail = (unsigned char*)(ws + off); off += align256(8);
    const bool probe = probe_out != nullptr;
    OzBuffers oz;
    if (emu) {
        oz.slices = (uint8_t*)(ws + off); off += align256(oz_slice_bytes(Np));
        oz.mtslices = (uint8_t*)(ws + off); off += align256(oz_mt_bytes(Np));
        double* r = (double*)(ws + off); off += align256((size_t)4 * Np * 8);
        oz.rowsig = r; oz.rowinv = r + Np; oz.gam = r + 2 * (size_t)Np; oz.gaminv = r + 3 * (size_t)Np;
        oz.mtsig = (double*)(ws + off); oz.mtmax = (unsigned long long*)(oz.mtsig + NB); oz.mtmax_stride = NB;
        oz.mtmax_bytes = (size_t)2 * NB * 8; off += align256((size_t)3 * NB * 8);
        oz.sdig = (uint8_t*)(ws + off); off += align256(oz_panel_bytes(Np));
        oz.ddig = (uint8_t*)(ws + off); off += align256(oz_ddig_bytes(Np));
        oz.dsc = (double*)(ws + off); off += align256((size_t)T * 2 * NB * 8);
        oz.cnt = (int*)(ws + off);
        if (cudaMemsetAsync(oz.cnt, 0, (4 * T + 8) * sizeof(int), st) != cudaSuccess) { set_error("memset counters failed"); return 1; }
        if (launch_oz_prep(e, N, Np, oz.rowsig, oz.rowinv, oz.gam, oz.gaminv, st)) return 1;
    }
    if (cudaMemsetAsync(info, 0, sizeof(int), st) != cudaSuccess) { set_error("memset info failed"); return 1; }
    if (launch_scale_points(X, ls, 1, N, Np, d, xs_out, sq_out, nullptr, st)) return 1;
    if (launch_cov_lower(xs_out, sq_out, e, 1, N, Np, d, W_out, probe ? pkv : nullptr, nullptr, st)) return 1;
    if (probe && launch_probe_sum(pkv, Np, T, 1, pu, nullptr, st)) return 1;
    if (factor_phase(W_out, Np, T, 1, dinv, dinvT, info, nullptr, DIAG_WRITE_INVERSE, oz, st)) return 1;
    if (emu) oz_set_groups(inverse_groups_of(precision));
    if (inverse_phase(W_out, Np, T, 1, dinvT, mt, nullptr, oz, st)) return 1;
    if (launch_wy(W_out, Np, N, 1, y, z, probe ? pu : nullptr, pt, nullptr, st)) return 1;
    if (launch_colnorm_alpha(W_out, Np, 1, z, cn, alpha_out, probe ? pt : nullptr, pw, nullptr, st)) return 1;
    if (probe) {
        if (launch_probe_value(pw, N, Np, 1, rho, fail, 1e300, nullptr, st)) return 1;
        if (cudaMemcpyAsync(probe_out, rho, 8, cudaMemcpyDeviceToDevice, st) != cudaSuccess) { set_error("copy of the probe value failed"); return 1; }
    }
    return 0;
}

size_t gpk_predict_workspace_bytes(int N, int chunk) {
    const size_t Np = round_up(N, NB), T = Np / NB, Bp = round_up(chunk, NB);
    return align256(Bp * Np * 8) + align256(T * Bp * 8);
}

int gpk_predict(const double* W, const double* alpha, const double* xs, const double* sq, const double* ls, int N,
                int d, const double* Xfit, long long M, double* mu_out, double* sd_out, int chunk, void* workspace,
                size_t workspace_bytes, void* stream) {
    if (bad_dims("gpk_predict", N, d, 0) || chunk < 1) return 2;
    if (M <= 0) return 0;
    if (workspace_bytes < gpk_predict_workspace_bytes(N, chunk)) { set_error("gpk_predict: workspace too small"); return 3; }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    const int Np = round_up(N, NB), T = Np / NB;
    const int BpMax = round_up(chunk, NB);
    double* KsT = (double*)workspace;
    double* partial = (double*)((char*)workspace + align256((size_t)BpMax * Np * 8));
    for (long long m0 = 0; m0 < M; m0 += chunk) {
        const int B = (int)((M - m0 < chunk) ? (M - m0) : chunk);
        const int Bp = round_up(B, NB);
        if (launch_cross_cov(xs, sq, ls, alpha, N, Np, d, Xfit, M, m0, B, Bp, KsT, mu_out, st)) return 1;
        if (launch_predict_gemm(W, Np, T, KsT, Bp, partial, st)) return 1;
        if (launch_predict_finalize(partial, T, Bp, B, sd_out + m0, st)) return 1;
    }
    return 0;
}

// ---- prediction without any allocation (no counterpart in the reference) ----
// Scratch of the native prediction chain for one chunk of GPK_STATIC_CHUNK grid points against at most GPK_STATIC_MAX_N
// known points, in a file-scope device array: K_*^T [chunk][Np] doubles, then the per-block partial sums [T][chunk].
// Every GPU that loads the library holds these 4.03 MiB until the process exits.
constexpr size_t STATIC_KST_BYTES = (size_t)GPK_STATIC_CHUNK * GPK_STATIC_MAX_N * 8;
constexpr size_t STATIC_PARTIAL_BYTES = (size_t)(GPK_STATIC_MAX_N / NB) * GPK_STATIC_CHUNK * 8;
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_KST_BYTES + STATIC_PARTIAL_BYTES];

// In host code `g_static_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[STATIC_MAX_DEVICES] = {};

static unsigned char* static_scratch_for_current_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= STATIC_MAX_DEVICES) {
        set_error("gpk_predict_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (g_static_ptr[dev] == nullptr) {
        unsigned char* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            set_error("gpk_predict_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        g_static_ptr[dev] = ptr;
    }
    return g_static_ptr[dev];
}

int gpk_predict_static(const double* W, const double* alpha, const double* xs, const double* sq, const double* ls, int N,
                       int d, const double* Xfit, long long M, double* mu_out, double* sd_out, void* stream) {
    if (bad_dims("gpk_predict_static", N, d, 0)) return 2;
    if (N > GPK_STATIC_MAX_N) {
        set_error("gpk_predict_static: N=%d exceeds GPK_STATIC_MAX_N=%d (the static scratch holds %d x %d doubles); "
                  "use gpk_predict with a caller-provided workspace", N, GPK_STATIC_MAX_N, GPK_STATIC_CHUNK, GPK_STATIC_MAX_N);
        return 2;
    }
    if (M <= 0) return 0;
    unsigned char* scratch = static_scratch_for_current_device();
    if (scratch == nullptr) return 1;
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    const int Np = round_up(N, NB), T = Np / NB;
    double* KsT = (double*)scratch;
    double* partial = (double*)(scratch + STATIC_KST_BYTES);
    // nothing has to start at zero: k_cross_cov writes all Bp x Np entries of K_*^T (zero rows for the padding points),
    // k_predict_gemm all T x Bp partial sums, before the next kernel of the chain reads them
    for (long long m0 = 0; m0 < M; m0 += GPK_STATIC_CHUNK) {
        const int B = (int)((M - m0 < GPK_STATIC_CHUNK) ? (M - m0) : GPK_STATIC_CHUNK);
        const int Bp = round_up(B, NB);
        if (launch_cross_cov(xs, sq, ls, alpha, N, Np, d, Xfit, M, m0, B, Bp, KsT, mu_out, st)) return 1;
        if (launch_predict_gemm(W, Np, T, KsT, Bp, partial, st)) return 1;
        if (launch_predict_finalize(partial, T, Bp, B, sd_out + m0, st)) return 1;
    }
    return 0;
}

size_t gpk_predict_state_bytes(int N) { return oz_pred_state_bytes(round_up(N, NB)); }

int gpk_predict_prepare(const double* W, const double* e, int N, void* state, size_t state_bytes, void* stream) {
    if (N < 1) { set_error("gpk_predict_prepare: N=%d", N); return 2; }
    if (state_bytes < gpk_predict_state_bytes(N)) { set_error("gpk_predict_prepare: state buffer too small"); return 3; }
    set_stream((cudaStream_t)stream);
    return launch_oz_pred_prepare(W, e, N, round_up(N, NB), (uint8_t*)state, (cudaStream_t)stream);
}

size_t gpk_predict_emu_workspace_bytes(int N, int chunk) {
    return oz_pred_workspace_bytes(round_up(N, NB), round_up(chunk, NB));
}

int gpk_predict_emu(const void* state, const double* alpha, const double* xs, const double* sq, const double* ls, int N,
                    int d, int precision, const double* Xfit, long long M, double* mu_out, double* sd_out, int chunk,
                    void* workspace, size_t workspace_bytes, void* stream) {
    if (bad_dims("gpk_predict_emu", N, d, 0) || chunk < 1) return 2;
    if (precision == GPK_PREC_FP64 || bad_precision("gpk_predict_emu", precision, N)) {
        if (precision == GPK_PREC_FP64) set_error("gpk_predict_emu: an emulated precision (GPK_PREC_EMU*) is required; gpk_predict is the native path");
        return 2;
    }
    oz_set_groups(inverse_groups_of(precision));
    if (M <= 0) return 0;
    if (workspace_bytes < gpk_predict_emu_workspace_bytes(N, chunk)) { set_error("gpk_predict_emu: workspace too small"); return 3; }
    cudaStream_t st = (cudaStream_t)stream;
    set_stream(st);
    const int Np = round_up(N, NB), T = Np / NB;
    for (long long m0 = 0; m0 < M; m0 += chunk) {
        const int B = (int)((M - m0 < chunk) ? (M - m0) : chunk);
        const int Bp = round_up(B, NB);
        double* partial = nullptr;
        if (launch_oz_predict_chunk((const uint8_t*)state, xs, sq, ls, alpha, N, Np, d, Xfit, M, m0, B, Bp,
                                    (uint8_t*)workspace, mu_out, &partial, st)) return 1;
        if (launch_predict_finalize(partial, 4 * T, Bp, B, sd_out + m0, st)) return 1;
    }
    return 0;
}

int gpk_kmeans_step(const double* X, int N, int d, const double* centers, const double* csq, int k, int* labels_inout,
                    int* changed_out, double* sums_out, double* counts_out, void* stream) {
    if (!X || !centers || !csq || !labels_inout || !changed_out || (sums_out && !counts_out)) { set_error("gpk_kmeans_step: null argument"); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_kmeans_step(X, N, d, centers, csq, k, labels_inout, changed_out, sums_out, counts_out, (cudaStream_t)stream);
}

size_t gpk_pso_layout(int P, int d, int S, int n_points, int max_iter, int max_evals, long long rng_len,
                      unsigned long long* offsets_out) {
    if (P < 1 || d < 1 || d > MAX_D || S < 2 || S > MAX_S || n_points < 1 || max_iter < 0 || max_evals < 1 || rng_len < 0) return 0;
    const PsoLayout L = pso_layout(P, d, S, n_points, max_iter, max_evals, rng_len);
    if (offsets_out) {
        const size_t o[GPK_PSO_N_OFFSETS] = {L.scalars, L.lower, L.upper, L.vmax, L.pos, L.vel, L.pbest_pos, L.pbest_sc, L.gbest_pos,
                                             L.scores, L.expected, L.d_sigma, L.measured_of, L.log_best, L.log_stall,
                                             L.log_restart, L.trace_pos, L.trace_sc, L.rng};
        for (int i = 0; i < GPK_PSO_N_OFFSETS; ++i) offsets_out[i] = o[i];
    }
    return L.total;
}

int gpk_pso_step(void* state, size_t state_bytes, int P, int d, int S, int n_points, int max_iter, int max_evals,
                 long long rng_len, const int* counts, const int* info, const unsigned char* fallback, double* positions_out,
                 unsigned char* active_out, void* stream) {
    const size_t need = gpk_pso_layout(P, d, S, n_points, max_iter, max_evals, rng_len, nullptr);
    if (need == 0) { set_error("gpk_pso_step: unsupported sizes P=%d d=%d S=%d", P, d, S); return 2; }
    if (state == nullptr || state_bytes < need) { set_error("gpk_pso_step: state buffer too small (%zu < %zu)", state_bytes, need); return 3; }
    if (!counts || !info || !positions_out || !active_out) { set_error("gpk_pso_step: null argument"); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_pso_step((unsigned char*)state, pso_layout(P, d, S, n_points, max_iter, max_evals, rng_len), P, S, counts, info,
                           fallback, positions_out, active_out, (cudaStream_t)stream);
}

int gpk_prob_surface(const double* table, long long R, int ncols, int npoints, double* y_out, double* p_out,
                     int* valid_out, void* stream) {
    if (npoints < 1) { set_error("gpk_prob_surface: npoints=%d", npoints); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_prob_surface(table, R, ncols, npoints, y_out, p_out, valid_out, (cudaStream_t)stream);
}

int gpk_surface_scan(const int* valid, long long R, long long* dst_out, long long* count_out, void* stream) {
    if (R < 0) { set_error("gpk_surface_scan: R=%lld", R); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_surface_scan(valid, R, dst_out, count_out, (cudaStream_t)stream);
}

int gpk_surface_gather(const int* valid, const long long* dst, long long R, int npoints, const double* y, const double* p,
                       double* y_compact, double* p_compact, long long* row_index_out, void* stream) {
    if (npoints < 1) { set_error("gpk_surface_gather: npoints=%d", npoints); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_surface_gather(valid, dst, R, npoints, y, p, y_compact, p_compact, row_index_out, (cudaStream_t)stream);
}

int gpk_sum_gaussians(const double* y, int n, const double* gaus, int ng, double half_dy, double* p_out,
                      void* stream) {
    set_stream((cudaStream_t)stream);
    return launch_sum_gaussians(y, n, gaus, ng, half_dy, p_out, (cudaStream_t)stream);
}

size_t gpk_oz_slice_bytes(int rows, int cols) { return (size_t)(rows / NB) * (cols / 32) * (oz_mt_bytes(32)); }

int gpk_oz_slice(const double* src, int ld, int rows, int cols, const double* rowinv, void* dst, void* stream) {
    if (rows < NB || rows % NB || cols < NB || cols % NB) { set_error("gpk_oz_slice: rows and cols must be positive multiples of %d", NB); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_oz_slice_generic(src, ld, rows, cols, rowinv, (uint8_t*)dst, (cudaStream_t)stream);
}

int gpk_oz_gemm_tn(const void* A, const void* B, int M, int N, int K, const double* sa, const double* sb, double* C,
                   int ldc, int* dbg, int groups, void* stream) {
    if (groups < 5 || groups > 8) { set_error("gpk_oz_gemm_tn: groups 5..8 (15 / 21 / 28 / 34 digit products)"); return 2; }
    oz_set_groups(groups);
    if (M < NB || M % NB || N < NB || N % NB || K < NB || K % NB) { set_error("gpk_oz_gemm_tn: M, N, K must be positive multiples of %d", NB); return 2; }
    set_stream((cudaStream_t)stream);
    return launch_oz_gemm_tn((const uint8_t*)A, (const uint8_t*)B, M, N, K, sa, sb, C, ldc, dbg, (cudaStream_t)stream);
}

int gpk_set_groups(int groups, int min_particles_per_group) {
    if (groups == 0 && min_particles_per_group == 0) {        // back to the built-in rule
        g_groups = 2; g_min_group = 16;
        return 0;
    }
    if (groups < 1 || groups > MAX_GROUPS || min_particles_per_group < 1) { set_error("gpk_set_groups: groups 1..%d, at least 1 particle per group", MAX_GROUPS); return 2; }
    g_groups = groups;
    g_min_group = min_particles_per_group;
    return 0;
}

// test hook: a kernel that traps, to show how a device-side fault (e.g. the pipeline watchdog of the digit GEMM, mbar_wait)
// reaches the caller: every later entry point returns a non-zero status and gpk_last_error_string() names the CUDA error
__global__ void k_debug_trap() { __trap(); }
int gpk_debug_trap(void* stream) {
    set_stream((cudaStream_t)stream);
    k_debug_trap<<<1, 32, 0, (cudaStream_t)stream>>>();
    if (check_launch("k_debug_trap")) return 1;
    cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
    if (e != cudaSuccess) { set_error("k_debug_trap: %s", cudaGetErrorString(e)); return 1; }
    return 0;
}

int gpk_timing_begin(void* stream) {
    g_cur_stream = (cudaStream_t)stream;
    for (auto& r : g_recs) g_pool.push_back(r.ev);
    g_recs.clear();
    g_cls_names.clear(); g_cls_ms.clear(); g_cls_n.clear();
    g_timing = true;
    timing_mark("(begin)");
    return 0;
}

int gpk_timing_end(void) {
    g_timing = false;
    if (g_recs.empty()) return 0;
    if (cudaEventSynchronize(g_recs.back().ev) != cudaSuccess) { set_error("timing: event sync failed"); return 1; }
    for (size_t i = 1; i < g_recs.size(); ++i) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, g_recs[i - 1].ev, g_recs[i].ev);
        g_cls_ms[g_recs[i].cls] += ms;
        g_cls_n[g_recs[i].cls] += 1;
    }
    return 0;
}

int gpk_timing_query(int idx, char* name_out, int name_cap, double* ms_out, long long* count_out) {
    if (idx < 0 || idx >= (int)g_cls_names.size()) return 1;
    snprintf(name_out, name_cap, "%s", g_cls_names[idx].c_str());
    *ms_out = g_cls_ms[idx];
    *count_out = g_cls_n[idx];
    return 0;
}

}  // extern "C"
