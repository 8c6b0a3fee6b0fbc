// Emulated-FP64 contractions on the 5th-generation tensor cores (tcgen05 / TMEM), sm_100a.
//
// sm_100a has no FP64 kind of tcgen05.mma; the FP64 tensor instruction is the warp-level DMMA (37 TFLOP/s, see
// gpk_gemm.cu).  The two contractions that make up 85 % of a loss evaluation (GP_func.py:22 Cholesky trailing
// updates and GP_func.py:38 the N-right-hand-side solve, restated as a triangular inverse) are instead computed
// here by an error-free split of each FP64 operand into OZ_NS = 7 signed 8-bit digits (Ozaki scheme):
//
//     x = sigma * u,  |u| <= 0.98,  sigma a power of two (row scale; a-priori bounds, see the slicers)
//     v = rint(u * 2^55) = sum_s b_s 2^(48 - 8 s),   b_s in [-128, 127]   (balanced digits: bytes of v + 0x..8080)
//     <x_a, x_b> = sigma_a sigma_b 2^-14 sum_g 2^(-8 g) acc_g,   acc_g = sum_k sum_{s+t=g} b_s[a,k] b_t[b,k]
//
// acc_g are exact int32 sums produced by tcgen05.mma.kind::i8; the 34 digit products with g = s+t <= 7 are kept
// (the dropped ones are mean-zero and of weight <= 2^-64 sigma_a sigma_b per term, below the rounding level of an
// FP64 dot product of this length; a 28-product mode, g <= 6, is selectable).  8 accumulators x 64 columns = all 512
// TMEM columns hold one 128 x 64 output tile; the epilogue rebuilds the FP64 value with two fused multiply-adds and
// applies it to the FP64 matrix in HBM.
//
// Operand storage ("sliced matrix"): for every 128-row block and every 32-wide K chunk one 28 KB record
//     [7 digits][16 row groups][2 k-halves][8 rows][16 bytes]
// which is exactly the shared-memory image tcgen05.mma reads through a no-swizzle K-major descriptor
// (core matrix = 8 rows x 16 B, LBO = 128 B between the k-halves, SBO = 256 B between row groups), so a pipeline
// stage is filled by plain bulk copies (cp.async.bulk, one elected thread) and needs no tensor map.
//
// Instruction shape: a 128 x 64 x 32 MMA is bound by its shared-memory operand reads (58 clk measured instead of 32,
// tools/umma_probe.cu).  The 64-row B digits t = 0.. of a stage are therefore stored back to back, so that ONE
// instruction with N = 64 * count multiplies the A digit s with `count` consecutive B digits and lands in the
// accumulators of the groups s + t0 .. s + t0 + count - 1, which are adjacent TMEM columns: 10-11 instructions of
// N <= 256 per K chunk instead of 28-34 of N = 64.
#include "gpk_ozdig.cuh"

#include <cstdlib>

namespace gpk {

constexpr int OZ_STAGES = 5;
#ifndef OZ_PREFETCH_CHUNKS
#define OZ_PREFETCH_CHUNKS 0
#endif
constexpr int OZ_PREFETCH = OZ_PREFETCH_CHUNKS;   // A chunks prefetched into L2 beyond the shared-memory ring (0 = off)
constexpr int OZ_THREADS = 320;                 // warp 0 producer, warp 1 MMA issuer, warps 2..9 epilogue
constexpr int OZ_TMEM_COLS = 512;
constexpr int OZ_TILE_N = 64;
constexpr size_t OZ_SMEM_BYTES = (size_t)OZ_STAGES * OZ_STAGE + 256;
// kind::i8, D = S32, A = B = signed 8 bit, both K-major, M = 128; the N field (bits 17..22, N >> 3) is added per MMA
constexpr uint32_t OZ_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((128u >> 4) << 24);
// no-swizzle K-major shared-memory descriptor without the start address: LBO = 128 B, SBO = 256 B, version 1
constexpr uint64_t OZ_SDESC = ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46);

// ---- PTX wrappers ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// A wait that can never hang the device: a barrier that does not complete within ~2 s traps the kernel.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) __trap();
    }
}
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile(
        "{\n.reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

// exact int64 -> double for |v| < 2^51 without the slow I2F path
__device__ __forceinline__ double ll2d_exact(long long v) {
    return __longlong_as_double(0x4338000000000000LL + v) - 6755399441055744.0;
}

// Exact double of the integer a0 2^16 + a1 2^8 + a2 (int32 accumulators, |value| < 2^48).  The register pair
// magic_hi : (a2 ^ 0x80000000), magic_hi = 0x43380000, is the double 2^52 + 2^51 + 2^31 + a2; the other two terms are added
// to it as 64-bit integers by two 32 x 32 + 64 multiply-adds, which keeps the mantissa field inside (0, 2^52); one FP64
// subtraction of 2^52 + 2^51 + 2^31 finishes the conversion.  magic_hi arrives as a run-time value: with the constant
// visible the compiler moves its addition behind the products and spends a carry chain per conversion.
__device__ __forceinline__ double oz_acc3(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t magic_hi) {
    long long w = (long long)(((unsigned long long)magic_hi << 32) | (unsigned long long)(a2 ^ 0x80000000u));
    w += (long long)(int)a1 * 256;
    w += (long long)(int)a0 * 65536;
    return __longlong_as_double(w) - 6755401588539392.0;      // 2^52 + 2^51 + 2^31
}
__device__ __forceinline__ double oz_acc2(uint32_t a0, uint32_t a1, uint32_t magic_hi) {
    long long w = (long long)(((unsigned long long)magic_hi << 32) | (unsigned long long)(a1 ^ 0x80000000u));
    w += (long long)(int)a0 * 256;
    return __longlong_as_double(w) - 6755401588539392.0;
}

enum OzEpi { OZ_SUB = 0, OZ_STORE = 1, OZ_STORE_NEG = 2, OZ_SUMSQ = 3 };

#ifdef OZ_PROFILE
// Instrumented build only (tools/oz_stall.py): where the MMA thread of every CTA spends its cycles, summed per EPI kind:
// [epi][0] total loop, [1] waiting for operands (full barriers), [2] waiting for the epilogue (freebar), [3] waiting for
// the tile queue, [4] issuing MMAs, [5] tiles, [6] K chunks, [7] CTAs
__device__ unsigned long long g_oz_prof[4][8];
#define OZ_PROF_T(var) const long long var = clock64()
#else
#define OZ_PROF_T(var)
#endif

struct OzTile {
    const uint8_t* a;     // first chunk of the A row block (128 rows), chunks are consecutive in K
    const uint8_t* b;     // first chunk of the B row block + half * OZ_DIGIT_HALF (64 rows of every digit)
    int nk;               // number of 32-wide K chunks
    double* c;            // &C[row0][col0] of the 128 x 64 FP64 tile
    int ldc;
    int store;            // 0: the FP64 result is not written (only its digits / maxima are wanted downstream)
    const double* sa;     // row scales sigma of the 128 A rows, or nullptr
    double sa_const;      // multiplies every row scale (use it alone when sa == nullptr)
    const double* sb;     // scales of the 64 B rows, or nullptr (= 1)
    int* dbg;             // optional raw accumulator dump [NG][128][64]
    // optional fused slicing of the tile's FP64 output (the next contraction's operand):
    uint8_t* dig;          // record of (output row block, K chunk of the tile's first column); nullptr = none
    const double* dig_rs;  // 128 row multipliers or nullptr
    double dig_rconst;     // u = value * dig_rs[row] * dig_rconst * dig_cs[col], |u| <= 0.98
    const double* dig_cs;  // 64 column multipliers or nullptr
    // optional row maxima of |output * amax_cs[col]| (the row scale of the next slicing pass): atomicMax on the bit
    // pattern of the non-negative double, amax[row]
    unsigned long long* amax;
    const double* amax_cs;
    // OZ_SUMSQ only: column sums of squares over each 32-row quadrant, red[q * red_ld + col] (no FP64 tile is stored)
    double* red;
    size_t red_ld;
};

// The MMAs of one K chunk.  NG = number of digit groups kept (g = s + t < NG): 7 -> 28 digit products, 8 -> 34.
// A digit s meets the B digits t = 0 .. min(7, NG - s) - 1 in pieces of at most 4 digits (N <= 256); the piece
// starting at t0 accumulates into the TMEM columns of the groups s + t0 ...  With NG = 8 the digit s = 4 goes first
// so that at the first chunk every accumulator is initialised by an instruction that covers it completely.
template <int NG, int TN>
__device__ __forceinline__ void oz_issue_chunk(uint32_t tmem, uint64_t da, uint64_t db, bool first) {
#pragma unroll
    for (int si = 0; si < OZ_NS; ++si) {
        const int s = (NG == 8) ? (si == 0 ? 4 : (si <= 4 ? si - 1 : si)) : si;
        const int total = (NG - s < OZ_NS) ? (NG - s) : OZ_NS;
#pragma unroll
        for (int t0 = 0; t0 < total; t0 += 4) {
            const int cnt = (total - t0 < 4) ? (total - t0) : 4;
            const bool init_piece = (NG == 8) ? (s == 4 || (s == 0 && t0 == 0)) : (s == 0);
            const uint32_t idesc = OZ_IDESC | ((uint32_t)((cnt * TN) >> 3) << 17);
            umma_i8(tmem + (s + t0) * TN, da + (uint64_t)(s * (OZ_DIGIT >> 4)),
                    db + (uint64_t)(t0 * ((TN * 32) >> 4)), idesc, (init_piece && first) ? 0u : 1u);
        }
    }
}

// Persistent tile loop with a dynamic scheduler: the producer thread draws 128 x 64 tiles from a global counter and
// hands the ids to the MMA thread and the epilogue warps through a small shared-memory queue, so CTAs that start late
// (other streams' kernels still on their SM) or draw long tiles simply take fewer.  f.make(tile, t) describes a tile
// (false = nothing to do).  The operand ring runs across tile boundaries: the next tile's first stages load while the
// epilogue drains the accumulators (all 512 TMEM columns hold one tile: MMA and epilogue of one CTA alternate).
constexpr int OZ_QDEPTH = 4;

// TN = tile width: 64 (one tile fills TMEM: 8 groups x 64 columns) or 32 (two accumulator sets of 256 columns: the epilogue
// of one tile overlaps the MMAs of the next; used by the 4-chunk panel contractions, whose tiles are too short to amortise
// the MMA / epilogue hand-off).
template <int EPI, int NG, int TN, class F>
__device__ __forceinline__ void oz_run(const F& f, int ntiles, int* counter, uint8_t* smem) {
    static_assert(TN == 64 || TN == 32, "tile width");
    static_assert(NG >= 5 && NG <= 8, "digit groups kept");
    static_assert(EPI != OZ_SUMSQ || TN == 64, "the sum-of-squares epilogue is written for 64-column tiles");
    constexpr uint32_t NSETS = (TN == 32) ? 2u : 1u;        // accumulator sets in TMEM
    constexpr uint32_t SET_COLS = NG * TN;
    constexpr uint32_t B_DIGIT = TN * 32;                   // bytes of one B digit of a stage (TN rows x 32 k)
    constexpr int ND = (NG < OZ_NS) ? NG : OZ_NS;           // leading digits of each operand that meet a partner (s + t < NG)
    constexpr uint32_t STAGE_TX = ND * OZ_DIGIT + ND * B_DIGIT;
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    uint8_t* ctrl = smem + (size_t)OZ_STAGES * OZ_STAGE;
    const uint32_t full0 = smem_u32(ctrl);                 // OZ_STAGES x 8 B
    const uint32_t empty0 = full0 + 8 * OZ_STAGES;         // OZ_STAGES x 8 B
    const uint32_t accbar = empty0 + 8 * OZ_STAGES;        // 2 x 8 B: accumulators of set s complete (MMA -> epilogue)
    const uint32_t freebar = accbar + 16;                  // 2 x 8 B: accumulators of set s drained (epilogue -> MMA)
    const uint32_t qfull0 = freebar + 16;                  // OZ_QDEPTH x 8 B: tile id written (producer -> consumers)
    const uint32_t qempty0 = qfull0 + 8 * OZ_QDEPTH;       // OZ_QDEPTH x 8 B: tile id read by every consumer
    volatile int* qtile = reinterpret_cast<volatile int*>(ctrl + 8 * (2 * OZ_STAGES + 4 + 2 * OZ_QDEPTH));
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(ctrl + 8 * (2 * OZ_STAGES + 4 + 2 * OZ_QDEPTH) + 4 * OZ_QDEPTH);
    const uint32_t stage0 = smem_u32(smem);

    if (tid == 0) {
        for (int s = 0; s < OZ_STAGES; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(accbar + 8 * b, 1);
            mbar_init(freebar + 8 * b, OZ_THREADS - 64);
        }
        for (int q = 0; q < OZ_QDEPTH; ++q) {
            mbar_init(qfull0 + 8 * q, 1);
            mbar_init(qempty0 + 8 * q, 1 + OZ_THREADS - 64);      // the MMA thread and every epilogue thread
        }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         smem_u32((const void*)tmem_slot)),
                     "r"((uint32_t)OZ_TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            OzTile t;
            uint32_t kg = 0;                     // chunk counter across tiles = position in the ring
            for (uint32_t n = 0;; ++n) {
                const uint32_t qs = n % OZ_QDEPTH;
                mbar_wait(qempty0 + 8 * qs, ((n / OZ_QDEPTH) & 1u) ^ 1u);
                int tile = atomicAdd(counter, 1);
                if (tile >= ntiles) tile = -1;
                qtile[qs] = tile;
                mbar_arrive(qfull0 + 8 * qs);
                if (tile < 0) break;
                if (!f.make(tile, t)) continue;
            if (OZ_PREFETCH > 0) {               // A chunks beyond the ring: HBM -> L2 ahead of the copy that needs them
                for (int k = OZ_STAGES; k < OZ_STAGES + OZ_PREFETCH && k < t.nk; ++k)
                    bulk_prefetch_l2(t.a + (size_t)k * OZ_CHUNK, OZ_CHUNK);
            }
            for (int k = 0; k < t.nk; ++k, ++kg) {
                const uint32_t s = kg % OZ_STAGES;
                const uint32_t ph = (kg / OZ_STAGES) & 1u;
                if (OZ_PREFETCH > 0 && k + OZ_STAGES + OZ_PREFETCH < t.nk)
                    bulk_prefetch_l2(t.a + (size_t)(k + OZ_STAGES + OZ_PREFETCH) * OZ_CHUNK, OZ_CHUNK);
                mbar_wait(empty0 + 8 * s, ph ^ 1u);
                mbar_expect_tx(full0 + 8 * s, STAGE_TX);
                const uint32_t dst = stage0 + s * OZ_STAGE;
                bulk_g2s(dst, t.a + (size_t)k * OZ_CHUNK, ND * OZ_DIGIT, full0 + 8 * s);      // digit-major record: a prefix
#pragma unroll
                for (int d = 0; d < ND; ++d)          // TN rows of every B digit, stored back to back
                    bulk_g2s(dst + OZ_CHUNK + d * B_DIGIT, t.b + (size_t)k * OZ_CHUNK + d * OZ_DIGIT, B_DIGIT,
                             full0 + 8 * s);
            }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            OzTile t;
            uint32_t kg = 0, it = 0;
#ifdef OZ_PROFILE
            long long pf_full = 0, pf_free = 0, pf_q = 0, pf_issue = 0, pf_tiles = 0, pf_chunks = 0;
            const long long pf_start = clock64();
#endif
            for (uint32_t n = 0;; ++n) {
                const uint32_t qs = n % OZ_QDEPTH;
                OZ_PROF_T(q0);
                mbar_wait(qfull0 + 8 * qs, (n / OZ_QDEPTH) & 1u);
                const int tile = qtile[qs];
                mbar_arrive(qempty0 + 8 * qs);
                if (tile < 0) break;
                if (!f.make(tile, t)) continue;
                OZ_PROF_T(q1);
                const uint32_t set = it % NSETS, use = it / NSETS;
                mbar_wait(freebar + 8 * set, (use & 1u) ^ 1u);      // the epilogue has drained this set's previous tile
                tc_fence_after();
                OZ_PROF_T(q2);
#ifdef OZ_PROFILE
                pf_q += q1 - q0; pf_free += q2 - q1; ++pf_tiles;
#endif
                ++it;
            for (int k = 0; k < t.nk; ++k, ++kg) {
                const uint32_t s = kg % OZ_STAGES;
                const uint32_t ph = (kg / OZ_STAGES) & 1u;
                OZ_PROF_T(c0);
                mbar_wait(full0 + 8 * s, ph);
                tc_fence_after();
                OZ_PROF_T(c1);
                const uint64_t da = OZ_SDESC | (uint64_t)(((stage0 + s * OZ_STAGE) >> 4) & 0x3FFF);
                const uint64_t db = OZ_SDESC | (uint64_t)(((stage0 + s * OZ_STAGE + OZ_CHUNK) >> 4) & 0x3FFF);
                oz_issue_chunk<NG, TN>(tmem + set * SET_COLS, da, db, k == 0);
                umma_commit(empty0 + 8 * s);     // frees the stage when these MMAs have read it
#ifdef OZ_PROFILE
                pf_full += c1 - c0; pf_issue += clock64() - c1; ++pf_chunks;
#endif
            }
            umma_commit(accbar + 8 * set);       // accumulators complete
            }
#ifdef OZ_PROFILE
            atomicAdd(&g_oz_prof[EPI][0], (unsigned long long)(clock64() - pf_start));
            atomicAdd(&g_oz_prof[EPI][1], (unsigned long long)pf_full);
            atomicAdd(&g_oz_prof[EPI][2], (unsigned long long)pf_free);
            atomicAdd(&g_oz_prof[EPI][3], (unsigned long long)pf_q);
            atomicAdd(&g_oz_prof[EPI][4], (unsigned long long)pf_issue);
            atomicAdd(&g_oz_prof[EPI][5], (unsigned long long)pf_tiles);
            atomicAdd(&g_oz_prof[EPI][6], (unsigned long long)pf_chunks);
            atomicAdd(&g_oz_prof[EPI][7], 1ull);
#endif
        }
    } else {
        // epilogue warps 2..9: TMEM lane quadrant = warp % 4, one output row and CW = TN / 2 of the tile's columns per thread
        constexpr int CW = TN / 2;
        const int q = warp & 3;
        const int ch = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        const uint32_t magic_hi = 0x43380000u + ((uint32_t)ntiles >> 31);     // = 0x43380000 (ntiles >= 0), opaque to the compiler
        OzTile t;
        uint32_t it = 0;
        for (uint32_t n = 0;; ++n) {
        const uint32_t qs = n % OZ_QDEPTH;
        mbar_wait(qfull0 + 8 * qs, (n / OZ_QDEPTH) & 1u);
        const int tile = qtile[qs];
        mbar_arrive(qempty0 + 8 * qs);
        if (tile < 0) break;
        if (!f.make(tile, t)) continue;
        const double rs = (t.sa ? t.sa[row] : 1.0) * t.sa_const * 0x1p-70;       // 2^-14 of the digit scaling, 2^-56 of x
        double* crow = t.c + (size_t)row * t.ldc + ch * CW;
        const double* sbp = t.sb ? t.sb + ch * CW : nullptr;
        const double drs = (t.dig_rs ? t.dig_rs[row] : 1.0) * t.dig_rconst * 0x1p55;
        const double* dcs = t.dig_cs ? t.dig_cs + ch * CW : nullptr;
        // digit record of the output: a 64-column tile covers two K chunks of the next contraction (one per column half), a
        // 32-column tile one chunk, in which the column half selects the 16-byte unit
        uint8_t* drow = t.dig ? t.dig + (TN == 64 ? (size_t)ch * OZ_CHUNK : (size_t)ch * 128) + (row >> 3) * 256 + (row & 7) * 16
                              : nullptr;
        // phase A (TMEM busy): accumulators -> FP64 output values in registers; phase B (after the accumulators are
        // released to the MMA thread): global stores and digit emission, overlapping the next tile's MMAs
        double val[CW];
        if (EPI == OZ_SUB) {                     // the tile is final input data: fetch it while the MMAs run
#pragma unroll
            for (int j = 0; j < CW / 4; ++j) ldg_f64x4(crow + 4 * j, val[4 * j], val[4 * j + 1], val[4 * j + 2], val[4 * j + 3]);
        }
        const uint32_t set = it % NSETS, use = it / NSETS;
        mbar_wait(accbar + 8 * set, use & 1u);
        tc_fence_after();
        ++it;
        const uint32_t taddr = tmem + set * SET_COLS + ((uint32_t)(q * 32) << 16) + ch * CW;
#pragma unroll
        for (int cc = 0; cc < CW; cc += 8) {
            uint32_t a[NG][8];
#pragma unroll
            for (int g = 0; g < NG; ++g) tmem_ld8(taddr + g * TN + cc, a[g]);
            tmem_wait_ld();
            if (t.dbg) {
#pragma unroll
                for (int g = 0; g < NG; ++g)
#pragma unroll
                    for (int j = 0; j < 8; ++j) t.dbg[(g * 128 + row) * TN + ch * CW + cc + j] = (int)a[g][j];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                // three exact integers hi = a0 2^16 + a1 2^8 + a2, mid = a3 2^16 + a4 2^8 + a5, lo = a6 2^8 + a7, each built
                // directly inside the bit pattern of a double (oz_acc3 / oz_acc2) and finished by one FP64 subtraction:
                // this phase is issue bound (8 warps on 4 schedulers hold TMEM while they convert)
                const double hi = oz_acc3(a[0][j], a[1][j], a[2][j], magic_hi);
                double x;                                                     // units of 2^-56 (rs carries the factor)
                if constexpr (NG >= 7) {
                    const double mid = oz_acc3(a[3][j], a[4][j], a[5][j], magic_hi);
                    const double lo = (NG == 8) ? oz_acc2(a[6][j], a[NG - 1][j], magic_hi) : oz_acc2(a[6][j], 0u, magic_hi);
                    x = fma(hi, 0x1p40, fma(mid, 0x1p16, lo));
                } else if constexpr (NG == 6) {
                    x = fma(hi, 0x1p40, oz_acc3(a[3][j], a[4][j], a[5][j], magic_hi) * 0x1p16);
                } else {
                    x = fma(hi, 0x1p40, oz_acc2(a[3][j], a[4][j], magic_hi) * 0x1p24);
                }
                const double v = x * (rs * (sbp ? sbp[cc + j] : 1.0));
                if (EPI == OZ_SUB) val[cc + j] -= v;
                else if (EPI == OZ_STORE) val[cc + j] = v;
                else if (EPI == OZ_STORE_NEG) val[cc + j] = -v;
                else val[cc + j] = v * v;
            }
        }
        tc_fence_before();
        mbar_arrive(freebar + 8 * set);          // this set of accumulators may be overwritten
        if (EPI != OZ_SUMSQ) {
            if (t.store) {
#pragma unroll
                for (int j = 0; j < CW / 4; ++j) stg_f64x4(crow + 4 * j, val[4 * j], val[4 * j + 1], val[4 * j + 2], val[4 * j + 3]);
            }
            if (t.amax) {
                double m = 0.0;
#pragma unroll
                for (int j = 0; j < CW; ++j) m = fmax(m, fabs(val[j] * t.amax_cs[ch * CW + j]));
                atomicMax(t.amax + row, (unsigned long long)__double_as_longlong(m));
            }
            if (drow) {                          // digits of the stored values: one 16-byte unit (16 consecutive k) per store
#pragma unroll
                for (int cc = 0; cc < CW; cc += 16) {
                    uint32_t lo[16], hi[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const double u = val[cc + j] * (drs * (dcs ? dcs[cc + j] : 1.0));
                        const long long w = (__double2ll_rn(u) + OZ_BIAS) ^ OZ_BIAS;
                        lo[j] = (uint32_t)w;
                        hi[j] = (uint32_t)(w >> 32);
                    }
                    uint8_t* d = drow + (cc >> 4) * 128;
#pragma unroll
                    for (int s = 0; s < ND; ++s) {
                        const int bytepos = 6 - s;
                        const uint32_t* w = (bytepos >= 4) ? hi : lo;
                        uint4 o4;
                        o4.x = pack_digit(w, bytepos & 3);
                        o4.y = pack_digit(w + 4, bytepos & 3);
                        o4.z = pack_digit(w + 8, bytepos & 3);
                        o4.w = pack_digit(w + 12, bytepos & 3);
                        *reinterpret_cast<uint4*>(d + s * OZ_DIGIT) = o4;
                    }
                }
            }
        }
        if constexpr (EPI == OZ_SUMSQ && TN == 64) {
            // butterfly: 32 columns x 32 rows (lanes) -> lane l holds the sum over the quadrant's rows of column l
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1) {
                const bool up = (lane & step) != 0;
#pragma unroll
                for (int i = 0; i < step; ++i) {
                    const double send = up ? val[i] : val[i + step];
                    const double keep = up ? val[i + step] : val[i];
                    val[i] = keep + __shfl_xor_sync(0xffffffffu, send, step);
                }
            }
            t.red[(size_t)q * t.red_ld + ch * 32 + lane] = val[0];
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"((uint32_t)OZ_TMEM_COLS)
                     : "memory");
    }
}

// ---- tile descriptions (functors) and the one persistent kernel ---------------------------------------------------
template <int EPI, int NG, int TN, class F>
__global__ void __launch_bounds__(OZ_THREADS, 1) k_oz(const F f, int ntiles, int* counter) {
    extern __shared__ __align__(128) uint8_t oz_smem[];
    oz_run<EPI, NG, TN, F>(f, ntiles, counter, oz_smem);
}

__device__ __forceinline__ void oz_no_digits(OzTile& t) {
    t.dbg = nullptr;
    t.dig = nullptr; t.dig_rs = nullptr; t.dig_rconst = 1.0; t.dig_cs = nullptr;
    t.amax = nullptr; t.amax_cs = nullptr;
    t.store = 1;
    t.red = nullptr; t.red_ld = 0;
}

// Left-looking Cholesky update of block column J: A_IJ -= sum_{K<J} L_IK L_JK^T from the digits of L.
// tile = p * nx + 2 * (I - J) + half.  sdig != nullptr: the epilogue also leaves the digits of
// S_IJ / (rowsig_i rowsig_j), I > J, the A operand of the panel solve.
struct OzSyrkTiles {
    double* mat; int Np, J, nx;
    const uint8_t* slices; size_t slice_stride;
    const double* rowsig; const double* rowinv;
    uint8_t* sdig; size_t sdig_stride;
    const unsigned char* active;
    __device__ __forceinline__ bool make(int tile, OzTile& t) const {
        const int p = tile / nx, bx = tile - p * nx;
        if (active && !active[p]) return false;
        const int I = J + (bx >> 1), h = bx & 1;
        const int kch = Np / 32;
        const uint8_t* S = slices + (size_t)p * slice_stride;
        t.a = S + (size_t)I * kch * OZ_CHUNK;
        t.b = S + (size_t)J * kch * OZ_CHUNK + (size_t)h * OZ_DIGIT_HALF;
        t.nk = 4 * J;
        t.c = mat + (size_t)p * Np * Np + (size_t)I * NB * Np + (size_t)J * NB + h * OZ_TILE_N;
        t.ldc = Np;
        t.sa = rowsig + I * NB;
        t.sa_const = 1.0;
        t.sb = rowsig + J * NB + h * OZ_TILE_N;
        oz_no_digits(t);
        if (sdig && I > J) {
            t.dig = sdig + (size_t)p * sdig_stride + ((size_t)I * 4 + 2 * h) * OZ_CHUNK;
            t.dig_rs = rowinv + I * NB;
            t.dig_cs = rowinv + J * NB + h * OZ_TILE_N;
            t.store = 0;       // S_IJ (I > J) is read again only as digits (by the TRSM); S_JJ goes to the diagonal kernel in FP64
        }
        return true;
    }
};

// Phase-2 update W_IJ = - sum_{k in [(J+1)NB, (I+1)NB)} (W gamma)[I-block][k] * (Mt / gamma)[:, k]^T,
// tile = p * nx + 2 * (T - 1 - I) + half (deepest contraction first).  emit: the epilogue also writes the digits of
// W_IJ gamma / 2 for the later steps J' < J.
struct OzInvuTiles {
    double* mat; int Np, T, J, nx;
    uint8_t* slices; size_t slice_stride;
    const uint8_t* mtslices; size_t mt_stride;
    const double* mtsig; const double* gam; int emit;
    const unsigned char* active;
    __device__ __forceinline__ bool make(int tile, OzTile& t) const {
        const int p = tile / nx, bx = tile - p * nx;
        if (active && !active[p]) return false;
        const int I = T - 1 - (bx >> 1), h = bx & 1;
        const int kch = Np / 32;
        t.a = slices + (size_t)p * slice_stride + ((size_t)I * kch + 4 * (J + 1)) * OZ_CHUNK;
        t.b = mtslices + (size_t)p * mt_stride + (size_t)(4 * (J + 1)) * OZ_CHUNK + (size_t)h * OZ_DIGIT_HALF;
        t.nk = 4 * (I - J);
        t.c = mat + (size_t)p * Np * Np + (size_t)I * NB * Np + (size_t)J * NB + h * OZ_TILE_N;
        t.ldc = Np;
        t.sa = nullptr;
        t.sa_const = 2.0;                              // W gamma is sliced with sigma = 2
        t.sb = mtsig + (size_t)p * NB + h * OZ_TILE_N;
        oz_no_digits(t);
        if (emit && J > 0) {
            t.dig = slices + (size_t)p * slice_stride + ((size_t)I * kch + 4 * J + 2 * h) * OZ_CHUNK;
            t.dig_rconst = 0.5;
            t.dig_cs = gam + J * NB + h * OZ_TILE_N;
        }
        return true;
    }
};

// Generic C = A B^T over sliced operands (test / library primitive): tile = mb * nx + nb, nx = N / 64.
struct OzGemmTiles {
    const uint8_t* A; const uint8_t* B; int kch, nx;
    const double* sa; const double* sb; double* C; int ldc; int* dbg; int ng;
    __device__ __forceinline__ bool make(int tile, OzTile& t) const {
        const int mb = tile / nx, nb = tile - mb * nx;
        t.a = A + (size_t)mb * kch * OZ_CHUNK;
        t.b = B + (size_t)(nb >> 1) * kch * OZ_CHUNK + (size_t)(nb & 1) * OZ_DIGIT_HALF;
        t.nk = kch;
        t.c = C + (size_t)mb * NB * ldc + (size_t)nb * OZ_TILE_N;
        t.ldc = ldc;
        t.sa = sa ? sa + mb * NB : nullptr;
        t.sa_const = 1.0;
        t.sb = sb ? sb + nb * OZ_TILE_N : nullptr;
        oz_no_digits(t);
        t.dbg = dbg ? dbg + (size_t)tile * (ng * 128 * OZ_TILE_N) : nullptr;
        return true;
    }
};

// particle-independent scales: row bound of L (|L_ik| <= sqrt(K_ii) = sqrt(1 + e_i^2)) and the column scaling
// gamma_k = 2^floor(log2 e_k) of W = L^-1 (|W_ik| <= 1 / e_k because K >= diag(e^2))
__global__ void k_oz_prep(const double* __restrict__ e, int N, int Np, double* __restrict__ rowsig,
                          double* __restrict__ rowinv, double* __restrict__ gam, double* __restrict__ gaminv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Np) return;
    double sig = 2.0, g = 1.0;
    if (i < N) {
        const double ei = fabs(e[i]);
        sig = oz_scale_for(sqrt(1.0 + ei * ei));
        g = (ei > 1e-290 && ei < 1e290) ? scalbn(1.0, ilogb(ei)) : 1.0;     // e = 0 is outside this mode's contract
    }
    rowsig[i] = sig;
    rowinv[i] = 1.0 / sig;
    gam[i] = g;
    gaminv[i] = 1.0 / g;
}

// digits of the block column J of L (rows below the diagonal block): blockIdx.x -> I = J + 1 + bx
__global__ void __launch_bounds__(256)
k_oz_slice_l(const double* __restrict__ mat, int Np, int J, const double* __restrict__ rowinv,
             uint8_t* __restrict__ slices, size_t slice_stride, const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int I = J + 1 + blockIdx.x;
    const int kch = Np / 32;
    oz_slice_tile(mat + (size_t)p * Np * Np + (size_t)I * NB * Np + (size_t)J * NB, Np, rowinv + I * NB, 1.0, nullptr,
                  slices + (size_t)p * slice_stride + ((size_t)I * kch + 4 * J) * OZ_CHUNK);
}

// digits of W gamma / 2: diag != 0 -> diagonal blocks I = 1 + bx; else block column J, rows I = J + 1 + bx
__global__ void __launch_bounds__(256)
k_oz_slice_w(const double* __restrict__ mat, int Np, int J, int diag, const double* __restrict__ gam,
             uint8_t* __restrict__ slices, size_t slice_stride, const unsigned char* __restrict__ active) {
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    const int I = diag ? 1 + blockIdx.x : J + 1 + blockIdx.x;
    const int Jc = diag ? I : J;
    const int kch = Np / 32;
    oz_slice_tile(mat + (size_t)p * Np * Np + (size_t)I * NB * Np + (size_t)Jc * NB, Np, nullptr, 0.5, gam + Jc * NB,
                  slices + (size_t)p * slice_stride + ((size_t)I * kch + 4 * Jc) * OZ_CHUNK);
}

// digits of Mt / gamma: blockIdx.x -> R = J + 1 + bx (128 columns of the panel).  The row scales come from the row maxima
// max_k |Mt[n][k] / gamma_k| the panel GEMM's epilogue collected (mtmax, one buffer per parity of J): every CTA derives the
// 128 scales it needs itself; the first CTA of a particle also publishes them for the inverse update (mtsig) and clears the
// other parity's maxima for the next block column (its last reader, the slicing of column J + 1, has finished).
__global__ void __launch_bounds__(256)
k_oz_slice_mt(const double* __restrict__ mt, int Np, int J, const double* __restrict__ gaminv,
              unsigned long long* __restrict__ mtmax, size_t mtmax_stride, double* __restrict__ mtsig,
              uint8_t* __restrict__ mtslices, size_t mt_stride, const unsigned char* __restrict__ active) {
    __shared__ double s_inv[NB];
    const int p = blockIdx.y;
    if (active && !active[p]) return;
    unsigned long long* cur = mtmax + (size_t)(J & 1) * mtmax_stride + (size_t)p * NB;
    if (threadIdx.x < NB) {
        const double sig = oz_scale_for(__longlong_as_double((long long)cur[threadIdx.x]));
        s_inv[threadIdx.x] = 1.0 / sig;
        if (blockIdx.x == 0) {
            mtsig[(size_t)p * NB + threadIdx.x] = sig;
            mtmax[(size_t)((J + 1) & 1) * mtmax_stride + (size_t)p * NB + threadIdx.x] = 0ull;
        }
    }
    __syncthreads();
    const int R = J + 1 + blockIdx.x;
    oz_slice_tile(mt + (size_t)p * NB * Np + (size_t)R * NB, Np, s_inv, 1.0, gaminv + R * NB,
                  mtslices + (size_t)p * mt_stride + (size_t)(4 * R) * OZ_CHUNK);
}

// generic slicer (test / library primitive): rows x cols FP64 (multiples of 128), grid (cols / 128, rows / 128)
__global__ void __launch_bounds__(256)
k_oz_slice_generic(const double* __restrict__ src, int ld, int kch, const double* __restrict__ rowinv,
                   uint8_t* __restrict__ dst) {
    const int cb = blockIdx.x, rb = blockIdx.y;
    oz_slice_tile(src + (size_t)rb * NB * ld + (size_t)cb * NB, ld, rowinv ? rowinv + rb * NB : nullptr, 1.0, nullptr,
                  dst + ((size_t)rb * kch + 4 * cb) * OZ_CHUNK);
}

// Panel solve on the digit GEMM: L_IJ = S_IJ D_J^T, tile = p * nx + (128 / TN) * (I - J - 1) + column part.  A = digits of
// S_IJ / (rowsig_i rowsig_j) left by the SYRK epilogue, B = digits of D_J rowsig_j; the epilogue stores L_IJ and its
// digits (L / rowsig_i).
template <int TN>
struct OzTrsmTiles {
    double* mat; int Np, T, J, nx;
    const uint8_t* sdig; size_t sdig_stride;
    const uint8_t* ddig; const double* dsc;
    const double* rowsig; const double* rowinv;
    uint8_t* slices; size_t slice_stride;
    const unsigned char* active;
    __device__ __forceinline__ bool make(int tile, OzTile& t) const {
        const int p = tile / nx, bx = tile - p * nx;
        if (active && !active[p]) return false;
        constexpr int PARTS = NB / TN;
        const int I = J + 1 + bx / PARTS, h = bx % PARTS;
        const int kch = Np / 32;
        t.a = sdig + (size_t)p * sdig_stride + (size_t)I * 4 * OZ_CHUNK;
        t.b = ddig + (((size_t)p * T + J) * 2 + 0) * (4 * (size_t)OZ_CHUNK) + (size_t)h * (TN * 32);
        t.nk = 4;
        t.c = mat + (size_t)p * Np * Np + (size_t)I * NB * Np + (size_t)J * NB + h * TN;
        t.ldc = Np;
        t.sa = rowsig + I * NB;
        t.sa_const = 1.0;
        t.sb = dsc + (((size_t)p * T + J) * 2 + 0) * NB + h * TN;
        oz_no_digits(t);
        t.dig = slices + (size_t)p * slice_stride + ((size_t)I * kch + 4 * J + (h * TN) / 32) * OZ_CHUNK;
        t.dig_rs = rowinv + I * NB;
        t.store = 0;           // L_IJ is consumed as digits only (later SYRKs, phase-2 panel); the inverse overwrites the block
        return true;
    }
};

// Phase-2 panel on the digit GEMM: Mt[n][r] = sum_j D_J^T[n][j] L[r][J NB + j], r in block R,
// tile = p * nx + (128 / TN) * (R - J - 1) + column part.
template <int TN>
struct OzInvpTiles {
    double* mt; int Np, T, J, nx;
    const uint8_t* ddig; const double* dsc;
    const uint8_t* slices; size_t slice_stride;
    const double* rowsig;
    unsigned long long* mtmax; const double* gaminv;      // row maxima of Mt / gamma for the panel's row scales
    const unsigned char* active;
    __device__ __forceinline__ bool make(int tile, OzTile& t) const {
        const int p = tile / nx, bx = tile - p * nx;
        if (active && !active[p]) return false;
        constexpr int PARTS = NB / TN;
        const int R = J + 1 + bx / PARTS, h = bx % PARTS;
        const int kch = Np / 32;
        t.a = ddig + (((size_t)p * T + J) * 2 + 1) * (4 * (size_t)OZ_CHUNK);
        t.b = slices + (size_t)p * slice_stride + ((size_t)R * kch + 4 * J) * OZ_CHUNK + (size_t)h * (TN * 32);
        t.nk = 4;
        t.c = mt + (size_t)p * NB * Np + (size_t)R * NB + h * TN;
        t.ldc = Np;
        t.sa = dsc + (((size_t)p * T + J) * 2 + 1) * NB;
        t.sa_const = 1.0;
        t.sb = rowsig + R * NB + h * TN;
        oz_no_digits(t);
        t.amax = mtmax + (size_t)p * NB;
        t.amax_cs = gaminv + R * NB + h * TN;
        return true;
    }
};

// ---- prediction on the digit GEMM (GP_func.py:38-39 for a chunk of grid points) ---------------------------------------
// sd^2 = 1 - |W k_*|^2: tile = I * nx + nb computes (W gamma / 2)[block I] x (k_* / gamma / sigma_B)[64 grid points]^T over
// k < (I+1) NB and reduces the squares over each 32-row quadrant: partial[(4 I + q) * Bp + n].
struct OzPredTiles {
    const uint8_t* wdig; const uint8_t* kdig; int Np, nx;
    const double* sigb; double* partial; size_t Bp;
    __device__ __forceinline__ bool make(int tile, OzTile& t) const {
        const int I = (Np / NB) - 1 - tile / nx, nb = tile % nx;          // deepest contraction first
        const int kch = Np / 32;
        t.a = wdig + (size_t)I * kch * OZ_CHUNK;
        t.b = kdig + (size_t)(nb >> 1) * kch * OZ_CHUNK + (size_t)(nb & 1) * OZ_DIGIT_HALF;
        t.nk = 4 * (I + 1);
        t.c = nullptr;
        t.ldc = 0;
        t.sa = nullptr;
        t.sa_const = 2.0 * sigb[0];
        t.sb = nullptr;
        oz_no_digits(t);
        t.red = partial + (size_t)(4 * I) * Bp + (size_t)nb * OZ_TILE_N;
        t.red_ld = Bp;
        return true;
    }
};

// gamma_k = 2^floor(log2 e_k), sigma_B = scale of max_k 1 / gamma_k, ginv_s[k] = 1 / (gamma_k sigma_B) (0 in the padding)
__global__ void __launch_bounds__(256) k_oz_pred_scales(const double* __restrict__ e, int N, int Np, double* __restrict__ gam,
                                                        double* __restrict__ ginv_s, double* __restrict__ sigb) {
    __shared__ double red[8];
    __shared__ double sig;
    double m = 0.0;
    for (int k = threadIdx.x; k < N; k += 256) {
        const double ek = fabs(e[k]);
        const double g = (ek > 1e-290 && ek < 1e290) ? scalbn(1.0, ilogb(ek)) : 1.0;
        m = fmax(m, 1.0 / g);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) m = fmax(m, red[w]);
        sig = oz_scale_for(m);
        sigb[0] = sig;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < Np; k += 256) {
        double g = 1.0, gi = 0.0;
        if (k < N) {
            const double ek = fabs(e[k]);
            g = (ek > 1e-290 && ek < 1e290) ? scalbn(1.0, ilogb(ek)) : 1.0;
            gi = 1.0 / (g * sig);
        }
        gam[k] = g;
        ginv_s[k] = gi;
    }
}

// digits of the whole lower block triangle of W (columns scaled by gamma, sigma = 2): grid (T, T), blocks J <= I
__global__ void __launch_bounds__(256) k_oz_slice_wall(const double* __restrict__ W, int Np, const double* __restrict__ gam,
                                                       uint8_t* __restrict__ wdig) {
    const int J = blockIdx.x, I = blockIdx.y;
    if (J > I) return;
    oz_slice_tile(W + (size_t)I * NB * Np + (size_t)J * NB, Np, nullptr, 0.5, gam + J * NB,
                  wdig + ((size_t)I * (Np / 32) + 4 * J) * OZ_CHUNK);
}

__device__ __forceinline__ double oz_rbf(double sq1, double sq2, double dot) {
    // kernel_func, GP_func.py:59-65, same operation order as rbf_from_parts in gpk_cov.cu
    double s = __dadd_rn(__dadd_rn(sq1, sq2), __dmul_rn(-2.0, dot));
    s = fmax(s, 0.0);
    return exp(__dmul_rn(-0.5, s));
}

// Cross covariance of a chunk of grid points straight into digit records (rows = grid points, K = training index),
// with the predictive mean mu_b = sum_j k(x_j, x_b) alpha_j (GP_func.py:32,36) fused in.  A CTA owns XC_PTS grid points
// (8 warps x XC_WPTS); a lane owns 16 consecutive j (one 16-byte unit of every digit).  The training-side arrays
// (scaled points, |x|^2, alpha, 1 / (gamma sigma_B)) are staged through shared memory in blocks of 512 j, transposed to
// [i = j % 16][lane = j / 16] with a row pitch of 33 so that both the coalesced fill and the per-lane reads are free of
// bank conflicts (a lane reading its 16 consecutive doubles straight from global memory touches 32 lines per request and
// is bound by the L1 tag rate: profiles/r01_ncu_predict_emu.md).
constexpr int XC_WPTS = 2;
constexpr int XC_PTS = 8 * XC_WPTS;
constexpr int XC_PITCH = 33;
constexpr int XC_ARR = 16 * XC_PITCH;       // doubles per staged array
template <int D>       // D = number of dimensions, compile time: a run-time bound costs MAX_D predicated pairs per element
__global__ void __launch_bounds__(256) k_oz_cross_cov(const double* __restrict__ xs, const double* __restrict__ sq,
                                                      const double* __restrict__ ls, const double* __restrict__ alpha,
                                                      const double* __restrict__ ginv_s, int N, int Np,
                                                      const double* __restrict__ Xfit, long long M, long long m0, int B,
                                                      int Bp, uint8_t* __restrict__ kdig, double* __restrict__ mu_out) {
    constexpr int d = D;
    extern __shared__ __align__(16) double xc_sm[];      // [d + 3][16][33]: xs rows, sq, alpha, ginv_s
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b0 = blockIdx.x * XC_PTS + warp * XC_WPTS;
    const int kch = Np / 32;
    double xb[XC_WPTS][D], sqb[XC_WPTS], acc[XC_WPTS];
#pragma unroll
    for (int q = 0; q < XC_WPTS; ++q) {
        const int b = b0 + q;
        sqb[q] = 0.0;
        acc[q] = 0.0;
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
            xb[q][dd] = 0.0;
            if (b < B) {
                xb[q][dd] = __ddiv_rn(Xfit[(size_t)dd * M + m0 + b], ls[dd]);
                const double v2 = __dmul_rn(xb[q][dd], xb[q][dd]);
                sqb[q] = (dd == 0) ? v2 : __dadd_rn(sqb[q], v2);
            }
        }
    }
    const int narr = d + 3;
    for (int jb = 0; jb < Np; jb += 512) {
        __syncthreads();
        for (int t = threadIdx.x; t < narr * 512; t += 256) {
            const int a = t >> 9, jl = t & 511, j = jb + jl;
            double v = 0.0;
            if (j < N) v = (a < d) ? xs[(size_t)a * Np + j] : (a == d) ? sq[j] : (a == d + 1) ? alpha[j] : ginv_s[j];
            xc_sm[a * XC_ARR + (jl & 15) * XC_PITCH + (jl >> 4)] = v;
        }
        __syncthreads();
        const int j0 = jb + lane * 16;
        if (j0 >= Np) continue;
#pragma unroll
        for (int q = 0; q < XC_WPTS; ++q) {
            const int b = b0 + q;
            if (b >= Bp) continue;
            uint32_t lo[16], hi[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const double* col = xc_sm + i * XC_PITCH + lane;
                double v = 0.0;
                if (j0 + i < N && b < B) {
                    double dot = __dmul_rn(col[0], xb[q][0]);
#pragma unroll
                    for (int dd = 1; dd < D; ++dd) dot = fma(col[dd * XC_ARR], xb[q][dd], dot);
                    v = oz_rbf(col[d * XC_ARR], sqb[q], dot);
                    acc[q] = fma(v, col[(d + 1) * XC_ARR], acc[q]);
                }
                const long long w = (__double2ll_rn(v * col[(d + 2) * XC_ARR] * 0x1p55) + OZ_BIAS) ^ OZ_BIAS;
                lo[i] = (uint32_t)w;
                hi[i] = (uint32_t)(w >> 32);
            }
            uint8_t* dd8 = kdig + (size_t)(b >> 7) * kch * OZ_CHUNK + ((b & 127) >> 3) * 256 + (b & 7) * 16 +
                           (size_t)(j0 >> 5) * OZ_CHUNK + ((j0 >> 4) & 1) * 128;
#pragma unroll
            for (int s = 0; s < OZ_NS; ++s) {
                const int bytepos = 6 - s;
                const uint32_t* w = (bytepos >= 4) ? hi : lo;
                uint4 o;
                o.x = pack_digit(w + 0, bytepos & 3);
                o.y = pack_digit(w + 4, bytepos & 3);
                o.z = pack_digit(w + 8, bytepos & 3);
                o.w = pack_digit(w + 12, bytepos & 3);
                *reinterpret_cast<uint4*>(dd8 + s * OZ_DIGIT) = o;
            }
        }
    }
#pragma unroll
    for (int q = 0; q < XC_WPTS; ++q) {
        double a = acc[q];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0 && b0 + q < B) mu_out[m0 + b0 + q] = a;
    }
}

// ---- launchers ------------------------------------------------------------------------------------------------
template <typename K>
static int oz_smem_attr(K kernel, bool& done) {
    if (!done) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)OZ_SMEM_BYTES);
        if (e != cudaSuccess) { set_error("oz smem attr: %s", cudaGetErrorString(e)); return 1; }
        done = true;
    }
    return 0;
}

size_t oz_slice_bytes(int Np) { return (size_t)(Np / NB) * (Np / 32) * OZ_CHUNK; }
size_t oz_mt_bytes(int Np) { return (size_t)(Np / 32) * OZ_CHUNK; }

int launch_oz_prep(const double* e, int N, int Np, double* rowsig, double* rowinv, double* gam, double* gaminv,
                   cudaStream_t st) {
    k_oz_prep<<<(Np + 255) / 256, 256, 0, st>>>(e, N, Np, rowsig, rowinv, gam, gaminv);
    return check_launch("k_oz_prep");
}

// digit groups kept (g = s + t < NG): 8 -> 34 digit products (default), 7 -> 28, 6 -> 21 (6 digits), 5 -> 15 (5 digits).
// Per calling thread: every C-ABI entry sets it from its own `precision` argument before launching.
static thread_local int g_oz_groups = 8;
void oz_set_groups(int ng) { g_oz_groups = (ng >= 5 && ng <= 8) ? ng : 8; }
int oz_get_groups() { return g_oz_groups; }

// Tile width of the 4-chunk panel contractions (TRSM, phase-2 panel): 64, one accumulator set fills TMEM, MMA and epilogue of
// a CTA alternate.  A 32-column variant with two accumulator sets (k_oz<.., TN = 32>) was measured in round 1
// (profiles/r01_ab_panel_tn.log): the hand-off wait drops from 4000 to 900 cycles per tile, but a 32-column tile costs 5700
// cycles against 10 000 for a 64-column one (MMAs of N <= 128 are bound by the A operand reads): 195.9 vs 192.5 ms per step,
// so it is not instantiated.

static int oz_grid(int ntiles) {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
        // experiment knob: leave some SMs to the panel kernels of the other particle groups (streams)
        if (const char* v = getenv("GPK_OZ_SMS")) { const int n = atoi(v); if (n > 0 && n < sms) sms = n; }
    }
    return ntiles < sms ? ntiles : sms;       // one persistent CTA per SM (215 KB of shared memory, all of TMEM)
}

// launches k_oz<EPI, 7 or 8, TN, F> on one persistent CTA per SM (one smem-attribute call per instantiation)
template <int EPI, int TN = OZ_TILE_N, class F>
static int oz_launch(const F& f, int ntiles, int* counter, cudaStream_t st) {
    static bool done[4] = {false, false, false, false};
    if (ntiles <= 0) return 0;
#define GPK_OZ_GO(NG_)                                                                                   \
    do {                                                                                                  \
        if (oz_smem_attr(k_oz<EPI, NG_, TN, F>, done[NG_ - 5])) return 1;                                 \
        k_oz<EPI, NG_, TN, F><<<oz_grid(ntiles), OZ_THREADS, OZ_SMEM_BYTES, st>>>(f, ntiles, counter);    \
    } while (0)
    switch (g_oz_groups) {
        case 5: GPK_OZ_GO(5); break;
        case 6: GPK_OZ_GO(6); break;
        case 7: GPK_OZ_GO(7); break;
        default: GPK_OZ_GO(8); break;
    }
#undef GPK_OZ_GO
    return 0;
}

size_t oz_panel_bytes(int Np) { return (size_t)(Np / NB) * 4 * OZ_CHUNK; }      // [T][4 chunks]: S digits of one block column
size_t oz_ddig_bytes(int Np) { return (size_t)(Np / NB) * 2 * 4 * OZ_CHUNK; }   // [T][2][4 chunks]: D_J and D_J^T digits

// sdig != nullptr: the epilogue also leaves the digits of S_IJ (I > J) for launch_oz_trsm
int launch_oz_syrk(double* mat, int Np, int T, int J, int P, const uint8_t* slices, const double* rowsig,
                   const double* rowinv, uint8_t* sdig, int* counter, const unsigned char* active, cudaStream_t st) {
    if (J == 0) return 0;
    const OzSyrkTiles f{mat, Np, J, 2 * (T - J), slices, oz_slice_bytes(Np), rowsig, rowinv, sdig, oz_panel_bytes(Np), active};
    if (oz_launch<OZ_SUB>(f, f.nx * P, counter, st)) return 1;
    return check_launch("k_oz_syrk");
}

__global__ void __launch_bounds__(256)
k_oz_slice_d(const double* __restrict__ dinv, const double* __restrict__ dinvT, int T, int J,
             const double* __restrict__ rowsig, uint8_t* __restrict__ dig, double* __restrict__ dsc,
             const unsigned char* __restrict__ active) {
    __shared__ double inv[NB];
    const int p = blockIdx.y, which = blockIdx.x;
    if (active && !active[p]) return;
    oz_slice_d_block(dinv, dinvT, T, J, p, which, rowsig, dig, dsc, inv);
}

int launch_oz_slice_d(const double* dinv, const double* dinvT, int T, int J, int P, const double* rowsig, uint8_t* ddig,
                      double* dsc, const unsigned char* active, cudaStream_t st) {
    k_oz_slice_d<<<dim3(2, P), 256, 0, st>>>(dinv, dinvT, T, J, rowsig, ddig, dsc, active);
    return check_launch("k_oz_slice_d");
}

int launch_oz_trsm(double* mat, int Np, int T, int J, int P, const uint8_t* sdig, const uint8_t* ddig, const double* dsc,
                   const double* rowsig, const double* rowinv, uint8_t* slices, int* counter, const unsigned char* active,
                   cudaStream_t st) {
    if (J >= T - 1) return 0;
    const OzTrsmTiles<64> f{mat, Np, T, J, 2 * (T - 1 - J), sdig, oz_panel_bytes(Np), ddig, dsc, rowsig, rowinv, slices,
                            oz_slice_bytes(Np), active};
    if (oz_launch<OZ_STORE, 64>(f, f.nx * P, counter, st)) return 1;
    return check_launch("k_oz_trsm");
}

int launch_oz_invp(double* mt, int Np, int T, int J, int P, const uint8_t* ddig, const double* dsc, const uint8_t* slices,
                   const double* rowsig, unsigned long long* mtmax, const double* gaminv, int* counter,
                   const unsigned char* active, cudaStream_t st) {
    if (J >= T - 1) return 0;
    const OzInvpTiles<64> f{mt, Np, T, J, 2 * (T - 1 - J), ddig, dsc, slices, oz_slice_bytes(Np), rowsig, mtmax, gaminv, active};
    if (oz_launch<OZ_STORE, 64>(f, f.nx * P, counter, st)) return 1;
    return check_launch("k_oz_invp");
}

int launch_oz_slice_l(const double* mat, int Np, int T, int J, int P, const double* rowinv, uint8_t* slices,
                      const unsigned char* active, cudaStream_t st) {
    if (J >= T - 1) return 0;
    k_oz_slice_l<<<dim3(T - 1 - J, P), 256, 0, st>>>(mat, Np, J, rowinv, slices, oz_slice_bytes(Np), active);
    return check_launch("k_oz_slice_l");
}

int launch_oz_slice_w(const double* mat, int Np, int T, int J, int diag, int P, const double* gam, uint8_t* slices,
                      const unsigned char* active, cudaStream_t st) {
    const int nblk = T - 1 - (diag ? 0 : J);
    if (nblk <= 0) return 0;
    k_oz_slice_w<<<dim3(nblk, P), 256, 0, st>>>(mat, Np, J, diag, gam, slices, oz_slice_bytes(Np), active);
    return check_launch("k_oz_slice_w");
}

int launch_oz_mt(const double* mt, int Np, int T, int J, int P, const double* gaminv, unsigned long long* mtmax,
                 size_t mtmax_stride, double* mtsig, uint8_t* mtslices, const unsigned char* active, cudaStream_t st) {
    if (J >= T - 1) return 0;
    k_oz_slice_mt<<<dim3(T - 1 - J, P), 256, 0, st>>>(mt, Np, J, gaminv, mtmax, mtmax_stride, mtsig, mtslices, oz_mt_bytes(Np), active);
    return check_launch("k_oz_slice_mt");
}

// emit != 0: the epilogue also writes the digits of W_IJ gamma / 2 (replaces launch_oz_slice_w for column J)
int launch_oz_invu(double* mat, int Np, int T, int J, int P, uint8_t* slices, const uint8_t* mtslices,
                   const double* mtsig, const double* gam, int emit, int* counter, const unsigned char* active,
                   cudaStream_t st) {
    if (J >= T - 1) return 0;
    const OzInvuTiles f{mat, Np, T, J, 2 * (T - 1 - J), slices, oz_slice_bytes(Np), mtslices, oz_mt_bytes(Np), mtsig, gam,
                        emit, active};
    if (oz_launch<OZ_STORE_NEG>(f, f.nx * P, counter, st)) return 1;
    return check_launch("k_oz_invu");
}

int launch_oz_slice_generic(const double* src, int ld, int rows, int cols, const double* rowinv, uint8_t* dst,
                            cudaStream_t st) {
    k_oz_slice_generic<<<dim3(cols / NB, rows / NB), 256, 0, st>>>(src, ld, cols / 32, rowinv, dst);
    return check_launch("k_oz_slice_generic");
}

int launch_oz_gemm_tn(const uint8_t* A, const uint8_t* B, int M, int N, int K, const double* sa, const double* sb,
                      double* C, int ldc, int* dbg, cudaStream_t st) {
    static int* counter = nullptr;           // the library primitive owns one tile counter
    if (!counter && cudaMalloc(&counter, sizeof(int)) != cudaSuccess) { set_error("oz counter alloc failed"); return 1; }
    if (cudaMemsetAsync(counter, 0, sizeof(int), st) != cudaSuccess) { set_error("oz counter memset failed"); return 1; }
    const OzGemmTiles f{A, B, K / 32, N / OZ_TILE_N, sa, sb, C, ldc, dbg, g_oz_groups};
    if (oz_launch<OZ_STORE>(f, f.nx * (M / NB), counter, st)) return 1;
    return check_launch("k_oz_gemm_tn");
}

// ---- prediction launchers -------------------------------------------------------------------------------------------
static size_t oz_align256(size_t x) { return (x + 255) & ~(size_t)255; }
// state of a fitted GP for the digit-GEMM prediction: [W digits][gamma Np][ginv_s Np][sigma_B]
size_t oz_pred_state_bytes(int Np) { return oz_align256(oz_slice_bytes(Np)) + oz_align256((size_t)2 * Np * 8) + 256; }
size_t oz_pred_workspace_bytes(int Np, int Bp) {
    return oz_align256((size_t)(Bp / NB) * (Np / 32) * OZ_CHUNK) + oz_align256((size_t)4 * (Np / NB) * Bp * 8) + 256;
}

int launch_oz_pred_prepare(const double* W, const double* e, int N, int Np, uint8_t* state, cudaStream_t st) {
    double* gam = (double*)(state + oz_align256(oz_slice_bytes(Np)));
    double* ginv_s = gam + Np;
    double* sigb = (double*)(state + oz_align256(oz_slice_bytes(Np)) + oz_align256((size_t)2 * Np * 8));
    k_oz_pred_scales<<<1, 256, 0, st>>>(e, N, Np, gam, ginv_s, sigb);
    if (check_launch("k_oz_pred_scales")) return 1;
    const int T = Np / NB;
    k_oz_slice_wall<<<dim3(T, T), 256, 0, st>>>(W, Np, gam, state);
    return check_launch("k_oz_slice_wall");
}

// one chunk of B grid points (Bp = B rounded up to 128): digits of K_*, mean, |W k_*|^2 partial sums
int launch_oz_predict_chunk(const uint8_t* state, const double* xs, const double* sq, const double* ls, const double* alpha,
                            int N, int Np, int d, const double* Xfit, long long M, long long m0, int B, int Bp,
                            uint8_t* workspace, double* mu_out, double** partial_out, cudaStream_t st) {
    const double* ginv_s = (const double*)(state + oz_align256(oz_slice_bytes(Np))) + Np;
    const double* sigb = (const double*)(state + oz_align256(oz_slice_bytes(Np)) + oz_align256((size_t)2 * Np * 8));
    uint8_t* kdig = workspace;
    double* partial = (double*)(workspace + oz_align256((size_t)(Bp / NB) * (Np / 32) * OZ_CHUNK));
    int* counter = (int*)((uint8_t*)partial + oz_align256((size_t)4 * (Np / NB) * Bp * 8));
    if (cudaMemsetAsync(counter, 0, sizeof(int), st) != cudaSuccess) { set_error("oz counter memset failed"); return 1; }
    const dim3 xgrid((Bp + XC_PTS - 1) / XC_PTS);
    const size_t xsmem = (size_t)(d + 3) * XC_ARR * sizeof(double);
#define GPK_XCOV(D) k_oz_cross_cov<D><<<xgrid, 256, xsmem, st>>>(xs, sq, ls, alpha, ginv_s, N, Np, Xfit, M, m0, B, Bp, kdig, mu_out)
    switch (d) {
        case 1: GPK_XCOV(1); break;
        case 2: GPK_XCOV(2); break;
        case 3: GPK_XCOV(3); break;
        case 4: GPK_XCOV(4); break;
        case 5: GPK_XCOV(5); break;
        case 6: GPK_XCOV(6); break;
        case 7: GPK_XCOV(7); break;
        default: GPK_XCOV(8); break;
    }
#undef GPK_XCOV
    if (check_launch("k_oz_cross_cov")) return 1;
    const OzPredTiles f{state, kdig, Np, Bp / OZ_TILE_N, sigb, partial, (size_t)Bp};
    if (oz_launch<OZ_SUMSQ>(f, f.nx * (Np / NB), counter, st)) return 1;
    *partial_out = partial;
    return check_launch("k_oz_predict");
}

}  // namespace gpk

#ifdef OZ_PROFILE
// instrumented build only: copies the 4 x 8 counters to the host and optionally clears them
extern "C" __attribute__((visibility("default"))) int gpk_debug_oz_prof(unsigned long long* out32, int reset) {
    if (cudaDeviceSynchronize() != cudaSuccess) return 1;
    if (cudaMemcpyFromSymbol(out32, gpk::g_oz_prof, sizeof(unsigned long long) * 32) != cudaSuccess) return 1;
    if (reset) {
        unsigned long long z[32] = {};
        if (cudaMemcpyToSymbol(gpk::g_oz_prof, z, sizeof z) != cudaSuccess) return 1;
    }
    return 0;
}
#endif
