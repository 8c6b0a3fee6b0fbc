// Device-resident particle-swarm loop of find_len_scales.len_scale_opt (find_len_scales.py:76-141).
//
// The reference evaluates a swarm, then updates it on the host, 501 + restarts times per experiment.  With the loss on the
// GPU a host round trip per iteration (copies, synchronise, finish of the scores, NumPy update) costs twice the kernel
// itself (profiles/r02_pso_host_profile_v2.log).  k_pso_scores + k_pso_step keep the loop on the device: after every swarm
// evaluation they
//   1. finishes the scores from the int32 counts: np.trapz(|counts / n - expected|, sigma_vals) with NumPy's operation
//      order and NumPy's pairwise summation (the same replica as the host routine gpk_finish_wasserstein), the proximity
//      penalty and the +1e13 of out-of-bounds particles                                  (find_len_scales.py:163-177,182);
//   2. absorbs them: personal / global bests by strict '<' and first-occurrence argmin, the stagnation counter, the
//      progress record                                                                    (find_len_scales.py:79-83,106-125);
//   3. prepares the next evaluation: the velocity / position update of the next iteration, or the soft restart
//      (find_len_scales.py:88-100,123-141), and writes the positions and the in-bounds mask the loss kernel reads.
// Random numbers come from a stream the host pre-generated with NumPy's own legacy generator, consumed in the reference's
// order (two rand(P, d) per iteration, one uniform(-noise, noise) per restart), so the trajectory is the host loop's bit
// for bit: every product and sum below is a separately rounded IEEE operation (__dmul_rn / __dadd_rn: no contraction), in
// NumPy's evaluation order.  Scores: one CTA per particle; the rest: one CTA.  The swarm state lives in one device buffer
// laid out by pso_layout().
#include "gpk_common.cuh"

namespace gpk {

// header words of the state buffer (int32), see include/gpk.h
enum PsoHdr { PH_P = 0, PH_D, PH_S, PH_NPOINTS, PH_MAX_ITER, PH_PATIENCE, PH_MAX_EVALS, PH_ITER, PH_PENDING, PH_STALL,
              PH_RESTARTS, PH_NEVALS, PH_ERROR, PH_RNG_CURSOR, PH_FALLBACKS, PH_RNG_LEN, PH_LOGGED, PH_WORDS = 32 };
enum PsoPending { PP_INIT = 0, PP_NORMAL = 1, PP_RESTART = 2, PP_DONE = 3 };

PsoLayout pso_layout(int P, int d, int S, int n_points, int max_iter, int max_evals, long long rng_len) {
    PsoLayout L{};
    size_t off = PH_WORDS * sizeof(int);
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 7) & ~(size_t)7; return o; };
    L.scalars = take(8 * 8);                         // [0] gbest score, [1] inertia decay
    L.lower = take((size_t)d * 8);
    L.upper = take((size_t)d * 8);
    L.vmax = take((size_t)d * 8);
    L.pos = take((size_t)P * d * 8);
    L.vel = take((size_t)P * d * 8);
    L.pbest_pos = take((size_t)P * d * 8);
    L.pbest_sc = take((size_t)P * 8);
    L.gbest_pos = take((size_t)d * 8);
    L.scores = take((size_t)P * 8);
    L.expected = take((size_t)S * 8);
    L.d_sigma = take((size_t)S * 8);
    L.measured_of = take((size_t)(n_points + 1) * 8);
    L.log_best = take((size_t)max_iter * 8);
    L.log_stall = take((size_t)max_iter * 4);
    L.log_restart = take((size_t)max_iter * 4);
    L.trace_pos = take((size_t)max_evals * P * d * 8);
    L.trace_sc = take((size_t)max_evals * P * 8);
    L.rng = take((size_t)rng_len * 8);
    L.total = off;
    return L;
}

// ---- np.trapz(|counts / n - expected|, sigma_vals) for one particle, by one warp ----------------------------------------
struct TrapzTerm {
    const int* c; const double* measured_of; const double* expected; const double* d_sigma; int n_points;
    // y_s = |counts[s] / n - expected[s]|: the quotient is looked up in the table NumPy's own division filled
    __device__ __forceinline__ double y(int s) const {
        int k = c[s];
        k = k < 0 ? 0 : (k > n_points ? n_points : k);
        return fabs(__dsub_rn(measured_of[k], expected[s]));
    }
    // t_s = d_s * (y_{s+1} + y_s) / 2.0   (np.trapz: d * (y[1:] + y[:-1]) / 2.0)
    __device__ __forceinline__ double operator()(int s) const {
        return __ddiv_rn(__dmul_rn(d_sigma[s], __dadd_rn(y(s + 1), y(s))), 2.0);
    }
};

// NumPy's pairwise sum of a block of n <= 128 terms starting at off: 8 running sums over the strided terms (lanes 0..7),
// combined as ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7)), then the n % 8 tail; n < 8: plain left-to-right sum.
// All 32 lanes call; the result is valid in every lane.
template <class Term>
__device__ __forceinline__ double pairwise_leaf(const Term& t, int off, int n) {
    const int lane = threadIdx.x & 31;
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; ++i) res = __dadd_rn(res, t(off + i));
        return res;
    }
    const int body = n - (n % 8);
    double r = 0.0;
    if (lane < 8) {
        r = t(off + lane);
        for (int i = 8; i < body; i += 8) r = __dadd_rn(r, t(off + i + lane));
    }
    const double r1 = __dadd_rn(r, __shfl_down_sync(0xffffffffu, r, 1));        // lanes 0, 2, 4, 6: r_k + r_{k+1}
    const double r2 = __dadd_rn(r1, __shfl_down_sync(0xffffffffu, r1, 2));      // lanes 0, 4: (..) + (..)
    double res = __dadd_rn(r2, __shfl_down_sync(0xffffffffu, r2, 4));           // lane 0
    res = __shfl_sync(0xffffffffu, res, 0);
    for (int i = body; i < n; ++i) res = __dadd_rn(res, t(off + i));
    return res;
}

// NumPy's pairwise summation: blocks of more than 128 terms are split at n / 2 rounded down to a multiple of 8
template <class Term>
__device__ double pairwise_sum_warp(const Term& t, int n) {
    int off[24], len[24], st[24];
    double left[24];
    int sp = 0;
    off[0] = 0; len[0] = n; st[0] = 0;
    double result = 0.0;
    while (true) {
        if (len[sp] <= 128) {
            result = pairwise_leaf(t, off[sp], len[sp]);
            while (true) {                               // return to the callers
                if (sp == 0) return result;
                --sp;
                if (st[sp] == 1) {                       // that was the left half: now the right one
                    left[sp] = result;
                    st[sp] = 2;
                    int n2 = len[sp] / 2;
                    n2 -= n2 % 8;
                    off[sp + 1] = off[sp] + n2; len[sp + 1] = len[sp] - n2; st[sp + 1] = 0;
                    ++sp;
                    break;
                }
                result = __dadd_rn(left[sp], result);    // both halves done
            }
        } else {
            st[sp] = 1;
            int n2 = len[sp] / 2;
            n2 -= n2 % 8;
            off[sp + 1] = off[sp]; len[sp + 1] = n2; st[sp + 1] = 0;
            ++sp;
        }
    }
}

struct SmemTerm {
    const double* t;
    __device__ __forceinline__ double operator()(int s) const { return t[s]; }
};

// Scores of the evaluated swarm, one CTA per particle: the S - 1 trapezoid terms are formed by all threads into shared
// memory, one warp then runs NumPy's pairwise summation over them (serial chains of ~16 additions per lane and leaf).
constexpr int PSO_SCORE_THREADS = 128;
__global__ void __launch_bounds__(PSO_SCORE_THREADS)
k_pso_scores(unsigned char* __restrict__ state, PsoLayout L, const int* __restrict__ counts) {
    extern __shared__ double s_terms[];
    const int* hdr = reinterpret_cast<const int*>(state);
    if (hdr[PH_PENDING] == PP_DONE) return;
    const int d = hdr[PH_D], S = hdr[PH_S];
    const int p = blockIdx.x, tid = threadIdx.x;
    const double* lower = reinterpret_cast<const double*>(state + L.lower);
    const double* upper = reinterpret_cast<const double*>(state + L.upper);
    const double* pos = reinterpret_cast<const double*>(state + L.pos) + (size_t)p * d;
    double* scores = reinterpret_cast<double*>(state + L.scores);
    bool oob = false;
    for (int k = 0; k < d; ++k) oob |= (pos[k] <= lower[k]) || (pos[k] >= upper[k]);
    if (oob) {                                           // np.where(oob, 1e13, scores)
        if (tid == 0) scores[p] = 1e13;
        return;
    }
    const TrapzTerm term{counts + (size_t)p * S, reinterpret_cast<const double*>(state + L.measured_of),
                         reinterpret_cast<const double*>(state + L.expected),
                         reinterpret_cast<const double*>(state + L.d_sigma), hdr[PH_NPOINTS]};
    for (int s = tid; s < S - 1; s += PSO_SCORE_THREADS) s_terms[s] = term(s);
    __syncthreads();
    if (tid < 32) {
        const double wass = pairwise_sum_warp(SmemTerm{s_terms}, S - 1);
        if (tid == 0) {
            double dmin = 0.0;
            for (int k = 0; k < d; ++k) {
                const double span = __dsub_rn(upper[k], lower[k]);
                const double dlo = __ddiv_rn(__dsub_rn(pos[k], lower[k]), span);
                const double dhi = __ddiv_rn(__dsub_rn(upper[k], pos[k]), span);
                const double m = fmin(dlo, dhi);
                dmin = (k == 0) ? m : fmin(dmin, m);
            }
            double prox = __dsub_rn(1.0, __dmul_rn(2.0, dmin));
            prox = fmin(fmax(prox, 0.0), 1.0);
            scores[p] = -__dsub_rn(-wass, __dmul_rn(0.01, prox));           // -(-wasserstein - (0.01 * proximity))
        }
    }
}

constexpr int PSO_THREADS = 256;

__global__ void __launch_bounds__(PSO_THREADS, 1)
k_pso_step(unsigned char* __restrict__ state, PsoLayout L, const int* __restrict__ info,
           const unsigned char* __restrict__ fallback, double* __restrict__ pos_out, unsigned char* __restrict__ active_out) {
    int* hdr = reinterpret_cast<int*>(state);
    const int pending = hdr[PH_PENDING];
    if (pending == PP_DONE) return;
    const int P = hdr[PH_P], d = hdr[PH_D];
    const int tid = threadIdx.x;
    double* scal = reinterpret_cast<double*>(state + L.scalars);
    const double* lower = reinterpret_cast<const double*>(state + L.lower);
    const double* upper = reinterpret_cast<const double*>(state + L.upper);
    const double* vmax = reinterpret_cast<const double*>(state + L.vmax);
    double* pos = reinterpret_cast<double*>(state + L.pos);
    double* vel = reinterpret_cast<double*>(state + L.vel);
    double* pbest_pos = reinterpret_cast<double*>(state + L.pbest_pos);
    double* pbest_sc = reinterpret_cast<double*>(state + L.pbest_sc);
    double* gbest_pos = reinterpret_cast<double*>(state + L.gbest_pos);
    double* scores = reinterpret_cast<double*>(state + L.scores);
    const double* rng = reinterpret_cast<const double*>(state + L.rng);
    __shared__ int s_next;          // what to prepare: 0 nothing (restart prepared / done), 1 the normal update of iteration s_iter
    __shared__ int s_iter, s_cursor, s_restart;
    __shared__ double s_w;

    // (1. the scores of the evaluated swarm were left in state.scores by k_pso_scores)

    // ---- 2. record, absorb, decide ---------------------------------------------------------------------------------------
    const int e = hdr[PH_NEVALS];
    if (e < hdr[PH_MAX_EVALS]) {
        double* tp = reinterpret_cast<double*>(state + L.trace_pos) + (size_t)e * P * d;
        double* ts = reinterpret_cast<double*>(state + L.trace_sc) + (size_t)e * P;
        for (int i = tid; i < P * d; i += PSO_THREADS) tp[i] = pos[i];
        for (int i = tid; i < P; i += PSO_THREADS) ts[i] = scores[i];
    }
    if (pending == PP_NORMAL) {                          // improved = scores < pbest_scores (before the argmin below)
        for (int p = tid; p < P; p += PSO_THREADS) {
            if (scores[p] < pbest_sc[p]) {
                pbest_sc[p] = scores[p];
                for (int k = 0; k < d; ++k) pbest_pos[p * d + k] = pos[p * d + k];
            }
        }
    } else {                                             // initial evaluation / re-evaluation after a restart
        for (int p = tid; p < P; p += PSO_THREADS) {
            pbest_sc[p] = scores[p];
            for (int k = 0; k < d; ++k) pbest_pos[p * d + k] = pos[p * d + k];
        }
    }
    __syncthreads();
    if (tid == 0) {
        int bad = 0, fb = 0;
        for (int p = 0; p < P; ++p) { bad |= (info[p] != 0); fb += fallback ? (int)fallback[p] : 0; }
        hdr[PH_FALLBACKS] += fb;
        hdr[PH_NEVALS] = e + 1;
        int g = 0;
        for (int p = 1; p < P; ++p) if (pbest_sc[p] < pbest_sc[g]) g = p;        // np.argmin: first occurrence
        int it = hdr[PH_ITER];
        s_next = 0; s_restart = 0;
        if (bad) {                                       // numpy.linalg.LinAlgError inside the swarm evaluation: stop here
            hdr[PH_ERROR] = 1;
            hdr[PH_PENDING] = PP_DONE;
        } else if (pending == PP_NORMAL) {
            int stall = hdr[PH_STALL];
            if (pbest_sc[g] < scal[0]) {
                scal[0] = pbest_sc[g];
                for (int k = 0; k < d; ++k) gbest_pos[k] = pbest_pos[g * d + k];
                stall = 0;
            } else {
                ++stall;
            }
            hdr[PH_STALL] = stall;
            reinterpret_cast<double*>(state + L.log_best)[it] = scal[0];
            reinterpret_cast<int*>(state + L.log_stall)[it] = stall;
            const int restart = stall >= hdr[PH_PATIENCE];
            reinterpret_cast<int*>(state + L.log_restart)[it] = restart;
            hdr[PH_LOGGED] = it + 1;
            if (restart) s_restart = 1;
            else { s_next = 1; s_iter = it + 1; }
        } else {
            scal[0] = pbest_sc[g];
            for (int k = 0; k < d; ++k) gbest_pos[k] = pbest_pos[g * d + k];
            hdr[PH_STALL] = 0;
            if (pending == PP_RESTART) { hdr[PH_RESTARTS] += 1; s_iter = it + 1; }
            else s_iter = 0;
            s_next = 1;
        }
        if (s_next && s_iter >= hdr[PH_MAX_ITER]) {      // the loop is over
            s_next = 0;
            hdr[PH_PENDING] = PP_DONE;
        }
        s_cursor = hdr[PH_RNG_CURSOR];
        const int need = s_restart ? P * d : (s_next ? 2 * P * d : 0);
        if (s_cursor + need > hdr[PH_RNG_LEN]) {         // more restarts than the host provided random numbers for
            hdr[PH_ERROR] = 2;
            hdr[PH_PENDING] = PP_DONE;
            s_next = 0; s_restart = 0;
        }
        if (s_next) {
            // w = max(0.4, 0.9 - i * inertia_decay)
            s_w = fmax(0.4, __dsub_rn(0.9, __dmul_rn((double)s_iter, scal[1])));
            hdr[PH_ITER] = s_iter;
            hdr[PH_PENDING] = PP_NORMAL;
            hdr[PH_RNG_CURSOR] = s_cursor + 2 * P * d;
        } else if (s_restart) {
            hdr[PH_PENDING] = PP_RESTART;
            hdr[PH_RNG_CURSOR] = s_cursor + P * d;
        }
    }
    __syncthreads();

    // ---- 3. the next evaluation's positions --------------------------------------------------------------------------------
    if (s_next) {
        const double* r1 = rng + s_cursor;
        const double* r2 = r1 + P * d;
        const double w = s_w;
        for (int i = tid; i < P * d; i += PSO_THREADS) {
            const int k = i % d;
            // velocities = (w * velocities + c1 * r1 * (pbest - positions) + c2 * r2 * (gbest - positions)), c1 = c2 = 1.4
            const double t1 = __dmul_rn(w, vel[i]);
            const double t4 = __dmul_rn(__dmul_rn(1.4, r1[i]), __dsub_rn(pbest_pos[i], pos[i]));
            const double t8 = __dmul_rn(__dmul_rn(1.4, r2[i]), __dsub_rn(gbest_pos[k], pos[i]));
            double v = __dadd_rn(__dadd_rn(t1, t4), t8);
            v = fmin(fmax(v, -vmax[k]), vmax[k]);                         // np.clip(velocities, -v_max, v_max)
            vel[i] = v;
            double x = __dadd_rn(pos[i], v);                              // positions += velocities
            x = fmin(fmax(x, lower[k]), upper[k]);                        // np.clip(positions, lower, upper)
            pos[i] = x;
        }
    } else if (s_restart) {
        const double* u = rng + s_cursor;
        for (int i = tid; i < P * d; i += PSO_THREADS) {
            const int k = i % d;
            // noise = 0.1 * (upper - lower); pbest += uniform(-noise, noise) = -noise + (noise - (-noise)) * u; clip
            const double noise = __dmul_rn(0.1, __dsub_rn(upper[k], lower[k]));
            const double range = __dsub_rn(noise, -noise);
            double x = __dadd_rn(pbest_pos[i], __dadd_rn(-noise, __dmul_rn(range, u[i])));
            x = fmin(fmax(x, lower[k]), upper[k]);
            pbest_pos[i] = x;
            pos[i] = x;                                                   // positions = pbest.copy(); velocities = 0
            vel[i] = 0.0;
        }
    } else {
        for (int p = tid; p < P; p += PSO_THREADS) active_out[p] = 0;     // finished (or failed): the remaining launches idle
        return;
    }
    __syncthreads();
    for (int i = tid; i < P * d; i += PSO_THREADS) pos_out[i] = pos[i];
    for (int p = tid; p < P; p += PSO_THREADS) {
        bool oob = false;
        for (int k = 0; k < d; ++k) oob |= (pos[p * d + k] <= lower[k]) || (pos[p * d + k] >= upper[k]);
        active_out[p] = oob ? 0 : 1;
    }
}

int launch_pso_step(unsigned char* state, const PsoLayout& L, int P, int S, const int* counts, const int* info,
                    const unsigned char* fallback, double* pos_out, unsigned char* active_out, cudaStream_t st) {
    k_pso_scores<<<P, PSO_SCORE_THREADS, (size_t)S * sizeof(double), st>>>(state, L, counts);
    if (check_launch("k_pso_scores")) return 1;
    k_pso_step<<<1, PSO_THREADS, 0, st>>>(state, L, info, fallback, pos_out, active_out);
    return check_launch("k_pso_step");
}

}  // namespace gpk
