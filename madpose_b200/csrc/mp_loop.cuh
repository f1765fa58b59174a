// K6 -- the hybrid LO-MSAC control loop on the device (hybrid_ransac.h:38-206) for a
// batch of pairs.  The reference's loop is sequential per pair; here every ROUND runs,
// for all pairs still iterating,
//   sample_kernel   : solver selection + minimal samples of the next <= B iterations
//                     (exact std::mt19937 streams, mp_rng.cuh)
//   minimal_kernel  : one thread per sample -> <= 15 fp64 models (K1/K2)
//   score_kernel    : fused fp32 truncated-MSAC scores of all new hypotheses (K3)
//   decide_kernel   : one CTA per pair replays the iterations IN ORDER on the scores:
//                     candidate bests are re-scored in fp64, best model / inliers /
//                     per-solver iteration caps updated, local optimisation (K5) run
//                     where the reference runs it, the break test taken per iteration.
// Iterations after an LO event inside a batch are discarded and their random numbers
// rewound (LO consumes stream B), so the sequence of (solver, sample) the device sees is
// the one the sequential CPU loop sees.  Termination is decided on the device; the host
// only relaunches rounds while the device-side count of unfinished pairs is non-zero.
#pragma once
#include "mp_block.cuh"
#include "mp_lm.cuh"
#include "mp_minimal.cuh"
#include "mp_rng.cuh"

namespace mp {

constexpr int LP_BMAX = 256;           // iterations per pair per round (upper bound; a pair starts at LP_B0 and doubles
                                       // its batch after every round without a new best model)
constexpr int LP_B0 = 64, LP_BDEF = 128; // measured on configs[1]: cap 64 -> 179-185 ms, 128 -> 178-181, 256 -> 183-187
constexpr int LP_MM = kMaxModelsAny;   // model slots per iteration
#ifndef MP_LP_THREADS
#define MP_LP_THREADS 128
#endif
constexpr int LP_THREADS = MP_LP_THREADS; // CTA of the scan / local-optimisation kernels
#ifndef MP_LO_CTAS
#define MP_LO_CTAS 2
#endif
constexpr int LP_CTAS_PER_SM = MP_LO_CTAS;
// Local optimisation of a pair with >= LP_WIDE_N matches runs on a CTA of LP_WIDE_THREADS threads (one per SM): the passes
// over the matches take half the time, the LM driver does not care.  The width is a function of the PAIR alone (never of
// the batch or the schedule), so a pair's result does not depend on what it is batched with.  Measured (ms / step, all
// pairs at 128 | all at 256 threads): configs[1] (2k matches) 124.9 | 133.5, configs[2] (5k) 266.7 | 255.3, configs[3]
// (10k) 154.0 | 125.3.
constexpr int LP_WIDE_N = 4096, LP_WIDE_THREADS = 256;
enum { PH_ITER = 0, PH_LO_PRE = 1, PH_LO_MAIN = 2, PH_FINISH = 3, PH_DONE = 4, PH_BUSY = 5 };

struct LoopParams {
    mp_options opt;
    mp_config cfg;
    int n_pairs, bmax;
    int b_after_lo; // batch length of the first round and of the first round after a local optimisation (doubles every
                    // quiet round up to bmax)
    double guard; // relative guard band of the fp32 score filter
    unsigned warp_from_it; // point-solver samples of iterations >= this are solved with one warp per sample (kinds that
                           // have such a solver); 0 = all of them, 0xffffffff = none
};

struct alignas(128) PairState { // 128-byte aligned: no cache line is shared between pairs (see sample_kernel)
    int phase, pre_done, B, batch_len;
    unsigned it;             // iterations completed = index of the next iteration
    unsigned iters[2], cap[2];
    unsigned total;          // num_iterations_total at loop exit
    int broke;
    double prior[2];
    double best_min_score, best_score;
    Model best_min, best;
    int best_solver, ev_solver;
    int number_lo, best_num_inliers, num_inl[3];
    double ratios[3];
    unsigned long long hyps_scored, evals;
    unsigned long long evals64, lm_blocks; // fp64 (model x match) evaluations; LM residual-block evaluations
    unsigned lm_solves, lm_iters;
    int sizes[2][3], non_min_sample_size, min_sample_size;
    int trace_n, trace_pad_; // events written to LoopBuffers::trace so far (mp_set_trace)
    unsigned long long clk[16]; // SM-clock cycles of thread 0 per phase: 0 scan 1 inliers 2 score64 3 lm eval 4 lm driver 5 shuffle 6 unflatten 7 LO total
};

// per-pair work buffers of one round
struct PairWork {
    int blen; // iterations of this round (0: the pair does not take part -- not iterating, or inside a local optimisation)
    int solver[LP_BMAX];
    int idx[LP_BMAX][8];
    unsigned long long posA[LP_BMAX], posB[LP_BMAX]; // stream positions AFTER iteration b
    int count[LP_BMAX];
    float s32[LP_BMAX * LP_MM];
};

struct LoopBuffers {
    const PairConst *pc;
    const Match64 *M;
    const float4 *S32;
    PairState *st;
    PairWork *wk;
    MtStream *rngA, *rngB;
    Model *models;  // [pair][LP_BMAX * LP_MM]
    HypF *hyps;     // same indexing
    int *inl_best;  // [3 * off] inlier lists of the global best (the caller's output layout)
    int *inl_cur;   // [3 * off] scratch: inliers of the model being refined
    int *flat;      // [3 * off + pad] scratch: flattened inlier list / shuffled sample
    int *sample;    // [3 * off + pad] scratch: un-flattened sample lists
    int *n_active;  // device counters of pairs not yet PH_DONE, one per group of gsz consecutive pairs
    int *n_iter;    // per group: pairs left in PH_ITER by the last scan kernel
    int *n_cand;    // per group: point-solver candidates of this round (reset every round)
    int *cand;      // [n_pairs * LP_BMAX * LP_MM] candidate list of the round: item * LP_MM + j (a group uses its own slice)
    unsigned char *cand_ok; // per model slot: did the candidate survive its tail
    int *n_ret;     // per group: pairs handed back to PH_ITER by lo_kernel launches so far (cumulative)
    int *n_items;   // per group: [2] number of samples of this round per minimal solver (reset every round)
    int *n_items_w; // per group: point-solver samples of this round routed to the warp-per-sample kernel (third list)
    int *items;     // [3][n_pairs * LP_BMAX] samples of this round, (pair - first) * LP_BMAX + b, grouped per solver so that
                    // the warps of the minimal kernel run ONE solver each (a group uses its own slice [first * LP_BMAX, ...))
    int n_pairs;
    int *lo_req;    // per pair: 0, or the round whose scan asked for a local optimisation / the final stage
                    // (negated while an lo_kernel CTA owns the pair)
    int gsz;
    unsigned long long *clk; // [8] phase cycle totals over all pairs (profiling aid)
    // parity forensics (mp_set_trace): per pair trace_cap records of 2 doubles, chronological:
    //   {it, score}   an iteration whose best candidate was re-scored in fp64 (score = the local best)
    //   {-1, score}   every ScoreModel inside LocalOptimization / the final least squares
    //   {-2, count}   every GetInliers (LeastSquaresFit, LO base, termination update): total inlier count
    //   {-3, iterations}, {-4, final cost}   every NonMinimalSolver / LeastSquares that ran
    double *trace;
    int trace_cap;
    unsigned long long *evals32; // per pair: (model x match) evaluations of the fused fp32 scoring kernel
};

// ---------------------------------------------------------------- block-level helpers
// (blockDim.x <= 512, all threads participate)

// exact ScoreModel (hybrid_ransac.h:265-282) of one model by the whole CTA
static __device__ __noinline__ double block_score64(const PairConst &pc, const Match64 *m, const Model &model,
                                                int score_type, double *red) {
    __builtin_assume(__isGlobal(&pc)); __builtin_assume(__isGlobal(m)); __builtin_assume(__isShared(red)); // address-space hints (mp_lm_warp.cuh)
    Eval64 e;
    make_eval64(pc, model, &e);
    double acc = 0;
    for (int i = threadIdx.x; i < pc.n; i += blockDim.x) {
        const Match64 mm = m[i];
        const double r0 = score_type == 1 ? DBL_MAX : eval64(e, mm, 0);
        const double r1 = score_type == 1 ? DBL_MAX : eval64(e, mm, 1);
        const double r2 = score_type == 2 ? DBL_MAX : eval64(e, mm, 2);
        acc += pc.w[0] * msac64(r0, pc.thr[0]) + pc.w[1] * msac64(r1, pc.thr[1]) + pc.w[2] * msac64(r2, pc.thr[2]);
    }
    return block_sum(acc, red);
}

// GetInliers (hybrid_ransac.h:289-349): three ascending index lists, out_idx[t * n + k];
// counts written to cnt_sh[3] (shared).  wcnt: shared int[3][16] (unused; kept for the callers' layout).
// Two passes per super-chunk of INL_CH * blockDim matches with no barrier inside the passes:
//   pass 1  residuals -> 3 flag bits per match kept in registers; per (chunk, warp, type) inlier counts to shared
//   scan    exclusive prefix of the counts in (chunk, warp) order, one warp per type
//   pass 2  ranks = prefix + popc(ballot below my lane) -> ordered, conflict-free stores
constexpr int INL_CH = 16, INL_MAXW = 8;
struct InlScratch { int cnt[3][INL_CH * INL_MAXW]; int base[3]; };
// score_out != nullptr: the same pass also returns ScoreModel(model) (hybrid_ransac.h:265-282) -- same residuals, same
// per-thread order of the sum and the same block_sum tree as block_score64, hence the same bits.
static __device__ __noinline__ void block_inliers(const PairConst &pc, const Match64 *m, const Model &model,
                                           const double *thr3, int *out_idx, int *cnt_sh, int (*wcnt)[16],
                                           double *score_out = nullptr, int score_type = 0, double *red = nullptr) {
    __builtin_assume(__isGlobal(&pc)); __builtin_assume(__isGlobal(m)); __builtin_assume(__isGlobal(out_idx));
    __builtin_assume(__isShared(cnt_sh)); __builtin_assume(__isShared(wcnt)); // address-space hints (mp_lm_warp.cuh)
    __shared__ InlScratch isc;
    Eval64 e;
    make_eval64(pc, model, &e);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5; // nw <= INL_MAXW
    const int n = pc.n;
    const double t0 = thr3[0], t1 = thr3[1], t2 = thr3[2];
    __syncthreads();
    if (threadIdx.x < 3) isc.base[threadIdx.x] = 0;
    __syncthreads();
    double acc = 0;
    for (int sbase = 0; sbase < n; sbase += INL_CH * blockDim.x) {
        unsigned long long flags = 0;
#pragma unroll 4
        for (int c = 0; c < INL_CH; c++) {
            const int i = sbase + c * blockDim.x + threadIdx.x;
            unsigned f = 0;
            if (i < n) {
                const Match64 mm = m[i];
                const double r0 = eval64(e, mm, 0), r1 = eval64(e, mm, 1), r2 = eval64(e, mm, 2);
                f = (r0 < t0 ? 1u : 0u) | (r1 < t1 ? 2u : 0u) | (r2 < t2 ? 4u : 0u);
                if (score_out) {
                    const double s0 = score_type == 1 ? DBL_MAX : r0, s1 = score_type == 1 ? DBL_MAX : r1,
                                 s2 = score_type == 2 ? DBL_MAX : r2;
                    acc += pc.w[0] * msac64(s0, pc.thr[0]) + pc.w[1] * msac64(s1, pc.thr[1]) + pc.w[2] * msac64(s2, pc.thr[2]);
                }
            }
            flags |= (unsigned long long)f << (3 * c);
            const unsigned b0 = __ballot_sync(0xffffffffu, f & 1u), b1 = __ballot_sync(0xffffffffu, f & 2u),
                           b2 = __ballot_sync(0xffffffffu, f & 4u);
            if (lane == 0) {
                isc.cnt[0][c * nw + warp] = __popc(b0); isc.cnt[1][c * nw + warp] = __popc(b1); isc.cnt[2][c * nw + warp] = __popc(b2);
            }
        }
        __syncthreads();
        for (int ty = warp; ty < 3; ty += nw) { // exclusive scan of the INL_CH * nw counts of type ty, from the running base
            int run = isc.base[ty];
            const int tot = INL_CH * nw;
            for (int o = 0; o < tot; o += 32) {
                const int k = o + lane;
                const int v = k < tot ? isc.cnt[ty][k] : 0;
                int inc = v;
                for (int d = 1; d < 32; d <<= 1) {
                    const int u = __shfl_up_sync(0xffffffffu, inc, d);
                    if (lane >= d) inc += u;
                }
                if (k < tot) isc.cnt[ty][k] = run + inc - v;
                run += __shfl_sync(0xffffffffu, inc, 31);
            }
            __syncwarp();
            if (lane == 0) isc.base[ty] = run;
        }
        __syncthreads();
#pragma unroll 4
        for (int c = 0; c < INL_CH; c++) {
            const int i = sbase + c * blockDim.x + threadIdx.x;
            const unsigned f = (unsigned)(flags >> (3 * c)) & 7u;
            const unsigned below = (1u << lane) - 1u;
            const unsigned b0 = __ballot_sync(0xffffffffu, f & 1u), b1 = __ballot_sync(0xffffffffu, f & 2u),
                           b2 = __ballot_sync(0xffffffffu, f & 4u);
            if (f & 1u) out_idx[isc.cnt[0][c * nw + warp] + __popc(b0 & below)] = i;
            if (f & 2u) out_idx[n + isc.cnt[1][c * nw + warp] + __popc(b1 & below)] = i;
            if (f & 4u) out_idx[2 * n + isc.cnt[2][c * nw + warp] + __popc(b2 & below)] = i;
        }
        __syncthreads();
    }
    if (threadIdx.x < 3) cnt_sh[threadIdx.x] = isc.base[threadIdx.x];
    __syncthreads();
    if (score_out) {
        const double sc = block_sum(acc, red);
        if (threadIdx.x == 0) *score_out = sc;
        __syncthreads();
    }
}

// utils::NumRequiredIterations (ext/RansacLib/RansacLib/utils.h:142-178), vector form
MP_HD unsigned num_required_iterations(const double *ratios, double prob_missing, const int *sizes, unsigned mn,
                                       unsigned mx) {
    double p = 1.0;
    for (int i = 0; i < 3; i++) p *= pow(ratios[i], (double)sizes[i]);
    if (p <= 0.0) return mx;
    if (p >= 1.0) return mn;
    const double q = 1.0 - p;
    if (q >= 0.99999999999999) return mx;
    const double it = ceil(log(prob_missing) / log(q) + 0.5);
    // the reference's static_cast<uint32_t>(double) is out of range for it >= 2^32; x86-64 gcc
    // truncates to 64 bits and keeps the low word -- reproduced so that caps agree with the CPU loop
    unsigned r = (unsigned)(unsigned long long)(long long)it;
    r = r < mx ? r : mx;
    return r > mn ? r : mn;
}

} // namespace mp
#include "mp_loop_lo.cuh"
