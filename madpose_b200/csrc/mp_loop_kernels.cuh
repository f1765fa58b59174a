// Kind-independent kernels of the device-side hybrid LO-MSAC loop (state initialisation,
// sampling); included by mp_estimate.cu only.  See mp_loop.cuh for the round structure.
#pragma once
#include "mp_loop.cuh"

namespace mp {

// ---- state initialisation: VerifyData (hybrid_ransac.h:552-577), ResetStatistics, RNG seeding
__global__ void init_state_kernel(LoopBuffers L, const LoopParams *prmp) {
    const LoopParams &prm = *prmp;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= prm.n_pairs) return;
    const PairConst &pc = L.pc[p];
    PairState &S = L.st[p];
    const int kind = pc.kind, k0 = kind == CALIB ? 3 : 4, k2 = kind == CALIB ? 5 : (kind == SHARED ? 6 : 7);
    S.sizes[0][0] = k0; S.sizes[0][1] = k0; S.sizes[0][2] = 0;
    S.sizes[1][0] = 0; S.sizes[1][1] = 0; S.sizes[1][2] = k2;
    S.non_min_sample_size = kind == CALIB ? 35 : 36; // hybrid_pose_estimator.h:47, ...shared_focal_estimator.h:47
    S.min_sample_size = k2;
    S.prior[0] = prm.cfg.solver_type == 1 ? 0.0 : 1.0; // EPI_ONLY: no depth solver (hybrid_pose_estimator.h:60-69)
    S.prior[1] = prm.cfg.solver_type == 2 ? 0.0 : 1.0; // MD_ONLY: no point solver
    for (int s = 0; s < 2; s++)
        for (int t = 0; t < 3; t++)
            if (S.sizes[s][t] > pc.n) { S.prior[s] = 0.0; break; }
    S.phase = (S.prior[0] > 0.0 || S.prior[1] > 0.0) ? PH_ITER : PH_DONE;
    S.pre_done = prm.opt.lo_starting_iterations == 0 ? 1 : 0;
    S.B = prm.b_after_lo; S.batch_len = 0; S.it = 0; S.total = 0; S.broke = 0;
    S.iters[0] = S.iters[1] = 0;
    const unsigned c = prm.opt.max_num_iterations_per_solver > prm.opt.min_num_iterations
                           ? prm.opt.max_num_iterations_per_solver : prm.opt.min_num_iterations;
    S.cap[0] = S.cap[1] = c;
    S.best_min_score = DBL_MAX; S.best_score = DBL_MAX; S.best_solver = -1; S.ev_solver = 0;
    Model d; // default model (pose.h:11,25,38,49): zero pose, scale 1, offsets 0, focal 1
    for (int i = 0; i < 9; i++) d.R[i] = 0;
    d.t[0] = d.t[1] = d.t[2] = 0; d.scale = 1; d.o0 = 0; d.o1 = 0; d.f0 = 1; d.f1 = 1;
    S.best = d; S.best_min = d;
    S.number_lo = 0; S.best_num_inliers = 0;
    for (int t = 0; t < 3; t++) { S.num_inl[t] = 0; S.ratios[t] = 0; }
    S.hyps_scored = 0; S.evals = 0; S.evals64 = 0; S.lm_blocks = 0; S.lm_solves = 0; S.lm_iters = 0;
    for (int q = 0; q < 16; q++) S.clk[q] = 0;
    S.trace_n = 0; S.trace_pad_ = 0;
    L.evals32[p] = 0;
    mt_seed(&L.rngA[p], prm.opt.random_seed);
    mt_seed(&L.rngB[p], prm.opt.random_seed);
    L.lo_req[p] = 0; L.wk[p].blen = 0;
    if (S.phase == PH_DONE) atomicSub(L.n_active + p / L.gsz, 1);
}

// ---- SelectMinimalSolver (hybrid_ransac.h:210-243) + HybridUniformSampling::Sample for the
// next iterations of every iterating pair.  One WARP per pair: the lanes extend the two
// generator rings far enough for the whole batch and copy the raw words the batch can
// consume into shared-memory windows; lane 0 then draws sequentially from the windows
// (the draws of one stream are inherently sequential: rejection sampling).
constexpr int SMP_WARPS = 2, SMP_WIN_A = 20 * LP_BMAX, SMP_WIN_B = 2 * LP_BMAX;

__device__ inline void warp_fill_window(MtStream *g, uint32_t *win, int need, int lane) {
    while (g->gen < g->pos + (unsigned long long)need) mt_refill(g, lane, 32, [] { __syncwarp(); });
    const unsigned base = (unsigned)(g->pos % MT_RING);
    for (int i = lane; i < need; i += 32) {
        unsigned r = base + (unsigned)i;
        if (r >= (unsigned)MT_RING) r -= MT_RING;
        win[i] = g->x[r];
    }
    __syncwarp();
}

__global__ void __launch_bounds__(SMP_WARPS * 32) sample_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    __shared__ uint32_t winA[SMP_WARPS][SMP_WIN_A], winB[SMP_WARPS][SMP_WIN_B];
    __shared__ unsigned char ssolver[SMP_WARPS][LP_BMAX];
    const LoopParams &prm = *prmp;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * SMP_WARPS + warp, p = first + q;
    if (q >= count) return;
    PairState &S = L.st[p];
    PairWork &W = L.wk[p];
    // the phase is published last by the kernel that owned the pair (lo_kernel: fence, then phase), so a pair seen
    // in PH_ITER has all its state visible; a pair inside a local optimisation simply sits this round out
    int ph = 0;
    if (lane == 0) { ph = *reinterpret_cast<volatile int *>(&S.phase); __threadfence(); }
    ph = __shfl_sync(0xffffffffu, ph, 0);
    if (ph != PH_ITER) {
        if (lane == 0) W.blen = 0;
        return;
    }
    MtStream *A = &L.rngA[p], *B = &L.rngB[p];
    const int n = L.pc[p].n;
    const unsigned max_it = prm.opt.max_num_iterations > prm.opt.min_num_iterations ? prm.opt.max_num_iterations
                                                                                   : prm.opt.min_num_iterations;
    const unsigned lo_start = prm.opt.lo_starting_iterations;
    unsigned len = (unsigned)S.B < max_it - S.it ? (unsigned)S.B : max_it - S.it;
    if (S.it < lo_start && len > lo_start - S.it) len = lo_start - S.it; // a batch never straddles lo_start
    {
        // iterations past the termination test of hybrid_ransac.h:169 are speculated for nothing: cut the batch at the
        // expected number of iterations until a solver reaches its cap (+10 % + 4; any length gives the same result)
        const double ps = S.prior[0] + S.prior[1];
        double est = 1e18;
        for (int s = 0; s < 2; s++)
            if (S.prior[s] > 0.0) {
                const double left = S.cap[s] > S.iters[s] ? (double)(S.cap[s] - S.iters[s]) : 1.0;
                est = fmin(est, left * ps / S.prior[s]);
            }
        est = est * 1.1 + 4.0;
        if (est < (double)len) len = (unsigned)est;
    }
    const int needA = 20 * (int)len, needB = 2 * (int)len; // words staged for this batch
    warp_fill_window(A, winA[warp], needA, lane);
    warp_fill_window(B, winB[warp], needB, lane);
    const unsigned long long posA0 = A->pos, posB0 = B->pos;
    const double prior0 = S.prior[0], prior1 = S.prior[1];
    int sizes[2][3], expect[2] = {0, 0}; // expect[s]: words of stream A one iteration of solver s takes without a retry
    for (int s = 0; s < 2; s++)
        for (int t = 0; t < 3; t++) {
            sizes[s][t] = S.sizes[s][t];
            expect[s] += mt_draw_words(n, sizes[s][t]);
        }
    // stream B: the solver of every iteration; generate_canonical takes exactly two words, so iteration b reads
    // words 2b, 2b + 1 of the window -- all iterations at once
    unsigned char *sol = ssolver[warp];
    for (unsigned b = lane; b < len; b += 32) {
        MtWindowRO sb{winB[warp], needB, (int)(2 * b), false};
        const double u = mt_uniform_real_t(sb, prior0 + prior1);
        int s = -1;
        double c = 0;
        for (int k = 0; k < 2; k++) {
            const double pk = k == 0 ? prior0 : prior1;
            if (pk == 0.0) continue;
            c += pk;
            if (u <= c) { s = k; break; }
        }
        if (s < 0) s = prior1 > 0.0 ? 1 : 0; // unreachable (u <= sum of priors)
        sol[b] = (unsigned char)s;
        W.solver[b] = s;
        W.posB[b] = posB0 + 2ull * (b + 1);
    }
    __syncwarp();
    // stream A: an iteration takes expect[s] words unless a draw repeats an index (or Lemire's method rejects a word).
    // 32 iterations are drawn at once at the offsets that assumes; the iterations up to and including the first one
    // that took a different number of words are right, the chunk restarts behind it with the true offset.
    auto draw_iteration = [&](auto &src, int s, unsigned b) {
        int tmp[8], scratch[16];
        for (int t = 0; t < 3; t++) {
            const int k = sizes[s][t];
            if (k <= 0) continue;
            mt_draw_sample_t(src, n, k, tmp, scratch); // type 1 of solver 0 is drawn and discarded (quirk Q1)
            if ((s == 0 && t == 0) || (s == 1 && t == 2))
                for (int i = 0; i < k; i++) W.idx[b][i] = tmp[i];
        }
    };
    unsigned b0 = 0;
    int o0 = 0; // words of stream A consumed by the iterations before b0
    bool serial_rest = false;
    while (b0 < len) {
        const unsigned b = b0 + lane;
        const bool on = b < len;
        const int s = on ? sol[b] : 0, ex = on ? expect[s] : 0;
        int pre = ex;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, pre, o);
            if (lane >= o) pre += v;
        }
        const int start = o0 + pre - ex;
        int used = ex;
        bool bad = false;
        if (on) {
            MtWindowRO sa{winA[warp], needA, start, false};
            draw_iteration(sa, s, b); // a wrong guess of `start` only writes W.idx[b], which the redo overwrites
            used = sa.used - start; bad = sa.overflow;
        }
        const unsigned dev = __ballot_sync(0xffffffffu, on && (used != ex || bad));
        const unsigned ovf = __ballot_sync(0xffffffffu, on && bad);
        const int f = dev ? __ffs(dev) - 1 : 32;
        if (on && (lane < f || (lane == f && !bad))) W.posA[b] = posA0 + (unsigned long long)(start + used);
        if (f == 32) {
            o0 += __shfl_sync(0xffffffffu, pre, 31);
            b0 += 32;
        } else {
            const int st_f = __shfl_sync(0xffffffffu, start, f), us_f = __shfl_sync(0xffffffffu, used, f);
            if ((ovf >> f) & 1u) { b0 += f; o0 = st_f; serial_rest = true; break; } // window used up: the rest in order
            b0 += f + 1; o0 = st_f + us_f;
        }
        __syncwarp(); // the redo of an iteration (another lane) writes W.idx[b] after the wrong guess did
    }
    __syncwarp();
    if (lane != 0) return;
    unsigned long long endA = posA0 + (unsigned long long)o0;
    if (serial_rest) {
        MtWindow sa{A, winA[warp], posA0, needA, o0, false}; // spills to the ring
        for (unsigned b = b0; b < len; b++) {
            draw_iteration(sa, (int)sol[b], b);
            W.posA[b] = sa.position();
        }
        endA = sa.position();
    }
    // Hand the samples to the minimal kernels: one list per solver, and the point-solver samples of the iterations
    // >= prm.warp_from_it in a third list that the warp-per-sample kernel solves.  Which form solves a sample depends
    // on its (pair, iteration) only -- never on the schedule.
    const unsigned it0i = S.it;
    int n1 = 0, n2 = 0;
    for (unsigned b = 0; b < len; b++) {
        if (sol[b] == 1) { if (it0i + b >= prm.warp_from_it) n2++; else n1++; }
    }
    {
        const int g = p / L.gsz, n0 = (int)len - n1 - n2;
        int o0i = n0 > 0 ? atomicAdd(L.n_items + 2 * g, n0) : 0, o1i = n1 > 0 ? atomicAdd(L.n_items + 2 * g + 1, n1) : 0;
        int o2i = n2 > 0 ? atomicAdd(L.n_items_w + g, n2) : 0;
        int *it0 = L.items + (size_t)first * LP_BMAX, *it1 = L.items + (size_t)(L.n_pairs + first) * LP_BMAX;
        int *it2 = L.items + (size_t)(2 * L.n_pairs + first) * LP_BMAX;
        for (unsigned b = 0; b < len; b++) {
            if (sol[b] == 1) { if (it0i + b >= prm.warp_from_it) it2[o2i++] = q * LP_BMAX + (int)b; else it1[o1i++] = q * LP_BMAX + (int)b; }
            else it0[o0i++] = q * LP_BMAX + (int)b;
        }
    }
    A->pos = endA; B->pos = posB0 + 2ull * len;
    S.batch_len = (int)len;
    W.blen = (int)len;
}

} // namespace mp
