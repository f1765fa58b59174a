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
    S.B = prm.bmax; S.batch_len = 0; S.it = 0; S.total = 0; S.broke = 0;
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
    for (int q = 0; q < 8; q++) S.clk[q] = 0;
    mt_seed(&L.rngA[p], prm.opt.random_seed);
    mt_seed(&L.rngB[p], prm.opt.random_seed);
    if (S.phase == PH_DONE) atomicSub(L.n_active + p / L.gsz, 1);
}

// ---- SelectMinimalSolver (hybrid_ransac.h:210-243) + HybridUniformSampling::Sample for the
// next iterations of every iterating pair; one thread per pair
__global__ void sample_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const LoopParams &prm = *prmp;
    const int q = blockIdx.x * blockDim.x + threadIdx.x, p = first + q;
    if (q >= count) return;
    PairState &S = L.st[p];
    if (S.phase != PH_ITER) { S.batch_len = 0; return; }
    PairWork &W = L.wk[p];
    MtStream *A = &L.rngA[p], *B = &L.rngB[p];
    const int n = L.pc[p].n;
    const unsigned max_it = prm.opt.max_num_iterations > prm.opt.min_num_iterations ? prm.opt.max_num_iterations
                                                                                   : prm.opt.min_num_iterations;
    const unsigned lo_start = prm.opt.lo_starting_iterations;
    unsigned len = (unsigned)S.B < max_it - S.it ? (unsigned)S.B : max_it - S.it;
    if (S.it < lo_start && len > lo_start - S.it) len = lo_start - S.it; // a batch never straddles lo_start
    for (unsigned b = 0; b < len; b++) {
        const double u = mt_uniform_real(B, S.prior[0] + S.prior[1]);
        int s = -1;
        double c = 0;
        for (int k = 0; k < 2; k++) {
            if (S.prior[k] == 0.0) continue;
            c += S.prior[k];
            if (u <= c) { s = k; break; }
        }
        if (s < 0) s = S.prior[1] > 0.0 ? 1 : 0; // unreachable (u <= sum of priors)
        int tmp[8], scratch[16];
        for (int t = 0; t < 3; t++) {
            const int k = S.sizes[s][t];
            if (k <= 0) continue;
            mt_draw_sample(A, n, k, tmp, scratch); // type 1 of solver 0 is drawn and discarded (quirk Q1)
            if ((s == 0 && t == 0) || (s == 1 && t == 2))
                for (int i = 0; i < k; i++) W.idx[b][i] = tmp[i];
        }
        W.solver[b] = s;
        W.posA[b] = A->pos; W.posB[b] = B->pos;
    }
    S.batch_len = (int)len;
}

} // namespace mp
