// Kernels of the device loop that are specialised per estimator kind (score + decide) and
// their launchers; instantiated once per kind in mp_kind{0,1,2}.cu so the three variants
// compile in parallel.
#pragma once
#include <cstdlib>
#include "mp_loop.cuh"

namespace mp {

// ---- K3 over all hypotheses of the round: one CTA per pair, tile staged once (bulk copy)
template <int KIND>
__global__ void __launch_bounds__(MP_CTA, 1) score_batch_kernel(LoopBuffers L, const LoopParams *prmp, int tile_cap, int first) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ unsigned short list[LP_BMAX * LP_MM];
    __shared__ int nlist;
    const LoopParams &prm = *prmp;
    const int p = first + blockIdx.x;
    PairWork &W = L.wk[p];
    const int blen = W.blen;
    if (blen == 0) return;
    const PairConst &pc = L.pc[p];
    if (threadIdx.x == 0) {
        int c = 0;
        for (int b = 0; b < blen; b++)
            for (int j = 0; j < W.count[b]; j++) list[c++] = (unsigned short)(b * LP_MM + j);
        nlist = c;
        mbar_init(&bar, 1);
    }
    __syncthreads();
    const int nh = nlist;
    if (nh == 0) return;
    if (threadIdx.x == 0) L.evals32[p] += (unsigned long long)nh * (unsigned long long)pc.n;
    const PairF pf = make_pairf(pc, prm.cfg.score_type);
    const float w0 = (float)pc.w[0], w2 = (float)pc.w[2];
    float4 *sP = reinterpret_cast<float4 *>(smem_raw), *sQ = sP + tile_cap;
    const float4 *gP = L.S32 + pc.off32, *gQ = gP + pc.n;
    const HypF *hyps = L.hyps + (size_t)p * LP_BMAX * LP_MM;
    uint32_t phase = 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int groups = (nh + MP_HT - 1) / MP_HT;
    for (int base = 0; base < pc.n; base += tile_cap) {
        const int len = min(tile_cap, pc.n - base);
        __syncthreads();
        stage_two(sP, gP + base, len * 16u, sQ, gQ + base, len * 16u, &bar, phase);
        TileF tile{sP, sQ, len};
        for (int g = warp; g < groups; g += MP_WARPS) {
            const int h0 = g * MP_HT, cnt = min(MP_HT, nh - h0);
            float out[MP_HT];
            warp_score_list<KIND>(tile, pf, w0, w2, hyps, list + h0, cnt, 0, len, out);
            if (lane == 0)
                for (int j = 0; j < cnt; j++) W.s32[list[h0 + j]] = (base == 0 ? 0.f : W.s32[list[h0 + j]]) + out[j];
        }
    }
}

// ---- the sequential part of the loop, one CTA per pair
template <int KIND> __device__ void block_update_termination(const BlockCtx &C, const Model &m, int *inl_best) {
    BlockScratch *sh = C.sh;
    PairState *S = C.S;
    block_inliers(*C.pc, C.M, m, C.pc->thr, inl_best, sh->cnt, sh->wcnt);
    if (threadIdx.x == 0) {
        const int n = C.pc->n;
        S->evals64 += n;
        S->best_num_inliers = sh->cnt[0] + sh->cnt[1] + sh->cnt[2];
        trace_event(C, -2.0, (double)sh->cnt[0] + 131072.0 * (double)sh->cnt[1] + 17179869184.0 * (double)sh->cnt[2]);
        for (int t = 0; t < 3; t++) { S->num_inl[t] = sh->cnt[t]; S->ratios[t] = n > 0 ? (double)sh->cnt[t] / n : 0.0; }
        for (int s = 0; s < 2; s++)
            S->cap[s] = num_required_iterations(S->ratios, 1.0 - C.prm->opt.success_probability, S->sizes[s],
                                                C.prm->opt.min_num_iterations, C.prm->opt.max_num_iterations_per_solver);
    }
    __syncthreads();
}

// scan_kernel: the iterations of this round, replayed in order (the PH_ITER part of the loop).  A pair that
// needs a local optimisation / the final stage leaves with its phase set and a ticket in lo_req; lo_kernel
// (usually on another stream, so that the other pairs keep iterating meanwhile) picks the ticket up.
template <int KIND> __global__ void __launch_bounds__(LP_THREADS, LP_CTAS_PER_SM) scan_kernel(LoopBuffers L, const LoopParams *prmp,
                                                                                const long long *off3, const int *caps, int first, int round) {
    __shared__ PairState S;
    __shared__ BlockScratch sh;
    const LoopParams &prm = *prmp;
    const int p = first + blockIdx.x;
    if (L.wk[p].blen == 0) return; // not iterating this round
    {
        const int *src = reinterpret_cast<const int *>(&L.st[p]);
        int *dst = reinterpret_cast<int *>(&S);
        for (int i = threadIdx.x; i < (int)(sizeof(PairState) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const PairConst &pc = L.pc[p];
    PairWork &W = L.wk[p];
    const Model *models = L.models + (size_t)p * LP_BMAX * LP_MM;
    MtStream *rA = &L.rngA[p], *rB = &L.rngB[p];
    const long long o3 = off3[p];
    int *inl_best = L.inl_best + 3 * pc.off;
    BlockCtx C;
    C.pc = &pc; C.M = L.M + pc.off; C.prm = &prm; C.S = &S; C.rngB = rB; C.sh = &sh;
    C.cap = caps[p];
    C.inl_cur = L.inl_cur + 3 * pc.off;
    C.flat = L.flat + 2 * o3; C.base = C.flat + C.cap; C.sample = L.sample + 3 * o3; // o3 = sum of caps of earlier pairs
    C.trace = L.trace ? L.trace + 2 * (size_t)p * L.trace_cap : nullptr; C.trace_cap = L.trace_cap;
    const unsigned lo_start = prm.opt.lo_starting_iterations;
    const unsigned max_it = prm.opt.max_num_iterations > prm.opt.min_num_iterations ? prm.opt.max_num_iterations
                                                                                   : prm.opt.min_num_iterations;
    const int n = pc.n;

    if (S.phase == PH_ITER) {
        const long long scan_t0 = clock64();
        bool lo_event = false;
        for (int b = 0; b < S.batch_len; b++) {
            const unsigned it = S.it;
            const int s = W.solver[b], cnt = W.count[b];
            __syncthreads();
            if (threadIdx.x == 0) { S.iters[s] += 1; S.hyps_scored += cnt; S.evals += (unsigned long long)cnt * n; }
            bool skip_break = false;
            if (cnt > 0) {
                const double bound = S.best_min_score < DBL_MAX ? S.best_min_score * (1.0 + prm.guard) : DBL_MAX;
                unsigned cand = 0;
                for (int j = 0; j < cnt; j++)
                    if (!((double)W.s32[b * LP_MM + j] > bound)) cand |= 1u << j;
                if (cand != 0 || it == lo_start) {
                    double loc = DBL_MAX;
                    int loc_id = -1;
                    for (int j = 0; j < cnt; j++) {
                        if (!((cand >> j) & 1u)) continue;
                        const double sc = block_score64(pc, C.M, models[b * LP_MM + j], prm.cfg.score_type, sh.red);
                        if (threadIdx.x == 0) S.evals64 += n;
                        if (sc < loc) { loc = sc; loc_id = j; }
                    }
                    if (threadIdx.x == 0 && loc_id >= 0) trace_event(C, (double)it, loc);
                    const bool newbest = loc_id >= 0 && loc < S.best_min_score;
                    if (newbest || it == lo_start) {
                        __syncthreads();
                        if (threadIdx.x == 0 && newbest) {
                            S.best_min_score = loc; S.best_min = models[b * LP_MM + loc_id];
                            if (loc < S.best_score) { S.best_score = loc; S.best = S.best_min; S.best_solver = s; }
                        }
                        __syncthreads();
                        const bool run_lo = it >= lo_start && S.best_min_score < DBL_MAX;
                        if (!newbest && !run_lo) skip_break = true; // `continue` of hybrid_ransac.h:140-141 (quirk Q13)
                        else if (run_lo) {
                            if (threadIdx.x == 0) {
                                S.phase = PH_LO_MAIN; S.ev_solver = s; S.it = it + 1;
                                rA->pos = W.posA[b]; rB->pos = W.posB[b]; // discard the speculated rest of the batch
                            }
                            __syncthreads();
                            lo_event = true;
                            break;
                        } else block_update_termination<KIND>(C, S.best, inl_best);
                    }
                }
            }
            __syncthreads();
            if (!skip_break && S.iters[s] >= S.cap[s]) { // hybrid_ransac.h:169
                if (threadIdx.x == 0) {
                    S.broke = 1; S.total = it; S.phase = PH_FINISH;
                    rA->pos = W.posA[b]; rB->pos = W.posB[b];
                }
                __syncthreads();
                break;
            }
            if (threadIdx.x == 0) S.it = it + 1;
            __syncthreads();
        }
        __syncthreads();
        if (S.phase == PH_ITER) {
            if (threadIdx.x == 0) {
                if (S.it >= max_it) { S.total = max_it; S.phase = PH_FINISH; }
                else if (S.it == lo_start && !S.pre_done) {
                    S.pre_done = 1;
                    if (S.best_min_score < DBL_MAX) S.phase = PH_LO_PRE; // hybrid_ransac.h:88-96
                }
                S.B = min(prm.bmax, S.B * 2);
            }
        } else if (lo_event && threadIdx.x == 0) S.B = prm.b_after_lo;
        if (threadIdx.x == 0) S.clk[0] += (unsigned long long)(clock64() - scan_t0);
        __syncthreads();
    }


    {
        int *dst = reinterpret_cast<int *>(&L.st[p]);
        const int *src = reinterpret_cast<const int *>(&S);
        for (int i = threadIdx.x; i < (int)(sizeof(PairState) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (S.phase == PH_ITER) atomicAdd(L.n_iter + p / L.gsz, 1);
        else *reinterpret_cast<volatile int *>(L.lo_req + p) = round; // PH_LO_PRE / PH_LO_MAIN / PH_FINISH
    }
}

// lo_kernel: local optimisations and the final stage of the pairs holding a ticket of a round <= `round`.
// THREADS: LP_THREADS (pairs below LP_WIDE_N matches, two CTAs per SM) or LP_WIDE_THREADS (the others, one per SM); a CTA
// of the wrong width for its pair leaves at once (the host launches a width only when the group holds such pairs).
template <int KIND, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == LP_THREADS ? LP_CTAS_PER_SM : 1) lo_kernel(LoopBuffers L, const LoopParams *prmp,
                                                                              const long long *off3, const int *caps, int first, int round,
                                                                              int heavy_cls, int heavy_from) {
    __shared__ PairState S;
    __shared__ BlockScratch sh;
    __shared__ double lo_score;
    __shared__ int claimed;
    const LoopParams &prm = *prmp;
    const int p = first + blockIdx.x;
    if ((L.pc[p].n >= LP_WIDE_N) != (THREADS == LP_WIDE_THREADS)) return;
    // heavy_cls: 0 = every pair; 1 / 2 = only the pairs that have run fewer than / at least heavy_from local optimisations
    // so far (the launch on the high-priority stream takes the pairs with the long chains first: scheduling only)
    if (heavy_cls != 0 && ((*reinterpret_cast<volatile int *>(&L.st[p].number_lo) >= heavy_from) != (heavy_cls == 2))) return;
    if (threadIdx.x == 0) {
        const int t = *reinterpret_cast<volatile int *>(L.lo_req + p);
        claimed = (t > 0 && t <= round && atomicCAS(L.lo_req + p, t, -t) == t) ? 1 : 0;
        __threadfence();
    }
    __syncthreads();
    if (!claimed) return;
    {
        const int *src = reinterpret_cast<const int *>(&L.st[p]);
        int *dst = reinterpret_cast<int *>(&S);
        for (int i = threadIdx.x; i < (int)(sizeof(PairState) / 4); i += blockDim.x) dst[i] = __ldcg(src + i);
    }
    __syncthreads();
    const PairConst &pc = L.pc[p];
    PairWork &W = L.wk[p];
    const Model *models = L.models + (size_t)p * LP_BMAX * LP_MM;
    MtStream *rA = &L.rngA[p], *rB = &L.rngB[p];
    const long long o3 = off3[p];
    int *inl_best = L.inl_best + 3 * pc.off;
    BlockCtx C;
    C.pc = &pc; C.M = L.M + pc.off; C.prm = &prm; C.S = &S; C.rngB = rB; C.sh = &sh;
    C.cap = caps[p];
    C.inl_cur = L.inl_cur + 3 * pc.off;
    C.flat = L.flat + 2 * o3; C.base = C.flat + C.cap; C.sample = L.sample + 3 * o3; // o3 = sum of caps of earlier pairs
    C.trace = L.trace ? L.trace + 2 * (size_t)p * L.trace_cap : nullptr; C.trace_cap = L.trace_cap;
    const unsigned lo_start = prm.opt.lo_starting_iterations;
    const unsigned max_it = prm.opt.max_num_iterations > prm.opt.min_num_iterations ? prm.opt.max_num_iterations
                                                                                   : prm.opt.min_num_iterations;
    const int n = pc.n;

    while (S.phase == PH_LO_PRE || S.phase == PH_LO_MAIN || S.phase == PH_FINISH) {
        if (S.phase == PH_LO_PRE) {
            if (threadIdx.x == 0) S.number_lo += 1;
            __syncthreads();
            block_local_optimization<KIND>(C, S.best_solver, &S.best, &S.best_score, &S.best_solver);
            block_update_termination<KIND>(C, S.best, inl_best);
            if (threadIdx.x == 0) S.phase = PH_ITER;
        } else if (S.phase == PH_LO_MAIN) {
            if (threadIdx.x == 0) { S.number_lo += 1; lo_score = S.best_min_score; }
            __syncthreads();
            block_local_optimization<KIND>(C, S.best_solver, &S.best_min, &lo_score, &S.best_solver);
            if (threadIdx.x == 0) {
                if (lo_score < S.best_score) { S.best_score = lo_score; S.best = S.best_min; S.best_solver = S.ev_solver; }
            }
            __syncthreads();
            block_update_termination<KIND>(C, S.best, inl_best);
            if (threadIdx.x == 0) {
                if (S.iters[S.ev_solver] >= S.cap[S.ev_solver]) { S.broke = 1; S.total = S.it - 1; S.phase = PH_FINISH; }
                else if (S.it >= max_it) { S.total = max_it; S.phase = PH_FINISH; }
                else S.phase = PH_ITER;
            }
        } else { // PH_FINISH: hybrid_ransac.h:177-203
            if (S.total <= lo_start && S.best_score < DBL_MAX) {
                if (threadIdx.x == 0) S.number_lo += 1;
                __syncthreads();
                block_local_optimization<KIND>(C, S.best_solver, &S.best, &S.best_score, &S.best_solver);
                block_update_termination<KIND>(C, S.best, inl_best);
            }
            if (prm.opt.final_least_squares) {
                __syncthreads();
                if (threadIdx.x == 0) sh.m_tmp = S.best;
                __syncthreads();
                // the LM below overwrites nothing it reads: lists = inlier lists of the global best
                block_lsq<KIND>(C, inl_best, S.num_inl[0], inl_best + n, S.num_inl[1], inl_best + 2 * n, S.num_inl[2], &sh.m_tmp);
                const double sc = block_score_counted(C, sh.m_tmp);
                __syncthreads();
                if (sc < S.best_score) {
                    __syncthreads();
                    if (threadIdx.x == 0) { S.best_score = sc; S.best = sh.m_tmp; }
                    __syncthreads();
                    block_update_termination<KIND>(C, S.best, inl_best);
                }
            }
            if (threadIdx.x == 0) S.phase = PH_DONE;
        }
        __syncthreads();
    }


    // publish: everything but the phase first (the copy carries PH_BUSY), then the ticket, then the phase
    __shared__ int final_phase;
    if (threadIdx.x == 0) { final_phase = S.phase; S.phase = PH_BUSY; }
    __syncthreads();
    {
        int *dst = reinterpret_cast<int *>(&L.st[p]);
        const int *src = reinterpret_cast<const int *>(&S);
        for (int i = threadIdx.x; i < (int)(sizeof(PairState) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        *reinterpret_cast<volatile int *>(L.lo_req + p) = 0;
        __threadfence();
        atomicExch(&L.st[p].phase, final_phase);
        if (final_phase == PH_ITER) atomicAdd(L.n_ret + p / L.gsz, 1);
        if (final_phase == PH_DONE) {
            atomicSub(L.n_active + p / L.gsz, 1);
            for (int q = 0; q < 16; q++) atomicAdd(L.clk + q, S.clk[q]);
        }
    }
}


template <int KIND>
cudaError_t launch_score_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, int tile_cap, size_t smem, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(score_batch_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    score_batch_kernel<KIND><<<count, MP_CTA, smem, s>>>(L, dprm, tile_cap, first);
    return cudaGetLastError();
}
template <int KIND>
cudaError_t launch_scan_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, const long long *off3,
                             const int *caps, int round, cudaStream_t s) {
    scan_kernel<KIND><<<count, LP_THREADS, 0, s>>>(L, dprm, off3, caps, first, round);
    return cudaGetLastError();
}
template <int KIND>
cudaError_t launch_lo_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, const long long *off3,
                           const int *caps, int round, int widths, cudaStream_t s, cudaStream_t s_heavy, int heavy_from) {
    // CTA size of the local optimisation: 128 measured best on configs[1] (64: 4 CTAs per SM but 12 % slower per step,
    // 32: worse); MP_LO_THREADS overrides it for experiments.  widths: bit 0 = the group holds pairs below LP_WIDE_N
    // matches, bit 1 = pairs from LP_WIDE_N on (LP_WIDE_THREADS threads, mp_loop.cuh).
    // s_heavy != nullptr: the pairs with >= heavy_from local optimisations behind them go to that (high-priority) stream
    static const int lo_threads = [] { const char *e = getenv("MP_LO_THREADS"); const int v = e ? atoi(e) : LP_THREADS; return (v == 32 || v == 64 || v == 128) ? v : LP_THREADS; }();
    const int ncls = s_heavy ? 2 : 1;
    for (int k = 0; k < ncls; k++) {
        const int cls = s_heavy ? (k == 0 ? 2 : 1) : 0; // the heavy class is enqueued first
        cudaStream_t st = (cls == 2) ? s_heavy : s;
        if (widths & 1) lo_kernel<KIND, LP_THREADS><<<count, lo_threads, 0, st>>>(L, dprm, off3, caps, first, round, cls, heavy_from);
        if (widths & 2) lo_kernel<KIND, LP_WIDE_THREADS><<<count, LP_WIDE_THREADS, 0, st>>>(L, dprm, off3, caps, first, round, cls, heavy_from);
    }
    return cudaGetLastError();
}

// mp_lm_batch: one CTA per LM problem of one pair
template <int KIND>
__global__ void __launch_bounds__(LP_THREADS, LP_CTAS_PER_SM) lm_batch_kernel(const PairConst *pcp, const Match64 *M, const LoopParams *prmp,
                                                                  int n_problems, const Model *start, const int *idx,
                                                                  const int *off, Model *out, double *info, bool nonminimal) {
    __shared__ PairState S;
    __shared__ BlockScratch sh;
    const int j = blockIdx.x;
    if (threadIdx.x == 0) { S.lm_solves = 0; S.lm_iters = 0; S.lm_blocks = 0; for (int q = 0; q < 16; q++) S.clk[q] = 0; sh.m_tmp = start[j]; }
    __syncthreads();
    BlockCtx C;
    C.pc = pcp; C.M = M + pcp->off; C.prm = prmp; C.S = &S; C.rngB = nullptr; C.sh = &sh; C.cap = 0;
    C.inl_cur = nullptr; C.flat = nullptr; C.base = nullptr; C.sample = nullptr; C.trace = nullptr; C.trace_cap = 0;
    const int *l0 = idx + off[3 * j], *l1 = idx + off[3 * j + 1], *l2 = idx + off[3 * j + 2];
    const int n0 = off[3 * j + 1] - off[3 * j], n1 = off[3 * j + 2] - off[3 * j + 1], n2 = off[3 * j + 3] - off[3 * j + 2];
    __shared__ LmInfo inf;
    if (threadIdx.x == 0) { inf = LmInfo{0, 0, 0, 0}; sh.lm_go = lm_setup(*pcp, prmp->cfg, n0, n1, n2, &sh.lm_setup, nonminimal) ? 1 : 0; }
    __syncthreads();
    if (sh.lm_go) {
        BlockEval<KIND> ev;
        ev.C = &C; ev.L = &sh.lm_setup; ev.lst[0] = l0; ev.lst[1] = l1; ev.lst[2] = l2; ev.n[0] = n0; ev.n[1] = n1; ev.n[2] = n2;
        block_lm<KIND>(C, sh.lm_setup, ev, sh.m_tmp, &sh.m_tmp, &inf);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        out[j] = sh.m_tmp;
        info[4 * j] = inf.iterations; info[4 * j + 1] = inf.termination; info[4 * j + 2] = inf.cost0; info[4 * j + 3] = inf.cost;
    }
}
template <int KIND>
cudaError_t launch_lm_kind(const PairConst *pc, const Match64 *M, const LoopParams *dprm, int n_problems, const Model *start,
                           const int *idx, const int *off, Model *out, double *info, bool nonminimal, cudaStream_t s) {
    lm_batch_kernel<KIND><<<n_problems, LP_THREADS, 0, s>>>(pc, M, dprm, n_problems, start, idx, off, out, info, nonminimal);
    return cudaGetLastError();
}

} // namespace mp
