// LocalOptimization / LeastSquaresFit (hybrid_ransac.h:383-538) and the LM solves they
// issue, executed by one CTA.  Control flow is uniform over the CTA: every thread runs the
// scalar driver on identical data; the residual-block sums, inlier extraction, scoring and
// the tail of the Fisher-Yates shuffles are spread over the threads.
#pragma once

namespace mp {

#define MP_CLK_BEGIN() const long long clk_t0_ = clock64()
#define MP_CLK_END(C_, slot_) do { if (threadIdx.x == 0) (C_).S->clk[slot_] += (unsigned long long)(clock64() - clk_t0_); } while (0)

constexpr int SH_MAXHEAD = 1024, SH_TAB = 2048;
constexpr int RS_CH = 8, RS_LD = 33; // warp reduction of the LM sums through shared memory: values per pass, padded row

struct BlockScratch { // shared memory of the decide kernel
    double red[32];
    double part[LP_WIDE_THREADS / 32][LM_ACC_DOUBLES];
    double total[LM_ACC_DOUBLES];
    int cnt[3], wcnt[3][16];
    int flag[4], base_len, fit_len, fit_cnt[3];
    Model m_lo, m_init, m_tmp;
    // LM: the point the CTA evaluates next and the driver's verdict (written by thread 0)
    double lm_x[LM_NA];
    int lm_go;
    LmSetup lm_setup;
    LmDriver drv;
    union { // two phases that never overlap share this memory
        struct { // fast Fisher-Yates head: draws, prefetched values, bucket offsets / entries, predecessor tables
            int sh_j[SH_MAXHEAD], sh_vi[SH_MAXHEAD], sh_vj[SH_MAXHEAD];
            int sh_key[SH_TAB], sh_val[SH_TAB];
        } f;
        double rs_tile[LP_WIDE_THREADS / 32][RS_CH * RS_LD]; // per warp: the LM evaluation's sums, RS_CH values x 32 lanes at a time
    } u;
};

struct BlockCtx {
    const PairConst *pc;
    const Match64 *M;      // this pair's records
    const LoopParams *prm;
    PairState *S;          // shared-memory copy of the pair state
    MtStream *rngB;
    int *inl_cur, *flat, *base, *sample; // global scratch of this pair
    int cap;               // capacity of flat / base and of each of the 3 sample lists
    BlockScratch *sh;
    double *trace;         // this pair's trace records or nullptr
    int trace_cap;
};

// thread 0 only
__device__ inline void trace_event(const BlockCtx &C, double tag, double value) {
    if (!C.trace) return;
    const int k = C.S->trace_n;
    if (k < C.trace_cap) { C.trace[2 * k] = tag; C.trace[2 * k + 1] = value; }
    C.S->trace_n = k + 1;
}

// ---- collective evaluation of cost, J^T J, J^T r at sh->lm_x; the sums are left in sh->total
template <int KIND> struct BlockEval {
    const BlockCtx *C;
    const LmSetup *L;
    const int *lst[3];
    int n[3];
    __device__ __noinline__ void operator()() const {
        MP_CLK_BEGIN();
        BlockScratch *sh = C->sh;
        __builtin_assume(__isShared(sh)); __builtin_assume(__isShared(L)); // (address-space hints: see mp_lm_warp.cuh)
        LmPoint p;
        lm_point(KIND, sh->lm_x, &p);
        LmAcc a;
        a.clear();
        const int n0 = L->use_reproj ? n[0] : 0, n1 = L->use_reproj ? n[1] : 0, n2 = L->use_samp ? n[2] : 0;
        const int ntot = n0 + n1 + n2;
        for (int j = threadIdx.x; j < ntot; j += blockDim.x) {
            int t, k;
            if (j < n0) { t = 0; k = j; } else if (j < n0 + n1) { t = 1; k = j - n0; } else { t = 2; k = j - n0 - n1; }
            const Match64 m = C->M[lst[t][k]];
            lm_block<KIND>(*C->pc, m, t, p, L->W, &a);
        }
        if (threadIdx.x == 0) C->S->lm_blocks += ntot;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
        // warp sums of the 78 accumulators through shared memory, RS_CH values at a time: every lane stores its RS_CH
        // values (row = value, column = lane, padded rows), lane (c, q) = (lane % 8, lane / 8) adds the 8 columns of quarter
        // q of value c in order, two shuffles combine the quarters.  ~35 instructions per pass, 10 passes; the register
        // reduce-scatter this replaces (80 -> 40 -> 20 -> 10 -> 5 values per lane with selects around 160 64-bit shuffles)
        // compiled to 2 200 instructions -- a quarter of the evaluation's instruction stream for ~2 000 residual blocks.
        // The order of the additions is fixed and does not depend on the CTA width.
        {
            const double *av = reinterpret_cast<const double *>(&a);
            double *tile = sh->u.rs_tile[warp];
            const int c = lane & (RS_CH - 1), q = lane >> 3;
#pragma unroll
            for (int ch = 0; ch < (LM_ACC_DOUBLES + RS_CH - 1) / RS_CH; ch++) {
#pragma unroll
                for (int k = 0; k < RS_CH; k++)
                    if (ch * RS_CH + k < LM_ACC_DOUBLES) tile[k * RS_LD + lane] = av[ch * RS_CH + k];
                __syncwarp();
                const double *col = tile + c * RS_LD + 8 * q;
                double s = col[0];
#pragma unroll
                for (int k = 1; k < 8; k++) s += col[k];
                s += __shfl_xor_sync(0xffffffffu, s, 8);
                s += __shfl_xor_sync(0xffffffffu, s, 16);
                if (lane < RS_CH && ch * RS_CH + lane < LM_ACC_DOUBLES) sh->part[warp][ch * RS_CH + lane] = s;
                __syncwarp();
            }
        }
        __syncthreads();
        for (int v = threadIdx.x; v < LM_ACC_DOUBLES; v += blockDim.x) { // any CTA size
            double s = 0;
            for (int w = 0; w < nw; w++) s += sh->part[w][v];
            sh->total[v] = s;
        }
        __syncthreads();
        MP_CLK_END(*C, 3);
    }
};
static_assert(sizeof(LmAcc) == LM_ACC_DOUBLES * sizeof(double), "LmAcc is a flat array of doubles");

// One LM solve by the CTA: warp 0 runs the driver (LmDriver state, mp_lm_warp.cuh), everybody evaluates.
// x0 -> x (final parameters, valid on thread 0 only); returns info on thread 0.
template <int KIND> __device__ __noinline__ void block_lm(const BlockCtx &C, const LmSetup &L, const BlockEval<KIND> &ev,
                                             const Model &start, Model *out, LmInfo *info_out) {
    MP_CLK_BEGIN();
    BlockScratch *sh = C.sh;
    __builtin_assume(__isShared(sh)); __builtin_assume(__isShared(&L)); __builtin_assume(__isShared(&start)); __builtin_assume(__isShared(out));
    LmDriver &drv = sh->drv; // driven by thread 0 only; shared memory keeps its arrays off the local-memory path
    __syncthreads();
    if (threadIdx.x == 0) {
        double x0[LM_NA];
        lm_model_to_x(start, x0);
        sh->lm_go = drv.begin(&L, x0) ? 1 : 0;
        for (int q = 0; q < 8; q++) drv.prof[q] = 0;
        for (int k = 0; k < LM_NA; k++) sh->lm_x[k] = drv.x[k];
    }
    __syncthreads();
    if (sh->lm_go) {
        const int lane = threadIdx.x & 31;
        bool first = true;
        while (true) { // ONE call site for each part of the driver: its code exists once in the kernel
            ev();
            if (threadIdx.x < 32) { // warp 0 drives (mp_lm_warp.cuh); the other warps wait at the barrier
                const bool go = (first ? warp_lm_first(drv, sh->total, lane) : warp_lm_feed(drv, sh->total, lane)) &&
                                warp_lm_propose(drv, lane);
                if (lane == 0) sh->lm_go = go ? 1 : 0;
                if (lane < LM_NA) sh->lm_x[lane] = drv.xc[lane];
            }
            first = false;
            __syncthreads();
            if (!sh->lm_go) break;
        }
    }
    if (threadIdx.x == 0) {
        Model m = start;
        lm_x_to_model(KIND, drv.x, &m);
        *out = m;
        *info_out = drv.info;
        for (int q = 0; q < 8; q++) C.S->clk[8 + q] += drv.prof[q];
    }
    __syncthreads();
    MP_CLK_END(C, 4);
}

// NonMinimalSolver / LeastSquares on explicit lists; *m is a shared-memory model.
template <int KIND> __device__ void block_lsq(const BlockCtx &C, const int *l0, int n0, const int *l1, int n1,
                                              const int *l2, int n2, Model *m, bool nonminimal = false) {
    BlockScratch *sh = C.sh;
    __syncthreads();
    if (threadIdx.x == 0) sh->lm_go = lm_setup(*C.pc, C.prm->cfg, n0, n1, n2, &sh->lm_setup, nonminimal) ? 1 : 0;
    __syncthreads();
    if (!sh->lm_go) return;
    BlockEval<KIND> ev;
    ev.C = &C; ev.L = &sh->lm_setup; ev.lst[0] = l0; ev.lst[1] = l1; ev.lst[2] = l2; ev.n[0] = n0; ev.n[1] = n1; ev.n[2] = n2;
    __shared__ LmInfo info;
    block_lm<KIND>(C, sh->lm_setup, ev, *m, m, &info);
    if (threadIdx.x == 0) {
        C.S->lm_solves += 1; C.S->lm_iters += info.iterations;
        trace_event(C, -3.0, (double)info.iterations); trace_event(C, -4.0, info.cost);
    }
    __syncthreads();
}

// utils::RandomShuffleAndResize (ext/RansacLib/RansacLib/utils.h:48-73) on v[0..*len) with
// stream B: for i in 0..n-2: swap(v[i], v[uniform_int(i, n-1)]), then resize(target).
// Only the first `target` outputs survive the resize and position i is final after step i, so
//  (1) ALL n-1 draws are evaluated in parallel straight from the generator ring (draw i uses
//      word pos+i unless an earlier draw hit a Lemire rejection, which is detected in parallel
//      and resolved sequentially for that one draw);
//  (2) the first head = min(target, n-1) swaps are replayed in parallel from prefetched
//      original values (see "Parallel replay" below).
static __device__ __noinline__ void block_shuffle_resize(const BlockCtx &C, int target, int *v, int *len_sh) {
    MP_CLK_BEGIN();
    MtStream *g = C.rngB;
    BlockScratch *sh = C.sh;
    __builtin_assume(__isShared(sh)); __builtin_assume(__isShared(len_sh)); __builtin_assume(__isGlobal(g)); __builtin_assume(__isGlobal(v));
    __syncthreads();
    const int n = *len_sh;
    __syncthreads();
    const int head = n - 1 < target ? (n - 1 > 0 ? n - 1 : 0) : target;
    const bool fast = head <= SH_MAXHEAD && n < (1 << 21);
    if (!fast) {
        if (threadIdx.x == 0)
            for (int i = 0; i < head; i++) {
                const int j = mt_uniform_int(g, i, n - 1);
                const int t = v[i]; v[i] = v[j]; v[j] = t;
            }
        __syncthreads();
    }
    int i = fast ? 0 : head; // draws still to take: uniform_int(i, n-1) for i in [i, n-1)
    while (i < n - 1) {
        if (g->pos == g->gen) mt_refill(g, (int)threadIdx.x, (int)blockDim.x, [] { __syncthreads(); });
        const unsigned long long pos = g->pos;
        const int avail = (int)(g->gen - pos), m = avail < n - 1 - i ? avail : n - 1 - i;
        int first = m;
        const unsigned rbase = (unsigned)(pos % MT_RING); // m <= 624 < MT_RING: one conditional wrap per index
        auto ring_at = [&](int j) { const unsigned r = rbase + (unsigned)j; return g->x[r >= (unsigned)MT_RING ? r - MT_RING : r]; };
        for (int j = threadIdx.x; j < m; j += blockDim.x) {
            const uint32_t w = mt_temper(ring_at(j));
            if (mt_lemire_rejects(w, (uint32_t)(n - (i + j)))) { first = j; break; }
        }
        __syncthreads();
        if (threadIdx.x == 0) sh->flag[0] = m;
        __syncthreads();
        if (first < m) atomicMin(&sh->flag[0], first);
        __syncthreads();
        first = sh->flag[0];
        if (fast) // draws [i, i+first) are valid: record those of the head
            for (int j = threadIdx.x; j < first && i + j < head; j += blockDim.x) {
                const uint32_t w = mt_temper(ring_at(j));
                const int ii = i + j, jj = ii + (int)(((unsigned long long)w * (uint32_t)(n - ii)) >> 32);
                sh->u.f.sh_j[ii] = jj; sh->u.f.sh_vi[ii] = v[ii]; sh->u.f.sh_vj[ii] = v[jj];
            }
        __syncthreads();
        if (threadIdx.x == 0) {
            g->pos = pos + first;
            if (first < m) { // the draw with a rejection, sequentially
                const int ii = i + first, jj = mt_uniform_int(g, ii, n - 1);
                if (fast && ii < head) { sh->u.f.sh_j[ii] = jj; sh->u.f.sh_vi[ii] = v[ii]; sh->u.f.sh_vj[ii] = v[jj]; }
            }
        }
        i += first + (first < m ? 1 : 0);
        __syncthreads();
    }
    if (fast && head > 0) {
        // Parallel replay.  Swap i exchanges positions i and j_i >= i, so position k < head is final after
        // step k and holds the value that sat at position j_k just before step k:
        //   a(i)   = value at position i just before step i = a(i') for the latest i' < i with j_i' == i, else v0[i]
        //   out[k] = a(k) if j_k == k, else a(q) for the latest q < k with j_q == j_k, else v0[j_k]
        // "latest earlier swap with the same target" = the largest time i < k with j_i == (j_k or k).  The swaps are
        // grouped by target through 1024 hash buckets (count, scan, scatter: the order inside a bucket is arbitrary and
        // does not matter, a bucket holds ~head / 1024 <= 1 swaps on average) and each query scans one bucket for the
        // maximum -- 8 barriers instead of the 55 stages of a bitonic sort of the (target, time) keys.
        constexpr int NB = 1024;
        int *start = sh->u.f.sh_key, *ent = sh->u.f.sh_key + NB;           // bucket offsets (exclusive scan), swap times by bucket
        int *ps = sh->u.f.sh_val, *pt = sh->u.f.sh_val + SH_MAXHEAD;       // pt doubles as the scatter cursors first
        for (int b = threadIdx.x; b < NB; b += blockDim.x) start[b] = 0;
        __syncthreads();
        for (int k = threadIdx.x; k < head; k += blockDim.x) atomicAdd(&start[sh->u.f.sh_j[k] & (NB - 1)], 1);
        __syncthreads();
        { // exclusive scan of the 1024 counts: a contiguous run per thread, warp scan, warp totals through sh->wcnt
            const int per = NB / (int)blockDim.x, b0 = threadIdx.x * per; // blockDim is 32, 64 or 128
            int run = 0;
            for (int q = 0; q < per; q++) run += start[b0 + q];
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            int inc = run;
            for (int dd = 1; dd < 32; dd <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc, dd); if (lane >= dd) inc += u; }
            if (lane == 31) sh->wcnt[0][warp] = inc;
            __syncthreads();
            int base = inc - run;
            for (int w = 0; w < warp; w++) base += sh->wcnt[0][w];
            for (int q = 0; q < per; q++) { const int c = start[b0 + q]; start[b0 + q] = base; pt[b0 + q] = base; base += c; }
        }
        __syncthreads();
        for (int k = threadIdx.x; k < head; k += blockDim.x) ent[atomicAdd(&pt[sh->u.f.sh_j[k] & (NB - 1)], 1)] = k;
        __syncthreads();
        auto latest_before = [&](int target, int before) { // max i < before with j_i == target, or -1
            const int b = target & (NB - 1), e1 = b + 1 < NB ? start[b + 1] : head;
            int best = -1;
            for (int e = start[b]; e < e1; e++) { const int i = ent[e]; if (i < before && i > best && sh->u.f.sh_j[i] == target) best = i; }
            return best;
        };
        // (the cursors in pt are dead after the barrier above; the queries read start / ent / sh_j only)
        for (int k = threadIdx.x; k < head; k += blockDim.x) { ps[k] = latest_before(sh->u.f.sh_j[k], k); pt[k] = latest_before(k, k); }
        const int last_hit = (head == n - 1 && n - 1 < target) ? latest_before(n - 1, head) : -1; // (every swap time is < head)
        __syncthreads();
        auto a_of = [&](int c) {
            while (pt[c] >= 0) c = pt[c];
            return sh->u.f.sh_vi[c];
        };
        for (int k = threadIdx.x; k < head; k += blockDim.x) {
            const int j = sh->u.f.sh_j[k];
            int out;
            if (j == k) out = a_of(k);
            else { const int q = ps[k]; out = q >= 0 ? a_of(q) : sh->u.f.sh_vj[k]; }
            v[k] = out;
        }
        if (threadIdx.x == 0 && last_hit >= 0) v[n - 1] = a_of(last_hit); // the list grows: its last element is kept too
    }
    __syncthreads();
    // resize: shrink, or grow with zeros (quirk Q3)
    for (int k = n + threadIdx.x; k < target; k += blockDim.x) v[k] = 0;
    __syncthreads();
    if (threadIdx.x == 0) *len_sh = target;
    __syncthreads();
    MP_CLK_END(C, 5);
}

// un-flatten (hybrid_ransac.h:443-449, 528-534): entry e -> (type e / n, index e % n), order kept.
// Stable three-way partition by ballot + prefix counts, one chunk of blockDim entries at a time.
static __device__ __noinline__ void block_unflatten(const BlockCtx &C, const int *flat, int len, int *cnt3_sh) {
    MP_CLK_BEGIN();
    BlockScratch *sh = C.sh;
    __builtin_assume(__isShared(sh)); __builtin_assume(__isShared(cnt3_sh)); __builtin_assume(__isGlobal(flat)); __builtin_assume(__isGlobal(C.sample));
    const int n = C.pc->n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    __syncthreads();
    if (threadIdx.x < 3) cnt3_sh[threadIdx.x] = 0;
    __syncthreads();
    for (int start = 0; start < len; start += blockDim.x) {
        const int k = start + threadIdx.x;
        int e = 0, t = -1;
        if (k < len) {
            e = flat[k]; t = 0;
            while (e >= n && t < 2) { e -= n; t++; }
        }
        unsigned bal[3];
        for (int q = 0; q < 3; q++) {
            bal[q] = __ballot_sync(0xffffffffu, t == q);
            if (lane == 0) sh->wcnt[q][warp] = __popc(bal[q]);
        }
        __syncthreads();
        if (t >= 0) {
            int off = cnt3_sh[t];
            for (int w = 0; w < warp; w++) off += sh->wcnt[t][w];
            C.sample[t * C.cap + off + __popc(bal[t] & ((1u << lane) - 1u))] = e;
        }
        __syncthreads();
        if (threadIdx.x < 3) {
            int c = 0;
            for (int w = 0; w < nw; w++) c += sh->wcnt[threadIdx.x][w];
            cnt3_sh[threadIdx.x] += c;
        }
        __syncthreads();
    }
    MP_CLK_END(C, 6);
}

// LeastSquaresFit (hybrid_ransac.h:484-538)
// have_inliers: inl_cur / sh->cnt already hold GetInliers(*m, thr) (block_score_inliers ran on the same model)
template <int KIND> __device__ __noinline__ void block_lsq_fit(const BlockCtx &C, const double *thr, int solver, Model *m,
                                                  bool all_inliers, bool have_inliers = false) {
    BlockScratch *sh = C.sh;
    __builtin_assume(__isShared(sh)); __builtin_assume(__isShared(m)); __builtin_assume(__isShared(C.S));
    __builtin_assume(__isGlobal(C.inl_cur)); __builtin_assume(__isGlobal(C.flat)); __builtin_assume(__isGlobal(C.pc));
    const int n = C.pc->n;
    if (!have_inliers) { MP_CLK_BEGIN(); block_inliers(*C.pc, C.M, *m, thr, C.inl_cur, sh->cnt, sh->wcnt); MP_CLK_END(C, 1); }
    if (threadIdx.x == 0) { C.S->evals64 += n; trace_event(C, -2.0, (double)sh->cnt[0] + 131072.0 * (double)sh->cnt[1] + 17179869184.0 * (double)sh->cnt[2]); }
    const int c0 = sh->cnt[0], c1 = sh->cnt[1], c2 = sh->cnt[2];
    const int *sz = C.S->sizes[solver];
    if (c0 < sz[0] || c1 < sz[1] || c2 < sz[2]) return;
    if (all_inliers) {
        block_lsq<KIND>(C, C.inl_cur, c0, C.inl_cur + n, c1, C.inl_cur + 2 * n, c2, m);
        return;
    }
    const int mult = C.prm->opt.min_sample_multiplicator;
    const int k0 = min(sz[0] * mult, c0), k1 = min(sz[1] * mult, c1), k2 = min(sz[2] * mult, c2);
    const int total = (k0 + k1 + k2) * mult; // x7 applied twice (quirk Q4)
    {
        int *flat = C.flat; const int *cur = C.inl_cur; // (local copies carry the address-space hints)
        __builtin_assume(__isGlobal(flat)); __builtin_assume(__isGlobal(cur));
        for (int k = threadIdx.x; k < c0; k += blockDim.x) flat[k] = cur[k];
        for (int k = threadIdx.x; k < c1; k += blockDim.x) flat[c0 + k] = cur[n + k] + n;
        for (int k = threadIdx.x; k < c2; k += blockDim.x) flat[c0 + c1 + k] = cur[2 * n + k] + 2 * n;
    }
    __syncthreads();
    if (threadIdx.x == 0) sh->fit_len = c0 + c1 + c2;
    block_shuffle_resize(C, total, C.flat, &sh->fit_len);
    block_unflatten(C, C.flat, total, sh->fit_cnt);
    const int s0 = sh->fit_cnt[0], s1 = sh->fit_cnt[1], s2 = sh->fit_cnt[2];
    block_lsq<KIND>(C, C.sample, s0, C.sample + C.cap, s1, C.sample + 2 * C.cap, s2, m);
}

// ScoreModel in fp64 + the work counters of the loop
__device__ inline double block_score_counted(const BlockCtx &C, const Model &m) {
    MP_CLK_BEGIN();
    const double s = block_score64(*C.pc, C.M, m, C.prm->cfg.score_type, C.sh->red);
    MP_CLK_END(C, 2);
    if (threadIdx.x == 0) { C.S->hyps_scored += 1; C.S->evals += C.pc->n; C.S->evals64 += C.pc->n; trace_event(C, -1.0, s); }
    return s;
}

// ScoreModel(m) and GetInliers(m, thr_inl) in ONE pass over the matches (the reference calls them back to back on the
// same model all through LocalOptimization); inliers are left in inl_cur / sh->cnt for the next block_lsq_fit.
// The work counters count both, as the reference evaluates both.
__device__ inline double block_score_inliers(const BlockCtx &C, const Model &m, const double *thr_inl) {
    MP_CLK_BEGIN();
    __shared__ double sc_sh;
    block_inliers(*C.pc, C.M, m, thr_inl, C.inl_cur, C.sh->cnt, C.sh->wcnt, &sc_sh, C.prm->cfg.score_type, C.sh->red);
    MP_CLK_END(C, 1);
    if (threadIdx.x == 0) { C.S->hyps_scored += 1; C.S->evals += C.pc->n; C.S->evals64 += C.pc->n; trace_event(C, -1.0, sc_sh); }
    return sc_sh;
}

// LocalOptimization (hybrid_ransac.h:383-482).  best / best_score / best_solver live in shared memory.
template <int KIND> __device__ __noinline__ void block_local_optimization(const BlockCtx &C, int solver, Model *best,
                                                             double *best_score, int *best_solver) {
    MP_CLK_BEGIN();
    BlockScratch *sh = C.sh;
    __builtin_assume(__isShared(sh)); __builtin_assume(__isShared(best)); __builtin_assume(__isShared(best_score)); __builtin_assume(__isShared(best_solver));
    __builtin_assume(__isShared(C.S)); __builtin_assume(__isGlobal(C.inl_cur)); __builtin_assume(__isGlobal(C.base)); __builtin_assume(__isGlobal(C.pc));
    const PairConst &pc = *C.pc;
    const mp_options &o = C.prm->opt;
    const int n = pc.n;
    double thr[3], step[3];
    for (int t = 0; t < 3; t++) {
        step[t] = (o.threshold_multiplier - 1.0) * pc.thr[t] / (double)(o.num_lsq_iterations - 1);
        thr[t] = pc.thr[t] * o.threshold_multiplier;
    }
    auto update_best = [&](double sc, const Model *m) {
        __syncthreads();
        if (threadIdx.x == 0 && sc < *best_score) { *best_score = sc; *best = *m; *best_solver = solver; }
        __syncthreads();
    };
    __syncthreads();
    if (threadIdx.x == 0) sh->m_init = *best;
    __syncthreads();
    block_lsq_fit<KIND>(C, thr, solver, &sh->m_init, true);
    // score of m_init + its base inliers at the nominal thresholds (flattened below with type offsets)
    double sc = block_score_inliers(C, sh->m_init, pc.thr);
    update_best(sc, &sh->m_init);
    if (threadIdx.x == 0) { C.S->evals64 += n; trace_event(C, -2.0, (double)sh->cnt[0] + 131072.0 * (double)sh->cnt[1] + 17179869184.0 * (double)sh->cnt[2]); }
    const int c0 = sh->cnt[0], c1 = sh->cnt[1], c2 = sh->cnt[2];
    {
        int *base = C.base; const int *cur = C.inl_cur;
        __builtin_assume(__isGlobal(base)); __builtin_assume(__isGlobal(cur));
        for (int k = threadIdx.x; k < c0; k += blockDim.x) base[k] = cur[k];
        for (int k = threadIdx.x; k < c1; k += blockDim.x) base[c0 + k] = cur[n + k] + n;
        for (int k = threadIdx.x; k < c2; k += blockDim.x) base[c0 + c1 + k] = cur[2 * n + k] + 2 * n;
    }
    __syncthreads();
    if (threadIdx.x == 0) sh->base_len = c0 + c1 + c2; // current length of the base list
    __syncthreads();
    const int nonmin = max(C.S->non_min_sample_size, min(C.S->min_sample_size * o.non_min_sample_multiplier, (c0 + c1 + c2) / 2));
    for (int r = 0; r < o.num_lo_steps; ++r) {
        // the sample is the list BEFORE this step's shuffle (quirk Q2)
        // (step 0: all base inliers = the three lists still held in inl_cur; later steps: the
        // nonmin-element list left by the previous shuffle)
        const int blen = sh->base_len;
        int s0 = c0, s1 = c1, s2 = c2;
        const int *l0 = C.inl_cur, *l1 = C.inl_cur + n, *l2 = C.inl_cur + 2 * n;
        __syncthreads();
        if (r > 0) {
            block_unflatten(C, C.base, blen, sh->fit_cnt);
            s0 = sh->fit_cnt[0]; s1 = sh->fit_cnt[1]; s2 = sh->fit_cnt[2];
            l0 = C.sample; l1 = C.sample + C.cap; l2 = C.sample + 2 * C.cap;
        }
        __syncthreads();
        block_shuffle_resize(C, nonmin, C.base, &sh->base_len);
        LmSetup Lg;
        if (!lm_setup(pc, C.prm->cfg, s0, s1, s2, &Lg, true)) continue; // NonMinimalSolver returned 0
        if (threadIdx.x == 0) sh->m_lo = sh->m_init;
        __syncthreads();
        block_lsq<KIND>(C, l0, s0, l1, s1, l2, s2, &sh->m_lo, true);
        sc = block_score_inliers(C, sh->m_lo, pc.thr); // the inliers feed the fit below
        update_best(sc, &sh->m_lo);
        block_lsq_fit<KIND>(C, pc.thr, solver, &sh->m_lo, false, true);
        double cur[3] = {thr[0], thr[1], thr[2]};
        for (int i = 0; i < o.num_lsq_iterations; ++i) {
            block_lsq_fit<KIND>(C, cur, solver, &sh->m_lo, false, i > 0);
            for (int t = 0; t < 3; t++) cur[t] -= step[t];
            if (i + 1 < o.num_lsq_iterations) sc = block_score_inliers(C, sh->m_lo, cur); // inliers at the next thresholds
            else sc = block_score_counted(C, sh->m_lo);
            update_best(sc, &sh->m_lo);
        }
    }
    __syncthreads();
    MP_CLK_END(C, 7);
}

} // namespace mp
