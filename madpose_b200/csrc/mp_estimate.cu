// Whole-path translation unit: mp_estimate_batch (host side of the device loop) and mp_lm_batch.
#include "mp_estimate.cuh"
#include "mp_loop_kernels.cuh"
#include <algorithm>
#include <thread>
#include <chrono>
#include <cstdio>
#include <cstdlib>

namespace mp {

constexpr int MP_MAX_GROUPS = 1024;

// PairState -> mp_model / mp_stats; undoes the focal normalisation
// (...shared_focal_estimator.cpp:49, ...two_focal_estimator.cpp:72-73)
__global__ void finalize_kernel(LoopBuffers L, int n_pairs, mp_model *models, mp_stats *stats) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const PairState &S = L.st[p];
    const PairConst &pc = L.pc[p];
    Model m = S.best;
    if (pc.kind == SHARED) { m.f0 *= pc.norm_scale; m.f1 = m.f0; }
    if (pc.kind == TWO) { m.f0 *= pc.norm_scale; m.f1 *= pc.norm_scale; }
    *reinterpret_cast<Model *>(models + p) = m;
    mp_stats o;
    o.num_iterations_total = S.total;
    o.num_iterations_per_solver[0] = S.iters[0]; o.num_iterations_per_solver[1] = S.iters[1];
    o.best_num_inliers = S.best_num_inliers; o.best_solver_type = S.best_solver; o.number_lo_iterations = S.number_lo;
    for (int t = 0; t < 3; t++) { o.num_inliers[t] = S.num_inl[t]; o.inlier_ratios[t] = S.ratios[t]; }
    o.reserved_ = 0;
    o.best_model_score = S.best_score;
    o.hypotheses_scored = S.hyps_scored; o.residual_evals = S.evals; o.lm_solves = S.lm_solves; o.lm_iterations = S.lm_iters;
    o.evals_fp64 = S.evals64; o.lm_block_evals = S.lm_blocks; o.evals_fp32 = L.evals32[p];
    stats[p] = o;
}

static int run_estimate_impl(mp_ctx *c, LoopAlloc &A, const mp_options *opt, const mp_config *cfg, mp_model *out_models,
                             mp_stats *out_stats, int32_t *out_inliers, bool record_start) {
    if (opt->num_lsq_iterations < 0 || opt->num_lo_steps < 0 || opt->min_sample_multiplicator < 0)
        return fail(MP_ERR_ARG, "negative LO option");
    if (c->resident_pairs <= 0) return fail(MP_ERR_ARG, "no resident batch: call mp_upload_batch first");
    // thresholds and weights are baked into the resident PairConst records by the front-end (A.1 rewrite): a resident
    // run with other values would silently use the stale ones
    for (int k = 0; k < 2; k++)
        if (opt->squared_inlier_thresholds[k] != c->up_thr[k] || opt->data_type_weights[k] != c->up_w[k])
            return fail(MP_ERR_ARG, "mp_estimate_resident: squared_inlier_thresholds / data_type_weights differ from the "
                                    "ones the batch was uploaded with; upload again");
    const int kind = c->resident_kind == MP_KIND_SCALE_ONLY ? CALIB : c->resident_kind, n_pairs = c->resident_pairs;
    cudaStream_t s = c->stream;
    if (record_start) MP_CUDA(cudaEventRecord(c->ev[0], s));
    const long long N = c->total_n;
    // per-pair scratch capacities
    MP_CUDA(A.h_off3.reserve((n_pairs + 1) * sizeof(long long)));
    MP_CUDA(A.h_caps.reserve(n_pairs * sizeof(int)));
    long long *off3 = A.h_off3.as<long long>();
    int *caps = A.h_caps.as<int>();
    const int k0 = kind == CALIB ? 3 : 4, k2 = kind == CALIB ? 5 : (kind == SHARED ? 6 : 7);
    const long long mult = opt->min_sample_multiplicator;
    const long long total_upper = std::max<long long>(2 * k0, k2) * mult * mult;
    off3[0] = 0;
    int max_n = 32;
    for (int p = 0; p < n_pairs; p++) {
        const int np_ = (int)(c->off[p + 1] - c->off[p]);
        long long cap = 3ll * np_ + std::max<long long>(total_upper, 64) + 64;
        if (cap > 2000000000ll) return fail(MP_ERR_ARG, "pair too large");
        caps[p] = (int)cap;
        off3[p + 1] = off3[p] + cap;
        max_n = std::max(max_n, np_);
    }
    const long long CAP = off3[n_pairs];
    const size_t slots = (size_t)n_pairs * LP_BMAX * LP_MM;
    MP_CUDA(A.st.reserve(n_pairs * sizeof(PairState)));
    MP_CUDA(A.wk.reserve(n_pairs * sizeof(PairWork)));
    MP_CUDA(A.rngA.reserve(n_pairs * sizeof(MtStream)));
    MP_CUDA(A.rngB.reserve(n_pairs * sizeof(MtStream)));
    MP_CUDA(A.models.reserve(slots * sizeof(Model)));
    MP_CUDA(A.hyps.reserve(slots * sizeof(HypF)));
    MP_CUDA(A.inl_best.reserve((3 * (size_t)N + 4) * sizeof(int)));
    MP_CUDA(A.inl_cur.reserve((3 * (size_t)N + 4) * sizeof(int)));
    MP_CUDA(A.flat.reserve(2 * (size_t)CAP * sizeof(int)));
    MP_CUDA(A.sample.reserve(3 * (size_t)CAP * sizeof(int)));
    MP_CUDA(A.misc.reserve(128 + 4 * MP_MAX_GROUPS));
    MP_CUDA(A.off3.reserve((n_pairs + 1) * sizeof(long long)));
    MP_CUDA(A.caps.reserve(n_pairs * sizeof(int)));
    MP_CUDA(A.prm.reserve(sizeof(LoopParams)));
    MP_CUDA(A.out_models.reserve(n_pairs * sizeof(mp_model)));
    MP_CUDA(A.out_stats.reserve(n_pairs * sizeof(mp_stats)));
    MP_CUDA(A.h_prm.reserve(sizeof(LoopParams)));
    MP_CUDA(A.h_active.reserve(4 * MP_MAX_GROUPS));
    LoopParams *prm = A.h_prm.as<LoopParams>();
    prm->opt = *opt; prm->cfg = *cfg; prm->n_pairs = n_pairs; // speculated iterations per pair per round: 128, or 256 for long loops (an iteration budget of thousands: every round
    // costs a fixed launch + solve latency; measured on configs[3], 64 pairs x ~7600 iterations: 176 -> 164 ms/step)
    // Small calls (<= 256 pairs) are bound by the number of rounds, not by solver throughput: they speculate 256 iterations
    // per round and restart at 128 after a local optimisation (measured, ms per call of 1 / 4 / 16 / 64 configs[1] pairs:
    // 4.1 / 12.8 / 20.4 / 26.1 -> 3.8 / 11.9 / 18.9 / 24.7).  Scheduling only: results do not depend on it.
    const bool small_call = n_pairs <= 256;
    prm->bmax = c->batch_override > 0 ? c->batch_override
                                      : ((small_call || std::max(opt->max_num_iterations, opt->min_num_iterations) >= 4096) ? LP_BMAX : LP_BDEF);
    {
        const char *e = getenv("MP_B_AFTER_LO"); // experiments
        const int v = e ? atoi(e) : 0;
        prm->b_after_lo = std::min(v > 0 ? v : (small_call ? 2 * LP_B0 : LP_B0), prm->bmax); // measured on configs[1] (bmax 64): 16 -> 245 ms, 32 -> 235, 64 -> 232
    } prm->guard = 1e-3;
    // Groups of consecutive pairs, each with its own round pipeline (sample -> minimal -> score -> scan) on its own
    // stream.  Local optimisations and the final stage run in lo_kernel launches on side streams, so a round never
    // waits for them: a pair inside a local optimisation sits the rounds out until its lo_kernel publishes it again.
    // which point-solver samples take the warp-per-sample form (kinds that have one): all of them for small calls,
    // the late iterations (>= MP_WARP_FROM_IT, i.e. the pairs that make up the ramp-down of a large batch) otherwise
    const unsigned warp_from = c->minimal_mode == 2 ? 0u : (c->minimal_mode == 1 ? 0xffffffffu :
                               (n_pairs <= MP_WARP_SOLVER_MAX_PAIRS ? 0u : (c->warp_from_it >= 0 ? (unsigned)c->warp_from_it : MP_WARP_FROM_IT)));
    prm->warp_from_it = (kind == SHARED) ? warp_from : 0xffffffffu;
    const int warp_solvers = prm->warp_from_it != 0xffffffffu ? 1 : 0;
    const bool serial = c->groups_override == 1; // one stream, scan and lo back to back (per-class timing pass)
    // automatic: about 750 pairs per group, at least two groups from 16 pairs on (two round pipelines overlap each
    // other's launch gaps; measured on configs[3], 64 pairs: 187 -> 176 ms/step)
    int G = c->groups_override > 0 ? c->groups_override : std::max((n_pairs + 749) / 750, n_pairs >= 16 ? 2 : 1);
    G = std::max(1, std::min(G, c->groups_override > 0 ? MP_MAX_GROUPS : 8));
    const int gsz = (n_pairs + G - 1) / G;
    G = (n_pairs + gsz - 1) / gsz;
    constexpr int SLOTS = 4, NSIDE = 12;
    MP_CUDA(A.misc.reserve(128 + 28 * MP_MAX_GROUPS));
    MP_CUDA(A.cand.reserve(slots * sizeof(int)));
    MP_CUDA(A.cand_ok.reserve(slots));
    MP_CUDA(A.items.reserve(3 * (size_t)n_pairs * LP_BMAX * sizeof(int)));
    MP_CUDA(A.h_active.reserve(4 * (2 + 2 * SLOTS) * MP_MAX_GROUPS));
    MP_CUDA(A.lo_req.reserve(n_pairs * sizeof(int)));
    int *h_active = A.h_active.as<int>(), *h_iter = h_active + G, *h_side = h_active + 2 * G; // h_side[2 * (g * SLOTS + k)] = {pairs not done, pairs handed back so far}
    for (int g = 0; g < G; g++) { h_active[g] = std::min(gsz, n_pairs - g * gsz); h_iter[g] = 0; }
    // MP_PRIO (experiments): 1 = the round pipelines run at high priority (their CTAs take freed SM slots before the
    // local optimisations' CTAs), 2 = the other way round, 0 = equal
    static const int prio_mode = [] { const char *e = getenv("MP_PRIO"); return e ? atoi(e) : 0; }();
    int prio_lo = 0, prio_hi = 0;
    MP_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi)); // numerically lower = higher priority
    while ((int)c->gstream.size() < G) {
        cudaStream_t gs; cudaEvent_t ge;
        MP_CUDA(cudaStreamCreateWithPriority(&gs, cudaStreamNonBlocking, prio_mode == 1 ? prio_hi : prio_lo));
        MP_CUDA(cudaEventCreateWithFlags(&ge, cudaEventDisableTiming));
        c->gstream.push_back(gs); c->gev.push_back(ge);
    }
    while ((int)c->sstream.size() < NSIDE) {
        cudaStream_t gs;
        MP_CUDA(cudaStreamCreateWithPriority(&gs, cudaStreamNonBlocking, prio_mode == 2 ? prio_hi : prio_lo));
        c->sstream.push_back(gs);
    }
    // Local optimisations of pairs that already have 2 or more behind them are launched on high-priority streams: the
    // step ends with the pairs that need 8-11 of them (2.8 on average), each a ~5 ms chain, and their CTAs should not
    // queue behind the others'.  Scheduling only -- a pair's result does not depend on it.  Measured on three 1500-pair
    // ranges of configs[1]: 124.6 / 134.4 / 133.1 -> 125.0 / 132.0 / 131.2 ms per step.  MP_LO_HEAVY = k overrides the
    // threshold (0 = off).
    static const int heavy_from = [] { const char *e = getenv("MP_LO_HEAVY"); return e ? atoi(e) : 2; }();
    while (heavy_from > 0 && (int)c->hstream.size() < NSIDE) {
        cudaStream_t hs; cudaEvent_t he;
        MP_CUDA(cudaStreamCreateWithPriority(&hs, cudaStreamNonBlocking, prio_hi));
        MP_CUDA(cudaEventCreateWithFlags(&he, cudaEventDisableTiming));
        c->hstream.push_back(hs); c->hev.push_back(he);
    }
    while ((int)c->sev.size() < G * SLOTS) {
        cudaEvent_t ge;
        MP_CUDA(cudaEventCreateWithFlags(&ge, cudaEventDisableTiming));
        c->sev.push_back(ge);
    }
    if (!c->ev_init) MP_CUDA(cudaEventCreateWithFlags(&c->ev_init, cudaEventDisableTiming));
    MP_CUDA(cudaMemcpyAsync(A.prm.p, prm, sizeof(LoopParams), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(A.off3.p, off3, (n_pairs + 1) * sizeof(long long), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(A.caps.p, caps, n_pairs * sizeof(int), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(A.misc.as<char>() + 128, h_active, G * sizeof(int), cudaMemcpyHostToDevice, s));
    LoopBuffers L;
    L.pc = c->d_pc.as<PairConst>(); L.M = c->d_m64.as<Match64>(); L.S32 = c->d_s32.as<float4>();
    L.st = A.st.as<PairState>(); L.wk = A.wk.as<PairWork>(); L.rngA = A.rngA.as<MtStream>(); L.rngB = A.rngB.as<MtStream>();
    L.models = A.models.as<Model>(); L.hyps = A.hyps.as<HypF>();
    L.inl_best = A.inl_best.as<int>(); L.inl_cur = A.inl_cur.as<int>(); L.flat = A.flat.as<int>(); L.sample = A.sample.as<int>();
    L.n_active = reinterpret_cast<int *>(A.misc.as<char>() + 128);
    L.n_iter = reinterpret_cast<int *>(A.misc.as<char>() + 128 + 4 * MP_MAX_GROUPS);
    L.n_items = reinterpret_cast<int *>(A.misc.as<char>() + 128 + 8 * MP_MAX_GROUPS);
    L.n_ret = reinterpret_cast<int *>(A.misc.as<char>() + 128 + 16 * MP_MAX_GROUPS);
    L.n_items_w = reinterpret_cast<int *>(A.misc.as<char>() + 128 + 20 * MP_MAX_GROUPS);
    L.n_cand = reinterpret_cast<int *>(A.misc.as<char>() + 128 + 24 * MP_MAX_GROUPS);
    L.cand = A.cand.as<int>(); L.cand_ok = A.cand_ok.as<unsigned char>();
    MP_CUDA(cudaMemsetAsync(L.n_ret, 0, G * sizeof(int), s));
    L.items = A.items.as<int>(); L.n_pairs = n_pairs;
    L.lo_req = A.lo_req.as<int>();
    L.gsz = gsz;
    MP_CUDA(A.evals32.reserve(n_pairs * sizeof(unsigned long long)));
    L.evals32 = A.evals32.as<unsigned long long>();
    L.trace = nullptr; L.trace_cap = c->trace_cap;
    if (c->trace_cap > 0) {
        MP_CUDA(A.trace.reserve((size_t)n_pairs * c->trace_cap * 2 * sizeof(double)));
        L.trace = A.trace.as<double>();
    }
    L.clk = reinterpret_cast<unsigned long long *>(A.misc.as<char>());
    MP_CUDA(cudaMemsetAsync(L.clk, 0, 128, s));
    const LoopParams *dprm = A.prm.as<LoopParams>();
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    init_state_kernel<<<(n_pairs + 63) / 64, 64, 0, s>>>(L, dprm);
    MP_CUDA(cudaGetLastError());
    int launches = 2;
    // dynamic tile + the kernel's static shared memory (candidate list of LP_BMAX * LP_MM shorts, barrier) <= opt-in limit
    int tile_cap = (c->smem_optin - (int)(LP_BMAX * LP_MM * sizeof(unsigned short)) - 2048) / 32;
    tile_cap = std::min(tile_cap, (max_n + 31) / 32 * 32);
    tile_cap = tile_cap / 32 * 32;
    const size_t smem = (size_t)tile_cap * 32;
    const unsigned max_it = std::max(opt->max_num_iterations, opt->min_num_iterations);
    const long long max_rounds = 64ll * max_it + 4096;
    long long rounds = 0;
    c->n_tick = 0;
    auto tick = [&](int cls) -> cudaError_t { // event before each kernel class; cls = -1 closes the last interval
        if (c->n_tick >= (int)c->tick_ev.size()) {
            cudaEvent_t ev;
            cudaError_t e_ = cudaEventCreate(&ev);
            if (e_ != cudaSuccess) return e_;
            c->tick_ev.push_back(ev); c->tick_cls.push_back(0);
        }
        c->tick_cls[c->n_tick] = cls;
        return cudaEventRecord(c->tick_ev[c->n_tick++], s);
    };
    MP_CUDA(cudaEventRecord(c->ev_init, s));
    struct GroupRt {
        int act = 0, iter = 1;        // latest known: pairs not done; pairs left iterating by the last scan
        int ret_seen = 0;             // pairs handed back to PH_ITER by lo_kernel launches, as last seen
        bool main_pending = false, lo_returned = true;
        bool slot_busy[SLOTS] = {false, false, false, false};
        long long round = 0;
    };
    std::vector<GroupRt> gr(G);
    for (int g = 0; g < G; g++) {
        gr[g].act = h_active[g];
        if (!serial) MP_CUDA(cudaStreamWaitEvent(c->gstream[g], c->ev_init, 0));
    }
    if (!serial)
        for (int k = 0; k < NSIDE; k++) MP_CUDA(cudaStreamWaitEvent(c->sstream[k], c->ev_init, 0));
    int side_rr = 0;
    // CTA widths of the local optimisation each group needs (bit 0: pairs below LP_WIDE_N matches, bit 1: the others)
    std::vector<int> lo_widths(G, 0);
    for (int p = 0; p < n_pairs; p++) lo_widths[p / gsz] |= (c->off[p + 1] - c->off[p]) >= LP_WIDE_N ? 2 : 1;
    const long long *d_off3 = A.off3.as<long long>();
    const int *d_caps = A.caps.as<int>();
    auto enqueue_round = [&](int g, int slot) -> int {
        cudaStream_t gs = serial ? s : c->gstream[g];
        const int first = g * gsz, count = std::min(gsz, n_pairs - first);
        const int round = (int)(++gr[g].round);
        MP_CUDA(cudaMemsetAsync(L.n_iter + g, 0, sizeof(int), gs));
        MP_CUDA(cudaMemsetAsync(L.n_items + 2 * g, 0, 2 * sizeof(int), gs));
        MP_CUDA(cudaMemsetAsync(L.n_items_w + g, 0, sizeof(int), gs));
        MP_CUDA(cudaMemsetAsync(L.n_cand + g, 0, sizeof(int), gs));
        if (serial) MP_CUDA(tick(0));
        sample_kernel<<<(count + SMP_WARPS - 1) / SMP_WARPS, SMP_WARPS * 32, 0, gs>>>(L, dprm, first, count);
        if (serial) MP_CUDA(tick(1));
        MP_CUDA(launch_minimal_batch(L, dprm, first, count, (long long)std::min(count, std::max(gr[g].act, 0)) * prm->bmax, kind, c->sm_count, warp_solvers, gs));
        if (serial) MP_CUDA(tick(2));
        if (kind == CALIB) MP_CUDA(launch_score_kind<CALIB>(L, dprm, first, count, tile_cap, smem, gs));
        else if (kind == SHARED) MP_CUDA(launch_score_kind<SHARED>(L, dprm, first, count, tile_cap, smem, gs));
        else MP_CUDA(launch_score_kind<TWO>(L, dprm, first, count, tile_cap, smem, gs));
        if (serial) MP_CUDA(tick(3));
        if (kind == CALIB) MP_CUDA(launch_scan_kind<CALIB>(L, dprm, first, count, d_off3, d_caps, round, gs));
        else if (kind == SHARED) MP_CUDA(launch_scan_kind<SHARED>(L, dprm, first, count, d_off3, d_caps, round, gs));
        else MP_CUDA(launch_scan_kind<TWO>(L, dprm, first, count, d_off3, d_caps, round, gs));
        MP_CUDA(cudaMemcpyAsync(h_iter + g, L.n_iter + g, sizeof(int), cudaMemcpyDeviceToHost, gs));
        MP_CUDA(cudaEventRecord(c->gev[g], gs));
        cudaStream_t ss = gs, hs = nullptr;
        int hk = -1;
        if (!serial) {
            const int k = side_rr++ % NSIDE;
            ss = c->sstream[k];
            MP_CUDA(cudaStreamWaitEvent(ss, c->gev[g], 0));
            if (heavy_from > 0) { hk = k; hs = c->hstream[k]; MP_CUDA(cudaStreamWaitEvent(hs, c->gev[g], 0)); }
        }
        if (kind == CALIB) MP_CUDA(launch_lo_kind<CALIB>(L, dprm, first, count, d_off3, d_caps, round, lo_widths[g], ss, hs, heavy_from));
        else if (kind == SHARED) MP_CUDA(launch_lo_kind<SHARED>(L, dprm, first, count, d_off3, d_caps, round, lo_widths[g], ss, hs, heavy_from));
        else MP_CUDA(launch_lo_kind<TWO>(L, dprm, first, count, d_off3, d_caps, round, lo_widths[g], ss, hs, heavy_from));
        if (serial) MP_CUDA(tick(-1));
        if (hs) { MP_CUDA(cudaEventRecord(c->hev[hk], hs)); MP_CUDA(cudaStreamWaitEvent(ss, c->hev[hk], 0)); } // the slot is done when both are
        MP_CUDA(cudaMemcpyAsync(h_side + 2 * (g * SLOTS + slot), L.n_active + g, sizeof(int), cudaMemcpyDeviceToHost, ss));
        MP_CUDA(cudaMemcpyAsync(h_side + 2 * (g * SLOTS + slot) + 1, L.n_ret + g, sizeof(int), cudaMemcpyDeviceToHost, ss));
        MP_CUDA(cudaEventRecord(c->sev[g * SLOTS + slot], ss));
        launches += 7;
        return MP_OK;
    };
    const bool trace = getenv("MP_TRACE") != nullptr; // host-side timeline of the rounds on stderr (debugging aid)
    const auto t_start = std::chrono::steady_clock::now();
    auto now_ms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count(); };
    while (true) {
        bool any = false, progress = false;
        for (int g = 0; g < G; g++) {
            GroupRt &R = gr[g];
            bool side_busy = false;
            for (int k = 0; k < SLOTS; k++) {
                if (!R.slot_busy[k]) continue;
                cudaError_t q = cudaEventQuery(c->sev[g * SLOTS + k]);
                if (q == cudaErrorNotReady) { side_busy = true; continue; }
                MP_CUDA(q);
                R.slot_busy[k] = false;
                R.act = std::min(R.act, h_side[2 * (g * SLOTS + k)]);
                if (h_side[2 * (g * SLOTS + k) + 1] > R.ret_seen) { R.ret_seen = h_side[2 * (g * SLOTS + k) + 1]; R.lo_returned = true; }
                if (trace) fprintf(stderr, "[mp] %8.2f ms  g%d  lo slot %d done   act %d\n", now_ms(), g, k, R.act);
                progress = true;
            }
            if (R.main_pending) {
                cudaError_t q = cudaEventQuery(c->gev[g]);
                if (q != cudaErrorNotReady) {
                    MP_CUDA(q);
                    R.main_pending = false;
                    R.iter = h_iter[g];
                    progress = true;
                    if (trace) fprintf(stderr, "[mp] %8.2f ms  g%d  round %lld done  iterating %d  act %d\n", now_ms(), g, R.round, R.iter, R.act);
                }
            }
            if (R.act <= 0 && !R.main_pending && !side_busy) continue; // group finished
            any = true;
            if (R.main_pending || R.act <= 0) continue;
            if (R.iter <= 0 && !R.lo_returned) continue; // nobody iterates; wait for a local optimisation to hand pairs back
            int slot = -1;
            for (int k = 0; k < SLOTS; k++)
                if (!R.slot_busy[k]) { slot = k; break; }
            if (slot < 0) continue;
            if (R.round > max_rounds) return fail(MP_ERR_CUDA, "device loop did not terminate");
            R.lo_returned = false;
            int rc2 = enqueue_round(g, slot);
            if (rc2) return rc2;
            R.main_pending = true; R.slot_busy[slot] = true;
            progress = true;
        }
        if (!any) break;
        if (!progress) std::this_thread::yield();
    }
    std::vector<long long> grounds(G);
    for (int g = 0; g < G; g++) grounds[g] = gr[g].round;
    for (int g = 0; g < G; g++) rounds = std::max(rounds, grounds[g]);
    // every group event has been observed complete by the host: the streams are drained
    finalize_kernel<<<(n_pairs + 63) / 64, 64, 0, s>>>(L, n_pairs, A.out_models.as<mp_model>(), A.out_stats.as<mp_stats>());
    MP_CUDA(cudaGetLastError());
    launches += 1;
    MP_CUDA(cudaEventRecord(c->ev[2], s));
    MP_CUDA(cudaMemcpyAsync(out_models, A.out_models.p, n_pairs * sizeof(mp_model), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(out_stats, A.out_stats.p, n_pairs * sizeof(mp_stats), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(c->phase_clk, L.clk, 128, cudaMemcpyDeviceToHost, s));
    c->last_groups = G;
    // The caller's inlier buffer is pageable (and, when freshly allocated, not yet backed by pages): a direct copy goes
    // through the driver's bounce buffers on one host thread.  Large lists land in a pinned buffer at link speed and are
    // moved -- and the pages touched -- by several host threads.
    const size_t inl_bytes = (out_inliers && N > 0) ? 3 * (size_t)N * sizeof(int) : 0;
    const bool staged = inl_bytes >= (size_t)(4u << 20);
    if (staged) {
        MP_CUDA(c->h_inl.reserve(inl_bytes));
        MP_CUDA(cudaMemcpyAsync(c->h_inl.p, A.inl_best.p, inl_bytes, cudaMemcpyDeviceToHost, s));
    } else if (inl_bytes)
        MP_CUDA(cudaMemcpyAsync(out_inliers, A.inl_best.p, inl_bytes, cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaEventRecord(c->ev[3], s));
    MP_CUDA(cudaStreamSynchronize(s));
    if (staged) {
        const int nt = (int)std::max<size_t>(1, std::min<size_t>(std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency())), inl_bytes >> 22));
        const size_t per = ((inl_bytes + nt - 1) / nt + 4095) & ~(size_t)4095;
        std::vector<std::thread> th;
        for (int t = 0; t < nt; t++) {
            const size_t o = (size_t)t * per;
            if (o >= inl_bytes) break;
            th.emplace_back([=] { std::memcpy(reinterpret_cast<char *>(out_inliers) + o, c->h_inl.as<char>() + o, std::min(per, inl_bytes - o)); });
        }
        for (auto &x : th) x.join();
    }
    float k = 0, t = 0;
    MP_CUDA(cudaEventElapsedTime(&k, c->ev[1], c->ev[2]));
    MP_CUDA(cudaEventElapsedTime(&t, c->ev[0], c->ev[3]));
    c->last_kernel_ms = k; c->last_total_ms = t; c->last_launches = launches;
    if (trace && c->phase_clk[8] != 0) { // sections of the LM driver: recorded by builds with -DMP_LM_PROF only
        const char *nm[8] = {"feed", "propose:setup", "cholesky", "forward", "backward", "model-change", "plus", "first"};
        for (int i = 0; i < 8; i++) fprintf(stderr, "[mp] LM driver %-14s %8.3f Gcycles\n", nm[i], c->phase_clk[8 + i] / 1e9);
    }
    c->last_rounds = (int)rounds;
    for (int i = 0; i < 4; i++) { c->class_ms[i] = 0; c->class_launches[i] = 0; }
    for (int i = 0; i + 1 < c->n_tick; i++) {
        const int cls = c->tick_cls[i];
        if (cls < 0) continue;
        float ms = 0;
        MP_CUDA(cudaEventElapsedTime(&ms, c->tick_ev[i], c->tick_ev[i + 1]));
        c->class_ms[cls] += ms; c->class_launches[cls] += 1;
    }
    return MP_OK;
}

// An error inside the round loop returns with kernels and copies still in flight on the group and side streams, and the
// next call would reuse their buffers: drain the device before handing the error to the caller.
static int run_estimate(mp_ctx *c, LoopAlloc &A, const mp_options *opt, const mp_config *cfg, mp_model *out_models,
                        mp_stats *out_stats, int32_t *out_inliers, bool record_start) {
    const int rc = run_estimate_impl(c, A, opt, cfg, out_models, out_stats, out_inliers, record_start);
    if (rc != MP_OK) {
        const std::string msg = g_err; // (keep the first error's text)
        cudaDeviceSynchronize();
        cudaGetLastError();
        g_err = msg;
    }
    return rc;
}

} // namespace mp

extern "C" {
int mp_lm_batch(mp_ctx *c, int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, int n_problems,
                const mp_model *start, const int32_t *idx, const int32_t *off, mp_model *out_models, double *out_info,
                int nonminimal) {
    using namespace mp;
    if (!c || !pair || !opt || !cfg || !start || !idx || !off || !out_models || !out_info || n_problems <= 0)
        return fail(MP_ERR_ARG, "mp_lm_batch: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    int rc = upload_pairs(c, kind, 1, pair, opt);
    if (rc) return rc;
    const int nidx = off[3 * n_problems];
    for (int i = 0; i < nidx; i++)
        if (idx[i] < 0 || idx[i] >= pair->n) return fail(MP_ERR_ARG, "mp_lm_batch: index out of range");
    if (!c->loop_alloc) c->loop_alloc = new LoopAlloc();
    LoopAlloc &A = *static_cast<LoopAlloc *>(c->loop_alloc);
    MP_CUDA(A.prm.reserve(sizeof(LoopParams)));
    MP_CUDA(A.h_prm.reserve(sizeof(LoopParams)));
    LoopParams *prm = A.h_prm.as<LoopParams>();
    prm->opt = *opt; prm->cfg = *cfg; prm->n_pairs = 1; prm->bmax = LP_BMAX; prm->b_after_lo = LP_BMAX / 4; prm->guard = 1e-3;
    prm->warp_from_it = 0xffffffffu;
    MP_CUDA(c->d_models.reserve(2 * (size_t)n_problems * sizeof(Model)));
    MP_CUDA(c->d_idx.reserve(((size_t)nidx + 3 * (size_t)n_problems + 8) * sizeof(int)));
    MP_CUDA(c->d_scores64.reserve(4 * (size_t)n_problems * sizeof(double)));
    Model *d_start = c->d_models.as<Model>(), *d_out = d_start + n_problems;
    int *d_idx = c->d_idx.as<int>(), *d_off = d_idx + nidx + 1;
    MP_CUDA(cudaMemcpyAsync(A.prm.p, prm, sizeof(LoopParams), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(d_start, start, n_problems * sizeof(Model), cudaMemcpyHostToDevice, s));
    if (nidx > 0) MP_CUDA(cudaMemcpyAsync(d_idx, idx, nidx * sizeof(int), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(d_off, off, (3 * n_problems + 1) * sizeof(int), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    const PairConst *pc = c->d_pc.as<PairConst>();
    const Match64 *M = c->d_m64.as<Match64>();
    const LoopParams *dprm = A.prm.as<LoopParams>();
    double *d_info = c->d_scores64.as<double>();
    const bool nm = nonminimal != 0;
    if (kind == CALIB || kind == MP_KIND_SCALE_ONLY) MP_CUDA(launch_lm_kind<CALIB>(pc, M, dprm, n_problems, d_start, d_idx, d_off, d_out, d_info, nm, s));
    else if (kind == SHARED) MP_CUDA(launch_lm_kind<SHARED>(pc, M, dprm, n_problems, d_start, d_idx, d_off, d_out, d_info, nm, s));
    else MP_CUDA(launch_lm_kind<TWO>(pc, M, dprm, n_problems, d_start, d_idx, d_off, d_out, d_info, nm, s));
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    MP_CUDA(cudaMemcpyAsync(out_models, d_out, n_problems * sizeof(Model), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(out_info, d_info, 4 * (size_t)n_problems * sizeof(double), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    float ms = 0;
    MP_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->last_kernel_ms = ms; c->last_total_ms = ms; c->last_launches = 2;
    return MP_OK;
}
int mp_upload_batch(mp_ctx *c, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt) {
    using namespace mp;
    if (!c || !pairs || !opt || n_pairs <= 0) return fail(MP_ERR_ARG, "mp_upload_batch: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    c->resident_pairs = 0;
    int rc = upload_pairs(c, kind, n_pairs, pairs, opt);
    if (rc) return rc;
    MP_CUDA(cudaStreamSynchronize(c->stream));
    c->resident_kind = kind; c->resident_pairs = n_pairs;
    return MP_OK;
}
int mp_estimate_resident(mp_ctx *c, const mp_options *opt, const mp_config *cfg, mp_model *out_models,
                         mp_stats *out_stats, int32_t *out_inliers) {
    using namespace mp;
    if (!c || !opt || !cfg || !out_models || !out_stats) return fail(MP_ERR_ARG, "mp_estimate_resident: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    if (!c->loop_alloc) c->loop_alloc = new LoopAlloc();
    return run_estimate(c, *static_cast<LoopAlloc *>(c->loop_alloc), opt, cfg, out_models, out_stats, out_inliers, true);
}
int mp_estimate_batch(mp_ctx *c, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt,
                      const mp_config *cfg, mp_model *out_models, mp_stats *out_stats, int32_t *out_inliers) {
    using namespace mp;
    if (!c || !pairs || !opt || !cfg || !out_models || !out_stats || n_pairs <= 0)
        return fail(MP_ERR_ARG, "mp_estimate_batch: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    if (!c->loop_alloc) c->loop_alloc = new LoopAlloc();
    c->resident_pairs = 0;
    MP_CUDA(cudaEventRecord(c->ev[0], c->stream));
    int rc = upload_pairs(c, kind, n_pairs, pairs, opt);
    if (rc) return rc;
    c->resident_kind = kind; c->resident_pairs = n_pairs;
    return run_estimate(c, *static_cast<LoopAlloc *>(c->loop_alloc), opt, cfg, out_models, out_stats, out_inliers, false);
}
int mp_set_trace(mp_ctx *c, int events_per_pair) {
    using namespace mp;
    if (!c || events_per_pair < 0 || events_per_pair > (1 << 22)) return fail(MP_ERR_ARG, "mp_set_trace: 0 (off) .. 4194304 events per pair");
    c->trace_cap = events_per_pair;
    return MP_OK;
}
int mp_read_trace(mp_ctx *c, int pair, double *out, int cap_events, int *n_events) {
    using namespace mp;
    if (!c || !out || !n_events || pair < 0 || cap_events < 0) return fail(MP_ERR_ARG, "mp_read_trace: bad arguments");
    LoopAlloc *A = static_cast<LoopAlloc *>(c->loop_alloc);
    if (!A || c->trace_cap <= 0 || !A->trace.p || pair >= c->resident_pairs) return fail(MP_ERR_ARG, "mp_read_trace: no trace recorded for this pair (mp_set_trace before the estimate call)");
    MP_CUDA(cudaSetDevice(c->device));
    PairState st;
    MP_CUDA(cudaMemcpy(&st, A->st.as<PairState>() + pair, sizeof(PairState), cudaMemcpyDeviceToHost));
    *n_events = st.trace_n; // may exceed the capacity: the records beyond it were dropped
    const int n = std::min(std::min(st.trace_n, c->trace_cap), cap_events);
    if (n > 0) MP_CUDA(cudaMemcpy(out, A->trace.as<double>() + 2 * (size_t)pair * c->trace_cap, 2 * (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
    return MP_OK;
}
int mp_set_minimal_mode(mp_ctx *c, int mode) {
    using namespace mp;
    if (!c || mode < 0 || mode > 2) return fail(MP_ERR_ARG, "mp_set_minimal_mode: 0 (automatic), 1 (thread per sample), 2 (warp per sample)");
    c->minimal_mode = mode;
    return MP_OK;
}
int mp_set_warp_from_iteration(mp_ctx *c, int iteration) {
    using namespace mp;
    if (!c || iteration < -1) return fail(MP_ERR_ARG, "mp_set_warp_from_iteration: -1 (default) or an iteration index >= 0");
    c->warp_from_it = iteration;
    return MP_OK;
}
int mp_set_batch(mp_ctx *c, int iterations) {
    using namespace mp;
    if (!c || iterations < 0 || iterations > LP_BMAX) return fail(MP_ERR_ARG, "mp_set_batch: 0 (default) .. 256");
    c->batch_override = iterations;
    return MP_OK;
}
int mp_set_groups(mp_ctx *c, int groups) {
    using namespace mp;
    if (!c || groups < 0 || groups > MP_MAX_GROUPS) return fail(MP_ERR_ARG, "mp_set_groups: 0 (automatic) .. 1024");
    c->groups_override = groups;
    return MP_OK;
}
int mp_last_phase_cycles(mp_ctx *c, unsigned long long *cycles8) {
    using namespace mp;
    if (!c || !cycles8) return fail(MP_ERR_ARG, "null argument");
    for (int i = 0; i < 8; i++) cycles8[i] = c->phase_clk[i];
    return MP_OK;
}
int mp_last_breakdown(mp_ctx *c, double *class_ms4, int *class_launches4, int *rounds) {
    using namespace mp;
    if (!c) return fail(MP_ERR_ARG, "null context");
    for (int i = 0; i < 4; i++) {
        if (class_ms4) class_ms4[i] = c->class_ms[i];
        if (class_launches4) class_launches4[i] = c->class_launches[i];
    }
    if (rounds) *rounds = c->last_rounds;
    return MP_OK;
}
}
