// ORACLE (test infrastructure only).  The hybrid LO-MSAC control loop, restated from
//   /root/reference/src/hybrid_ransac.h:38-206 (EstimateModel), :210-243 (SelectMinimalSolver),
//   :245-287 (scoring), :351-378 (termination update), :383-482 (LocalOptimization),
//   :484-538 (LeastSquaresFit), :540-577 (UpdateBestModel / VerifyData)
//   /root/reference/ext/RansacLib/RansacLib/hybrid_sampling.h:46-140 (HybridUniformSampling)
//   /root/reference/ext/RansacLib/RansacLib/utils.h:48-73,142-178 (shuffle, NumRequiredIterations)
// including the quirks Q1-Q14 of SURVEY.md A.7.  std::mt19937 and the libstdc++
// distributions are used directly, exactly as the reference does on this platform.
// oracle/_ref/ (ref_glue.cpp) compiles the reference's OWN loop headers around the same
// solver functions; tests/test_oracle_loop.py checks that both give identical results.
#pragma once
#include "om_lm.h"
#include "om_minimal.h"
#include <random>

namespace om {

struct Stats {
    uint32_t num_iterations_total = 0, num_iterations_per_solver[2] = {0, 0};
    int best_num_inliers = 0, best_solver_type = -1, number_lo_iterations = 0;
    double best_model_score = kMax;
    double inlier_ratios[3] = {0, 0, 0};
    std::vector<std::vector<int>> inlier_indices;
    // extra counters (work accounting for the CPU baseline)
    uint64_t hypotheses_scored = 0, residual_evals = 0;
    uint32_t lm_solves = 0, lm_iterations = 0;
};

// chronological records of 4 doubles, for parity forensics (tests/test_estimate_gpu.py::explain_fork):
//   {it, solver, n_models, best_local_score}   one per main-loop iteration
//   {it, -1, 0, score}                         every ScoreModel inside LocalOptimization / the final least squares
//   {it, -2, inlier count, min |r^2 - thr| / thr}   every GetInliers (LeastSquaresFit, LO base, termination update)
//   {it, -3, LM iterations, final cost}        every NonMinimalSolver / LeastSquares that ran
typedef std::vector<double> Trace;

struct Ransac {
    const Problem &P;
    Stats &st;
    Trace *trace;
    std::mt19937 rng, srng; // stream B (loop) and stream A (sampler)
    uint32_t cap[2];
    double prior[2];

    Ransac(const Problem &p, Stats &s, Trace *t) : P(p), st(s), trace(t) {}

    double score(const Model &m, bool lo_event = false) {
        st.hypotheses_scored++; st.residual_evals += P.n;
        const double sc = score_model(P, m);
        if (trace && lo_event) { trace->push_back(st.num_iterations_total); trace->push_back(-1); trace->push_back(0); trace->push_back(sc); }
        return sc;
    }
    void trace4(double kind, double a, double b) {
        if (trace) { trace->push_back(st.num_iterations_total); trace->push_back(kind); trace->push_back(a); trace->push_back(b); }
    }
    int inliers_traced(const Model &m, const double *thr, std::vector<std::vector<int>> *inl) {
        double mr = 0;
        const int c = get_inliers(P, m, thr, inl, trace ? &mr : nullptr);
        // the three per-type counts packed exactly into one double (17 bits each)
        trace4(-2, (double)(*inl)[0].size() + 131072.0 * (double)(*inl)[1].size() + 17179869184.0 * (double)(*inl)[2].size(), mr);
        return c;
    }
    void lsq(const std::vector<std::vector<int>> &sample, Model *m) {
        LmInfo info;
        if (lm_solve(P, sample.data(), m, &info)) { st.lm_solves++; st.lm_iterations += info.iterations; trace4(-3, info.iterations, info.cost); }
    }
    static void update_best(double sc, const Model &m, int solver, double *best_sc, Model *best, int *best_solver) {
        if (sc < *best_sc) { *best_sc = sc; *best = m; *best_solver = solver; }
    }
    // RansacLib utils.h:142-178 (vector form)
    static uint32_t num_required_iterations(const double *ratios, double prob_missing, const int *sizes, uint32_t mn, uint32_t mx) {
        double p = 1.0;
        for (int i = 0; i < 3; i++) p *= std::pow(ratios[i], (double)sizes[i]);
        if (p <= 0.0) return mx;
        if (p >= 1.0) return mn;
        double q = 1.0 - p;
        if (q >= 0.99999999999999) return mx;
        double it = std::ceil(std::log(prob_missing) / std::log(q) + 0.5);
        // static_cast<uint32_t>(double) in the reference (RansacLib utils.h:176) is out of range for it >= 2^32;
        // x86-64 gcc truncates to 64 bits and keeps the low word -- made explicit here
        uint32_t r = std::min((uint32_t)(uint64_t)(int64_t)it, mx);
        return std::max(mn, r);
    }
    void update_termination(const Model &m) {
        st.best_num_inliers = inliers_traced(m, P.thr, &st.inlier_indices);
        for (int t = 0; t < 3; t++) st.inlier_ratios[t] = P.n > 0 ? (double)st.inlier_indices[t].size() / P.n : 0.0;
        for (int s = 0; s < 2; s++)
            cap[s] = num_required_iterations(st.inlier_ratios, 1.0 - P.opt.success_probability, P.sizes[s],
                                             P.opt.min_num_iterations, P.opt.max_num_iterations_per_solver);
    }
    static void shuffle_resize(int target, std::mt19937 *g, std::vector<int> *v) {
        const int n = (int)v->size();
        for (int i = 0; i < n - 1; ++i) {
            std::uniform_int_distribution<int> dist(i, n - 1);
            std::swap((*v)[i], (*v)[dist(*g)]);
        }
        v->resize(target); // grows with zeros when target > n (quirk Q3)
    }
    void unflatten(std::vector<int> flat, std::vector<std::vector<int>> *sample) const {
        sample->assign(3, std::vector<int>());
        for (int idx : flat) {
            int t = 0;
            while (idx >= P.n) { idx -= P.n; t++; }
            (*sample)[t].push_back(idx);
        }
    }
    void least_squares_fit(const double *thr, int solver, Model *m, bool all_inliers) {
        std::vector<std::vector<int>> inl;
        inliers_traced(*m, thr, &inl);
        int sz[3];
        for (int t = 0; t < 3; t++) {
            if ((int)inl[t].size() < P.sizes[solver][t]) return;
            sz[t] = std::min(P.sizes[solver][t] * P.opt.min_sample_multiplicator, (int)inl[t].size());
        }
        if (all_inliers) { lsq(inl, m); return; }
        int total = (sz[0] + sz[1] + sz[2]) * P.opt.min_sample_multiplicator; // x7 twice (quirk Q4)
        std::vector<int> flat;
        for (int t = 0; t < 3; t++) for (int i : inl[t]) flat.push_back(i + t * P.n);
        shuffle_resize(total, &rng, &flat);
        std::vector<std::vector<int>> sample;
        unflatten(flat, &sample);
        lsq(sample, m);
    }
    void local_optimization(int solver, Model *best, double *best_score, int *best_solver) {
        double thr[3], step[3];
        for (int t = 0; t < 3; t++) {
            step[t] = (P.opt.threshold_multiplier - 1.0) * P.thr[t] / (int)(P.opt.num_lsq_iterations - 1);
            thr[t] = P.thr[t] * P.opt.threshold_multiplier;
        }
        Model m_init = *best;
        least_squares_fit(thr, solver, &m_init, true);
        double sc = score(m_init, true);
        update_best(sc, m_init, solver, best_score, best, best_solver);
        std::vector<std::vector<int>> base;
        inliers_traced(m_init, P.thr, &base);
        std::vector<int> base_flat;
        for (int t = 0; t < 3; t++) for (int i : base[t]) base_flat.push_back(i + t * P.n);
        const int nonmin = std::max(P.non_min_sample_size,
                                    std::min(P.min_sample_size * P.opt.non_min_sample_multiplier, (int)base_flat.size() / 2));
        for (int r = 0; r < P.opt.num_lo_steps; ++r) {
            std::vector<int> cur = base_flat;               // copy BEFORE the shuffle (quirk Q2)
            shuffle_resize(nonmin, &rng, &base_flat);
            std::vector<std::vector<int>> sample;
            unflatten(cur, &sample);
            Model m = m_init;
            LmInfo info;
            if (!lm_solve(P, sample.data(), &m, &info, true)) continue;
            st.lm_solves++; st.lm_iterations += info.iterations; trace4(-3, info.iterations, info.cost);
            sc = score(m, true);
            update_best(sc, m, solver, best_score, best, best_solver);
            least_squares_fit(P.thr, solver, &m, false);
            double cur_thr[3] = {thr[0], thr[1], thr[2]};
            for (int i = 0; i < P.opt.num_lsq_iterations; ++i) {
                least_squares_fit(cur_thr, solver, &m, false);
                sc = score(m, true);
                update_best(sc, m, solver, best_score, best, best_solver);
                for (int t = 0; t < 3; t++) cur_thr[t] -= step[t];
            }
        }
    }
    // HybridUniformSampling::Sample
    void draw(const int *sizes, std::vector<std::vector<int>> *sample) {
        sample->resize(3);
        for (int t = 0; t < 3; t++) {
            std::vector<int> &s = (*sample)[t];
            s.clear();
            const int k = sizes[t];
            if (k <= 0) continue;
            const double coeff = (double)P.n / (double)(P.n - k);
            if (coeff < M_E) {
                s.resize(k);
                std::uniform_int_distribution<int> dist(0, P.n - 1);
                for (int i = 0; i < k; ++i) {
                    bool dup = true;
                    while (dup) {
                        dup = false;
                        s[i] = dist(srng);
                        for (int j = 0; j < i; ++j) if (s[j] == s[i]) { dup = true; break; }
                    }
                }
            } else {
                s.resize(P.n);
                for (int i = 0; i < P.n; i++) s[i] = i;
                if (k == P.n) continue;
                for (int i = 0; i < P.n - 1; ++i) {
                    std::uniform_int_distribution<int> dist(i, P.n - 1);
                    std::swap(s[i], s[dist(srng)]);
                }
                s.resize(k);
            }
        }
    }
    int select_solver() {
        std::uniform_real_distribution<double> dist(0.0, prior[0] + prior[1]);
        const double u = dist(rng);
        double c = 0;
        for (int s = 0; s < 2; s++) {
            if (prior[s] == 0.0) continue;
            c += prior[s];
            if (u <= c) return s;
        }
        return -1;
    }
    int run(Model *best_model) {
        st = Stats();
        st.inlier_indices.assign(3, std::vector<int>());
        prior[0] = P.cfg.solver_type == 1 ? 0.0 : 1.0; // EPI_ONLY disables the depth solver
        prior[1] = P.cfg.solver_type == 2 ? 0.0 : 1.0; // MD_ONLY disables the point solver
        for (int s = 0; s < 2; s++)
            for (int t = 0; t < 3; t++) if (P.sizes[s][t] > P.n) { prior[s] = 0.0; break; }
        if (!(prior[0] > 0.0) && !(prior[1] > 0.0)) return 0;
        srng.seed(P.opt.random_seed); rng.seed(P.opt.random_seed);
        const uint32_t max_it = std::max(P.opt.max_num_iterations, P.opt.min_num_iterations);
        cap[0] = cap[1] = std::max(P.opt.max_num_iterations_per_solver, P.opt.min_num_iterations);
        Model best_min;
        double best_min_score = kMax;
        std::vector<std::vector<int>> sample;
        std::vector<Model> models;
        const uint32_t lo_start = P.opt.lo_starting_iterations;
        for (st.num_iterations_total = 0; st.num_iterations_total < max_it; ++st.num_iterations_total) {
            const uint32_t it = st.num_iterations_total;
            if (it == lo_start && best_min_score < kMax) {
                ++st.number_lo_iterations;
                local_optimization(st.best_solver_type, best_model, &st.best_model_score, &st.best_solver_type);
                update_termination(*best_model);
            }
            const int s = select_solver();
            st.num_iterations_per_solver[s] += 1;
            draw(P.sizes[s], &sample);
            minimal_solver(P, s, s == 0 ? sample[0].data() : sample[2].data(), &models);
            double loc = kMax;
            int loc_id = 0;
            for (size_t k = 0; k < models.size(); k++) {
                double sc = score(models[k]);
                if (sc < loc) { loc = sc; loc_id = (int)k; }
            }
            if (trace) { trace->push_back(it); trace->push_back(s); trace->push_back((double)models.size()); trace->push_back(loc); }
            if (!models.empty() && (loc < best_min_score || it == lo_start)) {
                const bool newbest = loc < best_min_score;
                if (newbest) {
                    best_min_score = loc; best_min = models[loc_id];
                    update_best(best_min_score, best_min, s, &st.best_model_score, best_model, &st.best_solver_type);
                }
                const bool run_lo = it >= lo_start && best_min_score < kMax;
                if (!newbest && !run_lo) continue; // skips the break test (quirk Q13)
                if (run_lo) {
                    ++st.number_lo_iterations;
                    double sc = best_min_score;
                    local_optimization(st.best_solver_type, &best_min, &sc, &st.best_solver_type);
                    update_best(sc, best_min, s, &st.best_model_score, best_model, &st.best_solver_type);
                }
                update_termination(*best_model);
            }
            if (st.num_iterations_per_solver[s] >= cap[s]) break;
        }
        if (st.num_iterations_total <= lo_start && st.best_model_score < kMax) {
            ++st.number_lo_iterations;
            local_optimization(st.best_solver_type, best_model, &st.best_model_score, &st.best_solver_type);
            update_termination(*best_model);
        }
        if (P.opt.final_least_squares) {
            Model refined = *best_model;
            lsq(st.inlier_indices, &refined);
            double sc = score(refined, true);
            if (sc < st.best_model_score) { st.best_model_score = sc; *best_model = refined; update_termination(*best_model); }
        }
        return st.best_num_inliers;
    }
};

// the three front-ends: run the loop, undo the focal normalisation
inline int estimate(int kind, const mp_pair &pair, const mp_options &opt, const mp_config &cfg, Model *out, Stats *st, Trace *trace = nullptr) {
    Problem P;
    make_problem(kind, pair, opt, cfg, &P);
    Ransac R(P, *st, trace);
    Model m; // default model: zero pose, scale 1, offsets 0, focal 1 (pose.h:11,25,38,49)
    int r = R.run(&m);
    if (kind == SHARED) { m.f0 *= P.norm_scale; m.f1 = m.f0; }
    if (kind == TWO) { m.f0 *= P.norm_scale; m.f1 *= P.norm_scale; }
    *out = m;
    return r;
}

} // namespace om
