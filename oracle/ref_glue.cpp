// ORACLE (test infrastructure only).  Builds oracle/_ref/liboracle_ref.so: the
// REFERENCE'S OWN control loop -- madpose::HybridLOMSAC from
// /root/reference/src/hybrid_ransac.h with RansacLib's sampler / utils
// (/root/reference/ext/RansacLib/RansacLib/{hybrid_ransac,hybrid_sampling,utils}.h),
// compiled where those headers lie (-I, nothing copied) -- instantiated around an
// adapter that exposes the oracle's restated solver math through the "HybridSolver"
// concept of /root/reference/src/hybrid_pose_estimator.h:41-85.  The rest of the
// reference (Eigen / Ceres / PoseLib / OpenCV code) cannot be built here.
// Used to validate oracle/om_ransac.h (same results, iteration for iteration) and as the
// "reference" CPU baseline of bench.py when present.
#include "om_lm.h"
#include "om_minimal.h"
#include "om_ransac.h"

#include <hybrid_ransac.h> // /root/reference/src

#include <thread>

namespace {

struct Adapter {
    const om::Problem &P;
    mutable om::Stats *st;
    explicit Adapter(const om::Problem &p, om::Stats *s) : P(p), st(s) {}
    int num_minimal_solvers() const { return 2; }
    int min_sample_size() const { return P.min_sample_size; }
    int non_minimal_sample_size() const { return P.non_min_sample_size; }
    int num_data_types() const { return 3; }
    void min_sample_sizes(std::vector<std::vector<int>> *s) const {
        s->resize(2);
        for (int k = 0; k < 2; k++) (*s)[k] = {P.sizes[k][0], P.sizes[k][1], P.sizes[k][2]};
    }
    void num_data(std::vector<int> *n) const { n->assign(3, P.n); }
    void solver_probabilities(std::vector<double> *p) const {
        p->assign(2, 1.0);
        if (P.cfg.solver_type == 1) (*p)[0] = 0.0; else if (P.cfg.solver_type == 2) (*p)[1] = 0.0;
    }
    int MinimalSolver(const std::vector<std::vector<int>> &sample, const int solver_idx, std::vector<om::Model> *models) const {
        om::minimal_solver(P, solver_idx, solver_idx == 0 ? sample[0].data() : sample[2].data(), models);
        return (int)models->size();
    }
    int NonMinimalSolver(const std::vector<std::vector<int>> &sample, const int, om::Model *m) const {
        om::LmInfo info;
        if (!om::lm_solve(P, sample.data(), m, &info, true)) return 0;
        st->lm_solves++; st->lm_iterations += info.iterations;
        return 1;
    }
    double EvaluateModelOnPoint(const om::Model &m, int t, int i, bool is_for_inlier = false) const {
        if (t == 0 && i == 0 && !is_for_inlier) { st->hypotheses_scored++; st->residual_evals += P.n; }
        return om::evaluate(P, m, t, i, is_for_inlier);
    }
    void LeastSquares(const std::vector<std::vector<int>> &sample, const int, om::Model *m) const {
        om::LmInfo info;
        if (om::lm_solve(P, sample.data(), m, &info)) { st->lm_solves++; st->lm_iterations += info.iterations; }
    }
};

} // namespace

extern "C" int omref_estimate_batch(int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt, const mp_config *cfg,
                                    mp_model *out_models, mp_stats *out_stats, int *out_inliers, int threads, double *, int) {
    std::vector<long long> off(n_pairs + 1, 0);
    for (int p = 0; p < n_pairs; p++) off[p + 1] = off[p] + 3ll * pairs[p].n;
    auto work = [&](int tid) {
        for (int p = tid; p < n_pairs; p += threads) {
            om::Problem P;
            om::make_problem(kind, pairs[p], *opt, *cfg, &P);
            om::Stats extra;
            Adapter solver(P, &extra);
            madpose::ExtendedHybridLORansacOptions ro;
            ro.min_num_iterations_ = opt->min_num_iterations; ro.max_num_iterations_ = opt->max_num_iterations;
            ro.max_num_iterations_per_solver_ = opt->max_num_iterations_per_solver;
            ro.success_probability_ = opt->success_probability; ro.random_seed_ = opt->random_seed;
            ro.squared_inlier_thresholds_ = {P.thr[0], P.thr[1], P.thr[2]}; // after the front-end rewrite
            ro.data_type_weights_ = {P.w[0], P.w[1], P.w[2]};
            ro.num_lo_steps_ = opt->num_lo_steps; ro.threshold_multiplier_ = opt->threshold_multiplier;
            ro.num_lsq_iterations_ = opt->num_lsq_iterations; ro.min_sample_multiplicator_ = opt->min_sample_multiplicator;
            ro.non_min_sample_multiplier_ = opt->non_min_sample_multiplier;
            ro.lo_starting_iterations_ = opt->lo_starting_iterations; ro.final_least_squares_ = opt->final_least_squares != 0;
            om::Model best;
            ransac_lib::HybridRansacStatistics rs;
            madpose::HybridLOMSAC<om::Model, std::vector<om::Model>, Adapter> lomsac;
            lomsac.EstimateModel(ro, solver, &best, &rs);
            if (kind == om::SHARED) { best.f0 *= P.norm_scale; best.f1 = best.f0; }
            if (kind == om::TWO) { best.f0 *= P.norm_scale; best.f1 *= P.norm_scale; }
            std::memcpy(out_models + p, &best, sizeof(mp_model));
            mp_stats *o = out_stats + p;
            std::memset(o, 0, sizeof(*o));
            o->num_iterations_total = rs.num_iterations_total;
            for (int s = 0; s < 2 && s < (int)rs.num_iterations_per_solver.size(); s++) o->num_iterations_per_solver[s] = rs.num_iterations_per_solver[s];
            o->best_num_inliers = rs.best_num_inliers; o->best_solver_type = rs.best_solver_type;
            o->number_lo_iterations = rs.number_lo_iterations; o->best_model_score = rs.best_model_score;
            for (int t = 0; t < 3 && t < (int)rs.inlier_ratios.size(); t++) {
                o->inlier_ratios[t] = rs.inlier_ratios[t];
                o->num_inliers[t] = (int)rs.inlier_indices[t].size();
                if (out_inliers) for (int k = 0; k < o->num_inliers[t]; k++) out_inliers[off[p] + (long long)t * pairs[p].n + k] = rs.inlier_indices[t][k];
            }
            o->hypotheses_scored = extra.hypotheses_scored; o->residual_evals = extra.residual_evals;
            o->lm_solves = extra.lm_solves; o->lm_iterations = extra.lm_iterations;
        }
    };
    if (threads <= 1) { threads = 1; work(0); return 0; }
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, t);
    for (auto &t : th) t.join();
    return 0;
}
