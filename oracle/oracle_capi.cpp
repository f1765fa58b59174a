// ORACLE (test infrastructure only).  C entry points so tests/ can drive the
// CPU restatement through ctypes.  Nothing here is linked into the product.
#include "om_affine.h"
#include "om_estimator.h"
#include "om_minimal.h"
#include "om_lm.h"
#include "om_ransac.h"
#include <thread>

extern "C" {

// kind: 0 calibrated (3 pts), 1 shared focal (4 pts), 2 two focal (4 pts)
// x1,x2: npts x 3 row-major homogeneous points.  out: up to 8 rows of
// (b1, a2, b2*a2, f1, f2)  -- same convention as solver.cpp:117-121 (1,b1,a2,b2*a2[,f..]).
int om_solve_affine(int kind, const double *x1, const double *x2, const double *d1, const double *d2, double *out) {
    std::vector<om::AffineSol> s;
    if (kind == 0) {
        double a[3][3], b[3][3];
        std::memcpy(a, x1, sizeof(a)); std::memcpy(b, x2, sizeof(b));
        s = om::solve_scale_and_shift(a, b, d1, d2);
    } else {
        double a[4][3], b[4][3];
        std::memcpy(a, x1, sizeof(a)); std::memcpy(b, x2, sizeof(b));
        s = kind == 1 ? om::solve_scale_and_shift_shared_focal(a, b, d1, d2)
                      : om::solve_scale_and_shift_two_focal(a, b, d1, d2);
    }
    for (size_t i = 0; i < s.size(); i++) {
        out[5 * i + 0] = s[i].b1; out[5 * i + 1] = s[i].a2; out[5 * i + 2] = s[i].b2 * s[i].a2;
        out[5 * i + 3] = s[i].f1; out[5 * i + 4] = s[i].f2;
    }
    return (int)s.size();
}

// ScoreModel for nm models of one pair (fp64, reference summation order)
int om_score(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, int nm, const mp_model *models,
             double *scores) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    for (int i = 0; i < nm; i++) scores[i] = om::score_model(P, *reinterpret_cast<const om::Model *>(models + i));
    return 0;
}
// per-match residuals r2[t*n + i] (inlier semantics, no gating) -- used to find near-threshold matches
int om_residuals(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, const mp_model *model,
                 double *r2, double *thr3) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    const om::Model &m = *reinterpret_cast<const om::Model *>(model);
    for (int t = 0; t < 3; t++)
        for (int i = 0; i < P.n; i++) r2[t * P.n + i] = om::evaluate(P, m, t, i, true);
    for (int t = 0; t < 3; t++) thr3[t] = P.thr[t];
    return 0;
}
// GetInliers; thr3 == NULL -> the problem's nominal thresholds
int om_inliers(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, const mp_model *model,
               const double *thr3, int *out_idx, int *out_counts) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    std::vector<std::vector<int>> inl;
    om::get_inliers(P, *reinterpret_cast<const om::Model *>(model), thr3 ? thr3 : P.thr, &inl);
    for (int t = 0; t < 3; t++) {
        out_counts[t] = (int)inl[t].size();
        for (size_t k = 0; k < inl[t].size(); k++) out_idx[t * P.n + k] = inl[t][k];
    }
    return 0;
}
// front-end constants: {norm_scale, thr0, thr1, thr2, w0, w1, w2, loss_scale}
int om_frontend(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, double *out8) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    out8[0] = P.norm_scale; out8[1] = P.thr[0]; out8[2] = P.thr[1]; out8[3] = P.thr[2];
    out8[4] = P.w[0]; out8[5] = P.w[1]; out8[6] = P.w[2]; out8[7] = P.loss_scale;
    return 0;
}

// MinimalSolver on explicit samples; out_models: n_samples x max_models
int om_minimal_solve(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, int solver_idx,
                     int n_samples, const int *idx, int max_models, mp_model *out_models, int *out_counts) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    int k = solver_idx == 0 ? P.sizes[0][0] : P.sizes[1][2];
    std::vector<om::Model> models;
    for (int s = 0; s < n_samples; s++) {
        om::minimal_solver(P, solver_idx, idx + (size_t)s * k, &models);
        int c = std::min((int)models.size(), max_models);
        out_counts[s] = c;
        for (int j = 0; j < c; j++) std::memcpy(out_models + (size_t)s * max_models + j, &models[j], sizeof(mp_model));
    }
    return 0;
}

// NonMinimalSolver / LeastSquares: problem j uses idx[off[3j+t] .. off[3j+t+1]) for type t
int om_lm(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, int n_problems,
          const mp_model *start, const int *idx, const int *off, mp_model *out_models, double *out_info, int nonminimal) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    for (int j = 0; j < n_problems; j++) {
        std::vector<int> s[3];
        for (int t = 0; t < 3; t++) s[t].assign(idx + off[3 * j + t], idx + off[3 * j + t + 1]);
        om::Model m = *reinterpret_cast<const om::Model *>(start + j);
        om::LmInfo info;
        om::lm_solve(P, s, &m, &info, nonminimal != 0);
        std::memcpy(out_models + j, &m, sizeof(mp_model));
        out_info[4 * j] = info.iterations; out_info[4 * j + 1] = info.termination;
        out_info[4 * j + 2] = info.cost0; out_info[4 * j + 3] = info.cost;
    }
    return 0;
}

// ---- direct access to the restated third-party solvers (cross-checked against OpenCV / constraint identities /
// scipy in tests/test_oracle_pins.py).  x0, x1: n x 3 row-major.
// 5-point: unit bearings -> essential matrices (row-major 9 each, up to 10) and the poses relpose_5pt keeps
int om_relpose_5pt(const double *x0, const double *x1, double *Es, int *nE, double *poses12, int *nP) {
    double a[5][3], b[5][3];
    std::memcpy(a, x0, sizeof(a)); std::memcpy(b, x1, sizeof(b));
    std::vector<std::vector<double>> E;
    om::relpose_5pt_essential(a, b, &E);
    *nE = (int)E.size();
    for (size_t i = 0; i < E.size(); i++) std::memcpy(Es + 9 * i, E[i].data(), 72);
    std::vector<om::Pose> P;
    om::relpose_5pt(a, b, &P);
    *nP = (int)P.size();
    for (size_t i = 0; i < P.size(); i++) { std::memcpy(poses12 + 12 * i, P[i].R, 72); std::memcpy(poses12 + 12 * i + 9, P[i].t, 24); }
    return 0;
}
// 7-point: homogeneous points -> fundamental matrices (up to 3)
int om_relpose_7pt(const double *x0, const double *x1, double *Fs, int *nF) {
    double a[7][3], b[7][3];
    std::memcpy(a, x0, sizeof(a)); std::memcpy(b, x1, sizeof(b));
    std::vector<std::vector<double>> F;
    om::relpose_7pt(a, b, &F);
    *nF = (int)F.size();
    for (size_t i = 0; i < F.size(); i++) std::memcpy(Fs + 9 * i, F[i].data(), 72);
    return 0;
}
// 6-point shared focal: (x, y, 1) points -> (R, t, f) candidates (13 doubles each, up to 15)
int om_relpose_6pt_focal(const double *x0, const double *x1, double *out13, int *n) {
    double a[6][3], b[6][3];
    std::memcpy(a, x0, sizeof(a)); std::memcpy(b, x1, sizeof(b));
    std::vector<om::FocalPose> fp;
    om::relpose_6pt_shared_focal(a, b, &fp);
    *n = (int)fp.size();
    for (size_t i = 0; i < fp.size(); i++) {
        std::memcpy(out13 + 13 * i, fp[i].pose.R, 72); std::memcpy(out13 + 13 * i + 9, fp[i].pose.t, 24); out13[13 * i + 12] = fp[i].focal;
    }
    return 0;
}
// cv::recoverPose restatement: p0, p1 n x 2
int om_recover_pose(const double *E, int n, const double *p0, const double *p1, double dist, double *R, double *t) {
    om::recover_pose(E, n, reinterpret_cast<const double (*)[2]>(p0), reinterpret_cast<const double (*)[2]>(p1), dist, R, t);
    return 0;
}
int om_bougnoux(const double *F, double *f0sq, double *f1sq) { om::bougnoux_focals(F, f0sq, f1sq); return 0; }
// residual vector of the LM problem at a model (the functors of cost_functions.h in the order of optimizer.h:54-77)
// out: 2 n0 + 2 n1 + n2 doubles; returns the count
int om_lm_residuals(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, const mp_model *model,
                    const int *idx, const int *off, double *out) {
    om::Problem P;
    om::make_problem(kind, *pair, *opt, *cfg, &P);
    const om::Model &m = *reinterpret_cast<const om::Model *>(model);
    om::LmData D; D.P = &P;
    const double sw = std::sqrt(P.sampson_w);
    D.W = P.kind == om::CALIB ? sw / (1.0 / (P.K0[0] + P.K0[4]) + 1.0 / (P.K1[0] + P.K1[4])) : sw;
    double x[om::NA];
    x[0] = m.o0; om::rot_to_quat(m.R, x + 1); std::memcpy(x + 5, m.t, 24); x[8] = m.scale; x[9] = m.o1; x[10] = m.f0; x[11] = m.f1;
    double f0 = x[10], f1 = P.kind == om::TWO ? x[11] : x[10];
    int k = 0;
    double r[2];
    if (P.cfg.lo_type != 1) {
        for (int j = off[0]; j < off[1]; j++) { om::res_reproj0<double>(D, idx[j], x, f0, f1, r); out[k++] = r[0]; out[k++] = r[1]; }
        for (int j = off[1]; j < off[2]; j++) { om::res_reproj1<double>(D, idx[j], x, f0, f1, r); out[k++] = r[0]; out[k++] = r[1]; }
    }
    if (P.cfg.lo_type != 2) for (int j = off[2]; j < off[3]; j++) { om::res_sampson<double>(D, idx[j], x, f0, f1, r); out[k++] = r[0]; }
    return k;
}

// utils::NumRequiredIterations restatement (KAT-1b of SURVEY.md Appendix C)
unsigned om_num_required_iterations(const double *ratios3, double prob_missing, const int *sizes3, unsigned mn, unsigned mx) {
    return om::Ransac::num_required_iterations(ratios3, prob_missing, sizes3, mn, mx);
}
// utils::RandomShuffleAndResize restatement on std::mt19937(seed): v has room for max(n, target) ints
int om_shuffle_resize(unsigned seed, int target, int n, int *v, int second_target, int second_n, int *v2) {
    std::mt19937 g(seed);
    std::vector<int> a(v, v + n);
    om::Ransac::shuffle_resize(target, &g, &a);
    for (int i = 0; i < target; i++) v[i] = a[i];
    std::vector<int> b(v2, v2 + second_n);
    om::Ransac::shuffle_resize(second_target, &g, &b);
    for (int i = 0; i < second_target; i++) v2[i] = b[i];
    return 0;
}

static void fill_stats(const om::Stats &st, int n, mp_stats *o, int *inl) {
    std::memset(o, 0, sizeof(*o));
    o->num_iterations_total = st.num_iterations_total;
    o->num_iterations_per_solver[0] = st.num_iterations_per_solver[0];
    o->num_iterations_per_solver[1] = st.num_iterations_per_solver[1];
    o->best_num_inliers = st.best_num_inliers; o->best_solver_type = st.best_solver_type;
    o->number_lo_iterations = st.number_lo_iterations; o->best_model_score = st.best_model_score;
    for (int t = 0; t < 3; t++) {
        o->inlier_ratios[t] = st.inlier_ratios[t];
        o->num_inliers[t] = t < (int)st.inlier_indices.size() ? (int)st.inlier_indices[t].size() : 0;
        if (inl) for (int k = 0; k < o->num_inliers[t]; k++) inl[t * n + k] = st.inlier_indices[t][k];
    }
    o->hypotheses_scored = st.hypotheses_scored; o->residual_evals = st.residual_evals;
    o->lm_solves = st.lm_solves; o->lm_iterations = st.lm_iterations;
}

// whole path for a batch of pairs on `threads` host threads (pair p -> thread p % threads).
// out_inliers layout = mp_estimate_batch.  trace (optional, pair 0 only): up to trace_cap doubles, -1 terminated.
int om_estimate_batch(int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt, const mp_config *cfg,
                      mp_model *out_models, mp_stats *out_stats, int *out_inliers, int threads, double *trace,
                      int trace_cap) {
    std::vector<long long> off(n_pairs + 1, 0);
    for (int p = 0; p < n_pairs; p++) off[p + 1] = off[p] + 3ll * pairs[p].n;
    auto work = [&](int tid) {
        for (int p = tid; p < n_pairs; p += threads) {
            om::Model m; om::Stats st; om::Trace tr;
            om::estimate(kind, pairs[p], *opt, *cfg, &m, &st, (p == 0 && trace) ? &tr : nullptr);
            std::memcpy(out_models + p, &m, sizeof(mp_model));
            fill_stats(st, pairs[p].n, out_stats + p, out_inliers ? out_inliers + off[p] : nullptr);
            if (p == 0 && trace) for (int i = 0; i < trace_cap; i++) trace[i] = i < (int)tr.size() ? tr[i] : -1.0;
        }
    };
    if (threads <= 1) { threads = 1; work(0); return 0; }
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, t);
    for (auto &t : th) t.join();
    return 0;
}

} // extern "C"
