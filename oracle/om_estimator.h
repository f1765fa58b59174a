// ORACLE (test infrastructure only - never linked into the product path).
// Front-ends, residuals, scoring and inlier extraction of the three hybrid
// estimators, restated without Eigen from
//   /root/reference/src/hybrid_pose_estimator.cpp:8-35,216-261 (+ .h:14-37)
//   /root/reference/src/hybrid_pose_shared_focal_estimator.cpp:8-51,160-202
//   /root/reference/src/hybrid_pose_two_focal_estimator.cpp:34-75,213-257
//   /root/reference/src/utils.h:56-83, /root/reference/src/solver.cpp:1188-1206
//   /root/reference/src/hybrid_ransac.h:265-349 (ScoreModel / GetInliers).
#pragma once
#include "../include/madpose_b200.h"
#include "om_linalg.h"
#include <cfloat>
#include <vector>

namespace om {

enum { CALIB = 0, SHARED = 1, TWO = 2 };
static const double kMax = DBL_MAX;

struct Model {
    double R[9], t[3], scale, o0, o1, f0, f1;
    Model() : scale(1), o0(0), o1(0), f0(1), f1(1) {
        for (double &v : R) v = 0;
        for (double &v : t) v = 0;
    }
};
static_assert(sizeof(Model) == sizeof(mp_model), "model layout");

struct Problem {
    int kind = 0, n = 0;
    bool scale_only = false;     // HybridEstimatePoseAndScale (hybrid_pose_estimator.cpp:37-63,297-442): calibrated, no offsets
    std::vector<double> x0, x1; // 2n, pixels (calibrated) or pp-subtracted / norm_scale (focal kinds)
    std::vector<double> d0, d1;
    double K0[9], K1[9], K0i[9], K1i[9];
    double min_depth[2];
    double thr[3], w[3];         // after the "three data types" rewrite
    double sampson_w = 1;        // sampson_squared_weight
    double loss_scale = 1;       // sampson_loss_scale_ (calibrated)
    double norm_scale = 1;       // focal kinds
    mp_options opt;
    mp_config cfg;
    int sizes[2][3];             // min_sample_sizes
    int non_min_sample_size = 35, min_sample_size = 5;
};

// poselib::normalize_points(x1, x2, T1, T2, normalize_scale=true, normalize_centroid=false,
// shared_scale=true) -- PoseLib v2.0.4 is NOT under /root/reference (fetched by
// CMakeLists.txt:26-30).  Published behaviour restated from memory (SURVEY.md A.1,
// UNVERIFIED): no centring, scale = sum_k(|x1_k| + |x2_k|) / (sqrt(2) * N), shared.
inline double normalize_points_scale(const std::vector<double> &a, const std::vector<double> &b, int n) {
    double s = 0;
    for (int i = 0; i < n; i++)
        s += std::sqrt(a[2 * i] * a[2 * i] + a[2 * i + 1] * a[2 * i + 1]) +
             std::sqrt(b[2 * i] * b[2 * i] + b[2 * i + 1] * b[2 * i + 1]);
    return s / (std::sqrt(2.0) * n);
}

inline void make_problem(int kind, const mp_pair &pr, const mp_options &opt, const mp_config &cfg, Problem *P) {
    P->scale_only = kind == 3; // MP_KIND_SCALE_ONLY: the calibrated estimator's scale-only twin
    if (kind == 3) kind = CALIB;
    P->kind = kind; P->n = pr.n; P->opt = opt; P->cfg = cfg;
    int n = pr.n;
    P->x0.assign(pr.x0, pr.x0 + 2 * n); P->x1.assign(pr.x1, pr.x1 + 2 * n);
    P->d0.assign(pr.d0, pr.d0 + n); P->d1.assign(pr.d1, pr.d1 + n);
    P->min_depth[0] = pr.min_depth[0]; P->min_depth[1] = pr.min_depth[1];
    double tr = opt.squared_inlier_thresholds[0], te = opt.squared_inlier_thresholds[1];
    double wr = opt.data_type_weights[0], we = opt.data_type_weights[1];
    double eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (kind == CALIB) {
        std::memcpy(P->K0, pr.K0, sizeof(eye)); std::memcpy(P->K1, pr.K1, sizeof(eye));
        mat3_inv(P->K0, P->K0i); mat3_inv(P->K1, P->K1i);
        double l = 1.0 / (P->K0[0] + P->K0[4]) + 1.0 / (P->K1[0] + P->K1[4]);
        P->loss_scale = 1.0 / (l * l); // hybrid_pose_estimator.h:35-36
        P->sizes[0][0] = 3; P->sizes[0][1] = 3; P->sizes[0][2] = 0;
        P->sizes[1][0] = 0; P->sizes[1][1] = 0; P->sizes[1][2] = 5;
        P->non_min_sample_size = 35; P->min_sample_size = 5;
    } else {
        for (int i = 0; i < n; i++) {
            P->x0[2 * i] -= pr.pp0[0]; P->x0[2 * i + 1] -= pr.pp0[1];
            P->x1[2 * i] -= pr.pp1[0]; P->x1[2 * i + 1] -= pr.pp1[1];
        }
        double ns = normalize_points_scale(P->x0, P->x1, n);
        P->norm_scale = ns;
        for (double &v : P->x0) v /= ns;
        for (double &v : P->x1) v /= ns;
        tr /= ns * ns; te /= ns * ns; // ...shared_focal_estimator.cpp:24-25
        std::memcpy(P->K0, eye, sizeof(eye)); std::memcpy(P->K1, eye, sizeof(eye));
        std::memcpy(P->K0i, eye, sizeof(eye)); std::memcpy(P->K1i, eye, sizeof(eye));
        int m = kind == SHARED ? 6 : 7;
        P->sizes[0][0] = 4; P->sizes[0][1] = 4; P->sizes[0][2] = 0;
        P->sizes[1][0] = 0; P->sizes[1][1] = 0; P->sizes[1][2] = m;
        P->non_min_sample_size = 36; P->min_sample_size = m;
    }
    // "three data types" rewrite, hybrid_pose_estimator.cpp:16-23
    double ws = we * (2 * tr / te);
    P->sampson_w = ws;
    P->w[0] = wr; P->w[1] = wr; P->w[2] = ws;
    P->thr[0] = tr; P->thr[1] = tr; P->thr[2] = te;
}

inline void model_K(const Problem &P, const Model &m, double *K0, double *K1, double *K0i, double *K1i) {
    if (P.kind == CALIB) {
        std::memcpy(K0, P.K0, 72); std::memcpy(K1, P.K1, 72); std::memcpy(K0i, P.K0i, 72); std::memcpy(K1i, P.K1i, 72);
        return;
    }
    double f0 = m.f0, f1 = P.kind == SHARED ? m.f0 : m.f1;
    double a[9] = {f0, 0, 0, 0, f0, 0, 0, 0, 1}, b[9] = {f1, 0, 0, 0, f1, 0, 0, 0, 1};
    double ai[9] = {1.0 / f0, 0, 0, 0, 1.0 / f0, 0, 0, 0, 1}, bi[9] = {1.0 / f1, 0, 0, 0, 1.0 / f1, 0, 0, 0, 1};
    std::memcpy(K0, a, 72); std::memcpy(K1, b, 72); std::memcpy(K0i, ai, 72); std::memcpy(K1i, bi, 72);
}

inline void essential(const double *R, const double *t, double *E) { // utils.h:56-62
    double Tx[9] = {0, -t[2], t[1], t[2], 0, -t[0], -t[1], t[0], 0};
    mat3_mul(Tx, R, E);
}

inline double sampson(const double *a, const double *b, const double *E) { // utils.h:64-83
    double e0 = E[0] * a[0] + E[1] * a[1] + E[2], e1 = E[3] * a[0] + E[4] * a[1] + E[5],
           e2 = E[6] * a[0] + E[7] * a[1] + E[8];
    double g0 = E[0] * b[0] + E[3] * b[1] + E[6], g1 = E[1] * b[0] + E[4] * b[1] + E[7];
    double C = b[0] * e0 + b[1] * e1 + e2;
    return C * C / (e0 * e0 + e1 * e1 + g0 * g0 + g1 * g1);
}

// poselib::check_cheirality, identical local copy at solver.cpp:1188-1206
inline bool cheirality(const double *R, const double *t, const double *x1, const double *x2, double min_depth) {
    double Rx1[3];
    mat3_vec(R, x1, Rx1);
    double a = -dot3(Rx1, x2), b1 = -dot3(Rx1, t), b2 = dot3(x2, t);
    double lambda1 = b1 - a * b2, lambda2 = -a * b1 + b2;
    min_depth = min_depth * (1 - a * a);
    return lambda1 > min_depth && lambda2 > min_depth;
}

// EvaluateModelOnPoint
inline double evaluate(const Problem &P, const Model &m, int t, int i, bool for_inlier) {
    if (!for_inlier && P.cfg.score_type == 1 && t != 2) return kMax;
    if (!for_inlier && P.cfg.score_type == 2 && t == 2) return kMax;
    double K0[9], K1[9], K0i[9], K1i[9];
    model_K(P, m, K0, K1, K0i, K1i);
    double h0[3] = {P.x0[2 * i], P.x0[2 * i + 1], 1}, h1[3] = {P.x1[2 * i], P.x1[2 * i + 1], 1};
    if (t == 0) {
        double a[3], q[3], p[3];
        mat3_vec(K0i, h0, a);
        double s = P.d0[i] + m.o0;
        for (double &v : a) v *= s;
        mat3_vec(m.R, a, q);
        for (int k = 0; k < 3; k++) q[k] += m.t[k];
        mat3_vec(K1, q, p);
        if (p[2] < 1e-2) return kMax;
        if (P.scale_only && P.d0[i] < 1e-2) return kMax; // hybrid_pose_estimator.cpp:386 (quirk Q16)
        double u = p[0] / p[2] - h1[0], v = p[1] / p[2] - h1[1];
        return u * u + v * v;
    } else if (t == 1) {
        double b[3], q[3], Rt_t[3], p[3];
        mat3_vec(K1i, h1, b);
        double s = P.d1[i] + m.o1;
        for (double &v : b) v = v * s * m.scale;
        mat3T_vec(m.R, b, q);
        mat3T_vec(m.R, m.t, Rt_t);
        for (int k = 0; k < 3; k++) q[k] -= Rt_t[k];
        mat3_vec(K0, q, p);
        if (p[2] < 1e-2) return kMax;
        if (P.scale_only && P.d1[i] < 1e-2) return kMax; // hybrid_pose_estimator.cpp:398
        double u = p[0] / p[2] - h0[0], v = p[1] / p[2] - h0[1];
        return u * u + v * v;
    } else {
        double E[9];
        essential(m.R, m.t, E);
        if (P.kind == CALIB) {
            double a[3], b[3], an[3], bn[3];
            mat3_vec(K0i, h0, a); mat3_vec(K1i, h1, b);
            double na = norm3(a), nb = norm3(b);
            for (int k = 0; k < 3; k++) { an[k] = a[k] / na; bn[k] = b[k] / nb; }
            if (!cheirality(m.R, m.t, an, bn, 1e-2)) return kMax;
            return sampson(a, b, E) * P.loss_scale;
        }
        double F[9], T[9], K1iT[9];
        mat3_T(K1i, K1iT);
        mat3_mul(K1iT, E, T);
        mat3_mul(T, K0i, F);
        return sampson(h0, h1, F);
    }
}

// ScoreModel, hybrid_ransac.h:265-282 (t-major, i ascending, fp64)
inline double score_model(const Problem &P, const Model &m) {
    double s = 0;
    for (int t = 0; t < 3; t++)
        for (int i = 0; i < P.n; i++) s += std::min(evaluate(P, m, t, i, false), P.thr[t]) * P.w[t];
    return s;
}

// GetInliers, hybrid_ransac.h:289-349 (common = false, all types)
// min_rel (optional, parity forensics): the smallest |r^2 - thr_t| / thr_t over all residuals of the pass, i.e. how
// close the pass came to a tie on the inlier test
inline int get_inliers(const Problem &P, const Model &m, const double *thr, std::vector<std::vector<int>> *inl,
                       double *min_rel = nullptr) {
    int cnt = 0;
    inl->resize(3);
    double mr = kMax;
    for (int t = 0; t < 3; t++) {
        (*inl)[t].clear();
        for (int i = 0; i < P.n; i++) {
            const double r = evaluate(P, m, t, i, true);
            if (r < thr[t]) { cnt++; (*inl)[t].push_back(i); }
            if (min_rel && r < kMax) mr = std::min(mr, std::fabs(r - thr[t]) / thr[t]);
        }
    }
    if (min_rel) *min_rel = mr;
    return cnt;
}

} // namespace om
