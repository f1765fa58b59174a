// ORACLE (test infrastructure only).  C entry points so tests/ can drive the
// CPU restatement through ctypes.  Nothing here is linked into the product.
#include "om_affine.h"
#include "om_estimator.h"

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

} // extern "C"
