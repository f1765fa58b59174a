// ORACLE (test infrastructure only).  MinimalSolver of the three hybrid estimators:
//   /root/reference/src/hybrid_pose_estimator.cpp:65-187
//   /root/reference/src/hybrid_pose_shared_focal_estimator.cpp:53-130
//   /root/reference/src/hybrid_pose_two_focal_estimator.cpp:77-185
// including the use_shift = false branches of the calibrated estimator (:87-120, :146-157) and the
// scale-only estimator's MinimalSolver (:322-367).
#pragma once
#include "om_affine.h"
#include "om_relpose.h"
#include "om_alt.h"

namespace om {

inline bool model_finite(const Model &m) {
    const double *p = reinterpret_cast<const double *>(&m);
    for (int k = 0; k < MP_MODEL_DOUBLES; k++) if (!std::isfinite(p[k])) return false;
    return true;
}

// solver 0: affine-depth solver + Kabsch (solver.cpp:482-534, 682-739, 986-1043) and the
// estimator's min-depth filter / offset1 conversion.  idx: 3 or 4 match indices.
inline void minimal_affine(const Problem &P, const int *idx, std::vector<Model> *models) {
    const int np = P.kind == CALIB ? 3 : 4;
    double x0[4][3], x1[4][3], d0[4], d1[4];
    for (int k = 0; k < np; k++) {
        int i = idx[k];
        double h0[3] = {P.x0[2 * i], P.x0[2 * i + 1], 1}, h1[3] = {P.x1[2 * i], P.x1[2 * i + 1], 1};
        if (P.kind == CALIB) { mat3_vec(P.K0i, h0, x0[k]); mat3_vec(P.K1i, h1, x1[k]); }
        else { std::memcpy(x0[k], h0, 24); std::memcpy(x1[k], h1, 24); }
        d0[k] = P.d0[i]; d1[k] = P.d1[i];
    }
    if (P.scale_only) { // hybrid_pose_estimator.cpp:325-331: Umeyama on the homogeneous rays (no depths), unit weights
        Model m;
        const double w[3] = {1, 1, 1};
        double sc;
        umeyama(3, x0, x1, w, m.R, m.t, &sc);
        m.scale = 1.0 / sc; // "scale now applies on the second camera"
        if (model_finite(m)) models->push_back(m);
        return;
    }
    if (P.kind == CALIB && !P.cfg.use_shift) { // hybrid_pose_estimator.cpp:87-120
        double p0[3][3], p1[3][3];
        for (int k = 0; k < 3; k++)
            for (int c = 0; c < 3; c++) { p0[k][c] = x0[k][c] * d0[k]; p1[k][c] = x1[k][c] * d1[k]; }
        auto dist = [](const double *a, const double *b) {
            double e[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
            return norm3(e);
        };
        const double v0[3] = {dist(p0[0], p0[1]), dist(p0[0], p0[2]), dist(p0[1], p0[2])};
        const double v1[3] = {dist(p1[0], p1[1]), dist(p1[0], p1[2]), dist(p1[1], p1[2])};
        const double scale = dot3(v1, v0) / dot3(v1, v1); // least-squares s with s * v1 = v0
        double Y[3][3];
        for (int k = 0; k < 3; k++)
            for (int c = 0; c < 3; c++) Y[k][c] = x1[k][c] * (d1[k] * scale);
        Model m;
        kabsch(3, p0, Y, m.R, m.t);
        m.scale = scale;
        if (model_finite(m)) models->push_back(m);
        return;
    }
    if (P.opt.use_ours || (P.kind == TWO && P.opt.use_4p4d)) { // alternative solvers (om_alt.h); use_ours wins over use_4p4d
        std::vector<Model> alt;
        if (P.kind == CALIB) { double a[3][3], b[3][3]; std::memcpy(a, x0, sizeof(a)); std::memcpy(b, x1, sizeof(b)); solve_pose_ours_calibrated(a, b, d0, d1, &alt); }
        else if (P.kind == SHARED) solve_pose_ours_shared_focal(x0, x1, d0, d1, &alt);
        else if (P.opt.use_ours) solve_pose_ours_two_focal(x0, x1, d0, d1, &alt);
        else solve_pose_4p4d_two_focal(x0, x1, d0, d1, &alt);
        for (Model m : alt) { // the estimator's filter and offset1 conversion, as for the default solver
            if (P.cfg.min_depth_constraint && !(m.o0 > -P.min_depth[0] && m.o1 > -P.min_depth[1] * m.scale)) continue;
            m.o1 /= m.scale;
            if (model_finite(m)) models->push_back(m);
        }
        return;
    }
    std::vector<AffineSol> sols;
    if (P.kind == CALIB) {
        double a[3][3], b[3][3];
        std::memcpy(a, x0, sizeof(a)); std::memcpy(b, x1, sizeof(b));
        sols = solve_scale_and_shift(a, b, d0, d1);
    } else if (P.kind == SHARED) sols = solve_scale_and_shift_shared_focal(x0, x1, d0, d1);
    else sols = solve_scale_and_shift_two_focal(x0, x1, d0, d1);
    for (const AffineSol &s : sols) {
        double X[4][3], Y[4][3];
        bool pos = true;
        for (int k = 0; k < np; k++) {
            double e1 = d0[k] + s.b1, e2 = d1[k] * s.a2 + s.b2 * s.a2;
            if (e1 <= 0 || e2 <= 0) pos = false; // solver.cpp:503
            double fa = P.kind == CALIB ? 1.0 : s.f1, fb = P.kind == CALIB ? 1.0 : s.f2;
            X[k][0] = x0[k][0] / fa * e1; X[k][1] = x0[k][1] / fa * e1; X[k][2] = x0[k][2] * e1;
            Y[k][0] = x1[k][0] / fb * e2; Y[k][1] = x1[k][1] / fb * e2; Y[k][2] = x1[k][2] * e2;
        }
        if (!pos) continue;
        Model m;
        kabsch(np, X, Y, m.R, m.t);
        m.scale = s.a2; m.o0 = s.b1; m.o1 = s.b2 * s.a2;
        m.f0 = P.kind == CALIB ? 1.0 : s.f1; m.f1 = P.kind == CALIB ? 1.0 : s.f2;
        if (P.cfg.min_depth_constraint && !(m.o0 > -P.min_depth[0] && m.o1 > -P.min_depth[1] * m.scale)) continue;
        m.o1 /= m.scale;
        if (model_finite(m)) models->push_back(m);
    }
}

// common tail of solver 1: triangulate with K0[I|0], K1[R|t], fit depth -> z affinely in
// both views (hybrid_pose_estimator.cpp:137-182)
inline bool depth_fit_tail(const Problem &P, int np, const int *idx, const double (*p0)[2], const double (*p1)[2],
                           double f0, double f1, Model *m) {
    double P0[12] = {f0, 0, 0, 0, 0, f0, 0, 0, 0, 0, 1, 0}, P1[12];
    for (int i = 0; i < 3; i++) {
        double k = i < 2 ? f1 : 1.0;
        for (int j = 0; j < 3; j++) P1[4 * i + j] = k * m->R[3 * i + j];
        P1[4 * i + 3] = k * m->t[i];
    }
    double X[7][3], z[7], d[7];
    for (int k = 0; k < np; k++) { triangulate(P0, P1, p0[k], p1[k], X[k]); z[k] = X[k][2]; d[k] = P.d0[idx[k]]; }
    if (P.scale_only || (P.kind == CALIB && !P.cfg.use_shift)) { // scale-only fits, no filter (:146-157, :351-363)
        auto ratio = [&](const double *dd, const double *zz) {
            double num = 0, den = 0;
            for (int k = 0; k < np; k++) { num += dd[k] * zz[k]; den += dd[k] * dd[k]; }
            return num / den;
        };
        const double s0 = ratio(d, z);
        for (int k = 0; k < 3; k++) m->t[k] /= s0;
        for (int k = 0; k < np; k++) {
            double Xs[3] = {X[k][0] / s0, X[k][1] / s0, X[k][2] / s0}, Y[3];
            mat3_vec(m->R, Xs, Y);
            z[k] = Y[2] + m->t[2]; d[k] = P.d1[idx[k]];
        }
        m->scale = ratio(d, z); m->o0 = 0; m->o1 = 0;
        return model_finite(*m);
    }
    double s0, c0;
    fit_affine(np, d, z, &s0, &c0);
    double off0 = c0 / s0;
    if (P.cfg.min_depth_constraint && off0 < -P.min_depth[0]) return false;
    for (int k = 0; k < 3; k++) m->t[k] /= s0;
    for (int k = 0; k < np; k++) {
        double Xs[3] = {X[k][0] / s0, X[k][1] / s0, X[k][2] / s0}, Y[3];
        mat3_vec(m->R, Xs, Y);
        z[k] = Y[2] + m->t[2]; d[k] = P.d1[idx[k]];
    }
    double s1, c1;
    fit_affine(np, d, z, &s1, &c1);
    double off1 = c1 / s1;
    if (P.cfg.min_depth_constraint && off1 < -P.min_depth[1]) return false;
    m->scale = s1; m->o0 = off0; m->o1 = off1;
    return model_finite(*m);
}

// solver 1: point-only relative pose + depth fits.  idx: 5 / 6 / 7 match indices.
inline void minimal_points(const Problem &P, const int *idx, std::vector<Model> *models) {
    const int np = P.sizes[1][2];
    double b0[7][3], b1[7][3], p0[7][2], p1[7][2];
    for (int k = 0; k < np; k++) {
        int i = idx[k];
        double h0[3] = {P.x0[2 * i], P.x0[2 * i + 1], 1}, h1[3] = {P.x1[2 * i], P.x1[2 * i + 1], 1}, a[3], b[3];
        if (P.kind == CALIB) { mat3_vec(P.K0i, h0, a); mat3_vec(P.K1i, h1, b); }
        else { std::memcpy(a, h0, 24); std::memcpy(b, h1, 24); }
        double na = norm3(a), nb = norm3(b);
        for (int c = 0; c < 3; c++) { b0[k][c] = a[c] / na; b1[k][c] = b[c] / nb; }
        p0[k][0] = a[0]; p0[k][1] = a[1]; p1[k][0] = b[0]; p1[k][1] = b[1];
    }
    if (P.kind == CALIB) {
        std::vector<Pose> poses;
        relpose_5pt(b0, b1, &poses);
        for (const Pose &ps : poses) {
            Model m;
            std::memcpy(m.R, ps.R, 72); std::memcpy(m.t, ps.t, 24);
            if (depth_fit_tail(P, np, idx, p0, p1, 1.0, 1.0, &m)) models->push_back(m);
        }
    } else if (P.kind == SHARED) {
        double h0[6][3], h1[6][3];
        for (int k = 0; k < 6; k++) { h0[k][0] = p0[k][0]; h0[k][1] = p0[k][1]; h0[k][2] = 1; h1[k][0] = p1[k][0]; h1[k][1] = p1[k][1]; h1[k][2] = 1; }
        std::vector<FocalPose> fps;
        relpose_6pt_shared_focal(h0, h1, &fps);
        for (const FocalPose &fp : fps) {
            Model m;
            std::memcpy(m.R, fp.pose.R, 72); std::memcpy(m.t, fp.pose.t, 24);
            m.f0 = fp.focal; m.f1 = fp.focal;
            if (depth_fit_tail(P, np, idx, p0, p1, fp.focal, fp.focal, &m)) models->push_back(m);
        }
    } else {
        std::vector<std::vector<double>> Fs;
        relpose_7pt(b0, b1, &Fs);
        for (const auto &F : Fs) {
            double f0sq, f1sq;
            bougnoux_focals(F.data(), &f0sq, &f1sq);
            double f0 = std::sqrt(std::fabs(f0sq)), f1 = std::sqrt(std::fabs(f1sq));
            if (!std::isfinite(f0) || !std::isfinite(f1) || f0 == 0 || f1 == 0) continue;
            double E[9]; // E = K1^T F K0
            for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) E[3 * i + j] = F[3 * i + j] * (i < 2 ? f1 : 1.0) * (j < 2 ? f0 : 1.0);
            Model m;
            recover_pose(E, np, p0, p1, 1e9, m.R, m.t); // NB: un-calibrated points with K = I, as the reference does
            m.f0 = f0; m.f1 = f1;
            if (depth_fit_tail(P, np, idx, p0, p1, f0, f1, &m)) models->push_back(m);
        }
    }
}

inline void minimal_solver(const Problem &P, int solver_idx, const int *idx, std::vector<Model> *models) {
    models->clear();
    if (solver_idx == 0) minimal_affine(P, idx, models);
    else minimal_points(P, idx, models);
}

} // namespace om
