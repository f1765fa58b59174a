// MinimalSolver of the three hybrid estimators on one sample, fp64, one thread:
//   solver 0: affine-depth solver + Kabsch + min-depth filter
//             (hybrid_pose_estimator.cpp:65-120, ...shared_focal....cpp:53-84, ...two_focal....cpp:77-113)
//   solver 1: point-only relative pose, DLT triangulation, affine depth fits
//             (hybrid_pose_estimator.cpp:121-187, ...shared_focal....cpp:85-130, ...two_focal....cpp:114-185)
// Models come out in the estimator's internal units, in the solver's root order.
#pragma once
#include "mp_affine.cuh"
#include "mp_relpose.cuh"
#include "mp_alt.cuh"

namespace mp {

// flags of the minimal solvers: EstimatorConfig::min_depth_constraint; EstimatorConfig::use_shift == false
// (honoured by the calibrated estimator only, hybrid_pose_estimator.cpp:73,146)
// HybridLORansacOptions::use_ours / use_4p4d: the alternative depth-aware solvers of mp_alt.cuh
enum { MF_MIN_DEPTH = 1u, MF_NO_SHIFT = 2u, MF_OURS = 4u, MF_4P4D = 8u };
MP_HD unsigned minimal_flags(const mp_config &cfg, const mp_options &opt) {
    return (cfg.min_depth_constraint ? MF_MIN_DEPTH : 0u) | (cfg.use_shift ? 0u : MF_NO_SHIFT) | (opt.use_ours ? MF_OURS : 0u) |
           (opt.use_4p4d ? MF_4P4D : 0u);
}

// calibrated estimator without the shift (EstimatorConfig::use_shift = false, hybrid_pose_estimator.cpp:87-120): scale
// from the three pairwise distances of the back-projected points, then Kabsch; one model, offsets 0, no filter.
// scale_only (HybridPoseEstimatorScaleOnly::MinimalSolver, :322-331): Umeyama on the homogeneous rays, unit weights.
MP_HD int minimal_affine_noshift(const PairConst &pc, const Match64 *M, const int *idx, Model *out) {
    double x0[3][3], x1[3][3], d0[3], d1[3];
    for (int k = 0; k < 3; k++) {
        const Match64 m = M[idx[k]];
        mat3_vec(pc.K0i, m.x0x, m.x0y, 1.0, x0[k]); mat3_vec(pc.K1i, m.x1x, m.x1y, 1.0, x1[k]);
        d0[k] = m.d0; d1[k] = m.d1;
    }
    Model m;
    m.o0 = 0; m.o1 = 0; m.f0 = 1; m.f1 = 1;
    if (pc.scale_only) {
        const double w[3] = {1, 1, 1};
        double sc;
        if (!umeyama(3, x0, x1, w, m.R, m.t, &sc)) return 0;
        m.scale = 1.0 / sc; // "scale now applies on the second camera" (:330)
    } else {
        double p0[3][3], p1[3][3];
        for (int k = 0; k < 3; k++)
            for (int c = 0; c < 3; c++) { p0[k][c] = x0[k][c] * d0[k]; p1[k][c] = x1[k][c] * d1[k]; }
        double v0[3], v1[3];
        const int pa[3] = {0, 0, 1}, pb[3] = {1, 2, 2};
        for (int e = 0; e < 3; e++) {
            const double a[3] = {p0[pa[e]][0] - p0[pb[e]][0], p0[pa[e]][1] - p0[pb[e]][1], p0[pa[e]][2] - p0[pb[e]][2]};
            const double b[3] = {p1[pa[e]][0] - p1[pb[e]][0], p1[pa[e]][1] - p1[pb[e]][1], p1[pa[e]][2] - p1[pb[e]][2]};
            v0[e] = norm3(a); v1[e] = norm3(b);
        }
        const double scale = dot3(v1, v0) / dot3(v1, v1);
        double Y[3][3];
        for (int k = 0; k < 3; k++)
            for (int c = 0; c < 3; c++) Y[k][c] = x1[k][c] * (d1[k] * scale);
        if (!kabsch<3>(p0, Y, m.R, m.t)) return 0;
        m.scale = scale;
    }
    if (!model_finite(m)) return 0;
    out[0] = m;
    return 1;
}

// solve_scale_shift_pose{,_shared_focal,_two_focal} (solver.cpp:482-534, 682-739, 986-1043, scale_on_x = false) on one
// sample: affine-depth solutions -> positive corrected depths (:503,702,1006) -> Kabsch.  Models carry the wrappers'
// convention: scale = a2, offset0 = b1, offset1 = b2 * a2 (NOT yet divided by the scale).  sols_out (optional): the raw
// solutions of solve_scale_and_shift* (:117-121: (b1, a2, b2 a2[, f...])) in the solver's order, count in *ns_out.
template <int KIND> MP_HD int affine_pose_models(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1,
                                                 Model *out, AffSol *sols_out = nullptr, int *ns_out = nullptr) {
    constexpr int NPT = KIND == CALIB ? 3 : 4;
    AffSol sols[8];
    int ns;
    if (KIND == CALIB) ns = affine_calibrated(x0, x1, d0, d1, sols);
    else if (KIND == SHARED) ns = affine_shared_focal(x0, x1, d0, d1, sols);
    else ns = affine_two_focal(x0, x1, d0, d1, sols);
    if (sols_out) { for (int q = 0; q < ns; q++) sols_out[q] = sols[q]; *ns_out = ns; }
    // pass 1 (cheap): solutions whose corrected depths are all positive, compacted in order, so that the expensive part
    // below runs over lists of similar length in every lane of a warp
    int keep[8], nk = 0;
    for (int q = 0; q < ns; q++) {
        const AffSol &s = sols[q];
        bool pos = true;
        for (int k = 0; k < NPT; k++) {
            const double e1 = d0[k] + s.b1, e2 = d1[k] * s.a2 + s.b2 * s.a2;
            if (e1 <= 0 || e2 <= 0) pos = false;
        }
        if (pos) keep[nk++] = q;
    }
    int n = 0;
    for (int qq = 0; qq < nk; qq++) {
        const AffSol &s = sols[keep[qq]];
        double X[NPT][3], Y[NPT][3];
        const double fa = KIND == CALIB ? 1.0 : s.f1, fb = KIND == CALIB ? 1.0 : s.f2;
        for (int k = 0; k < NPT; k++) {
            const double e1 = d0[k] + s.b1, e2 = d1[k] * s.a2 + s.b2 * s.a2;
            X[k][0] = x0[k][0] / fa * e1; X[k][1] = x0[k][1] / fa * e1; X[k][2] = x0[k][2] * e1;
            Y[k][0] = x1[k][0] / fb * e2; Y[k][1] = x1[k][1] / fb * e2; Y[k][2] = x1[k][2] * e2;
        }
        Model m;
        if (!kabsch<NPT>(X, Y, m.R, m.t)) continue;
        m.scale = s.a2; m.o0 = s.b1; m.o1 = s.b2 * s.a2; m.f0 = fa; m.f1 = fb;
        out[n++] = m;
    }
    return n;
}

template <int KIND> MP_HD int minimal_affine(const PairConst &pc, const Match64 *M, const int *idx, unsigned mflags,
                                             Model *out) {
    constexpr int NPT = KIND == CALIB ? 3 : 4;
    if (KIND == CALIB && (pc.scale_only || (mflags & MF_NO_SHIFT))) return minimal_affine_noshift(pc, M, idx, out);
    const bool min_depth_on = (mflags & MF_MIN_DEPTH) != 0;
    double x0[NPT][3], x1[NPT][3], d0[NPT], d1[NPT];
    for (int k = 0; k < NPT; k++) {
        const Match64 m = M[idx[k]];
        if (KIND == CALIB) { mat3_vec(pc.K0i, m.x0x, m.x0y, 1.0, x0[k]); mat3_vec(pc.K1i, m.x1x, m.x1y, 1.0, x1[k]); }
        else { x0[k][0] = m.x0x; x0[k][1] = m.x0y; x0[k][2] = 1.0; x1[k][0] = m.x1x; x1[k][1] = m.x1y; x1[k][2] = 1.0; }
        d0[k] = m.d0; d1[k] = m.d1;
    }
    int nm;
    if ((mflags & MF_OURS) || (KIND == TWO && (mflags & MF_4P4D))) { // use_ours wins over use_4p4d (...two_focal....cpp:87-93)
        if (KIND == CALIB) nm = alt_calibrated_ours(x0, x1, d0, d1, out);
        else if (KIND == SHARED) nm = alt_shared_focal_ours(x0, x1, d0, d1, out);
        else if (mflags & MF_OURS) nm = alt_two_focal_ours(x0, x1, d0, d1, out);
        else nm = alt_two_focal_4p4d(x0, x1, d0, d1, out);
    } else nm = affine_pose_models<KIND>(x0, x1, d0, d1, out);
    // the estimator's min-depth filter and offset1 conversion (hybrid_pose_estimator.cpp:79-85 and the focal twins)
    int n = 0;
    for (int q = 0; q < nm; q++) {
        Model m = out[q];
        if (min_depth_on && !(m.o0 > -pc.min_depth[0] && m.o1 > -pc.min_depth[1] * m.scale)) continue;
        m.o1 /= m.scale;
        if (model_finite(m)) out[n++] = m;
    }
    return n;
}

// triangulate the sample with K0 [I|0], K1 [R|t]; fit prior depth -> triangulated depth
// affinely in both views (hybrid_pose_estimator.cpp:137-182)
// scale_fit: the scale-only fits of hybrid_pose_estimator.cpp:146-157 / :351-363 (no offsets, no filter)
template <int NPT> MP_HD bool depth_fit_tail(const PairConst &pc, const double *d0s, const double *d1s,
                                             const double (*p0)[2], const double (*p1)[2], double f0, double f1,
                                             unsigned mflags, Model *m) {
    const bool min_depth_on = (mflags & MF_MIN_DEPTH) != 0;
    const bool scale_fit = pc.kind == CALIB && (pc.scale_only || (mflags & MF_NO_SHIFT));
    const double P0[12] = {f0, 0, 0, 0, 0, f0, 0, 0, 0, 0, 1, 0};
    double P1[12];
    for (int i = 0; i < 3; i++) {
        const double k = i < 2 ? f1 : 1.0;
        for (int j = 0; j < 3; j++) P1[4 * i + j] = k * m->R[3 * i + j];
        P1[4 * i + 3] = k * m->t[i];
    }
    double X[NPT][3], z[NPT];
    for (int k = 0; k < NPT; k++) {
        double Q[4];
        triangulate_h(P0, P1, p0[k], p1[k], Q);
        for (int c = 0; c < 3; c++) X[k][c] = Q[c] / Q[3];
        z[k] = X[k][2];
    }
    if (scale_fit) {
        double num = 0, den = 0;
        for (int k = 0; k < NPT; k++) { num += d0s[k] * z[k]; den += d0s[k] * d0s[k]; }
        const double s0 = num / den;
        for (int k = 0; k < 3; k++) m->t[k] /= s0;
        num = 0; den = 0;
        for (int k = 0; k < NPT; k++) {
            const double Xs[3] = {X[k][0] / s0, X[k][1] / s0, X[k][2] / s0};
            double Y[3];
            mat3_vecp(m->R, Xs, Y);
            num += d1s[k] * (Y[2] + m->t[2]); den += d1s[k] * d1s[k];
        }
        m->scale = num / den; m->o0 = 0; m->o1 = 0;
        return model_finite(*m);
    }
    double s0, c0;
    fit_affine(NPT, d0s, z, &s0, &c0);
    const double off0 = c0 / s0;
    if (min_depth_on && off0 < -pc.min_depth[0]) return false;
    for (int k = 0; k < 3; k++) m->t[k] /= s0;
    for (int k = 0; k < NPT; k++) {
        const double Xs[3] = {X[k][0] / s0, X[k][1] / s0, X[k][2] / s0};
        double Y[3];
        mat3_vecp(m->R, Xs, Y);
        z[k] = Y[2] + m->t[2];
    }
    double s1, c1;
    fit_affine(NPT, d1s, z, &s1, &c1);
    const double off1 = c1 / s1;
    if (min_depth_on && off1 < -pc.min_depth[1]) return false;
    m->scale = s1; m->o0 = off0; m->o1 = off1;
    return model_finite(*m);
}

// the sample of the point-only solvers: unit bearings b0/b1, homogeneous image points p0/p1 (K^-1-normalised when
// calibrated), prior depths
template <int KIND, int NPT> MP_HD void load_point_sample(const PairConst &pc, const Match64 *M, const int *idx, double (*b0)[3],
                                                         double (*b1)[3], double (*p0)[2], double (*p1)[2], double *d0, double *d1) {
    for (int k = 0; k < NPT; k++) {
        const Match64 m = M[idx[k]];
        double a[3], b[3];
        if (KIND == CALIB) { mat3_vec(pc.K0i, m.x0x, m.x0y, 1.0, a); mat3_vec(pc.K1i, m.x1x, m.x1y, 1.0, b); }
        else { a[0] = m.x0x; a[1] = m.x0y; a[2] = 1.0; b[0] = m.x1x; b[1] = m.x1y; b[2] = 1.0; }
        const double na = norm3(a), nb = norm3(b);
        for (int c = 0; c < 3; c++) { b0[k][c] = a[c] / na; b1[k][c] = b[c] / nb; }
        p0[k][0] = a[0]; p0[k][1] = a[1]; p1[k][0] = b[0]; p1[k][1] = b[1];
        d0[k] = m.d0; d1[k] = m.d1;
    }
}

// with_tail = false: the candidates (R, t, f0, f1) are returned WITHOUT the triangulation / depth-fit tail; the device
// loop runs that tail in minimal_tail_kernel, one thread per candidate (mp_minimal.cu)
// defer_pose (two focals, with_tail = false only): the candidate carries the factors of E (Ra in Model::R, t, and Rb in
// out[n + 4].R: `out` must hold 8 models) and no pose yet; the device loop runs recoverPose -- 28 triangulations per
// candidate -- in recover_pose_warp_kernel, one warp per candidate
template <int KIND> MP_HD int minimal_points(const PairConst &pc, const Match64 *M, const int *idx, unsigned mflags,
                                             Model *out, bool with_tail = true, bool defer_pose = false) {
    constexpr int NPT = KIND == CALIB ? 5 : (KIND == SHARED ? 6 : 7);
    double b0[NPT][3], b1[NPT][3], p0[NPT][2], p1[NPT][2], d0[NPT], d1[NPT];
    load_point_sample<KIND, NPT>(pc, M, idx, b0, b1, p0, p1, d0, d1);
    int n = 0;
    if (KIND == CALIB) {
        Pose poses[10];
        const int np = relpose_5pt(b0, b1, poses);
        for (int q = 0; q < np; q++) {
            Model m;
            for (int k = 0; k < 9; k++) m.R[k] = poses[q].R[k];
            for (int k = 0; k < 3; k++) m.t[k] = poses[q].t[k];
            m.f0 = 1.0; m.f1 = 1.0;
            if (!with_tail) { m.scale = 1; m.o0 = 0; m.o1 = 0; out[n++] = m; continue; }
            if (depth_fit_tail<NPT>(pc, d0, d1, p0, p1, 1.0, 1.0, mflags, &m)) out[n++] = m;
        }
    } else if (KIND == SHARED) {
        double h0[NPT][3], h1[NPT][3];
        for (int k = 0; k < NPT; k++) { h0[k][0] = p0[k][0]; h0[k][1] = p0[k][1]; h0[k][2] = 1; h1[k][0] = p1[k][0]; h1[k][1] = p1[k][1]; h1[k][2] = 1; }
        FocalPose fps[15];
        const int np = relpose_6pt_shared_focal(h0, h1, fps);
        for (int q = 0; q < np; q++) {
            Model m;
            for (int k = 0; k < 9; k++) m.R[k] = fps[q].pose.R[k];
            for (int k = 0; k < 3; k++) m.t[k] = fps[q].pose.t[k];
            m.f0 = fps[q].focal; m.f1 = fps[q].focal;
            if (!with_tail) { m.scale = 1; m.o0 = 0; m.o1 = 0; out[n++] = m; continue; }
            if (depth_fit_tail<NPT>(pc, d0, d1, p0, p1, m.f0, m.f1, mflags, &m)) out[n++] = m;
        }
    } else {
        double Fs[3][9];
        const int nf = fundamental_7pt(b0, b1, Fs);
        for (int q = 0; q < nf; q++) {
            double f0sq, f1sq;
            bougnoux_focals(Fs[q], &f0sq, &f1sq);
            const double f0 = sqrt(fabs(f0sq)), f1 = sqrt(fabs(f1sq)); // ...two_focal....cpp:121-122
            if (!finite_d(f0) || !finite_d(f1) || f0 == 0.0 || f1 == 0.0) continue;
            double E[9]; // E = K1^T F K0
            for (int i = 0; i < 3; i++)
                for (int j = 0; j < 3; j++) E[3 * i + j] = Fs[q][3 * i + j] * (i < 2 ? f1 : 1.0) * (j < 2 ? f0 : 1.0);
            Model m;
            if (defer_pose && !with_tail) { // factors of E: Ra, t here, Rb in the spare slot n + 4 (the kind uses <= 4 slots)
                Model &spare = out[n + 4];
                essential_factor(E, m.R, spare.R, m.t);
            } else {
                recover_pose<NPT>(E, p0, p1, 1e9, m.R, m.t); // un-calibrated points with K = I, as the reference does
            }
            m.f0 = f0; m.f1 = f1;
            if (!with_tail) { m.scale = 1; m.o0 = 0; m.o1 = 0; out[n++] = m; continue; }
            if (depth_fit_tail<NPT>(pc, d0, d1, p0, p1, f0, f1, mflags, &m)) out[n++] = m;
        }
    }
    return n;
}

// the tail of minimal_points on ONE candidate (R, t, f0, f1 set): triangulation + affine depth fits
template <int KIND> MP_HD bool minimal_points_tail(const PairConst &pc, const Match64 *M, const int *idx, unsigned mflags,
                                                   Model *m) {
    constexpr int NPT = KIND == CALIB ? 5 : (KIND == SHARED ? 6 : 7);
    double b0[NPT][3], b1[NPT][3], p0[NPT][2], p1[NPT][2], d0[NPT], d1[NPT];
    load_point_sample<KIND, NPT>(pc, M, idx, b0, b1, p0, p1, d0, d1);
    return depth_fit_tail<NPT>(pc, d0, d1, p0, p1, m->f0, m->f1, mflags, m);
}

// dispatch; out must hold kMaxModelsAny models
MP_HD int minimal_solver(const PairConst &pc, const Match64 *M, int solver_idx, const int *idx, unsigned mflags,
                         Model *out, bool with_tail = true) {
    if (solver_idx == 0) {
        if (pc.kind == CALIB) return minimal_affine<CALIB>(pc, M, idx, mflags, out);
        if (pc.kind == SHARED) return minimal_affine<SHARED>(pc, M, idx, mflags, out);
        return minimal_affine<TWO>(pc, M, idx, mflags, out);
    }
    if (pc.kind == CALIB) return minimal_points<CALIB>(pc, M, idx, mflags, out, with_tail);
    if (pc.kind == SHARED) return minimal_points<SHARED>(pc, M, idx, mflags, out, with_tail);
    return minimal_points<TWO>(pc, M, idx, mflags, out, with_tail);
}

} // namespace mp
