// ORACLE (test infrastructure only) -- continuation of om_lm.h: the trust-region
// Levenberg-Marquardt driver.  See the header of om_lm.h for what is restated.
#pragma once

namespace om {

constexpr int NT = 11; // tangent slots: o0, rot(3), t(3), scale, o1, f0, f1

struct LmInfo { int iterations = 0, termination = 0; double cost0 = 0, cost = 0; };
// termination: 0 not run (guard / no residuals), 1 convergence, 2 max iterations, 3 failure, 4 infeasible start

struct LmProblem {
    LmData D;
    const std::vector<int> *idx; // [3]
    bool use_reproj, use_samp;
    int slot[NT], nt;            // compact index of each tangent slot or -1
    double lower[NA]; bool has_lower[NA];
};

inline void rot_to_quat(const double *R, double *q) { // unit quaternion (w,x,y,z) of a rotation matrix
    double tr = R[0] + R[4] + R[8];
    if (tr > 0) {
        double s = std::sqrt(tr + 1.0) * 2; q[0] = 0.25 * s; q[1] = (R[7] - R[5]) / s; q[2] = (R[2] - R[6]) / s; q[3] = (R[3] - R[1]) / s;
    } else if (R[0] > R[4] && R[0] > R[8]) {
        double s = std::sqrt(1.0 + R[0] - R[4] - R[8]) * 2; q[0] = (R[7] - R[5]) / s; q[1] = 0.25 * s; q[2] = (R[1] + R[3]) / s; q[3] = (R[2] + R[6]) / s;
    } else if (R[4] > R[8]) {
        double s = std::sqrt(1.0 + R[4] - R[0] - R[8]) * 2; q[0] = (R[2] - R[6]) / s; q[1] = (R[1] + R[3]) / s; q[2] = 0.25 * s; q[3] = (R[5] + R[7]) / s;
    } else {
        double s = std::sqrt(1.0 + R[8] - R[0] - R[4]) * 2; q[0] = (R[3] - R[1]) / s; q[1] = (R[2] + R[6]) / s; q[2] = (R[5] + R[7]) / s; q[3] = 0.25 * s;
    }
    double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int k = 0; k < 4; k++) q[k] /= n;
}

// cost = 1/2 sum r^2 ; optionally A = J^T J (nt x nt, full), g = J^T r in tangent space
inline double lm_evaluate(const LmProblem &L, const double *x, bool want_jac, double *A, double *g) {
    const Problem &P = *L.D.P;
    const int nt = L.nt;
    if (want_jac) { for (int i = 0; i < nt * nt; i++) A[i] = 0; for (int i = 0; i < nt; i++) g[i] = 0; }
    double cost = 0;
    // ambient -> tangent map for the quaternion (QuaternionManifold::PlusJacobian)
    const double *q = x + 1;
    double PJ[4][3] = {{-q[1], -q[2], -q[3]}, {q[0], q[3], -q[2]}, {-q[3], q[0], q[1]}, {q[2], -q[1], q[0]}};
    auto accumulate = [&](const Jet *r, int nr) {
        for (int k = 0; k < nr; k++) {
            cost += 0.5 * r[k].a * r[k].a;
            if (!want_jac) continue;
            double jt[NT];
            jt[0] = r[k].v[0];
            for (int c = 0; c < 3; c++) jt[1 + c] = r[k].v[1] * PJ[0][c] + r[k].v[2] * PJ[1][c] + r[k].v[3] * PJ[2][c] + r[k].v[4] * PJ[3][c];
            for (int c = 0; c < 3; c++) jt[4 + c] = r[k].v[5 + c];
            jt[7] = r[k].v[8]; jt[8] = r[k].v[9]; jt[9] = r[k].v[10]; jt[10] = r[k].v[11];
            double jc[NT];
            for (int s = 0; s < NT; s++) if (L.slot[s] >= 0) jc[L.slot[s]] = jt[s];
            for (int a = 0; a < nt; a++) { g[a] += jc[a] * r[k].a; for (int b = 0; b < nt; b++) A[a * nt + b] += jc[a] * jc[b]; }
        }
    };
    Jet p[NA];
    for (int k = 0; k < NA; k++) p[k] = Jet(x[k], k);
    Jet f0 = p[10], f1 = P.kind == TWO ? p[11] : p[10];
    Jet r[2];
    if (L.use_reproj) {
        for (int i : L.idx[0]) { res_reproj0(L.D, i, p, f0, f1, r); accumulate(r, 2); }
        for (int i : L.idx[1]) { res_reproj1(L.D, i, p, f0, f1, r); accumulate(r, 2); }
    }
    if (L.use_samp) for (int i : L.idx[2]) { res_sampson(L.D, i, p, f0, f1, r); accumulate(r, 1); }
    return cost;
}

// x_plus = Plus(x, delta) with the quaternion manifold, then projection onto the bounds
inline void lm_plus(const LmProblem &L, const double *x, const double *d /*nt*/, double *xp) {
    std::memcpy(xp, x, sizeof(double) * NA);
    auto dv = [&](int s) { return L.slot[s] >= 0 ? d[L.slot[s]] : 0.0; };
    xp[0] += dv(0); xp[5] += dv(4); xp[6] += dv(5); xp[7] += dv(6); xp[8] += dv(7); xp[9] += dv(8); xp[10] += dv(9); xp[11] += dv(10);
    double w[3] = {dv(1), dv(2), dv(3)}, n = std::sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    if (n > 0) {
        double sn = std::sin(n) / n, e[4] = {std::cos(n), sn * w[0], sn * w[1], sn * w[2]};
        const double *q = x + 1;
        xp[1] = e[0] * q[0] - e[1] * q[1] - e[2] * q[2] - e[3] * q[3];
        xp[2] = e[0] * q[1] + e[1] * q[0] + e[2] * q[3] - e[3] * q[2];
        xp[3] = e[0] * q[2] - e[1] * q[3] + e[2] * q[0] + e[3] * q[1];
        xp[4] = e[0] * q[3] + e[1] * q[2] - e[2] * q[1] + e[3] * q[0];
    }
    for (int k = 0; k < NA; k++) if (L.has_lower[k] && xp[k] < L.lower[k]) xp[k] = L.lower[k];
}

inline bool cholesky_solve(int n, const double *A, const double *b, double *x) {
    double Lw[NT * NT];
    for (int i = 0; i < n; i++)
        for (int j = 0; j <= i; j++) {
            double s = A[i * n + j];
            for (int k = 0; k < j; k++) s -= Lw[i * n + k] * Lw[j * n + k];
            if (i == j) { if (!(s > 0) || !std::isfinite(s)) return false; Lw[i * n + i] = std::sqrt(s); }
            else Lw[i * n + j] = s / Lw[j * n + j];
        }
    double y[NT];
    for (int i = 0; i < n; i++) { double s = b[i]; for (int k = 0; k < i; k++) s -= Lw[i * n + k] * y[k]; y[i] = s / Lw[i * n + i]; }
    for (int i = n - 1; i >= 0; i--) { double s = y[i]; for (int k = i + 1; k < n; k++) s -= Lw[k * n + i] * x[k]; x[i] = s / Lw[i * n + i]; }
    for (int i = 0; i < n; i++) if (!std::isfinite(x[i])) return false;
    return true;
}

// NonMinimalSolver / LeastSquares.  Returns false when the guard rejects the sample
// (model untouched); true otherwise (like optimizer.h:113-125, whatever Ceres did).
// nonminimal: the call is NonMinimalSolver (true) or LeastSquares (false).  EstimatorConfig::use_shift reaches the
// optimiser only at hybrid_pose_estimator.cpp:206,281 (calibrated, both calls) and
// hybrid_pose_shared_focal_estimator.cpp:149 (shared focal, NonMinimalSolver only); the scale-only optimiser
// (optimizer.h:133-245) always holds the offsets constant and sets no offset bounds.
inline bool lm_use_shift(const Problem &P, bool nonminimal) {
    if (P.scale_only) return false;
    if (P.kind == CALIB) return P.cfg.use_shift != 0;
    if (P.kind == SHARED) return nonminimal ? P.cfg.use_shift != 0 : true;
    return true;
}
inline bool lm_solve(const Problem &P, const std::vector<int> *idx, Model *m, LmInfo *info, bool nonminimal = false) {
    LmInfo dummy; if (!info) info = &dummy; *info = LmInfo();
    const int k0 = P.sizes[0][0], k2 = P.sizes[1][2];
    if (((int)idx[0].size() < k0 && (int)idx[1].size() < k0) || (int)idx[2].size() < k2) return false;
    LmProblem L; L.D.P = &P; L.idx = idx;
    L.use_reproj = P.cfg.lo_type != 1; L.use_samp = P.cfg.lo_type != 2;
    double sw = std::sqrt(P.sampson_w);
    L.D.W = P.kind == CALIB ? sw / (1.0 / (P.K0[0] + P.K0[4]) + 1.0 / (P.K1[0] + P.K1[4])) : sw;
    const bool b0 = L.use_reproj && !idx[0].empty(), b1 = L.use_reproj && !idx[1].empty(), b2 = L.use_samp && !idx[2].empty();
    if (!b0 && !b1 && !b2) return false;
    bool act[NT];
    const bool shift = lm_use_shift(P, nonminimal);
    act[0] = b0 && shift; act[1] = act[2] = act[3] = act[4] = act[5] = act[6] = true;
    act[7] = b1; act[8] = b1 && shift; act[9] = P.kind != CALIB; act[10] = P.kind == TWO;
    L.nt = 0;
    for (int s = 0; s < NT; s++) L.slot[s] = act[s] ? L.nt++ : -1;
    for (int k = 0; k < NA; k++) { L.has_lower[k] = false; L.lower[k] = 0; }
    if (b1) { L.has_lower[8] = true; L.lower[8] = 1e-2; }
    if (!P.scale_only && P.cfg.min_depth_constraint && b0) { L.has_lower[0] = true; L.lower[0] = -P.min_depth[0] + 1e-2; }
    if (!P.scale_only && P.cfg.min_depth_constraint && b1) { L.has_lower[9] = true; L.lower[9] = -P.min_depth[1] + 1e-2; }
    if (P.kind == TWO) { L.has_lower[10] = L.has_lower[11] = true; L.lower[10] = L.lower[11] = 1e-6; }
    double x[NA];
    x[0] = m->o0; rot_to_quat(m->R, x + 1); std::memcpy(x + 5, m->t, 24); x[8] = m->scale; x[9] = m->o1; x[10] = m->f0; x[11] = m->f1;
    auto write_back = [&]() {
        double qn[4]; quat_normalize(x + 1, qn); quat_to_R(qn, m->R);
        m->o0 = x[0]; std::memcpy(m->t, x + 5, 24); m->scale = x[8]; m->o1 = x[9]; m->f0 = x[10]; m->f1 = P.kind == TWO ? x[11] : (P.kind == SHARED ? x[10] : m->f1);
    };
    for (int k = 0; k < NA; k++) if (L.has_lower[k] && x[k] < L.lower[k]) { info->termination = 4; write_back(); return true; }
    const int nt = L.nt;
    const bool constrained = true; // the scale bound (or a focal bound) is always present when b1 / TWO; treat uniformly
    double A[NT * NT], g[NT], An[NT * NT], gn[NT], scale[NT];
    double cost = lm_evaluate(L, x, true, A, g);
    info->cost0 = cost; info->cost = cost;
    for (int a = 0; a < nt; a++) scale[a] = 1.0 / (1.0 + std::sqrt(A[a * nt + a]));
    auto grad_max = [&](const double *xx, const double *gg) {
        double d[NT], xp[NA], mx = 0;
        for (int a = 0; a < nt; a++) d[a] = -gg[a];
        if (constrained) { lm_plus(L, xx, d, xp); for (int k = 0; k < NA; k++) mx = std::max(mx, std::fabs(xp[k] - xx[k])); }
        else for (int a = 0; a < nt; a++) mx = std::max(mx, std::fabs(gg[a]));
        return mx;
    };
    if (!std::isfinite(cost)) { info->termination = 3; write_back(); return true; }
    if (grad_max(x, g) <= P.cfg.ceres_gradient_tolerance) { info->termination = 1; write_back(); return true; }
    double radius = 1e4, decrease = 2.0;
    // TrustRegionStepEvaluator state
    const int max_nonmono = P.cfg.ceres_use_nonmonotonic_steps ? 5 : 0;
    double min_cost = cost, cur_cost = cost, ref_cost = cost, cand_cost = cost, acc_ref = 0, acc_cand = 0;
    int n_nonmono = 0, n_invalid = 0;
    const int max_it = (int)P.cfg.ceres_max_num_iterations;
    info->termination = 2;
    for (int it = 1; it <= max_it; it++) {
        info->iterations = it;
        if (radius < 1e-32) { info->termination = 1; break; }
        // damped, column-scaled normal equations
        double As[NT * NT], gs[NT], y[NT];
        for (int a = 0; a < nt; a++) { gs[a] = g[a] * scale[a]; for (int b = 0; b < nt; b++) As[a * nt + b] = A[a * nt + b] * scale[a] * scale[b]; }
        double Ad[NT * NT];
        std::memcpy(Ad, As, sizeof(double) * nt * nt);
        for (int a = 0; a < nt; a++) { double dg = std::min(std::max(As[a * nt + a], 1e-6), 1e32); Ad[a * nt + a] += dg / radius; }
        bool ok = cholesky_solve(nt, Ad, gs, y);
        double mcc = 0;
        if (ok) {
            for (int a = 0; a < nt; a++) y[a] = -y[a];
            double dg = 0, dAd = 0;
            for (int a = 0; a < nt; a++) { dg += y[a] * gs[a]; for (int b = 0; b < nt; b++) dAd += y[a] * As[a * nt + b] * y[b]; }
            mcc = -(dg + 0.5 * dAd);
        }
        if (!ok || !(mcc > 0)) { // invalid step
            if (++n_invalid >= 5) { info->termination = 3; break; }
            radius /= decrease; decrease *= 2;
            continue;
        }
        n_invalid = 0;
        double d[NT], xc[NA];
        for (int a = 0; a < nt; a++) d[a] = y[a] * scale[a];
        lm_plus(L, x, d, xc);
        double cc = lm_evaluate(L, xc, true, An, gn);
        if (!std::isfinite(cc)) cc = kMax;
        // parameter / function tolerance (tested on the candidate, before acceptance)
        double sn = 0, xn = 0;
        for (int k = 0; k < NA; k++) { sn += (xc[k] - x[k]) * (xc[k] - x[k]); xn += x[k] * x[k]; }
        if (std::sqrt(sn) <= P.cfg.ceres_parameter_tolerance * (std::sqrt(xn) + P.cfg.ceres_parameter_tolerance)) { info->termination = 1; break; }
        if (std::fabs(cost - cc) <= P.cfg.ceres_function_tolerance * cost) { info->termination = 1; break; }
        double rel = cc >= kMax ? -kMax : std::max((cur_cost - cc) / mcc, (ref_cost - cc) / (acc_ref + mcc));
        if (rel > 1e-3) {
            std::memcpy(x, xc, sizeof(x)); cost = cc;
            std::memcpy(A, An, sizeof(double) * nt * nt); std::memcpy(g, gn, sizeof(double) * nt);
            radius = std::min(1e16, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * rel - 1.0, 3))); decrease = 2.0;
            cur_cost = cc; acc_cand += mcc; acc_ref += mcc;
            if (cur_cost < min_cost) { min_cost = cur_cost; n_nonmono = 0; cand_cost = cur_cost; acc_cand = 0; }
            else { ++n_nonmono; if (cur_cost > cand_cost) { cand_cost = cur_cost; acc_cand = 0; } }
            if (n_nonmono == max_nonmono) { ref_cost = cand_cost; acc_ref = acc_cand; }
            if (grad_max(x, g) <= P.cfg.ceres_gradient_tolerance) { info->termination = 1; break; }
        } else {
            radius /= decrease; decrease *= 2;
        }
    }
    info->cost = cost;
    write_back();
    return true;
}

} // namespace om
