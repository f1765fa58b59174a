// K5 (part 2) -- the trust-region Levenberg-Marquardt driver of NonMinimalSolver /
// LeastSquares (hybrid_pose_estimator.cpp:189-214,264-289 and the focal twins ->
// optimizer.h:9-131, 247-369, 371-499).  Ceres is not available; the specification is
// Jacobi scaling fixed at iteration 0, LM diagonal clamp [1e-6, 1e32],
// radius 1e4 (max 1e16), min_relative_decrease 1e-3, non-monotonic steps (5), bounds by
// projection inside Plus, infeasible start = no-op; damped normal equations by Cholesky.
// The driver is scalar code every thread of the CTA executes redundantly on identical
// data; only `ev` (the evaluation of cost / J^T J / J^T r over the residual blocks) is a
// collective that spreads the blocks over the threads.
#pragma once

namespace mp {

struct LmInfo { int iterations, termination; double cost0, cost; };
// termination: 0 not run (guard), 1 convergence, 2 max iterations, 3 failure, 4 infeasible start

struct LmSetup {
    int kind, nt;
    int slot[LM_NT];       // compact index of each tangent slot or -1
    bool has_lower[LM_NA];
    double lower[LM_NA];
    double W;              // weight of the Sampson residual
    bool use_reproj, use_samp;
    double ftol, gtol, ptol;
    int max_it, max_nonmono;
};

// Guard + parameter activation + bounds.  n0,n1,n2 = sizes of the three index lists.
// Returns false when the estimator's guard rejects the sample (model left untouched).
// nonminimal: the call is NonMinimalSolver (true) or LeastSquares (false).  EstimatorConfig::use_shift reaches the
// optimiser only at hybrid_pose_estimator.cpp:206,281 (calibrated, both calls) and
// hybrid_pose_shared_focal_estimator.cpp:149 (shared focal, NonMinimalSolver only); the scale-only optimiser
// (optimizer.h:133-245) always holds the offsets constant and sets no offset bounds.
MP_HD bool lm_setup(const PairConst &pc, const mp_config &cfg, int n0, int n1, int n2, LmSetup *L, bool nonminimal) {
    const int k0 = pc.kind == CALIB ? 3 : 4, k2 = pc.kind == CALIB ? 5 : (pc.kind == SHARED ? 6 : 7);
    if ((n0 < k0 && n1 < k0) || n2 < k2) return false;
    L->kind = pc.kind;
    L->use_reproj = cfg.lo_type != 1; L->use_samp = cfg.lo_type != 2;
    const double sw = sqrt(pc.sampson_w);
    L->W = pc.kind == CALIB ? sw / (1.0 / (pc.K0[0] + pc.K0[4]) + 1.0 / (pc.K1[0] + pc.K1[4])) : sw;
    const bool b0 = L->use_reproj && n0 > 0, b1 = L->use_reproj && n1 > 0, b2 = L->use_samp && n2 > 0;
    if (!b0 && !b1 && !b2) return false;
    bool act[LM_NT];
    const bool shift = pc.scale_only ? false
                     : (pc.kind == CALIB ? cfg.use_shift != 0 : (pc.kind == SHARED ? (nonminimal ? cfg.use_shift != 0 : true) : true));
    act[0] = b0 && shift; act[1] = act[2] = act[3] = act[4] = act[5] = act[6] = true;
    act[7] = b1; act[8] = b1 && shift; act[9] = pc.kind != CALIB; act[10] = pc.kind == TWO;
    L->nt = 0;
    for (int s = 0; s < LM_NT; s++) L->slot[s] = act[s] ? L->nt++ : -1;
    for (int k = 0; k < LM_NA; k++) { L->has_lower[k] = false; L->lower[k] = 0; }
    if (b1) { L->has_lower[8] = true; L->lower[8] = 1e-2; }                                   // optimizer.h:79-81
    if (!pc.scale_only && cfg.min_depth_constraint && b0) { L->has_lower[0] = true; L->lower[0] = -pc.min_depth[0] + 1e-2; }
    if (!pc.scale_only && cfg.min_depth_constraint && b1) { L->has_lower[9] = true; L->lower[9] = -pc.min_depth[1] + 1e-2; }
    if (pc.kind == TWO) { L->has_lower[10] = L->has_lower[11] = true; L->lower[10] = L->lower[11] = 1e-6; }
    L->ftol = cfg.ceres_function_tolerance; L->gtol = cfg.ceres_gradient_tolerance; L->ptol = cfg.ceres_parameter_tolerance;
    L->max_it = (int)cfg.ceres_max_num_iterations;
    L->max_nonmono = cfg.ceres_use_nonmonotonic_steps ? 5 : 0;
    return true;
}

MP_HD void lm_model_to_x(const Model &m, double *x) {
    x[0] = m.o0; rot_to_quat(m.R, x + 1);
    x[5] = m.t[0]; x[6] = m.t[1]; x[7] = m.t[2];
    x[8] = m.scale; x[9] = m.o1; x[10] = m.f0; x[11] = m.f1;
}
MP_HD void lm_x_to_model(int kind, const double *x, Model *m) {
    quat_to_rot(x + 1, m->R);
    m->o0 = x[0]; m->t[0] = x[5]; m->t[1] = x[6]; m->t[2] = x[7]; m->scale = x[8]; m->o1 = x[9];
    m->f0 = x[10];
    if (kind == TWO) m->f1 = x[11];
    else if (kind == SHARED) m->f1 = x[10];
}

// x_plus = Plus(x, delta), then projection onto the bounds
MP_HD void lm_plus(const LmSetup &L, const double *x, const double *d, double *xp) {
    for (int k = 0; k < LM_NA; k++) xp[k] = x[k];
    double dv[LM_NT];
    for (int s = 0; s < LM_NT; s++) dv[s] = L.slot[s] >= 0 ? d[L.slot[s]] : 0.0;
    xp[0] += dv[0]; xp[5] += dv[4]; xp[6] += dv[5]; xp[7] += dv[6]; xp[8] += dv[7]; xp[9] += dv[8];
    xp[10] += dv[9]; xp[11] += dv[10];
    const double n = sqrt(dv[1] * dv[1] + dv[2] * dv[2] + dv[3] * dv[3]);
    if (n > 0) {
#ifdef __CUDA_ARCH__
        double sv, e0;
        sincos(n, &sv, &e0); // one argument reduction for both
        const double sn = sv / n, e1 = sn * dv[1], e2 = sn * dv[2], e3 = sn * dv[3];
#else
        const double sn = sin(n) / n, e0 = cos(n), e1 = sn * dv[1], e2 = sn * dv[2], e3 = sn * dv[3];
#endif
        const double *q = x + 1;
        xp[1] = e0 * q[0] - e1 * q[1] - e2 * q[2] - e3 * q[3];
        xp[2] = e0 * q[1] + e1 * q[0] + e2 * q[3] - e3 * q[2];
        xp[3] = e0 * q[2] - e1 * q[3] + e2 * q[0] + e3 * q[1];
        xp[4] = e0 * q[3] + e1 * q[2] - e2 * q[1] + e3 * q[0];
    }
    for (int k = 0; k < LM_NA; k++)
        if (L.has_lower[k] && xp[k] < L.lower[k]) xp[k] = L.lower[k];
}

// reduced full-slot sums -> compact nt x nt system
MP_HD double lm_compress(const LmSetup &L, const LmAcc &acc, double *A, double *g) {
    int k = 0;
    for (int a = 0; a < LM_NT; a++)
        for (int b = a; b < LM_NT; b++, k++) {
            const int ia = L.slot[a], ib = L.slot[b];
            if (ia >= 0 && ib >= 0) { A[ia * L.nt + ib] = acc.A[k]; A[ib * L.nt + ia] = acc.A[k]; }
        }
    for (int a = 0; a < LM_NT; a++)
        if (L.slot[a] >= 0) g[L.slot[a]] = acc.g[a];
    return acc.cost;
}

// The LM iteration as a state machine, so that ONE thread of a CTA can drive it while the whole
// CTA evaluates (cost, J^T J, J^T r) at the points it asks for:
//   begin(x0) -> evaluate at x -> first(acc) -> { propose() -> evaluate at xc -> feed(acc) }*
struct LmDriver {
    const LmSetup *L;
    LmInfo info;
    double x[LM_NA], xc[LM_NA];
    double A[LM_NT * LM_NT], g[LM_NT], As[LM_NT * LM_NT], gs[LM_NT], scale[LM_NT];
    double Ad[LM_NT * LM_NT], Lc[LM_NT * LM_NT], An[LM_NT * LM_NT], gn[LM_NT], y[LM_NT], yw[LM_NT]; // workspace (kept
    // in the object so that a driver placed in shared memory does all its array work there)
    double wr[LM_NT];      // warp driver (mp_lm_warp.cuh): reciprocals of the L D L^T pivots
    short cpos[LM_NT * (LM_NT + 1) / 2], cposT[LM_NT * (LM_NT + 1) / 2]; // warp driver: slot -> compact position
    double cost, radius, decrease, mcc;
    double min_cost, cur_cost, ref_cost, cand_cost, acc_ref, acc_cand;
    int n_nonmono, n_invalid, it;
    unsigned long long prof[8]; // device profiling aid (mp_lm_warp.cuh): cycles per driver section

    MP_HD double grad_max(const double *xx, const double *gg) const {
        double d[LM_NT], xp[LM_NA], mx = 0;
        for (int a = 0; a < L->nt; a++) d[a] = -gg[a];
        lm_plus(*L, xx, d, xp);
        for (int k = 0; k < LM_NA; k++) mx = fmax(mx, fabs(xp[k] - xx[k]));
        return mx;
    }
    // false: infeasible start, nothing to do (termination 4)
    MP_HD bool begin(const LmSetup *Lp, const double *x0) {
        L = Lp;
        info.iterations = 0; info.termination = 0; info.cost0 = 0; info.cost = 0;
        for (int k = 0; k < LM_NA; k++) x[k] = x0[k];
        for (int k = 0; k < LM_NA; k++)
            if (L->has_lower[k] && x[k] < L->lower[k]) { info.termination = 4; return false; }
        return true;
    }
    // evaluation at the start point; false: finished
    MP_HD bool first(const LmAcc &acc) {
        const int nt = L->nt;
        cost = lm_compress(*L, acc, A, g);
        info.cost0 = cost; info.cost = cost;
        for (int a = 0; a < nt; a++) scale[a] = 1.0 / (1.0 + sqrt(A[a * nt + a]));
        if (!finite_d(cost)) { info.termination = 3; return false; }
        if (grad_max(x, g) <= L->gtol) { info.termination = 1; return false; }
        radius = 1e4; decrease = 2.0;
        min_cost = cur_cost = ref_cost = cand_cost = cost; acc_ref = 0; acc_cand = 0;
        n_nonmono = 0; n_invalid = 0; it = 0;
        info.termination = 2;
        return true;
    }
    // next candidate xc; false: finished (max iterations / radius / repeated invalid steps)
    MP_HD bool propose() {
        const int nt = L->nt;
        while (true) {
            if (++it > L->max_it) { info.cost = cost; return false; }
            info.iterations = it;
            if (radius < 1e-32) { info.termination = 1; info.cost = cost; return false; }
            for (int a = 0; a < nt; a++) {
                gs[a] = g[a] * scale[a];
                for (int b = 0; b < nt; b++) { As[a * nt + b] = A[a * nt + b] * scale[a] * scale[b]; Ad[a * nt + b] = As[a * nt + b]; }
            }
            for (int a = 0; a < nt; a++) Ad[a * nt + a] += fmin(fmax(As[a * nt + a], 1e-6), 1e32) / radius;
            const bool ok = cholesky_solve_ws(nt, Ad, gs, y, Lc, yw);
            mcc = 0;
            if (ok) {
                for (int a = 0; a < nt; a++) y[a] = -y[a];
                double dg = 0, dAd = 0;
                for (int a = 0; a < nt; a++) {
                    dg += y[a] * gs[a];
                    for (int b = 0; b < nt; b++) dAd += y[a] * As[a * nt + b] * y[b];
                }
                mcc = -(dg + 0.5 * dAd);
            }
            if (!ok || !(mcc > 0)) {
                if (++n_invalid >= 5) { info.termination = 3; info.cost = cost; return false; }
                radius /= decrease; decrease *= 2;
                continue;
            }
            n_invalid = 0;
            double d[LM_NT];
            for (int a = 0; a < nt; a++) d[a] = y[a] * scale[a];
            lm_plus(*L, x, d, xc);
            return true;
        }
    }
    // evaluation at xc; false: converged
    MP_HD bool feed(const LmAcc &acc) {
        const int nt = L->nt;
        double cc = lm_compress(*L, acc, An, gn);
        if (!finite_d(cc)) cc = DBL_MAX;
        double sn = 0, xn = 0;
        for (int k = 0; k < LM_NA; k++) { sn += (xc[k] - x[k]) * (xc[k] - x[k]); xn += x[k] * x[k]; }
        if (sqrt(sn) <= L->ptol * (sqrt(xn) + L->ptol)) { info.termination = 1; info.cost = cost; return false; }
        if (fabs(cost - cc) <= L->ftol * cost) { info.termination = 1; info.cost = cost; return false; }
        const double rel = cc >= DBL_MAX ? -DBL_MAX : fmax((cur_cost - cc) / mcc, (ref_cost - cc) / (acc_ref + mcc));
        if (rel > 1e-3) {
            for (int k = 0; k < LM_NA; k++) x[k] = xc[k];
            cost = cc;
            for (int a = 0; a < nt * nt; a++) A[a] = An[a];
            for (int a = 0; a < nt; a++) g[a] = gn[a];
            const double tt = 2.0 * rel - 1.0;
            radius = fmin(1e16, radius / fmax(1.0 / 3.0, 1.0 - tt * tt * tt));
            decrease = 2.0;
            cur_cost = cc; acc_cand += mcc; acc_ref += mcc;
            if (cur_cost < min_cost) { min_cost = cur_cost; n_nonmono = 0; cand_cost = cur_cost; acc_cand = 0; }
            else { ++n_nonmono; if (cur_cost > cand_cost) { cand_cost = cur_cost; acc_cand = 0; } }
            if (n_nonmono == L->max_nonmono) { ref_cost = cand_cost; acc_ref = acc_cand; }
            if (grad_max(x, g) <= L->gtol) { info.termination = 1; info.cost = cost; return false; }
        } else {
            radius /= decrease; decrease *= 2;
        }
        return true;
    }
};

// single-thread form (host check): Ev: void operator()(const double *x, LmAcc *sums)
template <class Ev> MP_HD void lm_minimize(const LmSetup &L, Ev &ev, double *x, LmInfo *info) {
    LmDriver drv;
    LmAcc acc;
    if (drv.begin(&L, x)) {
        ev(drv.x, &acc);
        if (drv.first(acc))
            while (drv.propose()) {
                ev(drv.xc, &acc);
                if (!drv.feed(acc)) break;
            }
    }
    for (int k = 0; k < LM_NA; k++) x[k] = drv.x[k];
    *info = drv.info;
}

} // namespace mp
