// Residuals of the three hybrid estimators: exact fp64 evaluation (inlier tests,
// LM, re-scoring of candidate bests) and the fp32 form used by the fused scoring
// kernel.  Semantics: SURVEY.md Appendix A.2, i.e. EvaluateModelOnPoint of
// /root/reference/src/hybrid_pose_estimator.cpp:216-261,
// ...shared_focal_estimator.cpp:160-202, ...two_focal_estimator.cpp:213-257.
#pragma once
#include "mp_common.cuh"

namespace mp {

// ---------------------------------------------------------------- fp64 (exact)
// Everything that depends on the model only, hoisted out of the per-match loop
// (the reference rebuilds K, K^-1 and E on every call).
struct Eval64 {
    double M0[9], t[3];  // q0 = (d0+o0) * M0*[x0;1] + t,  M0 = R*K0^-1
    double M1[9], v1[3]; // q1 = scale*(d1+o1) * M1*[x1;1] + v1,  M1 = R^T*K1^-1, v1 = -R^T t
    double K0[9], K1[9]; // projection p = K q
    double K0i[9], K1i[9];
    double R[9], F[9];   // F = K1^-T [t]x R K0^-1 (focal kinds); E for calibrated is t x (R a)
    double o0, o1, scale, loss_scale;
    int kind, depth_guard;
};

MP_HD void make_eval64(const PairConst &pc, const Model &m, Eval64 *e) {
    e->kind = pc.kind; e->depth_guard = pc.scale_only;
    if (pc.kind == CALIB) {
        for (int i = 0; i < 9; i++) { e->K0[i] = pc.K0[i]; e->K1[i] = pc.K1[i]; e->K0i[i] = pc.K0i[i]; e->K1i[i] = pc.K1i[i]; }
    } else {
        double f0 = m.f0, f1 = pc.kind == SHARED ? m.f0 : m.f1;
        for (int i = 0; i < 9; i++) { e->K0[i] = 0; e->K1[i] = 0; e->K0i[i] = 0; e->K1i[i] = 0; }
        e->K0[0] = f0; e->K0[4] = f0; e->K0[8] = 1; e->K1[0] = f1; e->K1[4] = f1; e->K1[8] = 1;
        e->K0i[0] = 1.0 / f0; e->K0i[4] = 1.0 / f0; e->K0i[8] = 1; e->K1i[0] = 1.0 / f1; e->K1i[4] = 1.0 / f1; e->K1i[8] = 1;
    }
    for (int i = 0; i < 9; i++) e->R[i] = m.R[i];
    for (int i = 0; i < 3; i++) e->t[i] = m.t[i];
    mat3_mul(m.R, e->K0i, e->M0);
    double Rt[9] = {m.R[0], m.R[3], m.R[6], m.R[1], m.R[4], m.R[7], m.R[2], m.R[5], m.R[8]};
    mat3_mul(Rt, e->K1i, e->M1);
    mat3_vec(Rt, -m.t[0], -m.t[1], -m.t[2], e->v1);
    e->o0 = m.o0; e->o1 = m.o1; e->scale = m.scale; e->loss_scale = pc.loss_scale;
    // F = K1i^T * ([t]x R) * K0i
    double Tx[9] = {0, -m.t[2], m.t[1], m.t[2], 0, -m.t[0], -m.t[1], m.t[0], 0}, E[9], K1iT[9], T[9];
    mat3_mul(Tx, m.R, E);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) K1iT[3 * i + j] = e->K1i[3 * j + i];
    mat3_mul(K1iT, E, T);
    mat3_mul(T, e->K0i, e->F);
}

// squared residual of data type t (0,1,2) on one match; DBL_MAX is the reference's
// "invalid" sentinel (behind the camera / cheirality).  No score-type gating here.
MP_HD double eval64(const Eval64 &e, const Match64 &m, int t) {
    if (t == 0) {
        double u[3], p[3];
        mat3_vec(e.M0, m.x0x, m.x0y, 1.0, u);
        double s = m.d0 + e.o0;
        mat3_vec(e.K1, s * u[0] + e.t[0], s * u[1] + e.t[1], s * u[2] + e.t[2], p);
        if (p[2] < 1e-2 || (e.depth_guard && m.d0 < 1e-2)) return DBL_MAX;
        double a = p[0] / p[2] - m.x1x, b = p[1] / p[2] - m.x1y;
        return a * a + b * b;
    }
    if (t == 1) {
        double w[3], p[3];
        mat3_vec(e.M1, m.x1x, m.x1y, 1.0, w);
        double s = (m.d1 + e.o1) * e.scale;
        mat3_vec(e.K0, s * w[0] + e.v1[0], s * w[1] + e.v1[1], s * w[2] + e.v1[2], p);
        if (p[2] < 1e-2 || (e.depth_guard && m.d1 < 1e-2)) return DBL_MAX;
        double a = p[0] / p[2] - m.x0x, b = p[1] / p[2] - m.x0y;
        return a * a + b * b;
    }
    if (e.kind == CALIB) {
        double a[3], b[3], u[3];
        mat3_vec(e.K0i, m.x0x, m.x0y, 1.0, a);
        mat3_vec(e.K1i, m.x1x, m.x1y, 1.0, b);
        double na = 1.0 / sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
        double nb = 1.0 / sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
        mat3_vec(e.R, a[0] * na, a[1] * na, a[2] * na, u); // R * unit bearing
        double bx = b[0] * nb, by = b[1] * nb, bz = b[2] * nb;
        double al = -(u[0] * bx + u[1] * by + u[2] * bz);
        double b1 = -(u[0] * e.t[0] + u[1] * e.t[1] + u[2] * e.t[2]);
        double b2 = bx * e.t[0] + by * e.t[1] + bz * e.t[2];
        double l1 = b1 - al * b2, l2 = -al * b1 + b2, md = 1e-2 * (1 - al * al);
        if (!(l1 > md && l2 > md)) return DBL_MAX;
        // E a = t x (R a) with the un-normalised a (z = 1), E^T b = v1 x (R^T b)
        double Ra[3], Rtb[3];
        mat3_vec(e.R, a[0], a[1], 1.0, Ra);
        mat3T_vec(e.R, b[0], b[1], 1.0, Rtb);
        double c0 = e.t[1] * Ra[2] - e.t[2] * Ra[1], c1 = e.t[2] * Ra[0] - e.t[0] * Ra[2],
               c2 = e.t[0] * Ra[1] - e.t[1] * Ra[0];
        double g0 = e.v1[1] * Rtb[2] - e.v1[2] * Rtb[1], g1 = e.v1[2] * Rtb[0] - e.v1[0] * Rtb[2];
        double C = b[0] * c0 + b[1] * c1 + c2;
        return C * C / (c0 * c0 + c1 * c1 + g0 * g0 + g1 * g1) * e.loss_scale;
    }
    const double *F = e.F;
    double e0 = F[0] * m.x0x + F[1] * m.x0y + F[2], e1 = F[3] * m.x0x + F[4] * m.x0y + F[5],
           e2 = F[6] * m.x0x + F[7] * m.x0y + F[8];
    double g0 = F[0] * m.x1x + F[3] * m.x1y + F[6], g1 = F[1] * m.x1x + F[4] * m.x1y + F[7];
    double C = m.x1x * e0 + m.x1y * e1 + e2;
    return C * C / (e0 * e0 + e1 * e1 + g0 * g0 + g1 * g1);
}

// MSAC term with the reference's std::min semantics (NaN propagates like std::min(err, thr))
MP_HD double msac64(double err, double thr) { return (thr < err) ? thr : err; }

// ---------------------------------------------------------------- fp32 (scoring)
// Per-pair constants of the fused scoring kernel.
struct PairF {
    float k0fx, k0sk, k0fy, k1fx, k1sk, k1fy; // 2x2 upper-triangular parts of K0, K1 (calibrated)
    float thr0, thr1, thr2, loss;
    int n, gate; // gate: 0 none, 1 EPI_ONLY (types 0,1 masked), 2 MD_ONLY (type 2 masked)
};

// Hypothesis parameters, fp32, 32 floats (8 x float4) so that a warp can fetch them
// with broadcast LDS.128.  Calibrated uses R,t,v1,o0,s,so1; the focal kinds carry the
// focal-folded rotations.
struct alignas(16) HypF {
    float A[9];  // u = A*[x0;1]:  R (calibrated, on normalised a) or R*diag(1/f0,1/f0,1)
    float t[3];
    float B[9];  // w = B*[x1;1]:  R^T (calibrated, on normalised b) or R^T*diag(1/f1,1/f1,1)
    float v1[3]; // -R^T t
    float o0, s, so1, f0, f1, if0, if1, pad;
};
static_assert(sizeof(HypF) == 128, "HypF is 8 float4");

MP_HD void pack_hyp(int kind, const Model &m, HypF *h) {
    double f0 = kind == CALIB ? 1.0 : m.f0, f1 = kind == CALIB ? 1.0 : (kind == SHARED ? m.f0 : m.f1);
    double if0 = 1.0 / f0, if1 = 1.0 / f1;
    for (int r = 0; r < 3; r++) {
        h->A[3 * r + 0] = (float)(m.R[3 * r + 0] * if0); h->A[3 * r + 1] = (float)(m.R[3 * r + 1] * if0);
        h->A[3 * r + 2] = (float)m.R[3 * r + 2];
        h->B[3 * r + 0] = (float)(m.R[0 + r] * if1); h->B[3 * r + 1] = (float)(m.R[3 + r] * if1);
        h->B[3 * r + 2] = (float)m.R[6 + r];
        h->t[r] = (float)m.t[r];
        h->v1[r] = (float)(-(m.R[0 + r] * m.t[0] + m.R[3 + r] * m.t[1] + m.R[6 + r] * m.t[2]));
    }
    h->o0 = (float)m.o0; h->s = (float)m.scale; h->so1 = (float)(m.scale * m.o1);
    h->f0 = (float)f0; h->f1 = (float)f1; h->if0 = (float)if0; h->if1 = (float)if1; h->pad = 0.f;
}

} // namespace mp
