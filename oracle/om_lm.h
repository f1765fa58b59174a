// ORACLE (test infrastructure only).  NonMinimalSolver / LeastSquares of the three
// estimators: the Levenberg-Marquardt problems of
//   /root/reference/src/optimizer.h:9-131 (calibrated), :247-369 (shared focal), :371-499 (two focal)
// with the residual functors of /root/reference/src/cost_functions.h:16-129,137-256,264-387,
// and the guards of hybrid_pose_estimator.cpp:189-214,264-289 (and the focal twins).
//
// Ceres (>= 2.0, system package, NOT under /root/reference) is absent: PARITY UNPINNED
// against a real Ceres.  The trust-region Levenberg-Marquardt loop below restates the
// published behaviour of ceres::TrustRegionMinimizer + LevenbergMarquardtStrategy +
// TrustRegionStepEvaluator with the options the reference sets:
// forward-mode autodiff (Jets) in the ambient parameters, quaternion manifold
// (Plus(q, d) = [cos|d|, sin|d| d/|d|] * q), Jacobi column scaling fixed at iteration 0,
// LM diagonal clamp [1e-6, 1e32], radius 1e4 / max 1e16, min_relative_decrease 1e-3,
// non-monotonic steps (5), bounds by projection in Plus, infeasible start = no-op.
// Two stated deviations (this file is the specification the CUDA kernel follows):
// the damped normal equations are solved by Cholesky instead of DENSE_QR, and the
// projected Armijo line search Ceres adds for bounded problems is not performed.
#pragma once
#include "om_estimator.h"

namespace om {

constexpr int NA = 12; // ambient parameters: o0, q(w,x,y,z), t(3), scale, o1, f0, f1

struct Jet {
    double a, v[NA];
    Jet() : a(0) { for (double &x : v) x = 0; }
    Jet(double s) : a(s) { for (double &x : v) x = 0; }
    Jet(double s, int k) : a(s) { for (double &x : v) x = 0; v[k] = 1; }
};
inline Jet operator+(const Jet &x, const Jet &y) { Jet r; r.a = x.a + y.a; for (int i = 0; i < NA; i++) r.v[i] = x.v[i] + y.v[i]; return r; }
inline Jet operator-(const Jet &x, const Jet &y) { Jet r; r.a = x.a - y.a; for (int i = 0; i < NA; i++) r.v[i] = x.v[i] - y.v[i]; return r; }
inline Jet operator-(const Jet &x) { Jet r; r.a = -x.a; for (int i = 0; i < NA; i++) r.v[i] = -x.v[i]; return r; }
inline Jet operator*(const Jet &x, const Jet &y) { Jet r; r.a = x.a * y.a; for (int i = 0; i < NA; i++) r.v[i] = x.a * y.v[i] + x.v[i] * y.a; return r; }
inline Jet operator/(const Jet &x, const Jet &y) {
    Jet r; double iy = 1.0 / y.a; r.a = x.a * iy;
    for (int i = 0; i < NA; i++) r.v[i] = (x.v[i] - r.a * y.v[i]) * iy;
    return r;
}
inline Jet sqrt(const Jet &x) { Jet r; r.a = std::sqrt(x.a); double h = 0.5 / r.a; for (int i = 0; i < NA; i++) r.v[i] = x.v[i] * h; return r; }
inline double sqrt(double x) { return std::sqrt(x); }
inline double val(double x) { return x; }
inline double val(const Jet &x) { return x.a; }

template <class T> inline void quat_normalize(const T *q, T *o) {
    T n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int k = 0; k < 4; k++) o[k] = q[k] / n;
}
// rotation by a unit quaternion (w,x,y,z): p + 2 w (u x p) + 2 u x (u x p)
template <class T> inline void quat_rotate_unit(const T *q, const T *p, T *o) {
    T c0 = q[2] * p[2] - q[3] * p[1], c1 = q[3] * p[0] - q[1] * p[2], c2 = q[1] * p[1] - q[2] * p[0];
    c0 = c0 + c0; c1 = c1 + c1; c2 = c2 + c2;
    o[0] = p[0] + q[0] * c0 + (q[2] * c2 - q[3] * c1);
    o[1] = p[1] + q[0] * c1 + (q[3] * c0 - q[1] * c2);
    o[2] = p[2] + q[0] * c2 + (q[1] * c1 - q[2] * c0);
}
template <class T> inline void quat_to_R(const T *qn, T *R) { // unit quaternion -> row-major R
    T w = qn[0], x = qn[1], y = qn[2], z = qn[3];
    T two(2.0), one(1.0);
    R[0] = one - two * (y * y + z * z); R[1] = two * (x * y - w * z); R[2] = two * (x * z + w * y);
    R[3] = two * (x * y + w * z); R[4] = one - two * (x * x + z * z); R[5] = two * (y * z - w * x);
    R[6] = two * (x * z - w * y); R[7] = two * (y * z + w * x); R[8] = one - two * (x * x + y * y);
}

// parameter pack: p[0]=o0, p[1..4]=q, p[5..7]=t, p[8]=scale, p[9]=o1, p[10]=f0, p[11]=f1
// (shared focal: f1 aliases f0, the caller passes the same Jet for both)
struct LmData {
    const Problem *P;
    double W; // sampson residual weight
};

// reprojection 0 -> 1 (cost_functions.h:27-43, 146-166, 273-293)
template <class T> inline void res_reproj0(const LmData &D, int i, const T *p, const T &f0, const T &f1, T *r) {
    const Problem &P = *D.P;
    double h[3] = {P.x0[2 * i], P.x0[2 * i + 1], 1.0}, a[3];
    T x3[3], y[3], qn[4];
    T s = T(P.d0[i]) + p[0];
    if (P.kind == CALIB) { mat3_vec(P.K0i, h, a); for (int k = 0; k < 3; k++) x3[k] = T(a[k]) * s; }
    else { x3[0] = T(h[0]) / f0 * s; x3[1] = T(h[1]) / f0 * s; x3[2] = s; }
    quat_normalize(p + 1, qn);
    quat_rotate_unit(qn, x3, y);
    for (int k = 0; k < 3; k++) y[k] = y[k] + p[5 + k];
    T u, v, w;
    if (P.kind == CALIB) {
        u = T(P.K1[0]) * y[0] + T(P.K1[1]) * y[1] + T(P.K1[2]) * y[2];
        v = T(P.K1[3]) * y[0] + T(P.K1[4]) * y[1] + T(P.K1[5]) * y[2];
        w = T(P.K1[6]) * y[0] + T(P.K1[7]) * y[1] + T(P.K1[8]) * y[2];
    } else { u = f1 * y[0]; v = f1 * y[1]; w = y[2]; }
    r[0] = u / w - T(P.x1[2 * i]);
    r[1] = v / w - T(P.x1[2 * i + 1]);
}
// reprojection 1 -> 0 (cost_functions.h:62-82, 182-205, 309-332)
template <class T> inline void res_reproj1(const LmData &D, int i, const T *p, const T &f0, const T &f1, T *r) {
    const Problem &P = *D.P;
    double h[3] = {P.x1[2 * i], P.x1[2 * i + 1], 1.0}, b[3];
    T x3[3], y[3], qn[4], qi[4];
    T s = (T(P.d1[i]) + p[9]) * p[8];
    if (P.kind == CALIB) { mat3_vec(P.K1i, h, b); for (int k = 0; k < 3; k++) x3[k] = T(b[k]) * s; }
    else { x3[0] = T(h[0]) / f1 * s; x3[1] = T(h[1]) / f1 * s; x3[2] = s; }
    quat_normalize(p + 1, qn);
    qi[0] = qn[0]; qi[1] = -qn[1]; qi[2] = -qn[2]; qi[3] = -qn[3];
    for (int k = 0; k < 3; k++) x3[k] = x3[k] - p[5 + k];
    quat_rotate_unit(qi, x3, y);
    T u, v, w;
    if (P.kind == CALIB) {
        u = T(P.K0[0]) * y[0] + T(P.K0[1]) * y[1] + T(P.K0[2]) * y[2];
        v = T(P.K0[3]) * y[0] + T(P.K0[4]) * y[1] + T(P.K0[5]) * y[2];
        w = T(P.K0[6]) * y[0] + T(P.K0[7]) * y[1] + T(P.K0[8]) * y[2];
    } else { u = f0 * y[0]; v = f0 * y[1]; w = y[2]; }
    r[0] = u / w - T(P.x0[2 * i]);
    r[1] = v / w - T(P.x0[2 * i + 1]);
}
// weighted Sampson residual (cost_functions.h:100-124, 223-251, 350-382)
template <class T> inline void res_sampson(const LmData &D, int i, const T *p, const T &f0, const T &f1, T *r) {
    const Problem &P = *D.P;
    T qn[4], R[9], E[9];
    quat_normalize(p + 1, qn);
    quat_to_R(qn, R);
    const T *t = p + 5;
    for (int j = 0; j < 3; j++) {
        E[j] = -t[2] * R[3 + j] + t[1] * R[6 + j];
        E[3 + j] = t[2] * R[j] - t[0] * R[6 + j];
        E[6 + j] = -t[1] * R[j] + t[0] * R[3 + j];
    }
    double h0[3] = {P.x0[2 * i], P.x0[2 * i + 1], 1.0}, h1[3] = {P.x1[2 * i], P.x1[2 * i + 1], 1.0}, a[3], b[3];
    if (P.kind == CALIB) { mat3_vec(P.K0i, h0, a); mat3_vec(P.K1i, h1, b); }
    else {
        std::memcpy(a, h0, 24); std::memcpy(b, h1, 24);
        // F = K1^-T E K0^-1 with K = diag(f, f, 1)
        for (int j = 0; j < 3; j++) { E[j] = E[j] / f1; E[3 + j] = E[3 + j] / f1; }
        for (int j = 0; j < 3; j++) { E[3 * j] = E[3 * j] / f0; E[3 * j + 1] = E[3 * j + 1] / f0; }
    }
    T e0 = E[0] * T(a[0]) + E[1] * T(a[1]) + E[2], e1 = E[3] * T(a[0]) + E[4] * T(a[1]) + E[5],
      e2 = E[6] * T(a[0]) + E[7] * T(a[1]) + E[8];
    T g0 = E[0] * T(b[0]) + E[3] * T(b[1]) + E[6], g1 = E[1] * T(b[0]) + E[4] * T(b[1]) + E[7];
    T C = T(b[0]) * e0 + T(b[1]) * e1 + e2;
    r[0] = C / sqrt(e0 * e0 + e1 * e1 + g0 * g0 + g1 * g1) * T(D.W);
}

} // namespace om
#include "om_lm2.h"
