// ORACLE (test infrastructure only).  Depth-aware minimal solvers: affine
// (scale+shift) depth correction for calibrated / shared-focal / two-focal
// cameras followed by Kabsch alignment.
// Restates /root/reference/src/solver.cpp:35-125 (calibrated), :127-293
// (shared focal), :295-480 (two focal) and the pose wrappers :482-534,
// :682-739, :986-1043.  The reference solves the same polynomial systems with
// generated elimination templates (12x12 / 36x36 / 40x40 QR + action-matrix
// eigenvalues); here the identical equations ("pairwise 3-D distances agree in
// both views") are reduced with b1 as hidden variable to
// one univariate polynomial (degree 4 / 8 / 4).  Pinned against the
// reference's own numpy solvers (solver_py/*.py) in tests/golden/.
#pragma once
#include "om_linalg.h"
#include "om_poly.h"

namespace om {

typedef std::vector<double> Poly;
inline Poly pmul(const Poly &a, const Poly &b) {
    Poly r(a.size() + b.size() - 1, 0.0);
    for (size_t i = 0; i < a.size(); i++)
        for (size_t j = 0; j < b.size(); j++) r[i + j] += a[i] * b[j];
    return r;
}
inline Poly padd(const Poly &a, const Poly &b, double sb = 1.0) {
    Poly r(std::max(a.size(), b.size()), 0.0);
    for (size_t i = 0; i < a.size(); i++) r[i] += a[i];
    for (size_t i = 0; i < b.size(); i++) r[i] += sb * b[i];
    return r;
}
inline Poly pscale(const Poly &a, double s) {
    Poly r(a);
    for (double &c : r) c *= s;
    return r;
}

// |(d_i+b) p_i - (d_j+b) p_j|^2 = A b^2 + B b + C over the first `dim` coords
inline void pair_quadratic(const double *pi, const double *pj, double di, double dj, int dim, double *ABC) {
    double A = 0, B = 0, C = 0;
    for (int k = 0; k < dim; k++) {
        double e = pi[k] - pj[k], g = di * pi[k] - dj * pj[k];
        A += e * e; B += 2 * e * g; C += g * g;
    }
    ABC[0] = A; ABC[1] = B; ABC[2] = C;
}

struct AffineSol { double b1, a2, b2, f1, f2; }; // depth1 + b1 ; a2*(depth2 + b2)

// x1,x2: 3 points as rows [3][3] (homogeneous, z=1); reference solver.cpp:35-125
inline std::vector<AffineSol> solve_scale_and_shift(const double x1[3][3], const double x2[3][3], const double *d1,
                                                    const double *d2) {
    static const int pr[3][2] = {{0, 1}, {0, 2}, {1, 2}};
    double P[9], Q[9], Qi[9], M[9];
    for (int r = 0; r < 3; r++) {
        pair_quadratic(x1[pr[r][0]], x1[pr[r][1]], d1[pr[r][0]], d1[pr[r][1]], 3, P + 3 * r);
        pair_quadratic(x2[pr[r][0]], x2[pr[r][1]], d2[pr[r][0]], d2[pr[r][1]], 3, Q + 3 * r);
    }
    std::vector<AffineSol> out;
    if (!mat3_inv(Q, Qi)) return out;
    mat3_mul(Qi, P, M); // [s b2^2, s b2, s] = M [b1^2, b1, 1]
    Poly k0 = {M[2], M[1], M[0]}, k1 = {M[5], M[4], M[3]}, k2 = {M[8], M[7], M[6]};
    Poly q = padd(pmul(k1, k1), pmul(k0, k2), -1.0);
    for (double b1 : poly_real_roots(q)) {
        double s = poly_eval(k2, b1), sb2 = poly_eval(k1, b1);
        if (!(s > 0) || !std::isfinite(s)) continue; // Q9: sqrt of non-positive a2^2 dropped
        out.push_back({b1, std::sqrt(s), sb2 / s, 1.0, 1.0});
    }
    return out;
}

// 4 points as rows [4][3]; shared focal; reference solver.cpp:127-293
inline std::vector<AffineSol> solve_scale_and_shift_shared_focal(const double x1in[4][3], const double x2in[4][3],
                                                                 const double *d1, const double *d2) {
    static const int pr[4][2] = {{0, 1}, {0, 2}, {1, 2}, {0, 3}};
    double x1[4][3], x2[4][3], m1 = 0, m2 = 0;
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 2; k++) { m1 += std::fabs(x1in[i][k]); m2 += std::fabs(x2in[i][k]); }
    double f0 = 0.5 * (m1 / 8 + m2 / 8); // solver.cpp:134-138
    for (int i = 0; i < 4; i++) {
        x1[i][0] = x1in[i][0] / f0; x1[i][1] = x1in[i][1] / f0; x1[i][2] = 1;
        x2[i][0] = x2in[i][0] / f0; x2[i][1] = x2in[i][1] / f0; x2[i][2] = 1;
    }
    std::vector<double> N(16), rhs(16); // N y = w p(b1) + D1 ; columns of rhs: [A1 B1 C1 D1]
    for (int r = 0; r < 4; r++) {
        int i = pr[r][0], j = pr[r][1];
        double q1[3], q2[3];
        pair_quadratic(x1[i], x1[j], d1[i], d1[j], 2, q1);
        pair_quadratic(x2[i], x2[j], d2[i], d2[j], 2, q2);
        N[4 * r + 0] = q2[0]; N[4 * r + 1] = q2[1]; N[4 * r + 2] = q2[2];
        N[4 * r + 3] = (d2[i] - d2[j]) * (d2[i] - d2[j]);
        rhs[4 * r + 0] = q1[0]; rhs[4 * r + 1] = q1[1]; rhs[4 * r + 2] = q1[2];
        rhs[4 * r + 3] = (d1[i] - d1[j]) * (d1[i] - d1[j]);
    }
    std::vector<AffineSol> out;
    if (!lu_solve(4, 4, N, rhs)) return out;
    // y_r = w * g_r(b1) + h_r, y = [v b2^2, v b2, v, sigma], v = sigma*w
    Poly g[4];
    double h[4];
    for (int r = 0; r < 4; r++) { g[r] = {rhs[4 * r + 2], rhs[4 * r + 1], rhs[4 * r + 0]}; h[r] = rhs[4 * r + 3]; }
    // (1) v = sigma*w:  g3 w^2 + (h3 - g2) w - h2 = 0
    Poly a = g[3], b = padd(Poly{h[3]}, g[2], -1.0), c = {-h[2]};
    // (2) y1^2 = y0*y2
    Poly a2 = padd(pmul(g[1], g[1]), pmul(g[0], g[2]), -1.0);
    Poly b2 = padd(padd(pscale(g[1], 2 * h[1]), pscale(g[0], h[2]), -1.0), pscale(g[2], h[0]), -1.0);
    Poly c2 = {h[1] * h[1] - h[0] * h[2]};
    Poly ac = padd(pmul(a, c2), pmul(a2, c), -1.0), ab = padd(pmul(a, b2), pmul(a2, b), -1.0),
         bc = padd(pmul(b, c2), pmul(b2, c), -1.0);
    Poly res = padd(pmul(ac, ac), pmul(ab, bc), -1.0);
    for (double b1 : poly_real_roots(res)) {
        double den = -poly_eval(ab, b1);
        if (den == 0) continue;
        double w = poly_eval(ac, b1) / den;
        // Newton polish of (b1, w) on the two quadratics-in-w (the expanded
        // degree-8 resultant is only a root *locator*; its coefficients lose digits)
        for (int it = 0; it < 6; it++) {
            double da, db, dc, da2, db2, dc2;
            double va = poly_eval_d(a, b1, &da), vb = poly_eval_d(b, b1, &db), vc = poly_eval_d(c, b1, &dc);
            double va2 = poly_eval_d(a2, b1, &da2), vb2 = poly_eval_d(b2, b1, &db2), vc2 = poly_eval_d(c2, b1, &dc2);
            double F1 = (va * w + vb) * w + vc, F2 = (va2 * w + vb2) * w + vc2;
            double J11 = (da * w + db) * w + dc, J12 = 2 * va * w + vb;
            double J21 = (da2 * w + db2) * w + dc2, J22 = 2 * va2 * w + vb2;
            double det = J11 * J22 - J12 * J21;
            if (det == 0 || !std::isfinite(det)) break;
            double s1 = (F1 * J22 - F2 * J12) / det, s2 = (J11 * F2 - J21 * F1) / det;
            if (!std::isfinite(s1) || !std::isfinite(s2)) break;
            b1 -= s1; w -= s2;
            if (std::fabs(s1) <= 1e-16 * std::fabs(b1) && std::fabs(s2) <= 1e-16 * std::fabs(w)) break;
        }
        double y1 = w * poly_eval(g[1], b1) + h[1], y2 = w * poly_eval(g[2], b1) + h[2];
        double s = w * poly_eval(g[3], b1) + h[3];
        if (w < 0 || !(s > 0) || !std::isfinite(w) || !std::isfinite(s)) continue;
        out.push_back({b1, std::sqrt(s), y1 / y2, f0 / std::sqrt(w), f0 / std::sqrt(w)});
    }
    return out;
}

// two independent focals; reference solver.cpp:295-480
inline std::vector<AffineSol> solve_scale_and_shift_two_focal(const double x1in[4][3], const double x2in[4][3],
                                                              const double *d1, const double *d2) {
    static const int pr[5][2] = {{0, 1}, {0, 2}, {1, 2}, {0, 3}, {1, 3}};
    double x1[4][3], x2[4][3], m1 = 0, m2 = 0;
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 2; k++) { m1 += std::fabs(x1in[i][k]); m2 += std::fabs(x2in[i][k]); }
    double f10 = m1 / 8, f20 = m2 / 8; // solver.cpp:302-305
    for (int i = 0; i < 4; i++) {
        x1[i][0] = x1in[i][0] / f10; x1[i][1] = x1in[i][1] / f10; x1[i][2] = 1;
        x2[i][0] = x2in[i][0] / f20; x2[i][1] = x2in[i][1] / f20; x2[i][2] = 1;
    }
    double N[5][4], R1[5][4];
    for (int r = 0; r < 5; r++) {
        int i = pr[r][0], j = pr[r][1];
        double q1[3], q2[3];
        pair_quadratic(x1[i], x1[j], d1[i], d1[j], 2, q1);
        pair_quadratic(x2[i], x2[j], d2[i], d2[j], 2, q2);
        N[r][0] = q2[0]; N[r][1] = q2[1]; N[r][2] = q2[2]; N[r][3] = (d2[i] - d2[j]) * (d2[i] - d2[j]);
        R1[r][0] = q1[0]; R1[r][1] = q1[1]; R1[r][2] = q1[2]; R1[r][3] = (d1[i] - d1[j]) * (d1[i] - d1[j]);
    }
    std::vector<AffineSol> out;
    // left null vector n of N (5x4): n^T N = 0  <=> null space of N^T (4x5)
    double Nt[20], nv[5];
    for (int r = 0; r < 5; r++)
        for (int c = 0; c < 4; c++) Nt[c * 5 + r] = N[r][c];
    nullspace(4, 5, Nt, nv);
    // w1 = -k / rho(b1)
    double k = 0;
    Poly rho(3, 0.0);
    for (int r = 0; r < 5; r++) { k += nv[r] * R1[r][3]; rho[2] += nv[r] * R1[r][0]; rho[1] += nv[r] * R1[r][1]; rho[0] += nv[r] * R1[r][2]; }
    // y = w1 g(b1) + h from the 4 best-conditioned rows: drop the row with largest |n_r|
    int drop = 0;
    for (int r = 1; r < 5; r++) if (std::fabs(nv[r]) > std::fabs(nv[drop])) drop = r;
    std::vector<double> Nm(16), rhs(16);
    for (int r = 0, rr = 0; r < 5; r++) {
        if (r == drop) continue;
        for (int c = 0; c < 4; c++) { Nm[4 * rr + c] = N[r][c]; rhs[4 * rr + c] = R1[r][c]; }
        rr++;
    }
    if (!lu_solve(4, 4, Nm, rhs)) return out;
    Poly g[4], e[4]; // e_r = -k g_r + h_r rho   (= rho * y_r)
    for (int r = 0; r < 4; r++) {
        g[r] = {rhs[4 * r + 2], rhs[4 * r + 1], rhs[4 * r + 0]};
        e[r] = padd(pscale(g[r], -k), pscale(rho, rhs[4 * r + 3]));
    }
    Poly quart = padd(pmul(e[1], e[1]), pmul(e[0], e[2]), -1.0);
    for (double b1 : poly_real_roots(quart)) {
        double rb = poly_eval(rho, b1);
        if (rb == 0) continue;
        double w1 = -k / rb;
        double y1 = poly_eval(e[1], b1) / rb, y2 = poly_eval(e[2], b1) / rb, s = poly_eval(e[3], b1) / rb;
        double w2 = y2 / s;
        if (w1 < 0 || w2 < 0 || !(s > 0) || !std::isfinite(w1) || !std::isfinite(w2) || !std::isfinite(s)) continue;
        out.push_back({b1, std::sqrt(s), y1 / y2, f10 / std::sqrt(w1), f20 / std::sqrt(w2)});
    }
    return out;
}

// Kabsch alignment Y ~ R X + t on n points (columns given as rows X[i][3]).
// reference solver.cpp:506-525 (and :704-724, :1008-1028)
inline void kabsch(int n, const double (*X)[3], const double (*Y)[3], double *R, double *t) {
    double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0};
    for (int i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) { cx[k] += X[i][k]; cy[k] += Y[i][k]; }
    for (int k = 0; k < 3; k++) { cx[k] /= n; cy[k] /= n; }
    double S[9] = {0};
    for (int i = 0; i < n; i++)
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) S[3 * a + b] += (Y[i][a] - cy[a]) * (X[i][b] - cx[b]);
    double U[9], Sg[3], V[9], Vt[9];
    jacobi_svd(3, S, U, Sg, V);
    if (mat3_det(U) * mat3_det(V) < 0) { U[2] = -U[2]; U[5] = -U[5]; U[8] = -U[8]; }
    mat3_T(V, Vt);
    mat3_mul(U, Vt, R);
    double Rc[3];
    mat3_vec(R, cx, Rc);
    for (int k = 0; k < 3; k++) t[k] = cy[k] - Rc[k];
}

// estimate_scale_and_pose (solver.cpp:5-33): weighted Umeyama.  Centroids and the covariance are weighted, the
// scale is the UNWEIGHTED ratio sum(Yc . R Xc) / sum(R Xc . R Xc) (solver.cpp:28-29), t = cY - scale R cX.
inline void umeyama(int n, const double (*X)[3], const double (*Y)[3], const double *W, double *R, double *t, double *scale) {
    double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0}, ws = 0;
    for (int i = 0; i < n; i++) {
        ws += W[i];
        for (int k = 0; k < 3; k++) { cx[k] += X[i][k] * W[i]; cy[k] += Y[i][k] * W[i]; }
    }
    for (int k = 0; k < 3; k++) { cx[k] /= ws; cy[k] /= ws; }
    double S[9] = {0};
    for (int i = 0; i < n; i++)
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) S[3 * a + b] += (Y[i][a] - cy[a]) * W[i] * (X[i][b] - cx[b]);
    double U[9], Sg[3], V[9], Vt[9];
    jacobi_svd(3, S, U, Sg, V);
    if (mat3_det(U) * mat3_det(V) < 0) { U[2] = -U[2]; U[5] = -U[5]; U[8] = -U[8]; }
    mat3_T(V, Vt);
    mat3_mul(U, Vt, R);
    double num = 0, den = 0;
    for (int i = 0; i < n; i++) {
        double xc[3] = {X[i][0] - cx[0], X[i][1] - cx[1], X[i][2] - cx[2]}, xr[3];
        mat3_vec(R, xc, xr);
        for (int k = 0; k < 3; k++) { num += (Y[i][k] - cy[k]) * xr[k]; den += xr[k] * xr[k]; }
    }
    *scale = num / den;
    double Rc[3];
    mat3_vec(R, cx, Rc);
    for (int k = 0; k < 3; k++) t[k] = cy[k] - *scale * Rc[k];
}

} // namespace om
