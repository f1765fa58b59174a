// ORACLE (test infrastructure only).  The alternative depth-aware minimal solvers behind
// HybridLORansacOptions::use_ours / use_4p4d (src/hybrid_ransac.h:23-24), selected at
// hybrid_pose_estimator.cpp:75-78, hybrid_pose_shared_focal_estimator.cpp:62-65, hybrid_pose_two_focal_estimator.cpp:86-93:
//   solve_scale_shift_pose_ours               solver.cpp:536-680  (calibrated; 3 points, scale + both shifts)
//   solve_scale_shift_pose_shared_focal_ours  solver.cpp:818-984  (3 points + the third depth in view 1; no shifts)
//   solve_scale_shift_pose_two_focal_ours     solver.cpp:1045-1153 (3 points; no shifts; linear)
//   solve_scale_shift_pose_two_focal_4p4d     solver.cpp:1287-1406 (4 points; F from depth ratios; no scale, no shifts)
// All four encode "3-D distances between back-projected points agree in both views" (or, 4p4d, F x1 = q (e x x2)).
// They are restated from that geometry, not from the reference's generated coefficient lists:
//   * calibrated: the quartic of om_affine.h (identical equations); what differs from the default solver is what the
//     reference does AROUND the roots: sigma = s^2 >= 0.01 (solver.cpp:609), positive corrected depths (:656-660), and
//     the pose from the two difference vectors and their cross product instead of Kabsch (:662-676).
//   * shared focal: with F = f^2, sigma = s^2, delta = unknown depth of point 2 in view 1's partner: pair (0,1) gives
//     sigma = G(F) / H(F); the difference of pairs (0,2) and (1,2) is linear in delta; substitution leaves one quartic
//     in F (the F^5 terms cancel identically) -- four solutions, like the reference's 4 x 4 action matrix; every root
//     is checked against the two remaining pair equations.  Filters of solver.cpp:936-950: delta > 0, F >= 0,
//     sigma >= 0.001.
//   * two focal: linear 3 x 3 system in (1/f1^2, s^2/f2^2, s^2), all three must be positive (:1113).
//   * 4p4d: 11 x 11 linear system (e_z = 1), Bougnoux focals (NaN -> no solution), E = K2 F K1, motion_from_essential
//     with all four points; scale 1, offsets 0 (:1398-1400).  Quirk kept: the cheirality test inside
//     motion_from_essential runs on the UN-calibrated unit vectors x / |x| (solver.cpp:1301-1302) against the calibrated
//     E, so a correct pose can be rejected (about one noise-free scene in five returns nothing).
// Known reference quirk (D): solver_p3p_mono_3d writes solution k to column k of a matrix that is then shrunk to the
// number of kept solutions (solver.cpp:605-620), so a rejected root before a kept one leaves an uninitialised column and
// drops a valid one.  That is undefined behaviour; the restatement keeps exactly the valid solutions.
// No reference test or numpy counterpart exists for these solvers (solver_py/ covers the default three): PARITY
// UNPINNED; tests/test_alt_solvers.py checks the defining equations and ground-truth recovery instead.
#pragma once
#include "om_affine.h"
#include "om_relpose.h"

namespace om {

// pose from three 3-D point pairs through the difference vectors and their cross product (solver.cpp:662-676):
// R = [v1 v2 v1 x v2] [u1 u2 u1 x u2]^-1, t = Y0 - R X0
inline bool pose_from_triplet(const double X[3][3], const double Y[3][3], double *R, double *t) {
    double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3];
    for (int k = 0; k < 3; k++) { u1[k] = X[0][k] - X[1][k]; u2[k] = X[0][k] - X[2][k]; v1[k] = Y[0][k] - Y[1][k]; v2[k] = Y[0][k] - Y[2][k]; }
    cross3(u1, u2, u3); cross3(v1, v2, v3);
    double Xm[9] = {u1[0], u2[0], u3[0], u1[1], u2[1], u3[1], u1[2], u2[2], u3[2]};
    double Ym[9] = {v1[0], v2[0], v3[0], v1[1], v2[1], v3[1], v1[2], v2[2], v3[2]};
    double Xi[9];
    if (!mat3_inv(Xm, Xi)) return false;
    mat3_mul(Ym, Xi, R);
    double RX0[3];
    mat3_vec(R, X[0], RX0);
    for (int k = 0; k < 3; k++) t[k] = Y[0][k] - RX0[k];
    return true;
}

// calibrated, use_ours.  x0, x1: 3 homogeneous points (K^-1 x).  Models: scale = s, o0 = u, o1 = s v (wrapper convention)
inline void solve_pose_ours_calibrated(const double x0[3][3], const double x1[3][3], const double *d0, const double *d1,
                                       std::vector<Model> *out) {
    double a[3][3], b[3][3];
    std::memcpy(a, x0, sizeof(a)); std::memcpy(b, x1, sizeof(b));
    for (const AffineSol &s : solve_scale_and_shift(a, b, d0, d1)) {
        const double sigma = s.a2 * s.a2;
        if (sigma < 0.01) continue;                                   // solver.cpp:609
        bool pos = true;
        for (int k = 0; k < 3; k++) if (d0[k] + s.b1 <= 0 || d1[k] + s.b2 <= 0) pos = false; // :656-660
        if (!pos) continue;
        double X[3][3], Y[3][3];
        for (int k = 0; k < 3; k++)
            for (int c = 0; c < 3; c++) { X[k][c] = (d0[k] + s.b1) * x0[k][c]; Y[k][c] = s.a2 * (d1[k] + s.b2) * x1[k][c]; }
        Model m;
        if (!pose_from_triplet(X, Y, m.R, m.t)) continue;
        m.scale = s.a2; m.o0 = s.b1; m.o1 = s.a2 * s.b2;
        out->push_back(m);
    }
}

// shared focal, use_ours.  x0, x1: >= 3 points (x, y, 1) in normalised pixels; the first three are used; d1[2] is not.
inline void solve_pose_ours_shared_focal(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1,
                                         std::vector<Model> *out) {
    auto dist_xy = [](const double *p, double dp, const double *q, double dq) {
        const double ex = dp * p[0] - dq * q[0], ey = dp * p[1] - dq * q[1];
        return ex * ex + ey * ey;
    };
    const double A01 = dist_xy(x0[0], d0[0], x0[1], d0[1]), B01 = (d0[0] - d0[1]) * (d0[0] - d0[1]);
    const double A02 = dist_xy(x0[0], d0[0], x0[2], d0[2]), B02 = (d0[0] - d0[2]) * (d0[0] - d0[2]);
    const double A12 = dist_xy(x0[1], d0[1], x0[2], d0[2]), B12 = (d0[1] - d0[2]) * (d0[1] - d0[2]);
    const double C01 = dist_xy(x1[0], d1[0], x1[1], d1[1]), D01 = (d1[0] - d1[1]) * (d1[0] - d1[1]);
    const double q = x1[2][0] * x1[2][0] + x1[2][1] * x1[2][1];
    const double m0 = x1[0][0] * x1[2][0] + x1[0][1] * x1[2][1], m1 = x1[1][0] * x1[2][0] + x1[1][1] * x1[2][1];
    const double n0 = x1[0][0] * x1[0][0] + x1[0][1] * x1[0][1], n1 = x1[1][0] * x1[1][0] + x1[1][1] * x1[1][1];
    const double e0 = d1[0], e1 = d1[1];
    // polynomials in F = f^2 (ascending)
    const Poly G = {A01, B01}, H = {C01, D01};
    const Poly L = {e0 * m0 - e1 * m1, e0 - e1};                       // d2_0 (m0 + F) - d2_1 (m1 + F)
    const Poly U = {e0 * e0 * n0 - e1 * e1 * n1, e0 * e0 - e1 * e1};   // d2_0^2 (n0 + F) - d2_1^2 (n1 + F)
    const Poly dAB = {A02 - A12, B02 - B12};
    const Poly N = padd(pmul(G, U), pmul(dAB, H), -1.0);               // delta = N / (2 G L)
    const Poly AB02 = {A02, B02}, qF = {q, 1.0}, m0F = {m0, 1.0}, n0F = {n0, 1.0};
    const Poly GL = pmul(G, L), L2 = pmul(L, L);
    // 4 (A02 + F B02) H G L^2 - (q + F) N^2 + 4 d2_0 (m0 + F) N G L - 4 d2_0^2 (n0 + F) G^2 L^2 = 0
    Poly P = pscale(pmul(pmul(AB02, H), pmul(G, L2)), 4.0);
    P = padd(P, pmul(qF, pmul(N, N)), -1.0);
    P = padd(P, pmul(m0F, pmul(N, GL)), 4.0 * e0);
    P = padd(P, pmul(n0F, pmul(GL, GL)), -4.0 * e0 * e0);
    P.resize(5); // the F^5 terms cancel identically (what is left there is rounding noise): a quartic, four solutions like
                 // the reference's 4 x 4 action matrix
    for (double F : poly_real_roots(P)) {
        if (!(F >= 0) || !std::isfinite(F)) continue;                  // A2(0) < 0 (solver.cpp:944)
        const double g = poly_eval(G, F), h = poly_eval(H, F), den = 2 * g * poly_eval(L, F);
        if (den == 0 || h == 0) continue;
        const double delta = poly_eval(N, F) / den, sigma = g / h;
        if (!(delta > 0) || !std::isfinite(delta)) continue;           // eigenvalue 1/d3 < 0 (:936)
        if (!(sigma >= 0.001)) continue;                               // :948
        // guard: the root must solve both remaining pair equations
        bool ok = true;
        const double Ai[2] = {A02, A12}, Bi[2] = {B02, B12}, ei[2] = {e0, e1}, mi[2] = {m0, m1}, ni[2] = {n0, n1};
        for (int i = 0; i < 2; i++) {
            const double lhs = Ai[i] + F * Bi[i];
            const double rhs = sigma * (delta * delta * (q + F) - 2 * delta * ei[i] * (mi[i] + F) + ei[i] * ei[i] * (ni[i] + F));
            if (!(std::fabs(lhs - rhs) <= 1e-7 * (std::fabs(lhs) + std::fabs(rhs)))) ok = false;
        }
        if (!ok) continue;
        const double f = std::sqrt(F), s = std::sqrt(sigma);
        double X[3][3], Y[3][3];
        const double dd[3] = {d1[0], d1[1], delta};
        for (int k = 0; k < 3; k++) {
            X[k][0] = d0[k] * x0[k][0] / f; X[k][1] = d0[k] * x0[k][1] / f; X[k][2] = d0[k];
            Y[k][0] = s * dd[k] * x1[k][0] / f; Y[k][1] = s * dd[k] * x1[k][1] / f; Y[k][2] = s * dd[k];
        }
        Model m;
        if (!pose_from_triplet(X, Y, m.R, m.t)) continue;
        m.scale = s; m.o0 = 0; m.o1 = 0; m.f0 = f; m.f1 = f;
        out->push_back(m);
    }
}

// two focal, use_ours: linear in (a, b, c) = (1/f1^2, s^2/f2^2, s^2)
inline void solve_pose_ours_two_focal(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1,
                                      std::vector<Model> *out) {
    static const int pr[3][2] = {{0, 1}, {0, 2}, {1, 2}};
    std::vector<double> A(9), B(3);
    for (int r = 0; r < 3; r++) {
        const int i = pr[r][0], j = pr[r][1];
        const double ax = d0[i] * x0[i][0] - d0[j] * x0[j][0], ay = d0[i] * x0[i][1] - d0[j] * x0[j][1], az = d0[i] - d0[j];
        const double bx = d1[i] * x1[i][0] - d1[j] * x1[j][0], by = d1[i] * x1[i][1] - d1[j] * x1[j][1], bz = d1[i] - d1[j];
        A[3 * r] = ax * ax + ay * ay; A[3 * r + 1] = -(bx * bx + by * by); A[3 * r + 2] = -bz * bz;
        B[r] = -az * az;
    }
    if (!lu_solve(3, 1, A, B)) return;
    if (!(B[0] > 0 && B[1] > 0 && B[2] > 0)) return;                   // solver.cpp:1113
    const double fi = std::sqrt(B[0]), s = std::sqrt(B[2]), wi = std::sqrt(B[1] / B[2]); // 1/f1, s, 1/f2
    double X[3][3], Y[3][3];
    for (int k = 0; k < 3; k++) {
        X[k][0] = d0[k] * x0[k][0] * fi; X[k][1] = d0[k] * x0[k][1] * fi; X[k][2] = d0[k];
        Y[k][0] = s * d1[k] * x1[k][0] * wi; Y[k][1] = s * d1[k] * x1[k][1] * wi; Y[k][2] = s * d1[k];
    }
    Model m;
    if (!pose_from_triplet(X, Y, m.R, m.t)) return;
    m.scale = s; m.o0 = 0; m.o1 = 0; m.f0 = 1.0 / fi; m.f1 = 1.0 / wi;
    out->push_back(m);
}

// two focal, use_4p4d: F x0_i = q_i (e x x1_i), q_i = d1_i / d0_i, e_z = 1; 11 of the 12 equations (the last one dropped)
inline void solve_pose_4p4d_two_focal(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1,
                                      std::vector<Model> *out) {
    std::vector<double> A(121, 0.0), B(11, 0.0);
    int row = 0;
    for (int i = 0; i < 4; i++) {
        const double u = x0[i][0] / x0[i][2], v = x0[i][1] / x0[i][2], u2 = x1[i][0] / x1[i][2], v2 = x1[i][1] / x1[i][2];
        const double q = d1[i] / d0[i];
        // unknowns: F row-major (0..8), e_x (9), e_y (10); e_z = 1 moves to the right-hand side
        // row 0 of F x0 - q (e x x1) = 0: F0 . x0 - q (e_y - v2 e_z) = 0        (signs as solver.cpp:1324-1337)
        A[11 * row + 0] = -u; A[11 * row + 1] = -v; A[11 * row + 2] = -1; A[11 * row + 10] = q; B[row] = q * v2; row++;
        A[11 * row + 3] = -u; A[11 * row + 4] = -v; A[11 * row + 5] = -1; A[11 * row + 9] = -q; B[row] = -q * u2; row++;
        if (i == 3) break;
        A[11 * row + 6] = -u; A[11 * row + 7] = -v; A[11 * row + 8] = -1; A[11 * row + 9] = q * v2; A[11 * row + 10] = -q * u2; B[row] = 0; row++;
    }
    if (!lu_solve(11, 1, A, B)) return;
    double F[9];
    for (int k = 0; k < 9; k++) F[k] = B[k];
    double f0sq, f1sq;
    bougnoux_focals(F, &f0sq, &f1sq);                                  // focals_from_fundamental with pp = 0 (:1155-1178)
    const double f0 = std::sqrt(f0sq), f1 = std::sqrt(f1sq);
    if (std::isnan(f0) || std::isnan(f1)) return;                      // :1376-1379
    double E[9];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) E[3 * i + j] = F[3 * i + j] * (i < 2 ? f1 : 1.0) * (j < 2 ? f0 : 1.0);
    double b0[4][3], b1[4][3];
    for (int i = 0; i < 4; i++) {
        const double n0 = norm3(x0[i]), n1 = norm3(x1[i]);
        for (int k = 0; k < 3; k++) { b0[i][k] = x0[i][k] / n0; b1[i][k] = x1[i][k] / n1; }
    }
    std::vector<Pose> poses;
    motion_from_essential(E, 4, b0, b1, &poses);
    for (const Pose &p : poses) {
        Model m;
        std::memcpy(m.R, p.R, 72); std::memcpy(m.t, p.t, 24);
        m.scale = 1.0; m.o0 = 0; m.o1 = 0; m.f0 = f0; m.f1 = f1;
        out->push_back(m);
    }
}

} // namespace om
