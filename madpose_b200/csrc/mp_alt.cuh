// K1, alternative depth-aware minimal solvers behind HybridLORansacOptions::use_ours / use_4p4d
// (src/hybrid_ransac.h:23-24; selected at hybrid_pose_estimator.cpp:75-78, hybrid_pose_shared_focal_estimator.cpp:62-65,
// hybrid_pose_two_focal_estimator.cpp:86-93).  They replace
//   solve_scale_shift_pose_ours               solver.cpp:536-680   calibrated, 3 points, scale + both shifts
//   solve_scale_shift_pose_shared_focal_ours  solver.cpp:818-984   3 points + the first two depths of view 1, no shifts
//   solve_scale_shift_pose_two_focal_ours     solver.cpp:1045-1153 3 points, no shifts, linear
//   solve_scale_shift_pose_two_focal_4p4d     solver.cpp:1287-1406 4 points, F from the depth ratios, no scale / shifts
// and are written from the geometry ("3-D distances between back-projected points agree in both views"; 4p4d:
// F x0 = q (e x x1)), not from the reference's generated coefficient lists: the calibrated one reuses the quartic of
// mp_affine.cuh, the shared-focal one eliminates to ONE quartic in f^2 (derivation in oracle/om_alt.h), the two-focal
// ones are linear systems.  Models carry the wrappers' convention (scale, offset0, offset1 = shift * scale); the
// estimator's min-depth filter and offset1 /= scale follow in minimal_affine.  MP_HD scalar code, fixed-size arrays.
#pragma once
#include "mp_affine.cuh"
#include "mp_relpose.cuh"

namespace mp {

// R = [v1 v2 v1 x v2] [u1 u2 u1 x u2]^-1, t = Y0 - R X0 (solver.cpp:662-676): the pose of three exactly congruent points
MP_HD bool pose_from_triplet(const double (*X)[3], const double (*Y)[3], double *R, double *t) {
    double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3];
    for (int k = 0; k < 3; k++) { u1[k] = X[0][k] - X[1][k]; u2[k] = X[0][k] - X[2][k]; v1[k] = Y[0][k] - Y[1][k]; v2[k] = Y[0][k] - Y[2][k]; }
    cross3(u1, u2, u3); cross3(v1, v2, v3);
    const double Xm[9] = {u1[0], u2[0], u3[0], u1[1], u2[1], u3[1], u1[2], u2[2], u3[2]};
    const double Ym[9] = {v1[0], v2[0], v3[0], v1[1], v2[1], v3[1], v1[2], v2[2], v3[2]};
    double Xi[9];
    if (!inv3(Xm, Xi)) return false;
    mat3_mul(Ym, Xi, R);
    double RX0[3];
    mat3_vecp(R, X[0], RX0);
    for (int k = 0; k < 3; k++) t[k] = Y[0][k] - RX0[k];
    return true;
}

// calibrated, use_ours: x0, x1 = 3 homogeneous points (K^-1 x).  <= 4 models
MP_HD int alt_calibrated_ours(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1, Model *out) {
    AffSol sols[4];
    const int ns = affine_calibrated(x0, x1, d0, d1, sols);
    int n = 0;
    for (int q = 0; q < ns; q++) {
        const AffSol &s = sols[q];
        if (s.a2 * s.a2 < 0.01) continue;                                                     // solver.cpp:609
        bool pos = true;
        for (int k = 0; k < 3; k++) if (d0[k] + s.b1 <= 0 || d1[k] + s.b2 <= 0) pos = false;  // :656-660
        if (!pos) continue;
        double X[3][3], Y[3][3];
        for (int k = 0; k < 3; k++)
            for (int c = 0; c < 3; c++) { X[k][c] = (d0[k] + s.b1) * x0[k][c]; Y[k][c] = s.a2 * (d1[k] + s.b2) * x1[k][c]; }
        Model m;
        if (!pose_from_triplet(X, Y, m.R, m.t)) continue;
        m.scale = s.a2; m.o0 = s.b1; m.o1 = s.a2 * s.b2; m.f0 = 1.0; m.f1 = 1.0;
        out[n++] = m;
    }
    return n;
}

// out[0..NA+NB-2] += s * a * b
template <int NA, int NB> MP_HD void pmul_acc(const double *a, const double *b, double s, double *out) {
    for (int i = 0; i < NA; i++)
        for (int j = 0; j < NB; j++) out[i + j] = fma(s * a[i], b[j], out[i + j]);
}

// shared focal, use_ours: points (x, y, 1) in normalised pixels, the first three are used, d1[2] is not.  <= 4 models
MP_HD int alt_shared_focal_ours(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1, Model *out) {
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
    // polynomials in F = f^2, ascending coefficients
    const double G[2] = {A01, B01}, H[2] = {C01, D01};
    const double L[2] = {e0 * m0 - e1 * m1, e0 - e1}, U[2] = {e0 * e0 * n0 - e1 * e1 * n1, e0 * e0 - e1 * e1};
    const double dAB[2] = {A02 - A12, B02 - B12}, AB02[2] = {A02, B02}, qF[2] = {q, 1.0}, m0F[2] = {m0, 1.0}, n0F[2] = {n0, 1.0};
    double N[3], t3[3], GL[3], L2[3];
    pmul<2, 2>(G, U, N); pmul<2, 2>(dAB, H, t3);
    for (int i = 0; i < 3; i++) N[i] -= t3[i];                      // delta = N / (2 G L)
    pmul<2, 2>(G, L, GL); pmul<2, 2>(L, L, L2);
    double ABH[3], GL2[4], NN[5], NGL[5], GLGL[5], P[6] = {0, 0, 0, 0, 0, 0};
    pmul<2, 2>(AB02, H, ABH); pmul<2, 3>(G, L2, GL2);
    pmul<3, 3>(N, N, NN); pmul<3, 3>(N, GL, NGL); pmul<3, 3>(GL, GL, GLGL);
    // 4 (A02 + F B02) H G L^2 - (q + F) N^2 + 4 d2_0 (m0 + F) N G L - 4 d2_0^2 (n0 + F) G^2 L^2 = 0
    pmul_acc<3, 4>(ABH, GL2, 4.0, P);
    pmul_acc<2, 5>(qF, NN, -1.0, P);
    pmul_acc<2, 5>(m0F, NGL, 4.0 * e0, P);
    pmul_acc<2, 5>(n0F, GLGL, -4.0 * e0 * e0, P);
    // the F^5 terms cancel identically: a quartic (four solutions, like the reference's 4 x 4 action matrix)
    double roots[4];
    const int nr = real_roots<4>(P, 4, roots);
    int n = 0;
    for (int k = 0; k < nr && n < 4; k++) {
        const double F = roots[k];
        if (!(F >= 0) || !finite_d(F)) continue;                    // A2(0) < 0 (solver.cpp:944)
        const double g = poly_val(G, 1, F), h = poly_val(H, 1, F), den = 2 * g * poly_val(L, 1, F);
        if (den == 0 || h == 0) continue;
        const double delta = poly_val(N, 2, F) / den, sigma = g / h;
        if (!(delta > 0) || !finite_d(delta)) continue;             // eigenvalue 1/d3 < 0 (:936)
        if (!(sigma >= 0.001)) continue;                            // :948
        bool ok = true;                                             // guard: both remaining pair equations hold
        const double Ai[2] = {A02, A12}, Bi[2] = {B02, B12}, ei[2] = {e0, e1}, mi[2] = {m0, m1}, ni[2] = {n0, n1};
        for (int i = 0; i < 2; i++) {
            const double lhs = Ai[i] + F * Bi[i];
            const double rhs = sigma * (delta * delta * (q + F) - 2 * delta * ei[i] * (mi[i] + F) + ei[i] * ei[i] * (ni[i] + F));
            if (!(fabs(lhs - rhs) <= 1e-7 * (fabs(lhs) + fabs(rhs)))) ok = false;
        }
        if (!ok) continue;
        const double f = sqrt(F), s = sqrt(sigma);
        double X[3][3], Y[3][3];
        const double dd[3] = {d1[0], d1[1], delta};
        for (int i = 0; i < 3; i++) {
            X[i][0] = d0[i] * x0[i][0] / f; X[i][1] = d0[i] * x0[i][1] / f; X[i][2] = d0[i];
            Y[i][0] = s * dd[i] * x1[i][0] / f; Y[i][1] = s * dd[i] * x1[i][1] / f; Y[i][2] = s * dd[i];
        }
        Model m;
        if (!pose_from_triplet(X, Y, m.R, m.t)) continue;
        m.scale = s; m.o0 = 0; m.o1 = 0; m.f0 = f; m.f1 = f;
        out[n++] = m;
    }
    return n;
}

// two focal, use_ours: linear in (1/f1^2, s^2/f2^2, s^2), all three must be positive (solver.cpp:1113).  <= 1 model
MP_HD int alt_two_focal_ours(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1, Model *out) {
    const int pa[3] = {0, 0, 1}, pb[3] = {1, 2, 2};
    double A[9], B[3];
    for (int r = 0; r < 3; r++) {
        const int i = pa[r], j = pb[r];
        const double ax = d0[i] * x0[i][0] - d0[j] * x0[j][0], ay = d0[i] * x0[i][1] - d0[j] * x0[j][1], az = d0[i] - d0[j];
        const double bx = d1[i] * x1[i][0] - d1[j] * x1[j][0], by = d1[i] * x1[i][1] - d1[j] * x1[j][1], bz = d1[i] - d1[j];
        A[3 * r] = ax * ax + ay * ay; A[3 * r + 1] = -(bx * bx + by * by); A[3 * r + 2] = -bz * bz;
        B[r] = -az * az;
    }
    if (!lu_solve_inplace<3, 1>(A, B)) return 0;
    if (!(B[0] > 0 && B[1] > 0 && B[2] > 0)) return 0;
    const double fi = sqrt(B[0]), s = sqrt(B[2]), wi = sqrt(B[1] / B[2]); // 1 / f1, s, 1 / f2
    double X[3][3], Y[3][3];
    for (int k = 0; k < 3; k++) {
        X[k][0] = d0[k] * x0[k][0] * fi; X[k][1] = d0[k] * x0[k][1] * fi; X[k][2] = d0[k];
        Y[k][0] = s * d1[k] * x1[k][0] * wi; Y[k][1] = s * d1[k] * x1[k][1] * wi; Y[k][2] = s * d1[k];
    }
    Model m;
    if (!pose_from_triplet(X, Y, m.R, m.t)) return 0;
    m.scale = s; m.o0 = 0; m.o1 = 0; m.f0 = 1.0 / fi; m.f1 = 1.0 / wi;
    out[0] = m;
    return 1;
}

// two focal, use_4p4d: F x0_i = q_i (e x x1_i), q_i = d1_i / d0_i, e_z = 1: 11 of the 12 equations; Bougnoux focals;
// E = K2 F K1; motion_from_essential on the un-calibrated unit vectors (reference quirk, solver.cpp:1301-1302).  <= 4 models
MP_HD int alt_two_focal_4p4d(const double (*x0)[3], const double (*x1)[3], const double *d0, const double *d1, Model *out) {
    double A[121], B[11];
    for (int i = 0; i < 121; i++) A[i] = 0.0;
    int row = 0;
    for (int i = 0; i < 4; i++) {
        const double u = x0[i][0] / x0[i][2], v = x0[i][1] / x0[i][2], u2 = x1[i][0] / x1[i][2], v2 = x1[i][1] / x1[i][2];
        const double q = d1[i] / d0[i];
        A[11 * row + 0] = -u; A[11 * row + 1] = -v; A[11 * row + 2] = -1; A[11 * row + 10] = q; B[row] = q * v2; row++;
        A[11 * row + 3] = -u; A[11 * row + 4] = -v; A[11 * row + 5] = -1; A[11 * row + 9] = -q; B[row] = -q * u2; row++;
        if (i == 3) break;
        A[11 * row + 6] = -u; A[11 * row + 7] = -v; A[11 * row + 8] = -1; A[11 * row + 9] = q * v2; A[11 * row + 10] = -q * u2; B[row] = 0; row++;
    }
    if (!lu_solve_inplace<11, 1>(A, B)) return 0;
    double F[9];
    for (int k = 0; k < 9; k++) F[k] = B[k];
    double f0sq, f1sq;
    bougnoux_focals(F, &f0sq, &f1sq);
    const double f0 = sqrt(f0sq), f1 = sqrt(f1sq);
    if (f0 != f0 || f1 != f1) return 0;                             // solver.cpp:1376-1379
    double E[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) E[3 * i + j] = F[3 * i + j] * (i < 2 ? f1 : 1.0) * (j < 2 ? f0 : 1.0);
    double b0[4][3], b1[4][3];
    for (int i = 0; i < 4; i++) {
        const double na = norm3(x0[i]), nb = norm3(x1[i]);
        for (int k = 0; k < 3; k++) { b0[i][k] = x0[i][k] / na; b1[i][k] = x1[i][k] / nb; }
    }
    Pose ps[4];
    const int np_ = motion_from_essential<4>(E, b0, b1, ps, 4);
    for (int k = 0; k < np_; k++) {
        Model m;
        for (int c = 0; c < 9; c++) m.R[c] = ps[k].R[c];
        for (int c = 0; c < 3; c++) m.t[c] = ps[k].t[c];
        m.scale = 1.0; m.o0 = 0; m.o1 = 0; m.f0 = f0; m.f1 = f1;
        out[k] = m;
    }
    return np_;
}

} // namespace mp
