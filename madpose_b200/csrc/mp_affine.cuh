// K1 -- depth-aware minimal solvers (affine depth correction + Kabsch), fp64, one
// thread per sample, everything in registers / local arrays.
// Replaces solve_scale_and_shift{,_shared_focal,_two_focal} (solver.cpp:35-125,
// 127-293, 295-480) and the pose wrappers solve_scale_shift_pose* (solver.cpp:482-534,
// 682-739, 986-1043).  The reference eliminates with 12x12 / 36x36 / 40x40 templates
// and an action-matrix eigen-solve; here the same equations -- pairwise 3-D distances
// between back-projected, affinely corrected points agree in both views -- are reduced
// with b1 as the hidden variable to ONE univariate polynomial of degree 4 / 8 / 4
// whose real roots are isolated by a Sturm chain (mp_poly.cuh).
#pragma once
#include "mp_poly.cuh"
#include "mp_small.cuh"

namespace mp {

// out[0..NA+NB-2] = a * b  (coefficient counts NA, NB)
template <int NA, int NB> MP_HD void pmul(const double *a, const double *b, double *out) {
    for (int i = 0; i < NA + NB - 1; i++) out[i] = 0.0;
    for (int i = 0; i < NA; i++)
        for (int j = 0; j < NB; j++) out[i + j] = fma(a[i], b[j], out[i + j]);
}

// |(d_i + b) p_i - (d_j + b) p_j|^2 = q[0] b^2 + q[1] b + q[2] over the first DIM coordinates
template <int DIM> MP_HD void pair_quadratic(const double *pi, const double *pj, double di, double dj, double *q) {
    double A = 0, B = 0, C = 0;
    for (int k = 0; k < DIM; k++) {
        double e = pi[k] - pj[k], g = di * pi[k] - dj * pj[k];
        A = fma(e, e, A); B = fma(2 * e, g, B); C = fma(g, g, C);
    }
    q[0] = A; q[1] = B; q[2] = C;
}

struct AffSol { double b1, a2, b2, f1, f2; }; // corrected depths: d1 + b1 and a2 * (d2 + b2)

// calibrated, 3 points: x1, x2 rows (x, y, 1) of K^-1-normalised coordinates.  <= 4 solutions
MP_HD int affine_calibrated(const double (*x1)[3], const double (*x2)[3], const double *d1, const double *d2,
                            AffSol *out) {
    const int pa[3] = {0, 0, 1}, pb[3] = {1, 2, 2};
    double P[9], Q[9], Qi[9], M[9];
    for (int r = 0; r < 3; r++) {
        pair_quadratic<3>(x1[pa[r]], x1[pb[r]], d1[pa[r]], d1[pb[r]], P + 3 * r);
        pair_quadratic<3>(x2[pa[r]], x2[pb[r]], d2[pa[r]], d2[pb[r]], Q + 3 * r);
    }
    if (!inv3(Q, Qi)) return 0;
    mat3_mul(Qi, P, M); // [s b2^2, s b2, s]^T = M [b1^2, b1, 1]^T,  s = a2^2
    const double k0[3] = {M[2], M[1], M[0]}, k1[3] = {M[5], M[4], M[3]}, k2[3] = {M[8], M[7], M[6]};
    double q[5], t[5], roots[4];
    pmul<3, 3>(k1, k1, q);
    pmul<3, 3>(k0, k2, t);
    for (int i = 0; i < 5; i++) q[i] -= t[i];
    const int nr = real_roots<4>(q, 4, roots);
    int n = 0;
    for (int k = 0; k < nr; k++) {
        const double b1 = roots[k], s = poly_val(k2, 2, b1), sb2 = poly_val(k1, 2, b1);
        if (!(s > 0) || !finite_d(s)) continue;
        out[n].b1 = b1; out[n].a2 = sqrt(s); out[n].b2 = sb2 / s; out[n].f1 = 1.0; out[n].f2 = 1.0;
        n++;
    }
    return n;
}

// shared unknown focal, 4 points (x, y, 1) in (normalised) pixel units.  <= 8 solutions
MP_HD int affine_shared_focal(const double (*x1in)[3], const double (*x2in)[3], const double *d1, const double *d2,
                              AffSol *out) {
    const int pa[4] = {0, 0, 1, 0}, pb[4] = {1, 2, 2, 3};
    double x1[4][2], x2[4][2], m1 = 0, m2 = 0;
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 2; k++) { m1 += fabs(x1in[i][k]); m2 += fabs(x2in[i][k]); }
    const double f0 = 0.5 * (m1 / 8 + m2 / 8); // pre-scaling of solver.cpp:134-138
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 2; k++) { x1[i][k] = x1in[i][k] / f0; x2[i][k] = x2in[i][k] / f0; }
    // per pair:  w P1(b1) + D1 = sigma (w P2(b2) + D2);  unknown vector y = [v b2^2, v b2, v, sigma], v = sigma w
    double N[16], rhs[16];
    for (int r = 0; r < 4; r++) {
        const int i = pa[r], j = pb[r];
        pair_quadratic<2>(x2[i], x2[j], d2[i], d2[j], N + 4 * r);
        N[4 * r + 3] = (d2[i] - d2[j]) * (d2[i] - d2[j]);
        pair_quadratic<2>(x1[i], x1[j], d1[i], d1[j], rhs + 4 * r);
        rhs[4 * r + 3] = (d1[i] - d1[j]) * (d1[i] - d1[j]);
    }
    if (!lu_solve_inplace<4, 4>(N, rhs)) return 0;
    // y_r = w g_r(b1) + h_r
    double g[4][3], h[4];
    for (int r = 0; r < 4; r++) { g[r][0] = rhs[4 * r + 2]; g[r][1] = rhs[4 * r + 1]; g[r][2] = rhs[4 * r]; h[r] = rhs[4 * r + 3]; }
    // (1) v = sigma w       :  a w^2 + b w + c = 0,  a = g3, b = h3 - g2, c = -h2
    // (2) y1^2 = y0 y2      :  a2 w^2 + b2 w + c2 = 0
    double a[3], b[3], c0 = -h[2], a2[5], b2[3], c2, t5[5];
    for (int i = 0; i < 3; i++) { a[i] = g[3][i]; b[i] = -g[2][i]; }
    b[0] += h[3];
    pmul<3, 3>(g[1], g[1], a2);
    pmul<3, 3>(g[0], g[2], t5);
    for (int i = 0; i < 5; i++) a2[i] -= t5[i];
    for (int i = 0; i < 3; i++) b2[i] = 2 * h[1] * g[1][i] - h[2] * g[0][i] - h[0] * g[2][i];
    c2 = h[1] * h[1] - h[0] * h[2];
    // resultant in w of the two quadratics: (a c2 - a2 c)^2 - (a b2 - a2 b)(b c2 - b2 c)
    double ac[5], ab[7], bc[3], tmp7[7], res[9], t9[9];
    for (int i = 0; i < 5; i++) ac[i] = (i < 3 ? a[i] * c2 : 0.0) - a2[i] * c0;
    pmul<3, 3>(a, b2, ab);         // degree 4 in ab[0..4]
    ab[5] = 0; ab[6] = 0;
    pmul<5, 3>(a2, b, tmp7);       // degree 6
    for (int i = 0; i < 7; i++) ab[i] -= tmp7[i];
    for (int i = 0; i < 3; i++) bc[i] = b[i] * c2 - b2[i] * c0;
    pmul<5, 5>(ac, ac, res);
    pmul<7, 3>(ab, bc, t9);
    for (int i = 0; i < 9; i++) res[i] -= t9[i];
    double roots[8];
    const int nr = real_roots<8>(res, 8, roots);
    int n = 0;
    for (int k = 0; k < nr; k++) {
        double b1 = roots[k];
        const double den = -poly_val(ab, 6, b1);
        if (den == 0.0) continue;
        double w = poly_val(ac, 4, b1) / den;
        // Newton polish of (b1, w) on the two quadratics-in-w: the expanded octic only locates the root
        for (int it = 0; it < 6; it++) {
            double da, db, da2, db2;
            const double va = poly_val_d(a, 2, b1, &da), vb = poly_val_d(b, 2, b1, &db);
            const double va2 = poly_val_d(a2, 4, b1, &da2), vb2 = poly_val_d(b2, 2, b1, &db2);
            const double F1 = (va * w + vb) * w + c0, F2 = (va2 * w + vb2) * w + c2;
            const double J11 = (da * w + db) * w, J12 = 2 * va * w + vb;
            const double J21 = (da2 * w + db2) * w, J22 = 2 * va2 * w + vb2;
            const double det = J11 * J22 - J12 * J21;
            if (det == 0.0 || !finite_d(det)) break;
            const double s1 = (F1 * J22 - F2 * J12) / det, s2 = (J11 * F2 - J21 * F1) / det;
            if (!finite_d(s1) || !finite_d(s2)) break;
            b1 -= s1; w -= s2;
            if (fabs(s1) <= 1e-16 * fabs(b1) && fabs(s2) <= 1e-16 * fabs(w)) break;
        }
        const double y1 = w * poly_val(g[1], 2, b1) + h[1], y2 = w * poly_val(g[2], 2, b1) + h[2];
        const double s = w * poly_val(g[3], 2, b1) + h[3];
        if (w < 0 || !(s > 0) || !finite_d(w) || !finite_d(s)) continue;
        const double f = f0 / sqrt(w);
        out[n].b1 = b1; out[n].a2 = sqrt(s); out[n].b2 = y1 / y2; out[n].f1 = f; out[n].f2 = f;
        n++;
    }
    return n;
}

// two independent unknown focals, 4 points.  <= 4 solutions
MP_HD int affine_two_focal(const double (*x1in)[3], const double (*x2in)[3], const double *d1, const double *d2,
                           AffSol *out) {
    const int pa[5] = {0, 0, 1, 0, 1}, pb[5] = {1, 2, 2, 3, 3};
    double x1[4][2], x2[4][2], m1 = 0, m2 = 0;
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 2; k++) { m1 += fabs(x1in[i][k]); m2 += fabs(x2in[i][k]); }
    const double f10 = m1 / 8, f20 = m2 / 8; // solver.cpp:302-305
    for (int i = 0; i < 4; i++)
        for (int k = 0; k < 2; k++) { x1[i][k] = x1in[i][k] / f10; x2[i][k] = x2in[i][k] / f20; }
    // per pair: w1 P1(b1) + D1 = N_r . [u b2^2, u b2, u, sigma],  u = sigma w2
    double N[5][4], R1[5][4];
    for (int r = 0; r < 5; r++) {
        const int i = pa[r], j = pb[r];
        pair_quadratic<2>(x2[i], x2[j], d2[i], d2[j], N[r]);
        N[r][3] = (d2[i] - d2[j]) * (d2[i] - d2[j]);
        pair_quadratic<2>(x1[i], x1[j], d1[i], d1[j], R1[r]);
        R1[r][3] = (d1[i] - d1[j]) * (d1[i] - d1[j]);
    }
    // left null vector of the 5x4 matrix N eliminates y: nv . (w1 P1(b1) + D1) = 0  =>  w1 = -k / rho(b1)
    double Nt[20], nv[5];
    for (int r = 0; r < 5; r++)
        for (int c = 0; c < 4; c++) Nt[c * 5 + r] = N[r][c];
    nullspace_qr<4, 5>(Nt, nv);
    double k = 0, rho[3] = {0, 0, 0};
    for (int r = 0; r < 5; r++) {
        k = fma(nv[r], R1[r][3], k);
        rho[2] = fma(nv[r], R1[r][0], rho[2]); rho[1] = fma(nv[r], R1[r][1], rho[1]); rho[0] = fma(nv[r], R1[r][2], rho[0]);
    }
    int drop = 0; // solve y from the four rows with the smallest |nv| weight
    for (int r = 1; r < 5; r++) if (fabs(nv[r]) > fabs(nv[drop])) drop = r;
    double Nm[16], rhs[16];
    for (int r = 0, rr = 0; r < 5; r++) {
        if (r == drop) continue;
        for (int c = 0; c < 4; c++) { Nm[4 * rr + c] = N[r][c]; rhs[4 * rr + c] = R1[r][c]; }
        rr++;
    }
    if (!lu_solve_inplace<4, 4>(Nm, rhs)) return 0;
    // rho(b1) * y_r = e_r(b1) = -k g_r(b1) + h_r rho(b1)
    double e[4][3];
    for (int r = 0; r < 4; r++) {
        const double gr[3] = {rhs[4 * r + 2], rhs[4 * r + 1], rhs[4 * r]};
        for (int i = 0; i < 3; i++) e[r][i] = -k * gr[i] + rhs[4 * r + 3] * rho[i];
    }
    double quart[5], t5[5], roots[4];
    pmul<3, 3>(e[1], e[1], quart);
    pmul<3, 3>(e[0], e[2], t5);
    for (int i = 0; i < 5; i++) quart[i] -= t5[i];
    const int nr = real_roots<4>(quart, 4, roots);
    int n = 0;
    for (int q = 0; q < nr; q++) {
        const double b1 = roots[q], rb = poly_val(rho, 2, b1);
        if (rb == 0.0) continue;
        const double w1 = -k / rb;
        const double y1 = poly_val(e[1], 2, b1) / rb, y2 = poly_val(e[2], 2, b1) / rb, s = poly_val(e[3], 2, b1) / rb;
        const double w2 = y2 / s;
        if (w1 < 0 || w2 < 0 || !(s > 0) || !finite_d(w1) || !finite_d(w2) || !finite_d(s)) continue;
        out[n].b1 = b1; out[n].a2 = sqrt(s); out[n].b2 = y1 / y2; out[n].f1 = f10 / sqrt(w1); out[n].f2 = f20 / sqrt(w2);
        n++;
    }
    return n;
}

// Kabsch: Y ~ R X + t on NP points (solver.cpp:506-531 and the focal twins)
template <int NP> MP_HD bool kabsch(const double (*X)[3], const double (*Y)[3], double *R, double *t) {
    double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0}, S[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = 0; i < NP; i++)
        for (int k = 0; k < 3; k++) { cx[k] += X[i][k]; cy[k] += Y[i][k]; }
    for (int k = 0; k < 3; k++) { cx[k] /= NP; cy[k] /= NP; }
    for (int i = 0; i < NP; i++)
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) S[3 * a + b] = fma(Y[i][a] - cy[a], X[i][b] - cx[b], S[3 * a + b]);
    if (!procrustes_rotation(S, R)) return false;
    double Rc[3];
    mat3_vecp(R, cx, Rc);
    for (int k = 0; k < 3; k++) t[k] = cy[k] - Rc[k];
    return true;
}

// estimate_scale_and_pose (solver.cpp:5-33): weighted Umeyama, Y ~ scale R X + t.  Centroids and covariance are
// weighted; the scale is the unweighted ratio sum(Yc . R Xc) / sum(R Xc . R Xc) (solver.cpp:28-29).
MP_HD bool umeyama(int n, const double (*X)[3], const double (*Y)[3], const double *W, double *R, double *t, double *scale) {
    double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0}, S[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, ws = 0;
    for (int i = 0; i < n; i++) {
        ws += W[i];
        for (int k = 0; k < 3; k++) { cx[k] += X[i][k] * W[i]; cy[k] += Y[i][k] * W[i]; }
    }
    for (int k = 0; k < 3; k++) { cx[k] /= ws; cy[k] /= ws; }
    for (int i = 0; i < n; i++)
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) S[3 * a + b] += (Y[i][a] - cy[a]) * W[i] * (X[i][b] - cx[b]);
    if (!procrustes_rotation(S, R)) return false;
    double num = 0, den = 0;
    for (int i = 0; i < n; i++) {
        const double xc[3] = {X[i][0] - cx[0], X[i][1] - cx[1], X[i][2] - cx[2]};
        double xr[3];
        mat3_vecp(R, xc, xr);
        for (int k = 0; k < 3; k++) { num += (Y[i][k] - cy[k]) * xr[k]; den += xr[k] * xr[k]; }
    }
    *scale = num / den;
    double Rc[3];
    mat3_vecp(R, cx, Rc);
    for (int k = 0; k < 3; k++) t[k] = cy[k] - *scale * Rc[k];
    return true;
}

} // namespace mp
