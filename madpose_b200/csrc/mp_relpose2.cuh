// K2 (part 2) -- six-point shared-focal solver, DLT triangulation, Bougnoux focals,
// recoverPose and the 2-parameter depth fits.  PoseLib's relpose_6pt_shared_focal
// (...shared_focal_estimator.cpp:87) and cv::recoverPose (...two_focal_estimator.cpp:143)
// are third-party code absent from the reference tree; the published algorithms are
// used (hidden-variable formulation of Li 2006 / Kukelova et al. 2008 with w = 1/f^2).
#pragma once

namespace mp {

// ---- forms in (a, b): linear [a b 1]; quadratic [aa ab bb a b 1];
// cubic [aaa aab abb bbb aa ab bb a b 1]
MP_HD void mul_ll2(const double *p, const double *q, double *o) {
    o[0] = p[0] * q[0]; o[1] = p[0] * q[1] + p[1] * q[0]; o[2] = p[1] * q[1];
    o[3] = p[0] * q[2] + p[2] * q[0]; o[4] = p[1] * q[2] + p[2] * q[1]; o[5] = p[2] * q[2];
}
MP_HD void fma_ql2(const double *Q, const double *l, double s, double *c) { // c += s * Q * l
    const double l0 = l[0] * s, l1 = l[1] * s, l2 = l[2] * s;
    c[0] += Q[0] * l0;
    c[1] += Q[0] * l1 + Q[1] * l0;
    c[2] += Q[1] * l1 + Q[2] * l0;
    c[3] += Q[2] * l1;
    c[4] += Q[0] * l2 + Q[3] * l0;
    c[5] += Q[1] * l2 + Q[3] * l1 + Q[4] * l0;
    c[6] += Q[2] * l2 + Q[4] * l1;
    c[7] += Q[3] * l2 + Q[5] * l0;
    c[8] += Q[4] * l2 + Q[5] * l1;
    c[9] += Q[5] * l2;
}

// determinant of the complex 10x10 matrix C0 + w C1 + w^2 C2 at w = (wr, wi)
MP_HD void det10_complex(const double (*Cw)[100], double wr, double wi, double *dre, double *dim) {
    double Mr[100], Mi[100];
    const double w2r = wr * wr - wi * wi, w2i = 2 * wr * wi;
    for (int i = 0; i < 100; i++) {
        Mr[i] = Cw[0][i] + wr * Cw[1][i] + w2r * Cw[2][i];
        Mi[i] = wi * Cw[1][i] + w2i * Cw[2][i];
    }
    double dr = 1.0, di = 0.0;
    for (int c = 0; c < 10; c++) {
        int piv = c;
        double pm = Mr[c * 10 + c] * Mr[c * 10 + c] + Mi[c * 10 + c] * Mi[c * 10 + c];
        for (int r = c + 1; r < 10; r++) {
            double m = Mr[r * 10 + c] * Mr[r * 10 + c] + Mi[r * 10 + c] * Mi[r * 10 + c];
            if (m > pm) { pm = m; piv = r; }
        }
        if (!(pm > 0.0)) { *dre = 0; *dim = 0; return; }
        if (piv != c) {
            for (int k = 0; k < 10; k++) {
                double t = Mr[c * 10 + k]; Mr[c * 10 + k] = Mr[piv * 10 + k]; Mr[piv * 10 + k] = t;
                t = Mi[c * 10 + k]; Mi[c * 10 + k] = Mi[piv * 10 + k]; Mi[piv * 10 + k] = t;
            }
            dr = -dr; di = -di;
        }
        const double pr = Mr[c * 10 + c], pi = Mi[c * 10 + c];
        { double t = dr * pr - di * pi; di = dr * pi + di * pr; dr = t; }
        const double ir = pr / pm, ii = -pi / pm; // 1 / pivot
        for (int r = c + 1; r < 10; r++) {
            const double ar = Mr[r * 10 + c], ai = Mi[r * 10 + c];
            const double fr = ar * ir - ai * ii, fi = ar * ii + ai * ir;
            if (fr == 0.0 && fi == 0.0) continue;
            for (int k = c + 1; k < 10; k++) {
                const double br = Mr[c * 10 + k], bi = Mi[c * 10 + k];
                Mr[r * 10 + k] -= fr * br - fi * bi;
                Mi[r * 10 + k] -= fr * bi + fi * br;
            }
        }
    }
    *dre = dr; *dim = di;
}

struct FocalPose { Pose pose; double focal; };

// Is E an essential matrix?  || 2 E E^T E - tr(E E^T) E ||_F <= tol on the unit-norm matrix.  The hidden-variable
// determinant has real roots that are not solutions of the ten equations (the polish then stops at a least-squares
// point with a residual); an exact solver (PoseLib's action matrix) never returns those, so they are dropped.
MP_HD bool essential_valid(const double *Ein, double tol) {
    double n = 0, E[9], EEt[9], T[9];
    for (int k = 0; k < 9; k++) n += Ein[k] * Ein[k];
    n = sqrt(n);
    if (!(n > 0) || !finite_d(n)) return false;
    for (int k = 0; k < 9; k++) E[k] = Ein[k] / n;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
    const double tr = EEt[0] + EEt[4] + EEt[8];
    mat3_mul(EEt, E, T);
    double r = 0;
    for (int k = 0; k < 9; k++) { const double v = 2 * T[k] - tr * E[k]; r += v * v; }
    return sqrt(r) <= tol;
}

// x1, x2: 6 points (x, y, 1) in (normalised) pixel units, one shared unknown focal.  <= 15 candidates
MP_HD int relpose_6pt_shared_focal(const double (*x1)[3], const double (*x2)[3], FocalPose *out) {
    double A[54], N[27];
    epipolar_rows<6>(x1, x2, A);
    nullspace_qr<6, 9>(A, N);
    double L[9][3]; // F_k = a N[k] + b N[9+k] + N[18+k]
    for (int k = 0; k < 9; k++) { L[k][0] = N[k]; L[k][1] = N[9 + k]; L[k][2] = N[18 + k]; }
    // F Q F^T = S0 + w S1 (Q = diag(1,1,w)), symmetric quadratic forms
    double S0[6][6], S1[6][6], tmp[6];
    const int gi[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
    for (int i = 0; i < 3; i++)
        for (int j = i; j < 3; j++) {
            mul_ll2(L[3 * i], L[3 * j], S0[gi[i][j]]);
            mul_ll2(L[3 * i + 1], L[3 * j + 1], tmp);
            for (int m = 0; m < 6; m++) S0[gi[i][j]][m] += tmp[m];
            mul_ll2(L[3 * i + 2], L[3 * j + 2], S1[gi[i][j]]);
        }
    // trace of G = F Q F^T Q:  t0 + w t1 + w^2 t2
    double t0[6], t1[6], t2[6];
    for (int m = 0; m < 6; m++) { t0[m] = S0[0][m] + S0[3][m]; t1[m] = S1[0][m] + S1[3][m] + S0[5][m]; t2[m] = S1[5][m]; }
    // Cw[k][10 r + c]: coefficient of w^k, equation r, cubic monomial c
    double Cw[3][100];
    for (int k = 0; k < 3; k++)
        for (int i = 0; i < 100; i++) Cw[k][i] = 0.0;
    {
        double q1[6], q2[6];
        mul_ll2(L[4], L[8], q1); mul_ll2(L[5], L[7], q2);
        for (int m = 0; m < 6; m++) q1[m] -= q2[m];
        fma_ql2(q1, L[0], 1.0, Cw[0]);
        mul_ll2(L[3], L[8], q1); mul_ll2(L[5], L[6], q2);
        for (int m = 0; m < 6; m++) q1[m] -= q2[m];
        fma_ql2(q1, L[1], -1.0, Cw[0]);
        mul_ll2(L[3], L[7], q1); mul_ll2(L[4], L[6], q2);
        for (int m = 0; m < 6; m++) q1[m] -= q2[m];
        fma_ql2(q1, L[2], 1.0, Cw[0]);
    }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            const int r = 10 * (1 + 3 * i + j);
            // 2 sum_k G_ik F_kj, G_ik = (S0_ik + w S1_ik) * (k == 2 ? w : 1)
            for (int k = 0; k < 2; k++) {
                fma_ql2(S0[gi[i][k]], L[3 * k + j], 2.0, Cw[0] + r);
                fma_ql2(S1[gi[i][k]], L[3 * k + j], 2.0, Cw[1] + r);
            }
            fma_ql2(S0[gi[i][2]], L[6 + j], 2.0, Cw[1] + r);
            fma_ql2(S1[gi[i][2]], L[6 + j], 2.0, Cw[2] + r);
            fma_ql2(t0, L[3 * i + j], -1.0, Cw[0] + r);
            fma_ql2(t1, L[3 * i + j], -1.0, Cw[1] + r);
            fma_ql2(t2, L[3 * i + j], -1.0, Cw[2] + r);
        }
    for (int r = 0; r < 10; r++) { // rows scaled to unit max coefficient
        double mx = 0;
        for (int c = 0; c < 10; c++)
            for (int k = 0; k < 3; k++) mx = fmax(mx, fabs(Cw[k][10 * r + c]));
        if (!(mx > 0)) return 0;
        const double im = 1.0 / mx;
        for (int c = 0; c < 10; c++)
            for (int k = 0; k < 3; k++) Cw[k][10 * r + c] *= im;
    }
    // det C(w) has degree 15: sample on the unit circle (conjugate symmetry: 9 of 16 points),
    // inverse DFT -> coefficients.  The polynomial is only a root locator (roots are polished below).
    const double PI = 3.14159265358979323846;
    double vr[9], vi[9];
    for (int k = 0; k <= 8; k++) det10_complex(Cw, cos(2 * PI * k / 16), sin(2 * PI * k / 16), &vr[k], &vi[k]);
    double coef[16], roots[15];
    for (int j = 0; j < 16; j++) {
        double s = vr[0] + ((j & 1) ? -vr[8] : vr[8]);
        for (int k = 1; k < 8; k++) {
            const double ang = 2 * PI * ((j * k) & 15) / 16;
            s += 2.0 * (vr[k] * cos(ang) + vi[k] * sin(ang));
        }
        coef[j] = s / 16.0;
    }
    int nr = real_roots<15>(coef, 15, roots);
    { // keep the admissible roots (w = 1 / f^2 > 0) only, in order: the lanes of a warp then walk lists of similar length
        int k = 0;
        for (int q = 0; q < nr; q++)
            if (roots[q] > 0) roots[k++] = roots[q];
        nr = k;
    }
    int n = 0;
    for (int q = 0; q < nr && n < 15; q++) {
        const double wr = roots[q];
        // null vector of C(w): elimination with full pivoting, free variable = 1
        double M[100];
        int cp[10];
        for (int i = 0; i < 100; i++) M[i] = Cw[0][i] + wr * (Cw[1][i] + wr * Cw[2][i]);
        for (int i = 0; i < 10; i++) cp[i] = i;
        for (int c = 0; c < 9; c++) {
            int pr = c, pc = c;
            double pm = fabs(M[10 * c + c]);
            for (int r = c; r < 10; r++)
                for (int k = c; k < 10; k++)
                    if (fabs(M[10 * r + k]) > pm) { pm = fabs(M[10 * r + k]); pr = r; pc = k; }
            for (int k = 0; k < 10; k++) { double t = M[10 * c + k]; M[10 * c + k] = M[10 * pr + k]; M[10 * pr + k] = t; }
            for (int r = 0; r < 10; r++) { double t = M[10 * r + c]; M[10 * r + c] = M[10 * r + pc]; M[10 * r + pc] = t; }
            { int t = cp[c]; cp[c] = cp[pc]; cp[pc] = t; }
            for (int r = c + 1; r < 10; r++) {
                const double f = M[10 * r + c] / M[10 * c + c];
                for (int k = c; k < 10; k++) M[10 * r + k] -= f * M[10 * c + k];
            }
        }
        double v[10], vv[10];
        v[9] = 1.0;
        for (int c = 8; c >= 0; c--) {
            double s = 0;
            for (int k = c + 1; k < 10; k++) s += M[10 * c + k] * v[k];
            v[c] = -s / M[10 * c + c];
        }
        for (int i = 0; i < 10; i++) vv[cp[i]] = v[i];
        if (vv[9] == 0.0 || !finite_d(vv[9])) continue;
        double a = vv[7] / vv[9], b = vv[8] / vv[9], w = wr;
        // Gauss-Newton polish of (a, b, w) on the ten (row-scaled) equations
        for (int it = 0; it < 5; it++) {
            const double mo[10] = {a * a * a, a * a * b, a * b * b, b * b * b, a * a, a * b, b * b, a, b, 1.0};
            const double ma[10] = {3 * a * a, 2 * a * b, b * b, 0, 2 * a, b, 0, 1, 0, 0};
            const double mb[10] = {0, a * a, 2 * a * b, 3 * b * b, 0, a, 2 * b, 0, 1, 0};
            double JtJ[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Jtr[3] = {0, 0, 0};
            for (int r = 0; r < 10; r++) {
                double f = 0, g[3] = {0, 0, 0};
                for (int c = 0; c < 10; c++) {
                    const double c0 = Cw[0][10 * r + c], c1 = Cw[1][10 * r + c], c2 = Cw[2][10 * r + c];
                    const double cw = c0 + w * (c1 + w * c2);
                    f += cw * mo[c]; g[0] += cw * ma[c]; g[1] += cw * mb[c]; g[2] += (c1 + 2 * w * c2) * mo[c];
                }
                for (int i = 0; i < 3; i++) {
                    Jtr[i] += g[i] * f;
                    for (int j = 0; j < 3; j++) JtJ[3 * i + j] += g[i] * g[j];
                }
            }
            double inv[9], st[3];
            if (!inv3(JtJ, inv)) break;
            mat3_vecp(inv, Jtr, st);
            if (!finite_d(st[0] + st[1] + st[2])) break;
            a -= st[0]; b -= st[1]; w -= st[2];
            if (fabs(st[0]) <= 1e-15 * fabs(a) && fabs(st[1]) <= 1e-15 * fabs(b) && fabs(st[2]) <= 1e-15 * fabs(w)) break;
        }
        if (!(w > 0) || !finite_d(a + b + w)) continue;
        const double f = 1.0 / sqrt(w);
        double E[9]; // E = K F K, K = diag(f, f, 1)
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++)
                E[3 * i + j] = (a * N[3 * i + j] + b * N[9 + 3 * i + j] + N[18 + 3 * i + j]) * (i < 2 ? f : 1.0) * (j < 2 ? f : 1.0);
        if (!essential_valid(E, 1e-8)) continue;
        double b1[6][3], b2[6][3];
        for (int i = 0; i < 6; i++) {
            const double p[3] = {x1[i][0] / f, x1[i][1] / f, 1.0}, qv[3] = {x2[i][0] / f, x2[i][1] / f, 1.0};
            const double np = norm3(p), nq = norm3(qv);
            for (int k = 0; k < 3; k++) { b1[i][k] = p[k] / np; b2[i][k] = qv[k] / nq; }
        }
        Pose ps[4];
        const int np = motion_from_essential<6>(E, b1, b2, ps, 4);
        for (int k = 0; k < np && n < 15; k++) { out[n].pose = ps[k]; out[n].focal = f; n++; }
    }
    return n;
}

// COLMAP-style DLT triangulation (utils.h:24-38): P1, P2 row-major 3x4, p pixel coordinates
MP_HD void triangulate_h(const double *P1, const double *P2, const double *p1, const double *p2, double *Xh /*[4]*/) {
    double A[16];
    for (int k = 0; k < 4; k++) {
        A[k] = p1[0] * P1[8 + k] - P1[k];
        A[4 + k] = p1[1] * P1[8 + k] - P1[4 + k];
        A[8 + k] = p2[0] * P2[8 + k] - P2[k];
        A[12 + k] = p2[1] * P2[8 + k] - P2[4 + k];
    }
    smallest_right_singular4(A, Xh);
}

// z ~ s d + c (normal equations, hybrid_pose_estimator.cpp:160-163)
MP_HD void fit_affine(int n, const double *d, const double *z, double *s, double *c) {
    double sdd = 0, sd = 0, sdz = 0, sz = 0;
    for (int i = 0; i < n; i++) { sdd += d[i] * d[i]; sd += d[i]; sdz += d[i] * z[i]; sz += z[i]; }
    const double det = sdd * n - sd * sd;
    *s = (sdz * n - sd * sz) / det;
    *c = (sdd * sz - sd * sdz) / det;
}

// Bougnoux focal lengths with principal points at the origin
// (...two_focal_estimator.cpp:11-32).  With p = (0,0,1) and II = diag(1,1,0):
// f1^2 = -[s(e2) II F p p^T F^T p] / [s(e2) II F II F^T p] evaluated as the (2,2)
// entries; written out on the few entries that survive the zeros.
MP_HD void bougnoux_focals(const double *F, double *f1sq, double *f2sq) {
    auto nullvec = [](const double *M, double *e) {
        double r[3][3], best = -1;
        cross3(M, M + 3, r[0]); cross3(M, M + 6, r[1]); cross3(M + 3, M + 6, r[2]);
        for (int k = 0; k < 3; k++) {
            const double nn = dot3(r[k], r[k]);
            if (nn > best) { best = nn; e[0] = r[k][0]; e[1] = r[k][1]; e[2] = r[k][2]; }
        }
    };
    double Ft[9], e1[3], e2[3];
    mat3_transpose(F, Ft);
    nullvec(F, e1);  // F e1 = 0
    nullvec(Ft, e2); // F^T e2 = 0
    e1[0] /= e1[2]; e1[1] /= e1[2]; e2[0] /= e2[2]; e2[1] /= e2[2];
    // row 2 of [e]x II = (-e_y, e_x, 0)  ->  T = (-e_y, e_x, 0) . Fa  (a row vector)
    auto ratio = [](const double *e, const double *Fa, const double *Fb) {
        const double T[3] = {-e[1] * Fa[0] + e[0] * Fa[3], -e[1] * Fa[1] + e[0] * Fa[4], -e[1] * Fa[2] + e[0] * Fa[5]};
        const double num = T[2] * Fb[8];                 // T pp^T Fb p   = T_z * Fb[2][2]
        const double den = T[0] * Fb[2] + T[1] * Fb[5];  // T II Fb p     = T_x Fb[0][2] + T_y Fb[1][2]
        return -num / den;
    };
    *f1sq = ratio(e2, F, Ft);
    *f2sq = ratio(e1, Ft, F);
}

// cv::recoverPose(E, pts0, pts1, I, R, t, dist): candidates (R1,t),(R2,t),(R1,-t),(R2,-t);
// every point is DLT-triangulated, a point counts when it lies in front of both cameras
// and closer than dist; the first candidate with the largest count wins.
// one (pose hypothesis c, point) test of recoverPose: c = 0..3 -> (Ra,t), (Rb,t), (Ra,-t), (Rb,-t)
MP_HD bool recover_pose_point_ok(const double *Ra, const double *Rb, const double *tt, int c, const double *p0, const double *p1,
                                 double dist) {
    const double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    const double *Rc = (c & 1) ? Rb : Ra;
    const double sg = c < 2 ? 1.0 : -1.0;
    double P1[12];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) P1[4 * i + j] = Rc[3 * i + j];
        P1[4 * i + 3] = sg * tt[i];
    }
    double Q[4];
    triangulate_h(P0, P1, p0, p1, Q);
    bool ok = Q[2] * Q[3] > 0;
    const double X[3] = {Q[0] / Q[3], Q[1] / Q[3], Q[2] / Q[3]};
    ok = ok && X[2] < dist;
    const double z1 = P1[8] * X[0] + P1[9] * X[1] + P1[10] * X[2] + P1[11];
    ok = ok && z1 > 0 && z1 < dist;
    return ok;
}
// the winner of the four counts: the first candidate with the largest count
MP_HD void recover_pose_select(const double *Ra, const double *Rb, const double *tt, const int *good4, double *R, double *t) {
    int best = -1, bestc = 0;
    for (int c = 0; c < 4; c++)
        if (best < 0 || good4[c] > bestc) { best = c; bestc = good4[c]; }
    const double *Rc = (best & 1) ? Rb : Ra;
    const double sg = best < 2 ? 1.0 : -1.0;
    for (int k = 0; k < 9; k++) R[k] = Rc[k];
    for (int k = 0; k < 3; k++) t[k] = sg * tt[k];
}
template <int NP> MP_HD void recover_pose(const double *E, const double (*p0)[2], const double (*p1)[2], double dist,
                                          double *R, double *t) {
    double Ra[9], Rb[9], tt[3];
    essential_factor(E, Ra, Rb, tt);
    int good[4];
    for (int c = 0; c < 4; c++) {
        good[c] = 0;
        for (int i = 0; i < NP; i++) good[c] += recover_pose_point_ok(Ra, Rb, tt, c, p0[i], p1[i], dist) ? 1 : 0;
    }
    recover_pose_select(Ra, Rb, tt, good, R, t);
}

} // namespace mp
