// K2 (part 1) -- point-only relative-pose solvers the hybrid estimators mix in, fp64,
// one thread per sample.  The reference calls third-party code here (PoseLib v2.0.4
// relpose_5pt / relpose_7pt: hybrid_pose_estimator.cpp:134, ...two_focal....cpp:116);
// the algorithms are the published ones (Nister 2004 five-point; classical seven-point),
// written for fixed-size local arrays: constraint polynomials are expanded with
// hand-indexed products of linear / quadratic forms instead of generic polynomial types.
// The (R, t) factorisation of E follows the construction the reference keeps in
// solver.cpp:1219-1285 (order of the four candidates included).
#pragma once
#include "mp_poly.cuh"
#include "mp_small.cuh"

namespace mp {

struct Pose { double R[9], t[3]; };

// poselib::check_cheirality (identical copy: solver.cpp:1188-1206)
MP_HD bool cheirality(const double *R, const double *t, const double *x1, const double *x2, double min_depth) {
    double Rx1[3];
    mat3_vecp(R, x1, Rx1);
    const double a = -dot3(Rx1, x2), b1 = -dot3(Rx1, t), b2 = dot3(x2, t);
    const double l1 = b1 - a * b2, l2 = -a * b1 + b2, md = min_depth * (1 - a * a);
    return l1 > md && l2 > md;
}

// E -> (Ra, Rb, t): candidates (Ra,t), (Ra,-t), (Rb,-t), (Rb,t)
MP_HD void essential_factor(const double *E, double *Ra, double *Rb, double *t) {
    const double c0[3] = {E[0], E[3], E[6]}, c1[3] = {E[1], E[4], E[7]}, c2[3] = {E[2], E[5], E[8]};
    double u12[3], u13[3], u23[3];
    cross3(c0, c1, u12); cross3(c0, c2, u13); cross3(c1, c2, u23);
    const double n12 = dot3(u12, u12), n13 = dot3(u13, u13), n23 = dot3(u23, u23);
    const double *a, *u;
    double nn;
    if (n12 > n13) { if (n12 > n23) { a = c0; u = u12; nn = n12; } else { a = c1; u = u23; nn = n23; } }
    else           { if (n13 > n23) { a = c0; u = u13; nn = n13; } else { a = c1; u = u23; nn = n23; } }
    double w0[3], w1[3], w2[3];
    const double na = norm3(a), su = sqrt(nn);
    for (int k = 0; k < 3; k++) { w1[k] = a[k] / na; w2[k] = u[k] / su; }
    cross3(w2, w1, w0);
    for (int k = 0; k < 3; k++) w0[k] = -w0[k];
    double v0[3], v1[3], v2[3];
    mat3T_vecp(E, w1, v0);
    mat3T_vecp(E, w0, v1);
    for (int k = 0; k < 3; k++) v1[k] = -v1[k];
    const double n0 = norm3(v0);
    for (int k = 0; k < 3; k++) v0[k] /= n0;
    const double d = dot3(v0, v1);
    for (int k = 0; k < 3; k++) v1[k] -= d * v0[k];
    const double n1 = norm3(v1);
    for (int k = 0; k < 3; k++) v1[k] /= n1;
    cross3(v0, v1, v2);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            Ra[3 * i + j] = w0[i] * v0[j] + w1[i] * v1[j] + w2[i] * v2[j];
            Rb[3 * i + j] = -w0[i] * v0[j] - w1[i] * v1[j] + w2[i] * v2[j];
        }
    for (int k = 0; k < 3; k++) t[k] = w2[k];
}

// appends the candidates of E for which ALL n bearing pairs pass cheirality (min depth 0)
template <int NP> MP_HD int motion_from_essential(const double *E, const double (*x1)[3], const double (*x2)[3],
                                                   Pose *out, int cap) {
    double Ra[9], Rb[9], t[3], tn[3];
    essential_factor(E, Ra, Rb, t);
    for (int k = 0; k < 3; k++) tn[k] = -t[k];
    int n = 0;
    for (int c = 0; c < 4 && n < cap; c++) {
        const double *R = c < 2 ? Ra : Rb, *tt = (c == 0 || c == 3) ? t : tn;
        if (!finite_d(R[0] + R[4] + R[8])) continue;
        bool ok = true;
        for (int i = 0; i < NP && ok; i++) ok = cheirality(R, tt, x1[i], x2[i], 0.0);
        if (!ok) continue;
        for (int k = 0; k < 9; k++) out[n].R[k] = R[k];
        for (int k = 0; k < 3; k++) out[n].t[k] = tt[k];
        n++;
    }
    return n;
}

template <int NP> MP_HD void epipolar_rows(const double (*x1)[3], const double (*x2)[3], double *A) {
    for (int r = 0; r < NP; r++)
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) A[9 * r + 3 * i + j] = x2[r][i] * x1[r][j];
}

// ---- forms in (x, y, z): linear [x y z 1]; quadratic [xx xy xz yy yz zz x y z 1];
// cubic [xxx xxy xxz xyy xyz xzz yyy yyz yzz zzz xx xy xz yy yz zz x y z 1]
MP_HD void mul_ll3(const double *a, const double *b, double *q) {
    q[0] = a[0] * b[0]; q[1] = a[0] * b[1] + a[1] * b[0]; q[2] = a[0] * b[2] + a[2] * b[0];
    q[3] = a[1] * b[1]; q[4] = a[1] * b[2] + a[2] * b[1]; q[5] = a[2] * b[2];
    q[6] = a[0] * b[3] + a[3] * b[0]; q[7] = a[1] * b[3] + a[3] * b[1]; q[8] = a[2] * b[3] + a[3] * b[2];
    q[9] = a[3] * b[3];
}
// c += s * q * l
MP_HD void fma_ql3(const double *q, const double *l, double s, double *c) {
    const double l0 = l[0] * s, l1 = l[1] * s, l2 = l[2] * s, l3 = l[3] * s;
    c[0] += q[0] * l0;
    c[1] += q[0] * l1 + q[1] * l0;
    c[2] += q[0] * l2 + q[2] * l0;
    c[3] += q[1] * l1 + q[3] * l0;
    c[4] += q[1] * l2 + q[2] * l1 + q[4] * l0;
    c[5] += q[2] * l2 + q[5] * l0;
    c[6] += q[3] * l1;
    c[7] += q[3] * l2 + q[4] * l1;
    c[8] += q[4] * l2 + q[5] * l1;
    c[9] += q[5] * l2;
    c[10] += q[0] * l3 + q[6] * l0;
    c[11] += q[1] * l3 + q[6] * l1 + q[7] * l0;
    c[12] += q[2] * l3 + q[6] * l2 + q[8] * l0;
    c[13] += q[3] * l3 + q[7] * l1;
    c[14] += q[4] * l3 + q[7] * l2 + q[8] * l1;
    c[15] += q[5] * l3 + q[8] * l2;
    c[16] += q[6] * l3 + q[9] * l0;
    c[17] += q[7] * l3 + q[9] * l1;
    c[18] += q[8] * l3 + q[9] * l2;
    c[19] += q[9] * l3;
}

// Five-point essential matrices.  x1, x2: 5 unit bearings.  Es: up to 10 row-major 3x3.
MP_HD int essential_5pt(const double (*x1)[3], const double (*x2)[3], double (*Es)[9]) {
    double A[45], N[36];
    epipolar_rows<5>(x1, x2, A);
    nullspace_qr<5, 9>(A, N);
    double L[9][4]; // E_k = x N[k] + y N[9+k] + z N[18+k] + N[27+k]
    for (int k = 0; k < 9; k++) { L[k][0] = N[k]; L[k][1] = N[9 + k]; L[k][2] = N[18 + k]; L[k][3] = N[27 + k]; }
    // symmetric G = E E^T (quadratic forms), index map for (i,j)
    double G[6][10], tmp[10];
    const int gi[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
    for (int i = 0; i < 3; i++)
        for (int j = i; j < 3; j++) {
            double *g = G[gi[i][j]];
            mul_ll3(L[3 * i], L[3 * j], g);
            for (int k = 1; k < 3; k++) {
                mul_ll3(L[3 * i + k], L[3 * j + k], tmp);
                for (int m = 0; m < 10; m++) g[m] += tmp[m];
            }
        }
    double tr[10];
    for (int m = 0; m < 10; m++) tr[m] = G[0][m] + G[3][m] + G[5][m];
    // cubic constraints, own monomial order; row 0 = det, rows 1..9 = 2 E E^T E - tr(E E^T) E
    double C[10][20];
    for (int r = 0; r < 10; r++)
        for (int m = 0; m < 20; m++) C[r][m] = 0.0;
    {
        double q1[10], q2[10];
        mul_ll3(L[4], L[8], q1); mul_ll3(L[5], L[7], q2);
        for (int m = 0; m < 10; m++) q1[m] -= q2[m];
        fma_ql3(q1, L[0], 1.0, C[0]);
        mul_ll3(L[3], L[8], q1); mul_ll3(L[5], L[6], q2);
        for (int m = 0; m < 10; m++) q1[m] -= q2[m];
        fma_ql3(q1, L[1], -1.0, C[0]);
        mul_ll3(L[3], L[7], q1); mul_ll3(L[4], L[6], q2);
        for (int m = 0; m < 10; m++) q1[m] -= q2[m];
        fma_ql3(q1, L[2], 1.0, C[0]);
    }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double *c = C[1 + 3 * i + j];
            for (int k = 0; k < 3; k++) fma_ql3(G[gi[i][k]], L[3 * k + j], 2.0, c);
            fma_ql3(tr, L[3 * i + j], -1.0, c);
        }
    // Nister's elimination order: [xxx yyy xxy xyy xxz xx yyz yy xyz xy | xzz xz x yzz yz y zzz zz z 1]
    const int ord[20] = {0, 6, 1, 3, 2, 10, 7, 13, 4, 11, 5, 12, 16, 8, 14, 17, 9, 15, 18, 19};
    double Lm[100], Rm[100];
    for (int r = 0; r < 10; r++)
        for (int c = 0; c < 10; c++) { Lm[10 * r + c] = C[r][ord[c]]; Rm[10 * r + c] = C[r][ord[10 + c]]; }
    if (!lu_solve_inplace<10, 10>(Lm, Rm)) return 0;
    // B(z): three rows (x-coefficient deg 3, y-coefficient deg 3, constant deg 4)
    double B[3][3][5];
    for (int q = 0; q < 3; q++) {
        const double *e = Rm + 10 * (4 + 2 * q), *f = Rm + 10 * (5 + 2 * q);
        B[q][0][0] = e[2]; B[q][0][1] = e[1] - f[2]; B[q][0][2] = e[0] - f[1]; B[q][0][3] = -f[0]; B[q][0][4] = 0;
        B[q][1][0] = e[5]; B[q][1][1] = e[4] - f[5]; B[q][1][2] = e[3] - f[4]; B[q][1][3] = -f[3]; B[q][1][4] = 0;
        B[q][2][0] = e[9]; B[q][2][1] = e[8] - f[9]; B[q][2][2] = e[7] - f[8]; B[q][2][3] = e[6] - f[7]; B[q][2][4] = -f[6];
    }
    auto pm = [](const double *a, int na, const double *b, int nb, double s, double *out) { // out += s*a*b
        for (int i = 0; i <= na; i++)
            for (int j = 0; j <= nb; j++) out[i + j] = fma(s * a[i], b[j], out[i + j]);
    };
    double m0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, m1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, m2[7] = {0, 0, 0, 0, 0, 0, 0};
    pm(B[0][1], 3, B[1][2], 4, 1.0, m0); pm(B[0][2], 4, B[1][1], 3, -1.0, m0);
    pm(B[0][2], 4, B[1][0], 3, 1.0, m1); pm(B[0][0], 3, B[1][2], 4, -1.0, m1);
    pm(B[0][0], 3, B[1][1], 3, 1.0, m2); pm(B[0][1], 3, B[1][0], 3, -1.0, m2);
    double det[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, roots[10];
    pm(m0, 7, B[2][0], 3, 1.0, det); pm(m1, 7, B[2][1], 3, 1.0, det); pm(m2, 6, B[2][2], 4, 1.0, det);
    const int nr = real_roots<10>(det, 10, roots);
    int n = 0;
    for (int k = 0; k < nr; k++) {
        double z = roots[k];
        const double dd = poly_val(m2, 6, z);
        if (dd == 0.0 || !finite_d(dd)) continue;
        double x = poly_val(m0, 7, z) / dd, y = poly_val(m1, 7, z) / dd;
        // Gauss-Newton polish of (x, y, z) on the ten cubic constraints (the expanded degree-10
        // determinant only locates the root)
        for (int it = 0; it < 5; it++) {
            const double mo[20] = {x * x * x, x * x * y, x * x * z, x * y * y, x * y * z, x * z * z, y * y * y, y * y * z,
                                   y * z * z, z * z * z, x * x, x * y, x * z, y * y, y * z, z * z, x, y, z, 1.0};
            const double mx[20] = {3 * x * x, 2 * x * y, 2 * x * z, y * y, y * z, z * z, 0, 0, 0, 0, 2 * x, y, z, 0, 0, 0, 1, 0, 0, 0};
            const double my[20] = {0, x * x, 0, 2 * x * y, x * z, 0, 3 * y * y, 2 * y * z, z * z, 0, 0, x, 0, 2 * y, z, 0, 0, 1, 0, 0};
            const double mz[20] = {0, 0, x * x, 0, x * y, 2 * x * z, 0, y * y, 2 * y * z, 3 * z * z, 0, 0, x, 0, y, 2 * z, 0, 0, 1, 0};
            double JtJ[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Jtr[3] = {0, 0, 0};
            for (int r = 0; r < 10; r++) {
                double f = 0, g[3] = {0, 0, 0};
                for (int c = 0; c < 20; c++) { f += C[r][c] * mo[c]; g[0] += C[r][c] * mx[c]; g[1] += C[r][c] * my[c]; g[2] += C[r][c] * mz[c]; }
                for (int i = 0; i < 3; i++) {
                    Jtr[i] += g[i] * f;
                    for (int j = 0; j < 3; j++) JtJ[3 * i + j] += g[i] * g[j];
                }
            }
            double inv[9], st[3];
            if (!inv3(JtJ, inv)) break;
            mat3_vecp(inv, Jtr, st);
            if (!finite_d(st[0] + st[1] + st[2])) break;
            x -= st[0]; y -= st[1]; z -= st[2];
            if (fabs(st[0]) <= 1e-15 * fabs(x) && fabs(st[1]) <= 1e-15 * fabs(y) && fabs(st[2]) <= 1e-15 * fabs(z)) break;
        }
        bool ok = true;
        for (int m = 0; m < 9; m++) {
            Es[n][m] = x * N[m] + y * N[9 + m] + z * N[18 + m] + N[27 + m];
            ok = ok && finite_d(Es[n][m]);
        }
        if (ok) n++;
    }
    return n;
}

MP_HD int relpose_5pt(const double (*x1)[3], const double (*x2)[3], Pose *poses /*[10]*/) {
    double Es[10][9];
    const int ne = essential_5pt(x1, x2, Es);
    int n = 0;
    for (int k = 0; k < ne && n < 10; k++) n += motion_from_essential<5>(Es[k], x1, x2, poses + n, 10 - n);
    return n;
}

// Seven-point fundamental matrices (unit Frobenius norm).  <= 3 solutions
MP_HD int fundamental_7pt(const double (*x1)[3], const double (*x2)[3], double (*Fs)[9]) {
    double A[63], N[18];
    epipolar_rows<7>(x1, x2, A);
    nullspace_qr<7, 9>(A, N);
    const double *F0 = N, *F1 = N + 9;
    double Fp[9], Fm[9];
    for (int k = 0; k < 9; k++) { Fp[k] = F1[k] + F0[k]; Fm[k] = F1[k] - F0[k]; }
    // det(x F0 + F1) = c3 x^3 + c2 x^2 + c1 x + c0 from four determinant evaluations
    const double c0 = det3(F1), c3 = det3(F0), dp = det3(Fp), dm = det3(Fm);
    const double cub[4] = {c0, 0.5 * (dp - dm) - c3, 0.5 * (dp + dm) - c0, c3};
    double roots[3];
    const int nr = real_roots<3>(cub, 3, roots);
    int n = 0;
    for (int k = 0; k < nr; k++) {
        double nn = 0;
        for (int m = 0; m < 9; m++) { Fs[n][m] = roots[k] * F0[m] + F1[m]; nn = fma(Fs[n][m], Fs[n][m], nn); }
        nn = sqrt(nn);
        if (!(nn > 0) || !finite_d(nn)) continue;
        for (int m = 0; m < 9; m++) Fs[n][m] /= nn;
        n++;
    }
    return n;
}

} // namespace mp
#include "mp_relpose2.cuh"
