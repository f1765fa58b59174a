// ORACLE (test infrastructure only).  Point-only relative pose solvers the hybrid
// estimators mix in, plus the geometry that follows them.
//
// The reference calls THIRD-PARTY code here that is NOT under /root/reference:
//   PoseLib v2.0.4 (CMakeLists.txt:26-30, FetchContent): relpose_5pt
//     (hybrid_pose_estimator.cpp:134), relpose_6pt_shared_focal
//     (hybrid_pose_shared_focal_estimator.cpp:87), relpose_7pt
//     (hybrid_pose_two_focal_estimator.cpp:116);
//   OpenCV calib3d (unpinned): cv::recoverPose (hybrid_pose_two_focal_estimator.cpp:143).
// Their published algorithms are restated (PARITY UNPINNED against those libraries):
//   5pt  = Nister, "An efficient solution to the five-point relative pose problem" (PAMI 2004):
//          4-D null space, 10 cubic constraints, Gauss-Jordan, det of the 3x3 matrix B(z) ->
//          degree-10 polynomial, real roots by Sturm sequences (as PoseLib does);
//   7pt  = classical: 2-D null space, det(x F0 + F1) = 0 cubic, unit-norm F;
//   6pt  = Li, "A simple solution to the six-point two-view focal-length problem" (ECCV 2006) /
//          Kukelova-Bujnak-Pajdla (BMVC 2008): 3-D null space, det + 9 trace constraints with
//          Q = diag(1,1,w), w = 1/f^2 as hidden variable, det C(w) of degree 15;
//   essential-matrix decomposition + all-point cheirality = the copy of PoseLib's
//          motion_from_essential kept in /root/reference/src/solver.cpp:1219-1285.
// Triangulation, Bougnoux focals and the depth fits follow the reference's own code
// (utils.h:24-53, hybrid_pose_two_focal_estimator.cpp:11-32, hybrid_pose_estimator.cpp:159-182).
#pragma once
#include "om_estimator.h"
#include "om_poly.h"
#include <complex>

namespace om {

// ---- polynomials in three variables, exponent of each variable <= 3 ---------
struct P3 {
    double c[4][4][4];
    P3() { std::memset(c, 0, sizeof(c)); }
    static P3 lin(double cx, double cy, double cz, double c1) {
        P3 p; p.c[1][0][0] = cx; p.c[0][1][0] = cy; p.c[0][0][1] = cz; p.c[0][0][0] = c1; return p;
    }
    static P3 constant(double v) { P3 p; p.c[0][0][0] = v; return p; }
};
inline P3 operator+(const P3 &a, const P3 &b) {
    P3 r;
    for (int i = 0; i < 64; i++) (&r.c[0][0][0])[i] = (&a.c[0][0][0])[i] + (&b.c[0][0][0])[i];
    return r;
}
inline P3 operator-(const P3 &a, const P3 &b) {
    P3 r;
    for (int i = 0; i < 64; i++) (&r.c[0][0][0])[i] = (&a.c[0][0][0])[i] - (&b.c[0][0][0])[i];
    return r;
}
inline P3 operator*(const P3 &a, double s) {
    P3 r;
    for (int i = 0; i < 64; i++) (&r.c[0][0][0])[i] = (&a.c[0][0][0])[i] * s;
    return r;
}
inline P3 operator*(const P3 &a, const P3 &b) {
    P3 r;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) for (int k = 0; k < 4; k++) {
        double v = a.c[i][j][k];
        if (v == 0) continue;
        for (int l = 0; i + l < 4; l++) for (int m = 0; j + m < 4; m++) for (int n = 0; k + n < 4; n++)
            r.c[i + l][j + m][k + n] += v * b.c[l][m][n];
    }
    return r;
}
inline double p3_eval(const P3 &p, double x, double y, double z, double *g = nullptr) {
    double v = 0, gx = 0, gy = 0, gz = 0, px[4] = {1, x, x * x, x * x * x}, py[4] = {1, y, y * y, y * y * y},
           pz[4] = {1, z, z * z, z * z * z};
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) for (int k = 0; k < 4; k++) {
        double c = p.c[i][j][k];
        if (c == 0) continue;
        v += c * px[i] * py[j] * pz[k];
        if (i) gx += c * i * px[i - 1] * py[j] * pz[k];
        if (j) gy += c * j * px[i] * py[j - 1] * pz[k];
        if (k) gz += c * k * px[i] * py[j] * pz[k - 1];
    }
    if (g) { g[0] = gx; g[1] = gy; g[2] = gz; }
    return v;
}

struct Pose { double R[9], t[3]; };

// all-point cheirality with min_depth = 0 (solver.cpp:1208-1217)
inline bool cheirality_all(const double *R, const double *t, int n, const double (*x1)[3], const double (*x2)[3]) {
    for (int i = 0; i < n; i++) if (!cheirality(R, t, x1[i], x2[i], 0.0)) return false;
    return true;
}

// The four (R, t) factorisations of E in the construction of solver.cpp:1219-1285
// (copy of PoseLib's motion_from_essential): order (Ra,t), (Ra,-t), (Rb,-t), (Rb,t).
inline void essential_factor(const double *E, double *Ra, double *Rb, double *t) {
    double c0[3] = {E[0], E[3], E[6]}, c1[3] = {E[1], E[4], E[7]}, c2[3] = {E[2], E[5], E[8]};
    double u12[3], u13[3], u23[3];
    cross3(c0, c1, u12); cross3(c0, c2, u13); cross3(c1, c2, u23);
    double n12 = dot3(u12, u12), n13 = dot3(u13, u13), n23 = dot3(u23, u23);
    double w1[3], w2[3], w0[3];
    const double *a, *u; double nn;
    if (n12 > n13) { if (n12 > n23) { a = c0; u = u12; nn = n12; } else { a = c1; u = u23; nn = n23; } }
    else           { if (n13 > n23) { a = c0; u = u13; nn = n13; } else { a = c1; u = u23; nn = n23; } }
    double na = norm3(a), su = std::sqrt(nn);
    for (int k = 0; k < 3; k++) { w1[k] = a[k] / na; w2[k] = u[k] / su; }
    cross3(w2, w1, w0);
    for (int k = 0; k < 3; k++) w0[k] = -w0[k];
    double v0[3], v1[3], v2[3];
    mat3T_vec(E, w1, v0);               // row0 = w1^T E
    mat3T_vec(E, w0, v1);               // row1 = -w0^T E
    for (int k = 0; k < 3; k++) v1[k] = -v1[k];
    double n0 = norm3(v0);
    for (int k = 0; k < 3; k++) v0[k] /= n0;
    double d = dot3(v0, v1);
    for (int k = 0; k < 3; k++) v1[k] -= d * v0[k];
    double n1 = norm3(v1);
    for (int k = 0; k < 3; k++) v1[k] /= n1;
    cross3(v0, v1, v2);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            Ra[3 * i + j] = w0[i] * v0[j] + w1[i] * v1[j] + w2[i] * v2[j];
            Rb[3 * i + j] = -w0[i] * v0[j] - w1[i] * v1[j] + w2[i] * v2[j];
        }
    for (int k = 0; k < 3; k++) t[k] = w2[k];
}

inline void motion_from_essential(const double *E, int n, const double (*x1)[3], const double (*x2)[3],
                                  std::vector<Pose> *out) {
    double Ra[9], Rb[9], t[3], tn[3];
    essential_factor(E, Ra, Rb, t);
    for (int k = 0; k < 3; k++) tn[k] = -t[k];
    const double *Rs[4] = {Ra, Ra, Rb, Rb};
    const double *ts[4] = {t, tn, tn, t};
    for (int c = 0; c < 4; c++) {
        if (!std::isfinite(Rs[c][0] + Rs[c][4] + Rs[c][8])) continue;
        if (cheirality_all(Rs[c], ts[c], n, x1, x2)) {
            Pose p; std::memcpy(p.R, Rs[c], 72); std::memcpy(p.t, ts[c], 24); out->push_back(p);
        }
    }
}

// rows of the epipolar constraint x2^T E x1 = 0 for row-major vec(E)
inline void epipolar_rows(int n, const double (*x1)[3], const double (*x2)[3], double *A) {
    for (int r = 0; r < n; r++)
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) A[9 * r + 3 * i + j] = x2[r][i] * x1[r][j];
}

// ---- 5-point (calibrated) -----------------------------------------------------
// x1,x2: 5 unit bearings.  Returns essential matrices (row-major).
inline int relpose_5pt_essential(const double (*x1)[3], const double (*x2)[3], std::vector<std::vector<double>> *Es) {
    double A[45], N[36];
    epipolar_rows(5, x1, x2, A);
    nullspace(5, 9, A, N);
    P3 E[9];
    for (int k = 0; k < 9; k++) E[k] = P3::lin(N[k], N[9 + k], N[18 + k], N[27 + k]);
    // EEt, trace, det
    P3 EEt[9], tr;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
            if (i == j) tr = tr + EEt[3 * i + j];
        }
    P3 eq[10];
    eq[0] = E[0] * (E[4] * E[8] - E[5] * E[7]) - E[1] * (E[3] * E[8] - E[5] * E[6]) + E[2] * (E[3] * E[7] - E[4] * E[6]);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            P3 s = EEt[3 * i] * E[j] + EEt[3 * i + 1] * E[3 + j] + EEt[3 * i + 2] * E[6 + j];
            eq[1 + 3 * i + j] = s * 2.0 - tr * E[3 * i + j];
        }
    // Nister's monomial order
    static const int mono[20][3] = {{3,0,0},{0,3,0},{2,1,0},{1,2,0},{2,0,1},{2,0,0},{0,2,1},{0,2,0},{1,1,1},{1,1,0},
                                    {1,0,2},{1,0,1},{1,0,0},{0,1,2},{0,1,1},{0,1,0},{0,0,3},{0,0,2},{0,0,1},{0,0,0}};
    std::vector<double> L(100), Rr(100);
    for (int r = 0; r < 10; r++)
        for (int c = 0; c < 20; c++) {
            double v = eq[r].c[mono[c][0]][mono[c][1]][mono[c][2]];
            if (c < 10) L[10 * r + c] = v; else Rr[10 * r + c - 10] = v;
        }
    if (!lu_solve(10, 10, L, Rr)) return 0; // rows now: mono_r + sum_c Rr[r][c] mono_{10+c}
    // B(z): rows k = e - z f, l = g - z h, m = i - z j (rows 4..9)
    std::vector<double> B[3][3];
    for (int q = 0; q < 3; q++) {
        const double *e = &Rr[10 * (4 + 2 * q)], *f = &Rr[10 * (5 + 2 * q)];
        B[q][0] = {e[2], e[1] - f[2], e[0] - f[1], -f[0]};
        B[q][1] = {e[5], e[4] - f[5], e[3] - f[4], -f[3]};
        B[q][2] = {e[9], e[8] - f[9], e[7] - f[8], e[6] - f[7], -f[6]};
    }
    auto mul = [](const std::vector<double> &a, const std::vector<double> &b) {
        std::vector<double> r(a.size() + b.size() - 1, 0.0);
        for (size_t i = 0; i < a.size(); i++) for (size_t j = 0; j < b.size(); j++) r[i + j] += a[i] * b[j];
        return r;
    };
    auto sub = [](std::vector<double> a, const std::vector<double> &b) {
        if (a.size() < b.size()) a.resize(b.size(), 0.0);
        for (size_t i = 0; i < b.size(); i++) a[i] -= b[i];
        return a;
    };
    auto add = [](std::vector<double> a, const std::vector<double> &b) {
        if (a.size() < b.size()) a.resize(b.size(), 0.0);
        for (size_t i = 0; i < b.size(); i++) a[i] += b[i];
        return a;
    };
    std::vector<double> m0 = sub(mul(B[0][1], B[1][2]), mul(B[0][2], B[1][1]));
    std::vector<double> m1 = sub(mul(B[0][2], B[1][0]), mul(B[0][0], B[1][2]));
    std::vector<double> m2 = sub(mul(B[0][0], B[1][1]), mul(B[0][1], B[1][0]));
    std::vector<double> det = add(add(mul(m0, B[2][0]), mul(m1, B[2][1])), mul(m2, B[2][2]));
    det.resize(11, 0.0);
    int cnt = 0;
    for (double z : poly_real_roots(det)) {
        double d = poly_eval(m2, z);
        if (d == 0 || !std::isfinite(d)) continue;
        double x = poly_eval(m0, z) / d, y = poly_eval(m1, z) / d;
        // Gauss-Newton polish of (x, y, z) on the ten cubic constraints: the expanded degree-10
        // determinant only locates the root (its coefficients lose digits)
        for (int it = 0; it < 5; it++) {
            double JtJ[9] = {0}, Jtr[3] = {0};
            for (int r = 0; r < 10; r++) {
                double g[3], f = p3_eval(eq[r], x, y, z, g);
                for (int i = 0; i < 3; i++) { Jtr[i] += g[i] * f; for (int j = 0; j < 3; j++) JtJ[3 * i + j] += g[i] * g[j]; }
            }
            double inv[9], st[3];
            if (!mat3_inv(JtJ, inv)) break;
            mat3_vec(inv, Jtr, st);
            if (!std::isfinite(st[0] + st[1] + st[2])) break;
            x -= st[0]; y -= st[1]; z -= st[2];
            if (std::fabs(st[0]) <= 1e-15 * std::fabs(x) && std::fabs(st[1]) <= 1e-15 * std::fabs(y) && std::fabs(st[2]) <= 1e-15 * std::fabs(z)) break;
        }
        std::vector<double> Ev(9);
        for (int k = 0; k < 9; k++) Ev[k] = x * N[k] + y * N[9 + k] + z * N[18 + k] + N[27 + k];
        bool ok = true;
        for (double v : Ev) ok = ok && std::isfinite(v);
        if (!ok) continue;
        Es->push_back(Ev);
        cnt++;
    }
    return cnt;
}

inline int relpose_5pt(const double (*x1)[3], const double (*x2)[3], std::vector<Pose> *poses) {
    std::vector<std::vector<double>> Es;
    relpose_5pt_essential(x1, x2, &Es);
    for (auto &E : Es) motion_from_essential(E.data(), 5, x1, x2, poses);
    return (int)poses->size();
}

// ---- 7-point (uncalibrated) -----------------------------------------------------
inline int relpose_7pt(const double (*x1)[3], const double (*x2)[3], std::vector<std::vector<double>> *Fs) {
    double A[63], N[18];
    epipolar_rows(7, x1, x2, A);
    nullspace(7, 9, A, N);
    const double *F0 = N, *F1 = N + 9;
    double Fp[9], Fm[9];
    for (int k = 0; k < 9; k++) { Fp[k] = F1[k] + F0[k]; Fm[k] = F1[k] - F0[k]; }
    double c0 = mat3_det(F1), c3 = mat3_det(F0), dp = mat3_det(Fp), dm = mat3_det(Fm);
    double c2 = 0.5 * (dp + dm) - c0, c1 = 0.5 * (dp - dm) - c3;
    for (double x : poly_real_roots({c0, c1, c2, c3})) {
        std::vector<double> F(9);
        double nn = 0;
        for (int k = 0; k < 9; k++) { F[k] = x * F0[k] + F1[k]; nn += F[k] * F[k]; }
        nn = std::sqrt(nn);
        if (!(nn > 0) || !std::isfinite(nn)) continue;
        for (double &v : F) v /= nn;
        Fs->push_back(F);
    }
    return (int)Fs->size();
}

} // namespace om
#include "om_relpose2.h"
