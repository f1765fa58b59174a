// ORACLE (test infrastructure only) -- continuation of om_relpose.h: 6-point shared
// focal, triangulation, Bougnoux focals, recoverPose and the depth-affine fits.
#pragma once

namespace om {

typedef std::complex<double> cplx;

template <class T> inline T det_lu(int n, std::vector<T> A) {
    T det = 1.0;
    for (int c = 0; c < n; c++) {
        int piv = c;
        for (int r = c + 1; r < n; r++) if (std::abs(A[r * n + c]) > std::abs(A[piv * n + c])) piv = r;
        if (std::abs(A[piv * n + c]) == 0.0) return T(0.0);
        if (piv != c) { for (int k = 0; k < n; k++) std::swap(A[c * n + k], A[piv * n + k]); det = -det; }
        det *= A[c * n + c];
        for (int r = c + 1; r < n; r++) {
            T f = A[r * n + c] / A[c * n + c];
            for (int k = c; k < n; k++) A[r * n + k] -= f * A[c * n + k];
        }
    }
    return det;
}

struct FocalPose { Pose pose; double focal; };

// Is E an essential matrix?  || 2 E E^T E - tr(E E^T) E ||_F <= tol on the unit-norm matrix.  The hidden-variable
// determinant has real roots that are not solutions of the ten equations (the polish then stops at a least-squares
// point with a residual); an exact solver (PoseLib's action matrix) never returns those, so they are dropped.
inline bool essential_valid(const double *Ein, double tol) {
    double n = 0, E[9], EEt[9], T[9];
    for (int k = 0; k < 9; k++) n += Ein[k] * Ein[k];
    n = std::sqrt(n);
    if (!(n > 0) || !std::isfinite(n)) return false;
    for (int k = 0; k < 9; k++) E[k] = Ein[k] / n;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) EEt[3 * i + j] = E[3 * i] * E[3 * j] + E[3 * i + 1] * E[3 * j + 1] + E[3 * i + 2] * E[3 * j + 2];
    const double tr = EEt[0] + EEt[4] + EEt[8];
    mat3_mul(EEt, E, T);
    double r = 0;
    for (int k = 0; k < 9; k++) { const double v = 2 * T[k] - tr * E[k]; r += v * v; }
    return std::sqrt(r) <= tol;
}

// x1,x2: 6 points (x, y, 1) in (normalised) pixel units, shared unknown focal.
// Returns (pose, focal) candidates.  madpose passes unit-normalised bearings to PoseLib;
// the constraints are homogeneous per point, so the unnormalised (x,y,1) are used here.
inline int relpose_6pt_shared_focal(const double (*x1)[3], const double (*x2)[3], std::vector<FocalPose> *out) {
    double A[54], N[27];
    epipolar_rows(6, x1, x2, A);
    nullspace(6, 9, A, N);
    // F = a F0 + b F1 + F2 ; third variable of P3 is w = 1/f^2
    P3 F[9], w = P3::lin(0, 0, 1, 0), one = P3::constant(1.0);
    for (int k = 0; k < 9; k++) F[k] = P3::lin(N[k], N[9 + k], 0, N[18 + k]);
    const P3 *Q[3] = {&one, &one, &w};
    // G = F Q F^T Q
    P3 G[9], tr;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            P3 s;
            for (int k = 0; k < 3; k++) s = s + F[3 * i + k] * F[3 * j + k] * (*Q[k]);
            G[3 * i + j] = s * (*Q[j]);
            if (i == j) tr = tr + G[3 * i + j];
        }
    P3 eq[10];
    eq[0] = F[0] * (F[4] * F[8] - F[5] * F[7]) - F[1] * (F[3] * F[8] - F[5] * F[6]) + F[2] * (F[3] * F[7] - F[4] * F[6]);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            P3 s = G[3 * i] * F[j] + G[3 * i + 1] * F[3 + j] + G[3 * i + 2] * F[6 + j];
            eq[1 + 3 * i + j] = s * 2.0 - tr * F[3 * i + j];
        }
    static const int mono[10][2] = {{3,0},{2,1},{1,2},{0,3},{2,0},{1,1},{0,2},{1,0},{0,1},{0,0}};
    // C(w) = C0 + C1 w + C2 w^2, rows scaled to unit max coefficient
    double Cw[3][100];
    for (int r = 0; r < 10; r++) {
        double mx = 0;
        for (int c = 0; c < 10; c++) for (int k = 0; k < 3; k++) mx = std::max(mx, std::fabs(eq[r].c[mono[c][0]][mono[c][1]][k]));
        if (!(mx > 0)) return 0;
        for (int c = 0; c < 10; c++) for (int k = 0; k < 3; k++) Cw[k][10 * r + c] = eq[r].c[mono[c][0]][mono[c][1]][k] / mx;
    }
    auto detC = [&](auto wv) {
        typedef decltype(wv) T;
        std::vector<T> M(100);
        for (int i = 0; i < 100; i++) M[i] = Cw[0][i] + wv * (Cw[1][i] + wv * Cw[2][i]);
        return det_lu<T>(10, M);
    };
    // det C(w) has degree 15: sample on a circle, inverse DFT -> coefficients (root locator)
    const int NP = 16;
    const double rad = 1.0, PI = 3.14159265358979323846;
    cplx vals[NP];
    for (int k = 0; k < NP; k++) vals[k] = detC(cplx(rad * std::cos(2 * PI * k / NP), rad * std::sin(2 * PI * k / NP)));
    std::vector<double> coef(NP);
    for (int j = 0; j < NP; j++) {
        cplx s = 0;
        for (int k = 0; k < NP; k++) s += vals[k] * cplx(std::cos(2 * PI * j * k / NP), -std::sin(2 * PI * j * k / NP));
        coef[j] = (s / (double)NP).real() / std::pow(rad, j);
    }
    int cnt = 0;
    for (double wr : poly_real_roots(coef)) {
        if (!(wr > 0)) continue;
        // null vector of C(w): Gaussian elimination with full pivoting, free variable = 1
        double M[100];
        for (int i = 0; i < 100; i++) M[i] = Cw[0][i] + wr * (Cw[1][i] + wr * Cw[2][i]);
        int cp[10];
        for (int i = 0; i < 10; i++) cp[i] = i;
        for (int c = 0; c < 9; c++) {
            int pr = c, pc = c;
            for (int r = c; r < 10; r++) for (int k = c; k < 10; k++) if (std::fabs(M[10 * r + k]) > std::fabs(M[10 * pr + pc])) { pr = r; pc = k; }
            for (int k = 0; k < 10; k++) std::swap(M[10 * c + k], M[10 * pr + k]);
            for (int r = 0; r < 10; r++) std::swap(M[10 * r + c], M[10 * r + pc]);
            std::swap(cp[c], cp[pc]);
            for (int r = c + 1; r < 10; r++) {
                double f = M[10 * r + c] / M[10 * c + c];
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
        if (vv[9] == 0 || !std::isfinite(vv[9])) continue;
        double a = vv[7] / vv[9], b = vv[8] / vv[9], ww = wr;
        // Gauss-Newton polish of (a, b, w) on the 10 original equations
        for (int it = 0; it < 5; it++) {
            double JtJ[9] = {0}, Jtr[3] = {0};
            for (int r = 0; r < 10; r++) {
                double g[3], f = p3_eval(eq[r], a, b, ww, g);
                for (int i = 0; i < 3; i++) { Jtr[i] += g[i] * f; for (int j = 0; j < 3; j++) JtJ[3 * i + j] += g[i] * g[j]; }
            }
            double inv[9], st[3];
            if (!mat3_inv(JtJ, inv)) break;
            mat3_vec(inv, Jtr, st);
            if (!std::isfinite(st[0] + st[1] + st[2])) break;
            a -= st[0]; b -= st[1]; ww -= st[2];
            if (std::fabs(st[0]) <= 1e-15 * std::fabs(a) && std::fabs(st[1]) <= 1e-15 * std::fabs(b) && std::fabs(st[2]) <= 1e-15 * std::fabs(ww)) break;
        }
        if (!(ww > 0) || !std::isfinite(a + b + ww)) continue;
        double f = 1.0 / std::sqrt(ww), Fm[9], E[9];
        for (int k = 0; k < 9; k++) Fm[k] = a * N[k] + b * N[9 + k] + N[18 + k];
        // E = K F K, K = diag(f, f, 1)
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) E[3 * i + j] = Fm[3 * i + j] * (i < 2 ? f : 1.0) * (j < 2 ? f : 1.0);
        if (!essential_valid(E, 1e-8)) continue;
        double b1[6][3], b2[6][3];
        for (int i = 0; i < 6; i++) {
            double p[3] = {x1[i][0] / f, x1[i][1] / f, 1.0}, q[3] = {x2[i][0] / f, x2[i][1] / f, 1.0};
            double np = norm3(p), nq = norm3(q);
            for (int k = 0; k < 3; k++) { b1[i][k] = p[k] / np; b2[i][k] = q[k] / nq; }
        }
        std::vector<Pose> poses;
        motion_from_essential(E, 6, b1, b2, &poses);
        for (auto &p : poses) { out->push_back({p, f}); cnt++; }
    }
    return cnt;
}

// COLMAP-style DLT triangulation, utils.h:24-38.  P1, P2: 3x4 row-major.
inline void triangulate(const double *P1, const double *P2, const double *p1, const double *p2, double *X) {
    double A[16], U[16], S[4], V[16];
    for (int k = 0; k < 4; k++) {
        A[k] = p1[0] * P1[8 + k] - P1[k];
        A[4 + k] = p1[1] * P1[8 + k] - P1[4 + k];
        A[8 + k] = p2[0] * P2[8 + k] - P2[k];
        A[12 + k] = p2[1] * P2[8 + k] - P2[4 + k];
    }
    jacobi_svd(4, A, U, S, V);
    for (int k = 0; k < 3; k++) X[k] = V[4 * k + 3] / V[12 + 3];
}

// 2-parameter fit z ~ s*d + c  via the normal equations (hybrid_pose_estimator.cpp:160-163)
inline void fit_affine(int n, const double *d, const double *z, double *s, double *c) {
    double sdd = 0, sd = 0, sdz = 0, sz = 0;
    for (int i = 0; i < n; i++) { sdd += d[i] * d[i]; sd += d[i]; sdz += d[i] * z[i]; sz += z[i]; }
    double det = sdd * n - sd * sd;
    *s = (sdz * n - sd * sz) / det;
    *c = (sdd * sz - sd * sdz) / det;
}

// hybrid_pose_two_focal_estimator.cpp:11-32 with p1 = p2 = (0,0,1)
inline void bougnoux_focals(const double *F, double *f1sq, double *f2sq) {
    auto nullvec = [](const double *M, double *e) { // M e = 0: best cross product of two rows
        double r[3][3], best = -1;
        const double *a = M, *b = M + 3, *c = M + 6;
        cross3(a, b, r[0]); cross3(a, c, r[1]); cross3(b, c, r[2]);
        for (int k = 0; k < 3; k++) { double nn = dot3(r[k], r[k]); if (nn > best) { best = nn; std::memcpy(e, r[k], 24); } }
    };
    double Ft[9], e1[3] = {0, 0, 0}, e2[3] = {0, 0, 0};
    mat3_T(F, Ft);
    nullvec(F, e1);  // F e1 = 0   (V.col(2))
    nullvec(Ft, e2); // F^T e2 = 0 (U.col(2))
    for (int k = 0; k < 2; k++) { e1[k] /= e1[2]; e2[k] /= e2[2]; }
    e1[2] = 1; e2[2] = 1;
    double se1[9] = {0, -e1[2], e1[1], e1[2], 0, -e1[0], -e1[1], e1[0], 0};
    double se2[9] = {0, -e2[2], e2[1], e2[2], 0, -e2[0], -e2[1], e2[0], 0};
    double II[9] = {1, 0, 0, 0, 1, 0, 0, 0, 0}, pp[9] = {0, 0, 0, 0, 0, 0, 0, 0, 1};
    auto quad = [&](const double *S, const double *Fa, const double *Mid, const double *Fb) {
        // (p^T S II Fa Mid Fb p) with p = (0,0,1): entry (2,2) of the product
        double T1[9], T2[9], T3[9], T4[9];
        mat3_mul(S, II, T1); mat3_mul(T1, Fa, T2); mat3_mul(T2, Mid, T3); mat3_mul(T3, Fb, T4);
        return T4[8];
    };
    *f1sq = -quad(se2, F, pp, Ft) / quad(se2, F, II, Ft);
    *f2sq = -quad(se1, Ft, pp, F) / quad(se1, Ft, II, F);
}

// cv::recoverPose(E, pts0, pts1, I, R, t, distanceThresh) restated: four candidates
// (R1,t),(R2,t),(R1,-t),(R2,-t), DLT-triangulate every point, count points in front of
// both cameras and closer than the threshold, first candidate with the maximal count wins.
// OpenCV takes R1,R2,t from its SVD (sign conventions unpinned); here they come from the
// deterministic construction of essential_factor().
inline void recover_pose(const double *E, int n, const double (*p0)[2], const double (*p1)[2], double dist, double *R,
                         double *t) {
    double Ra[9], Rb[9], tt[3], tn[3];
    essential_factor(E, Ra, Rb, tt);
    for (int k = 0; k < 3; k++) tn[k] = -tt[k];
    const double *Rs[4] = {Ra, Rb, Ra, Rb};
    const double *ts[4] = {tt, tt, tn, tn};
    int best = -1, bestc = 0;
    double P0[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    for (int c = 0; c < 4; c++) {
        double P1[12];
        for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) P1[4 * i + j] = Rs[c][3 * i + j]; P1[4 * i + 3] = ts[c][i]; }
        int good = 0;
        for (int i = 0; i < n; i++) {
            double A[16], U[16], S[4], V[16];
            for (int k = 0; k < 4; k++) {
                A[k] = p0[i][0] * P0[8 + k] - P0[k]; A[4 + k] = p0[i][1] * P0[8 + k] - P0[4 + k];
                A[8 + k] = p1[i][0] * P1[8 + k] - P1[k]; A[12 + k] = p1[i][1] * P1[8 + k] - P1[4 + k];
            }
            jacobi_svd(4, A, U, S, V);
            double Q[4] = {V[3], V[7], V[11], V[15]};
            bool ok = Q[2] * Q[3] > 0;
            double X[3] = {Q[0] / Q[3], Q[1] / Q[3], Q[2] / Q[3]};
            ok = ok && X[2] < dist;
            double z1 = P1[8] * X[0] + P1[9] * X[1] + P1[10] * X[2] + P1[11];
            ok = ok && z1 > 0 && z1 < dist;
            good += ok;
        }
        if (best < 0 || good > bestc) { best = c; bestc = good; }
    }
    std::memcpy(R, Rs[best], 72);
    std::memcpy(t, ts[best], 24);
}

} // namespace om
