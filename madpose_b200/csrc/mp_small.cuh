// Small fixed-size dense linear algebra for the per-sample fp64 solvers (K1/K2) and
// the LM normal equations (K5).  Everything is MP_HD scalar code on fixed-size
// arrays: one GPU thread owns one problem; no heap, no STL.
// These stand where the reference calls Eigen (JacobiSVD solver.cpp:517 / utils.h:35,
// colPivHouseholderQr solver.cpp:100, partialPivLu solver.cpp:573, ldlt
// hybrid_pose_estimator.cpp:163) and PoseLib's QR null-space extraction.
#pragma once
#include "mp_common.cuh"

namespace mp {

MP_HD void cross3(const double *a, const double *b, double *c) {
    double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
    c[0] = x; c[1] = y; c[2] = z;
}
MP_HD double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
MP_HD double norm3(const double *a) { return sqrt(dot3(a, a)); }
MP_HD void mat3_vecp(const double *A, const double *x, double *y) { mat3_vec(A, x[0], x[1], x[2], y); }
MP_HD void mat3T_vecp(const double *A, const double *x, double *y) { mat3T_vec(A, x[0], x[1], x[2], y); }
MP_HD void mat3_transpose(const double *A, double *B) {
    double T[9] = {A[0], A[3], A[6], A[1], A[4], A[7], A[2], A[5], A[8]};
    for (int i = 0; i < 9; i++) B[i] = T[i];
}

// One-sided (Hestenes) Jacobi on the columns of the N x N row-major matrix A:
// on return the columns of A are mutually orthogonal (A_out = A_in * V), V orthogonal.
// Column norms are the singular values (unsorted).
template <int N> MP_HD void jacobi_cols(double *A, double *V) {
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) V[i * N + j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 40; sweep++) {
        double off = 0;
        for (int p = 0; p < N - 1; p++)
            for (int q = p + 1; q < N; q++) {
                double a = 0, b = 0, c = 0;
                for (int k = 0; k < N; k++) {
                    double x = A[k * N + p], y = A[k * N + q];
                    a = fma(x, x, a); b = fma(y, y, b); c = fma(x, y, c);
                }
                double lim = sqrt(a * b);
                // a converged pair is "rotated" by the identity instead of skipped: the lanes of a warp stay
                // together (1 * x - 0 * y == x exactly, so the result is the one of the skipping form)
                const bool skip = !(fabs(c) > 1e-300) || !(fabs(c) > 1e-17 * lim);
                double rel = fabs(c) / lim;
                off = (!skip && rel > off) ? rel : off;
                double zeta = (b - a) / (2.0 * c);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
                if (skip) { cs = 1.0; sn = 0.0; }
                for (int k = 0; k < N; k++) {
                    double x = A[k * N + p], y = A[k * N + q];
                    A[k * N + p] = cs * x - sn * y; A[k * N + q] = sn * x + cs * y;
                    x = V[k * N + p]; y = V[k * N + q];
                    V[k * N + p] = cs * x - sn * y; V[k * N + q] = sn * x + cs * y;
                }
            }
        if (off < 1e-16) break;
    }
}

// Rotation of the orthogonal Procrustes problem max tr(R^T S)... i.e. R = U diag(1,1,det) V^T
// for S = U Sigma V^T (Kabsch, solver.cpp:515-525).  Only the two dominant singular
// pairs are used: R = u1 v1^T + u2 v2^T + (u1 x u2)(v1 x v2)^T, which equals the
// det-corrected product also when S has rank 2 (three-point samples).
MP_HD bool procrustes_rotation(const double *S, double *R) {
    double A[9], V[9], nrm[3];
    for (int i = 0; i < 9; i++) A[i] = S[i];
    jacobi_cols<3>(A, V);
    for (int j = 0; j < 3; j++) nrm[j] = sqrt(A[j] * A[j] + A[3 + j] * A[3 + j] + A[6 + j] * A[6 + j]);
    int i0 = 0, i1 = 1, i2 = 2; // sort descending (stable)
    if (nrm[i1] > nrm[i0]) { int t = i0; i0 = i1; i1 = t; }
    if (nrm[i2] > nrm[i1]) { int t = i1; i1 = i2; i2 = t; }
    if (nrm[i1] > nrm[i0]) { int t = i0; i0 = i1; i1 = t; }
    if (!(nrm[i1] > 0.0)) return false;
    double u1[3], u2[3], u3[3], v1[3], v2[3], v3[3];
    for (int k = 0; k < 3; k++) {
        u1[k] = A[3 * k + i0] / nrm[i0]; u2[k] = A[3 * k + i1] / nrm[i1];
        v1[k] = V[3 * k + i0]; v2[k] = V[3 * k + i1];
    }
    cross3(u1, u2, u3); cross3(v1, v2, v3);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) R[3 * i + j] = u1[i] * v1[j] + u2[i] * v2[j] + u3[i] * v3[j];
    return true;
}

// Right singular vector of the smallest singular value of a 4x4 matrix (DLT triangulation, utils.h:24-38), Jacobi form
MP_HD void smallest_right_singular4_jacobi(const double *Ain, double *v) {
    double A[16], V[16];
    for (int i = 0; i < 16; i++) A[i] = Ain[i];
    jacobi_cols<4>(A, V);
    int best = 0;
    double bn = DBL_MAX;
    for (int j = 0; j < 4; j++) {
        double s = A[j] * A[j] + A[4 + j] * A[4 + j] + A[8 + j] * A[8 + j] + A[12 + j] * A[12 + j];
        if (s < bn) { bn = s; best = j; }
    }
    for (int k = 0; k < 4; k++) v[k] = V[4 * k + best];
}

// The same vector by Householder QR + inverse iteration on R^T R (the triangulation matrices have one singular value
// far below the others: the iteration gains (sigma_4 / sigma_3)^2 per step, 2-4 steps in practice) -- an eighth of the
// work of the Jacobi sweeps.  Falls back to the Jacobi form when the iteration has not settled to 1e-14 in 8 steps or
// meets a non-finite value (an exactly singular R).  The sign of v is arbitrary, as with the SVD.
MP_HD void smallest_right_singular4(const double *Ain, double *v) {
    double R[16];
    for (int i = 0; i < 16; i++) R[i] = Ain[i];
    bool ok = true;
    for (int c = 0; c < 3; c++) { // Householder QR, Q is not needed
        double nr = 0;
        for (int i = c; i < 4; i++) nr = fma(R[4 * i + c], R[4 * i + c], nr);
        nr = sqrt(nr);
        if (nr == 0.0) continue;
        const double alpha = R[4 * c + c] > 0 ? -nr : nr;
        double w[4] = {0, 0, 0, 0};
        for (int i = c; i < 4; i++) w[i] = R[4 * i + c];
        w[c] -= alpha;
        double wn = 0;
        for (int i = c; i < 4; i++) wn = fma(w[i], w[i], wn);
        if (wn == 0.0) continue;
        const double inv = 2.0 / wn;
        for (int j = c; j < 4; j++) {
            double t = 0;
            for (int i = c; i < 4; i++) t = fma(w[i], R[4 * i + j], t);
            t *= inv;
            for (int i = c; i < 4; i++) R[4 * i + j] = fma(-t, w[i], R[4 * i + j]);
        }
    }
    double rd[4];
    for (int i = 0; i < 4; i++) { rd[i] = 1.0 / R[5 * i]; ok = ok && finite_d(rd[i]); }
    double x[4] = {0.5, 0.5, 0.5, 0.5};
    bool done = false;
    for (int it = 0; ok && it < 8 && !done; it++) {
        double y[4], z[4];
        for (int i = 0; i < 4; i++) { // R^T y = x (forward)
            double t = x[i];
            for (int k = 0; k < i; k++) t = fma(-R[4 * k + i], y[k], t);
            y[i] = t * rd[i];
        }
        for (int i = 3; i >= 0; i--) { // R z = y (backward)
            double t = y[i];
            for (int k = i + 1; k < 4; k++) t = fma(-R[4 * i + k], z[k], t);
            z[i] = t * rd[i];
        }
        const double nz = sqrt(z[0] * z[0] + z[1] * z[1] + z[2] * z[2] + z[3] * z[3]);
        if (!(nz > 0.0) || !finite_d(nz)) { ok = false; break; }
        const double sg = (z[0] * x[0] + z[1] * x[1] + z[2] * x[2] + z[3] * x[3]) < 0 ? -1.0 / nz : 1.0 / nz;
        double dmax = 0;
        for (int i = 0; i < 4; i++) { const double xn = z[i] * sg; dmax = fmax(dmax, fabs(xn - x[i])); x[i] = xn; }
        done = dmax <= 1e-14;
    }
    if (ok && done) { for (int i = 0; i < 4; i++) v[i] = x[i]; return; }
    smallest_right_singular4_jacobi(Ain, v);
}

// A (N x N) X = B (N x M), Gaussian elimination with partial pivoting, in place
// (A destroyed, B <- X).  Returns false on an exactly singular pivot.
template <int N, int M> MP_HD bool lu_solve_inplace(double *A, double *B) {
    for (int c = 0; c < N; c++) {
        int piv = c;
        double pv = fabs(A[c * N + c]);
        for (int r = c + 1; r < N; r++) {
            double v = fabs(A[r * N + c]);
            if (v > pv) { pv = v; piv = r; }
        }
        if (!(pv > 0.0)) return false;
        if (piv != c) {
            for (int k = 0; k < N; k++) { double t = A[c * N + k]; A[c * N + k] = A[piv * N + k]; A[piv * N + k] = t; }
            for (int k = 0; k < M; k++) { double t = B[c * M + k]; B[c * M + k] = B[piv * M + k]; B[piv * M + k] = t; }
        }
        double ip = 1.0 / A[c * N + c];
        for (int r = c + 1; r < N; r++) {
            double f = A[r * N + c] * ip;
            if (f == 0.0) continue;
            for (int k = c + 1; k < N; k++) A[r * N + k] = fma(-f, A[c * N + k], A[r * N + k]);
            for (int k = 0; k < M; k++) B[r * M + k] = fma(-f, B[c * M + k], B[r * M + k]);
        }
    }
    for (int c = N - 1; c >= 0; c--) {
        double ip = 1.0 / A[c * N + c];
        for (int k = 0; k < M; k++) {
            double s = B[c * M + k];
            for (int j = c + 1; j < N; j++) s = fma(-A[c * N + j], B[j * M + k], s);
            B[c * M + k] = s * ip;
        }
    }
    return true;
}

// Orthonormal basis of the null space of A (M x N row-major, M < N, full row rank):
// Householder QR of A^T; the last N-M columns of Q.  basis: (N-M) x N row-major.
template <int M, int N> MP_HD void nullspace_qr(const double *A, double *basis) {
    double Q[N * N], R[N * M], v[N];
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) Q[i * N + j] = (i == j) ? 1.0 : 0.0;
    for (int i = 0; i < N; i++)
        for (int j = 0; j < M; j++) R[i * M + j] = A[j * N + i];
    for (int c = 0; c < M; c++) {
        double nr = 0;
        for (int i = c; i < N; i++) nr = fma(R[i * M + c], R[i * M + c], nr);
        nr = sqrt(nr);
        if (nr == 0.0) continue;
        double alpha = R[c * M + c] > 0 ? -nr : nr;
        for (int i = c; i < N; i++) v[i] = R[i * M + c];
        v[c] -= alpha;
        double vn = 0;
        for (int i = c; i < N; i++) vn = fma(v[i], v[i], vn);
        if (vn == 0.0) continue;
        double inv = 2.0 / vn;
        for (int j = c; j < M; j++) {
            double s = 0;
            for (int i = c; i < N; i++) s = fma(v[i], R[i * M + j], s);
            s *= inv;
            for (int i = c; i < N; i++) R[i * M + j] = fma(-s, v[i], R[i * M + j]);
        }
        for (int j = 0; j < N; j++) {
            double s = 0;
            for (int i = c; i < N; i++) s = fma(Q[j * N + i], v[i], s);
            s *= inv;
            for (int i = c; i < N; i++) Q[j * N + i] = fma(-s, v[i], Q[j * N + i]);
        }
    }
    for (int k = 0; k < N - M; k++)
        for (int i = 0; i < N; i++) basis[k * N + i] = Q[i * N + (M + k)];
}

// Cholesky solve of the SPD system A x = b (n <= NMAX, A row-major with leading dim n).
// (L: n*n doubles, y: n doubles of caller workspace)
MP_HD bool cholesky_solve_ws(int n, const double *A, const double *b, double *x, double *L, double *y) {
    for (int i = 0; i < n; i++)
        for (int j = 0; j <= i; j++) {
            double s = A[i * n + j];
            for (int k = 0; k < j; k++) s -= L[i * n + k] * L[j * n + k];
            if (i == j) {
                if (!(s > 0.0) || !finite_d(s)) return false;
                L[i * n + i] = sqrt(s);
            } else L[i * n + j] = s / L[j * n + j];
        }
    for (int i = 0; i < n; i++) {
        double s = b[i];
        for (int k = 0; k < i; k++) s -= L[i * n + k] * y[k];
        y[i] = s / L[i * n + i];
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = y[i];
        for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k];
        x[i] = s / L[i * n + i];
    }
    for (int i = 0; i < n; i++)
        if (!finite_d(x[i])) return false;
    return true;
}
template <int NMAX> MP_HD bool cholesky_solve(int n, const double *A, const double *b, double *x) {
    double L[NMAX * NMAX], y[NMAX];
    return cholesky_solve_ws(n, A, b, x, L, y);
}

} // namespace mp
