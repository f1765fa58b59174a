// ORACLE (test infrastructure only - never linked into the product path).
// Small dense linear-algebra kit, plain C++17, no Eigen.  Replaces the Eigen
// calls the reference makes on the hot path (JacobiSVD, colPivHouseholderQr,
// EigenSolver, ldlt: /root/reference/src/solver.cpp:100,104,517,
// /root/reference/src/utils.h:35, hybrid_pose_estimator.cpp:163).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace om {

// ---- 3x3 row-major helpers -------------------------------------------------
inline void mat3_mul(const double *A, const double *B, double *C) {
    double T[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) T[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    std::memcpy(C, T, sizeof(T));
}
inline void mat3_T(const double *A, double *B) {
    double T[9] = {A[0], A[3], A[6], A[1], A[4], A[7], A[2], A[5], A[8]};
    std::memcpy(B, T, sizeof(T));
}
inline void mat3_vec(const double *A, const double *x, double *y) {
    double a = A[0] * x[0] + A[1] * x[1] + A[2] * x[2];
    double b = A[3] * x[0] + A[4] * x[1] + A[5] * x[2];
    double c = A[6] * x[0] + A[7] * x[1] + A[8] * x[2];
    y[0] = a; y[1] = b; y[2] = c;
}
inline void mat3T_vec(const double *A, const double *x, double *y) {
    double a = A[0] * x[0] + A[3] * x[1] + A[6] * x[2];
    double b = A[1] * x[0] + A[4] * x[1] + A[7] * x[2];
    double c = A[2] * x[0] + A[5] * x[1] + A[8] * x[2];
    y[0] = a; y[1] = b; y[2] = c;
}
inline double mat3_det(const double *A) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
}
inline bool mat3_inv(const double *A, double *B) {
    double d = mat3_det(A);
    if (d == 0.0 || !std::isfinite(d)) return false;
    double id = 1.0 / d;
    double T[9];
    T[0] = (A[4] * A[8] - A[5] * A[7]) * id; T[1] = (A[2] * A[7] - A[1] * A[8]) * id; T[2] = (A[1] * A[5] - A[2] * A[4]) * id;
    T[3] = (A[5] * A[6] - A[3] * A[8]) * id; T[4] = (A[0] * A[8] - A[2] * A[6]) * id; T[5] = (A[2] * A[3] - A[0] * A[5]) * id;
    T[6] = (A[3] * A[7] - A[4] * A[6]) * id; T[7] = (A[1] * A[6] - A[0] * A[7]) * id; T[8] = (A[0] * A[4] - A[1] * A[3]) * id;
    std::memcpy(B, T, sizeof(T));
    return true;
}
inline void cross3(const double *a, const double *b, double *c) {
    double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
    c[0] = x; c[1] = y; c[2] = z;
}
inline double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline double norm3(const double *a) { return std::sqrt(dot3(a, a)); }

// ---- one-sided Jacobi SVD, n x n (n <= 9), A = U diag(S) V^T, S descending ---
// Same decomposition Eigen::JacobiSVD delivers (unique up to sign/rotation in
// degenerate subspaces).
inline void jacobi_svd(int n, const double *Ain, double *U, double *S, double *V) {
    double A[81];
    std::memcpy(A, Ain, sizeof(double) * n * n);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) V[i * n + j] = (i == j);
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < n - 1; p++)
            for (int q = p + 1; q < n; q++) {
                double a = 0, b = 0, c = 0;
                for (int k = 0; k < n; k++) {
                    a += A[k * n + p] * A[k * n + p];
                    b += A[k * n + q] * A[k * n + q];
                    c += A[k * n + p] * A[k * n + q];
                }
                if (c == 0.0 || std::fabs(c) <= 1e-300) continue;
                off = std::max(off, std::fabs(c) / std::sqrt(a * b + 1e-300));
                if (std::fabs(c) <= 1e-17 * std::sqrt(a * b)) continue;
                double zeta = (b - a) / (2.0 * c);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
                for (int k = 0; k < n; k++) {
                    double x = A[k * n + p], y = A[k * n + q];
                    A[k * n + p] = cs * x - sn * y;
                    A[k * n + q] = sn * x + cs * y;
                    x = V[k * n + p]; y = V[k * n + q];
                    V[k * n + p] = cs * x - sn * y;
                    V[k * n + q] = sn * x + cs * y;
                }
            }
        if (off < 1e-16) break;
    }
    int order[9];
    double nrm[9];
    for (int j = 0; j < n; j++) {
        double s = 0;
        for (int k = 0; k < n; k++) s += A[k * n + j] * A[k * n + j];
        nrm[j] = std::sqrt(s);
        order[j] = j;
    }
    std::stable_sort(order, order + n, [&](int a, int b) { return nrm[a] > nrm[b]; });
    double Vs[81];
    for (int jj = 0; jj < n; jj++) {
        int j = order[jj];
        S[jj] = nrm[j];
        for (int k = 0; k < n; k++) {
            Vs[k * n + jj] = V[k * n + j];
            if (U) U[k * n + jj] = nrm[j] > 0 ? A[k * n + j] / nrm[j] : 0.0;
        }
    }
    std::memcpy(V, Vs, sizeof(double) * n * n);
    if (U && n == 3) {
        // complete U to an orthonormal basis when the smallest singular value is ~0
        if (S[2] <= 1e-14 * S[0]) {
            double c[3], a[3] = {U[0], U[3], U[6]}, b[3] = {U[1], U[4], U[7]};
            cross3(a, b, c);
            double nn = norm3(c);
            if (nn > 0) { U[2] = c[0] / nn; U[5] = c[1] / nn; U[8] = c[2] / nn; }
        }
    }
}

// ---- generic dense solve with partial pivoting: A (n x n) X = B (n x m) ------
inline bool lu_solve(int n, int m, std::vector<double> A, std::vector<double> &B) {
    for (int c = 0; c < n; c++) {
        int piv = c;
        for (int r = c + 1; r < n; r++)
            if (std::fabs(A[r * n + c]) > std::fabs(A[piv * n + c])) piv = r;
        if (A[piv * n + c] == 0.0) return false;
        if (piv != c) {
            for (int k = 0; k < n; k++) std::swap(A[c * n + k], A[piv * n + k]);
            for (int k = 0; k < m; k++) std::swap(B[c * m + k], B[piv * m + k]);
        }
        double ip = 1.0 / A[c * n + c];
        for (int r = c + 1; r < n; r++) {
            double f = A[r * n + c] * ip;
            if (f == 0.0) continue;
            for (int k = c; k < n; k++) A[r * n + k] -= f * A[c * n + k];
            for (int k = 0; k < m; k++) B[r * m + k] -= f * B[c * m + k];
        }
    }
    for (int c = n - 1; c >= 0; c--) {
        for (int k = 0; k < m; k++) {
            double s = B[c * m + k];
            for (int j = c + 1; j < n; j++) s -= A[c * n + j] * B[j * m + k];
            B[c * m + k] = s / A[c * n + c];
        }
    }
    return true;
}

// ---- null space of A (m x n, m < n, full row rank) via Householder QR of A^T -
// basis: (n-m) vectors of length n, row-major [k][n], orthonormal.
inline void nullspace(int m, int n, const double *A, double *basis) {
    std::vector<double> Q(n * n, 0.0), R(n * m);
    for (int i = 0; i < n; i++) Q[i * n + i] = 1.0;
    for (int i = 0; i < n; i++)
        for (int j = 0; j < m; j++) R[i * m + j] = A[j * n + i];
    std::vector<double> v(n);
    for (int c = 0; c < m; c++) {
        double nr = 0;
        for (int i = c; i < n; i++) nr += R[i * m + c] * R[i * m + c];
        nr = std::sqrt(nr);
        if (nr == 0) continue;
        double alpha = R[c * m + c] > 0 ? -nr : nr;
        for (int i = 0; i < n; i++) v[i] = 0;
        for (int i = c; i < n; i++) v[i] = R[i * m + c];
        v[c] -= alpha;
        double vn = 0;
        for (int i = c; i < n; i++) vn += v[i] * v[i];
        if (vn == 0) continue;
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int i = c; i < n; i++) s += v[i] * R[i * m + j];
            s = 2 * s / vn;
            for (int i = c; i < n; i++) R[i * m + j] -= s * v[i];
        }
        for (int j = 0; j < n; j++) { // Q = Q * H
            double s = 0;
            for (int i = c; i < n; i++) s += Q[j * n + i] * v[i];
            s = 2 * s / vn;
            for (int i = c; i < n; i++) Q[j * n + i] -= s * v[i];
        }
    }
    for (int k = 0; k < n - m; k++)
        for (int i = 0; i < n; i++) basis[k * n + i] = Q[i * n + (m + k)];
}

} // namespace om
