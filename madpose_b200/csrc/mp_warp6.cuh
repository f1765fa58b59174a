// K2 for the shared-focal estimator with ONE WARP PER SAMPLE: the 6-point shared-focal relative pose solver
// (replaces poselib::relpose_6pt_shared_focal at hybrid_pose_shared_focal_estimator.cpp:87; same hidden-variable
// formulation as the scalar relpose_6pt_shared_focal of mp_relpose2.cuh, which stays as the host-checked statement
// of the arithmetic).  No thread-private arrays: the work is spread over the lanes and the operands live in
// registers and in a per-warp shared-memory workspace.
//   null space of the 6 x 9 epipolar system   Householder, lane = row of Q / column of R
//   coefficient matrices C0, C1, C2 (10 x 10) lane = equation
//   det C(w) at 9 points of the unit circle   3 matrices at a time, lane = (matrix, row): complex elimination with
//                                             partial pivoting, rows in registers, pivot row through shared memory
//   real roots of the degree-15 determinant   mp_warp_poly.cuh (lane = coefficient / evaluation point / root)
//   per admissible root w > 0                 3 roots at a time, lane = (root, row): 10 x 10 elimination with full
//                                             pivoting in shared memory, null vector, Gauss-Newton polish of (a, b, w)
//                                             (lane = equation), essential-matrix test, E -> (R, t) with the
//                                             cheirality test one point per lane
// All 32 lanes of the warp call warp_relpose_6pt together.
#pragma once
#include "mp_minimal.cuh"
#include "mp_warp_poly.cuh"

namespace mp {

constexpr int W6_LD = 11;   // padded row length of the 10 x 10 elimination matrices (conflict-free row access)
// cos / sin of 2 pi k / 16 (the unit-circle evaluation points and the inverse DFT)
__device__ const double W6_COS[16] = {1.0, 0.92387953251128674, 0.70710678118654757, 0.38268343236508984, 0.0,
                                      -0.38268343236508973, -0.70710678118654746, -0.92387953251128674, -1.0,
                                      -0.92387953251128685, -0.70710678118654768, -0.38268343236509034, 0.0,
                                      0.38268343236509, 0.70710678118654735, 0.92387953251128652};
__device__ const double W6_SIN[16] = {0.0, 0.38268343236508978, 0.70710678118654746, 0.92387953251128674, 1.0,
                                      0.92387953251128674, 0.70710678118654757, 0.38268343236508989, 0.0,
                                      -0.38268343236508967, -0.70710678118654746, -0.92387953251128652, -1.0,
                                      -0.92387953251128663, -0.70710678118654768, -0.38268343236509039};

struct alignas(16) Warp6Work {
    double x1[6][3], x2[6][3];         // the sample, (x, y, 1)
    double N[27];                      // null-space basis: F_k = a N[k] + b N[9 + k] + N[18 + k]
    double Cw[3][100];                 // C(w) = C0 + w C1 + w^2 C2, rows scaled to unit max coefficient
    double vr[9], vi[9];               // det C at the 9 unit-circle points
    double coef[16];                   // det C(w) = sum coef[j] w^j
    double roots[16];
    union {
        struct { double Q[81], R[54]; double S0[6][6], S1[6][6], T[3][6]; } a;  // stages 1-2
        struct { double mag[3][10]; double prow[3][2][10]; } d;                 // stage 3
        WarpPolyWork<15> poly;                                                  // stage 5
        struct {                                                                // stage 6
            double M[3][10 * W6_LD];
            double pv[3][10];  int pk[3][10];     // per virtual row: best |entry| and its column
            int cp[3][10], phys[3][10];           // column permutation; physical row at each virtual position
            double v[3][10];                      // null vector (permuted order)
            double red[3][10][9];                 // Gauss-Newton partial products
            double abw[3][3];
            int ncand[3];
        } r;
    } u;
};

// ---- stage 1: orthonormal null-space basis of the 6 x 9 epipolar system (Householder QR of A^T)
__device__ inline void warp6_nullspace(Warp6Work &W, int lane) {
    double *Q = W.u.a.Q, *R = W.u.a.R;
    for (int e = lane; e < 81; e += 32) Q[e] = (e / 9 == e % 9) ? 1.0 : 0.0;
    for (int e = lane; e < 54; e += 32) { // R[i][j] = A[j][i], A[j][3 a + b] = x2[j][a] * x1[j][b]
        const int i = e / 6, j = e % 6;
        R[e] = W.x2[j][i / 3] * W.x1[j][i % 3];
    }
    __syncwarp();
    for (int c = 0; c < 6; c++) {
        double v[9], nr = 0;
#pragma unroll
        for (int i = 0; i < 9; i++) { v[i] = i >= c ? R[i * 6 + c] : 0.0; nr = fma(v[i], v[i], nr); }
        nr = sqrt(nr);
        const double rcc = R[c * 6 + c];
        const double alpha = rcc > 0 ? -nr : nr;
        double vn = 0;
#pragma unroll
        for (int i = 0; i < 9; i++) { if (i == c) v[i] -= alpha; vn = fma(v[i], v[i], vn); }
        __syncwarp(); // everybody has read column c
        if (nr != 0.0 && vn != 0.0) {
            const double inv = 2.0 / vn;
            if (lane >= c && lane < 6) { // column `lane` of R
                double s = 0;
#pragma unroll
                for (int i = 0; i < 9; i++) s = fma(v[i], R[i * 6 + lane], s);
                s *= inv;
#pragma unroll
                for (int i = 0; i < 9; i++) if (i >= c) R[i * 6 + lane] = fma(-s, v[i], R[i * 6 + lane]);
            } else if (lane >= 8 && lane < 17) { // row `lane - 8` of Q
                const int j = lane - 8;
                double s = 0;
#pragma unroll
                for (int i = 0; i < 9; i++) s = fma(Q[j * 9 + i], v[i], s);
                s *= inv;
#pragma unroll
                for (int i = 0; i < 9; i++) if (i >= c) Q[j * 9 + i] = fma(-s, v[i], Q[j * 9 + i]);
            }
        }
        __syncwarp();
    }
    if (lane < 27) W.N[lane] = Q[(lane % 9) * 9 + 6 + lane / 9];
    __syncwarp();
}

// ---- stage 2: the ten cubic constraints in (a, b) with coefficients quadratic in w (row 0: det F = 0, rows 1-9:
// 2 G F - tr(G) F = 0 with G = F Q F^T Q, Q = diag(1, 1, w)); returns false when a row vanishes
__device__ inline bool warp6_build(Warp6Work &W, int lane) {
    auto Lf = [&](int k, double *l) { l[0] = W.N[k]; l[1] = W.N[9 + k]; l[2] = W.N[18 + k]; };
    const int gi[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
    if (lane < 6) { // S0 + w S1 = F Q F^T, symmetric: entry g = (i, j)
        const int i = lane < 3 ? 0 : (lane < 5 ? 1 : 2), j = lane < 3 ? lane : (lane < 5 ? lane - 2 : 2);
        double la[3], lb[3], q[6], tmp[6];
        Lf(3 * i, la); Lf(3 * j, lb); mul_ll2(la, lb, q);
        Lf(3 * i + 1, la); Lf(3 * j + 1, lb); mul_ll2(la, lb, tmp);
#pragma unroll
        for (int m = 0; m < 6; m++) W.u.a.S0[lane][m] = q[m] + tmp[m];
        Lf(3 * i + 2, la); Lf(3 * j + 2, lb); mul_ll2(la, lb, q);
#pragma unroll
        for (int m = 0; m < 6; m++) W.u.a.S1[lane][m] = q[m];
    }
    __syncwarp();
    if (lane < 6) { // trace of G: t0 + w t1 + w^2 t2
        W.u.a.T[0][lane] = W.u.a.S0[0][lane] + W.u.a.S0[3][lane];
        W.u.a.T[1][lane] = W.u.a.S1[0][lane] + W.u.a.S1[3][lane] + W.u.a.S0[5][lane];
        W.u.a.T[2][lane] = W.u.a.S1[5][lane];
    }
    __syncwarp();
    bool ok = true;
    if (lane < 10) {
        double c0[10], c1[10], c2[10];
#pragma unroll
        for (int m = 0; m < 10; m++) { c0[m] = 0; c1[m] = 0; c2[m] = 0; }
        if (lane == 0) {
            double l4[3], l8[3], l5[3], l7[3], l3[3], l6[3], l0[3], l1[3], l2[3], q1[6], q2[6];
            Lf(0, l0); Lf(1, l1); Lf(2, l2); Lf(3, l3); Lf(4, l4); Lf(5, l5); Lf(6, l6); Lf(7, l7); Lf(8, l8);
            mul_ll2(l4, l8, q1); mul_ll2(l5, l7, q2);
            for (int m = 0; m < 6; m++) q1[m] -= q2[m];
            fma_ql2(q1, l0, 1.0, c0);
            mul_ll2(l3, l8, q1); mul_ll2(l5, l6, q2);
            for (int m = 0; m < 6; m++) q1[m] -= q2[m];
            fma_ql2(q1, l1, -1.0, c0);
            mul_ll2(l3, l7, q1); mul_ll2(l4, l6, q2);
            for (int m = 0; m < 6; m++) q1[m] -= q2[m];
            fma_ql2(q1, l2, 1.0, c0);
        } else {
            const int i = (lane - 1) / 3, j = (lane - 1) % 3;
            double l[3];
            for (int k = 0; k < 2; k++) {
                Lf(3 * k + j, l);
                fma_ql2(W.u.a.S0[gi[i][k]], l, 2.0, c0);
                fma_ql2(W.u.a.S1[gi[i][k]], l, 2.0, c1);
            }
            Lf(6 + j, l);
            fma_ql2(W.u.a.S0[gi[i][2]], l, 2.0, c1);
            fma_ql2(W.u.a.S1[gi[i][2]], l, 2.0, c2);
            Lf(3 * i + j, l);
            fma_ql2(W.u.a.T[0], l, -1.0, c0);
            fma_ql2(W.u.a.T[1], l, -1.0, c1);
            fma_ql2(W.u.a.T[2], l, -1.0, c2);
        }
        double mx = 0;
#pragma unroll
        for (int m = 0; m < 10; m++) mx = fmax(mx, fmax(fabs(c0[m]), fmax(fabs(c1[m]), fabs(c2[m]))));
        ok = mx > 0;
        const double im = 1.0 / mx;
#pragma unroll
        for (int m = 0; m < 10; m++) {
            W.Cw[0][10 * lane + m] = c0[m] * im; W.Cw[1][10 * lane + m] = c1[m] * im; W.Cw[2][10 * lane + m] = c2[m] * im;
        }
    }
    ok = __all_sync(WP_FULL, ok);
    __syncwarp();
    return ok;
}

// ---- stage 3: det C(w) at w_k = exp(2 pi i k / 16), k = 0..8 (conjugate symmetry gives the other 7)
__device__ inline void warp6_dets(Warp6Work &W, int lane) {
    const int g = lane / 10, r = lane % 10;
    const bool act = lane < 30;
    for (int pass = 0; pass < 3; pass++) {
        const int k = pass * 3 + (act ? g : 0);
        const double wr = W6_COS[k], wi = W6_SIN[k];
        const double w2r = wr * wr - wi * wi, w2i = 2 * wr * wi;
        double Mr[10], Mi[10];
#pragma unroll
        for (int c = 0; c < 10; c++) {
            const double c0 = W.Cw[0][10 * r + c], c1 = W.Cw[1][10 * r + c], c2 = W.Cw[2][10 * r + c];
            Mr[c] = c0 + wr * c1 + w2r * c2;
            Mi[c] = wi * c1 + w2i * c2;
        }
        int pos = r;           // position of my row in the row order of the scalar elimination
        double dr = 1.0, di = 0.0;
        bool dead = false;     // a zero pivot column: the determinant is 0
        __syncwarp();
#pragma unroll
        for (int c = 0; c < 10; c++) {
            if (act) W.u.d.mag[g][pos] = pos >= c ? Mr[c] * Mr[c] + Mi[c] * Mi[c] : -1.0;
            __syncwarp();
            int piv = c;
            double pm = 0;
            if (act) {
                pm = W.u.d.mag[g][c];
                for (int q = c + 1; q < 10; q++) { const double m = W.u.d.mag[g][q]; if (m > pm) { pm = m; piv = q; } }
            }
            if (!(pm > 0.0)) dead = true;
            const bool mine = act && pos == piv;
            if (piv != c) { // the pivot row and the row at position c trade places
                if (pos == piv) pos = c; else if (pos == c) pos = piv;
                dr = -dr; di = -di;
            }
            if (mine) {
#pragma unroll
                for (int q = 0; q < 10; q++) if (q >= c) { W.u.d.prow[g][0][q] = Mr[q]; W.u.d.prow[g][1][q] = Mi[q]; }
            }
            __syncwarp();
            if (act && !dead) {
                const double pr = W.u.d.prow[g][0][c], pi = W.u.d.prow[g][1][c];
                { const double t = dr * pr - di * pi; di = dr * pi + di * pr; dr = t; }
                if (pos > c) {
                    const double rpm = 1.0 / pm, ir = pr * rpm, ii = -pi * rpm; // 1 / pivot
                    const double ar = Mr[c], ai = Mi[c];
                    const double fr = ar * ir - ai * ii, fi = ar * ii + ai * ir;
                    if (!(fr == 0.0 && fi == 0.0)) {
#pragma unroll
                        for (int q = 0; q < 10; q++)
                            if (q > c) {
                                const double br = W.u.d.prow[g][0][q], bi = W.u.d.prow[g][1][q];
                                Mr[q] -= fr * br - fi * bi;
                                Mi[q] -= fr * bi + fi * br;
                            }
                    }
                }
            }
            __syncwarp();
        }
        if (act && r == 0 && k <= 8) { W.vr[k] = dead ? 0.0 : dr; W.vi[k] = dead ? 0.0 : di; }
        __syncwarp();
    }
}

// ---- stage 4: inverse DFT -> coefficients of det C(w) (degree 15)
__device__ inline void warp6_coefs(Warp6Work &W, int lane) {
    if (lane < 16) {
        const int j = lane;
        double s = W.vr[0] + ((j & 1) ? -W.vr[8] : W.vr[8]);
        for (int k = 1; k < 8; k++) {
            const int a16 = (j * k) & 15;
            s += 2.0 * (W.vr[k] * W6_COS[a16] + W.vi[k] * W6_SIN[a16]);
        }
        W.coef[j] = s / 16.0;
    }
    __syncwarp();
}

// E -> up to 4 (R, t); the candidates for which all six bearing pairs pass cheirality are written to out (order kept).
// The 10 lanes of group g call it together; lanes r < 6 test one point each.  Returns the number written (uniform
// over the group); out may be nullptr to count only.
__device__ inline int group_motion_from_essential(const Warp6Work &W, const double *E, double f, int g, int r, bool act,
                                                  unsigned *mask_out, double *Ra, double *Rb, double *t) {
    essential_factor(E, Ra, Rb, t);
    unsigned ok4 = 0xfu;
    if (act && r < 6) {
        const double p[3] = {W.x1[r][0] / f, W.x1[r][1] / f, 1.0}, q[3] = {W.x2[r][0] / f, W.x2[r][1] / f, 1.0};
        const double np_ = norm3(p), nq = norm3(q);
        const double b1[3] = {p[0] / np_, p[1] / np_, p[2] / np_}, b2[3] = {q[0] / nq, q[1] / nq, q[2] / nq};
        const double tn[3] = {-t[0], -t[1], -t[2]};
        ok4 = 0;
        for (int c = 0; c < 4; c++) {
            const double *R = c < 2 ? Ra : Rb, *tt = (c == 0 || c == 3) ? t : tn;
            if (cheirality(R, tt, b1, b2, 0.0)) ok4 |= 1u << c;
        }
    }
    // AND over the six point lanes of the group: one ballot per candidate
    unsigned res = 0;
    for (int c = 0; c < 4; c++) {
        const unsigned bal = __ballot_sync(WP_FULL, !act || r >= 6 || ((ok4 >> c) & 1u));
        const unsigned grp = act ? (0x3ffu << (10 * g)) : 0u;
        if (act && (bal & grp) == grp) res |= 1u << c;
    }
    *mask_out = res;
    return __popc(res);
}

// The solver.  Sample points (x, y, 1) must be in W.x1 / W.x2.  Writes up to 15 candidates (R, t, focal; scale 1,
// offsets 0) to out (global or shared memory) and returns their number (uniform over the warp).
// LOCKSTEP: all warps of the CTA call it together (valid = this warp has a sample) and meet at a CTA barrier between
// the stages, so that they execute the same stage's code at the same time (the solver is ~170 KB of instructions).
template <bool LOCKSTEP> __device__ inline int warp_relpose_6pt(Warp6Work &W, Model *out, int lane, bool valid = true) {
#ifndef W6_BARRIERS
#define W6_BARRIERS 15 // which of the four stage barriers the lockstep form keeps (bit k = the barrier after stage k)
#endif
    if (valid) warp6_nullspace(W, lane);
    if (LOCKSTEP && (W6_BARRIERS & 1)) __syncthreads();
    bool ok = valid;
    if (ok) ok = warp6_build(W, lane);
    if (LOCKSTEP && (W6_BARRIERS & 2)) __syncthreads();
    if (ok) warp6_dets(W, lane);
    if (LOCKSTEP && (W6_BARRIERS & 4)) __syncthreads();
    int nr = 0;
    if (ok) {
        warp6_coefs(W, lane);
        nr = warp_real_roots<15>(W.coef, 15, W.roots, W.u.poly, true, lane);
    }
    if (LOCKSTEP && (W6_BARRIERS & 8)) __syncthreads();
    __syncwarp();
    int n = 0;
    const int g = lane / 10, r = lane % 10;
    for (int base = 0; base < nr && n < 15; base += 3) {
        const bool act = lane < 30 && base + g < nr;
        const double wr = act ? W.roots[base + g] : 1.0;
        __syncwarp();
        double *M = W.u.r.M[act ? g : 0];
        if (act) {
#pragma unroll
            for (int c = 0; c < 10; c++)
                M[r * W6_LD + c] = W.Cw[0][10 * r + c] + wr * (W.Cw[1][10 * r + c] + wr * W.Cw[2][10 * r + c]);
            W.u.r.cp[g][r] = r;
        }
        int pos = r;
        __syncwarp();
        // elimination with full pivoting: null vector of C(w)
        for (int c = 0; c < 9; c++) {
            if (act) {
                double pm = -1.0; int pk = c;
                if (pos >= c) {
                    pm = fabs(M[r * W6_LD + c]);
                    for (int k = c + 1; k < 10; k++) { const double m = fabs(M[r * W6_LD + k]); if (m > pm) { pm = m; pk = k; } }
                }
                W.u.r.pv[g][pos] = pm; W.u.r.pk[g][pos] = pk;
                W.u.r.phys[g][pos] = r;
            }
            __syncwarp();
            int pr = c, pc = c, prow = 0;
            if (act) {
                double pm = W.u.r.pv[g][c];
                pc = W.u.r.pk[g][c];
                // (the scalar code starts from entry (c, c) and takes strictly larger entries in (row, column) order)
                for (int q = c + 1; q < 10; q++) { const double m = W.u.r.pv[g][q]; if (m > pm) { pm = m; pr = q; pc = W.u.r.pk[g][q]; } }
                prow = W.u.r.phys[g][pr];
            }
            __syncwarp();
            if (act) {
                if (pr != c) { if (pos == pr) pos = c; else if (pos == c) pos = pr; }
                if (pc != c) { const double t = M[r * W6_LD + c]; M[r * W6_LD + c] = M[r * W6_LD + pc]; M[r * W6_LD + pc] = t; }
                if (r == 0 && pc != c) { const int t = W.u.r.cp[g][c]; W.u.r.cp[g][c] = W.u.r.cp[g][pc]; W.u.r.cp[g][pc] = t; }
            }
            __syncwarp();
            if (act && pos > c) {
                const double f = M[r * W6_LD + c] / M[prow * W6_LD + c];
                for (int k = c; k < 10; k++) M[r * W6_LD + k] -= f * M[prow * W6_LD + k];
            }
            __syncwarp();
        }
        if (act) W.u.r.phys[g][pos] = r; // final positions (the last step may have moved rows 8 / 9)
        __syncwarp();
        // back substitution, free variable = 1: the row at position c computes v[c] once v[c+1..9] are there
        if (act && r == 0) W.u.r.v[g][9] = 1.0;
        __syncwarp();
        for (int c = 8; c >= 0; c--) {
            if (act && pos == c) {
                double s = 0;
                for (int k = c + 1; k < 10; k++) s += M[r * W6_LD + k] * W.u.r.v[g][k];
                W.u.r.v[g][c] = -s / M[r * W6_LD + c];
            }
            __syncwarp();
        }
        double a = 0, b = 0, w = wr;
        bool good = act;
        if (act) {
            // vv[cp[i]] = v[i]; (a, b) = (vv[7], vv[8]) / vv[9]
            double v7 = 0, v8 = 0, v9 = 0;
            for (int i = 0; i < 10; i++) {
                const int ci = W.u.r.cp[g][i];
                const double vi = W.u.r.v[g][i];
                if (ci == 7) v7 = vi; else if (ci == 8) v8 = vi; else if (ci == 9) v9 = vi;
            }
            if (v9 == 0.0 || !finite_d(v9)) good = false;
            a = v7 / v9; b = v8 / v9;
        }
        // Gauss-Newton polish of (a, b, w) on the ten row-scaled equations: lane = equation.  Every lane runs the same
        // barrier sequence; a group that has finished (done) skips its updates.
        bool done = !act;
        for (int it = 0; it < 5; it++) {
            if (!done) {
                const double mo[10] = {a * a * a, a * a * b, a * b * b, b * b * b, a * a, a * b, b * b, a, b, 1.0};
                const double ma[10] = {3 * a * a, 2 * a * b, b * b, 0, 2 * a, b, 0, 1, 0, 0};
                const double mb[10] = {0, a * a, 2 * a * b, 3 * b * b, 0, a, 2 * b, 0, 1, 0};
                double f = 0, g0 = 0, g1 = 0, g2 = 0;
#pragma unroll
                for (int c = 0; c < 10; c++) {
                    const double c0 = W.Cw[0][10 * r + c], c1 = W.Cw[1][10 * r + c], c2 = W.Cw[2][10 * r + c];
                    const double cw = c0 + w * (c1 + w * c2);
                    f += cw * mo[c]; g0 += cw * ma[c]; g1 += cw * mb[c]; g2 += (c1 + 2 * w * c2) * mo[c];
                }
                double *o = W.u.r.red[g][r];
                o[0] = g0 * g0; o[1] = g0 * g1; o[2] = g0 * g2; o[3] = g1 * g1; o[4] = g1 * g2; o[5] = g2 * g2;
                o[6] = g0 * f; o[7] = g1 * f; o[8] = g2 * f;
            }
            __syncwarp();
            if (!done) {
                double s[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
                for (int q = 0; q < 10; q++)
#pragma unroll
                    for (int e = 0; e < 9; e++) s[e] += W.u.r.red[g][q][e];
                const double JtJ[9] = {s[0], s[1], s[2], s[1], s[3], s[4], s[2], s[4], s[5]}, Jtr[3] = {s[6], s[7], s[8]};
                double inv[9], st[3];
                if (!inv3(JtJ, inv)) done = true;
                else {
                    mat3_vecp(inv, Jtr, st);
                    if (!finite_d(st[0] + st[1] + st[2])) done = true;
                    else {
                        a -= st[0]; b -= st[1]; w -= st[2];
                        if (fabs(st[0]) <= 1e-15 * fabs(a) && fabs(st[1]) <= 1e-15 * fabs(b) && fabs(st[2]) <= 1e-15 * fabs(w)) done = true;
                    }
                }
            }
            __syncwarp();
            if (__all_sync(WP_FULL, done)) break;
        }
        __syncwarp();
        good = good && (w > 0) && finite_d(a + b + w);
        // E = K F K, K = diag(f, f, 1)
        double f = 1.0, E[9];
        if (good) {
            f = 1.0 / sqrt(w);
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int j = 0; j < 3; j++)
                    E[3 * i + j] = (a * W.N[3 * i + j] + b * W.N[9 + 3 * i + j] + W.N[18 + 3 * i + j]) * (i < 2 ? f : 1.0) * (j < 2 ? f : 1.0);
            good = essential_valid(E, 1e-8);
        }
        if (!good) {
#pragma unroll
            for (int k = 0; k < 9; k++) E[k] = (k % 4 == 0) ? 1.0 : 0.0; // harmless input for the idle groups
        }
        unsigned mask = 0;
        double Ra[9], Rb[9], t[3];
        group_motion_from_essential(W, E, f, g, r, act, &mask, Ra, Rb, t);
        if (!good) mask = 0;
        if (!finite_d(Ra[0] + Ra[4] + Ra[8])) mask &= ~0x3u; // candidates 0, 1 use Ra; 2, 3 use Rb
        if (!finite_d(Rb[0] + Rb[4] + Rb[8])) mask &= ~0xcu;
        const int cnt = __popc(mask);
        if (lane < 30 && r == 0) W.u.r.ncand[g] = act ? cnt : 0;
        __syncwarp();
        const int before = (g > 0 ? W.u.r.ncand[0] : 0) + (g > 1 ? W.u.r.ncand[1] : 0);
        const int total = W.u.r.ncand[0] + W.u.r.ncand[1] + W.u.r.ncand[2];
        if (act && r == 0 && cnt > 0) {
            int k = n + before;
            for (int c = 0; c < 4; c++) {
                if (!((mask >> c) & 1u) || k >= 15) continue;
                const double *R = c < 2 ? Ra : Rb;
                const double sg = (c == 0 || c == 3) ? 1.0 : -1.0;
                Model m;
                for (int q = 0; q < 9; q++) m.R[q] = R[q];
                for (int q = 0; q < 3; q++) m.t[q] = sg * t[q];
                m.scale = 1; m.o0 = 0; m.o1 = 0; m.f0 = f; m.f1 = f;
                out[k++] = m;
            }
        }
        n = min(n + total, 15);
        __syncwarp();
    }
    return n;
}

} // namespace mp
