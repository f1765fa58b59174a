// K5 (part 1) -- residual blocks of the local-optimisation problems with ANALYTIC
// Jacobians in the tangent space (the reference uses Ceres autodiff functors,
// cost_functions.h:16-129, 137-256, 264-387, on the problems of optimizer.h:9-131,
// 247-369, 371-499).  Parameters (ambient): x[0]=o0, x[1..4]=q (w,x,y,z), x[5..7]=t,
// x[8]=scale, x[9]=o1, x[10]=f0 (or the shared focal), x[11]=f1.
// Tangent slots: 0 o0 | 1-3 rotation (q <- [cos|d|, sin|d| d/|d|] * q, i.e. R <- R(2d) R)
// | 4-6 t | 7 scale | 8 o1 | 9 f0 (shared focal) | 10 f1.
#pragma once
#include "mp_small.cuh"

namespace mp {

constexpr int LM_NA = 12, LM_NT = 11;

MP_HD void quat_to_rot(const double *q, double *R) { // normalises q (like ceres::QuaternionRotatePoint)
    const double n = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const double w = q[0] * n, x = q[1] * n, y = q[2] * n, z = q[3] * n;
    R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
    R[3] = 2 * (x * y + w * z); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - w * x);
    R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = 1 - 2 * (x * x + y * y);
}
MP_HD void rot_to_quat(const double *R, double *q) { // Eigen::Quaterniond(R) convention, w first (utils.h:111-114)
    const double tr = R[0] + R[4] + R[8];
    if (tr > 0) {
        const double s = sqrt(tr + 1.0) * 2;
        q[0] = 0.25 * s; q[1] = (R[7] - R[5]) / s; q[2] = (R[2] - R[6]) / s; q[3] = (R[3] - R[1]) / s;
    } else if (R[0] > R[4] && R[0] > R[8]) {
        const double s = sqrt(1.0 + R[0] - R[4] - R[8]) * 2;
        q[0] = (R[7] - R[5]) / s; q[1] = 0.25 * s; q[2] = (R[1] + R[3]) / s; q[3] = (R[2] + R[6]) / s;
    } else if (R[4] > R[8]) {
        const double s = sqrt(1.0 + R[4] - R[0] - R[8]) * 2;
        q[0] = (R[2] - R[6]) / s; q[1] = (R[1] + R[3]) / s; q[2] = 0.25 * s; q[3] = (R[5] + R[7]) / s;
    } else {
        const double s = sqrt(1.0 + R[8] - R[0] - R[4]) * 2;
        q[0] = (R[3] - R[1]) / s; q[1] = (R[2] + R[6]) / s; q[2] = (R[5] + R[7]) / s; q[3] = 0.25 * s;
    }
    const double n = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    for (int k = 0; k < 4; k++) q[k] *= n;
}

// Everything the residual blocks need that depends on the parameters only.
struct LmPoint {
    double R[9], t[3], o0, scale, o1, f0, f1;
};
MP_HD void lm_point(int kind, const double *x, LmPoint *p) {
    quat_to_rot(x + 1, p->R);
    p->t[0] = x[5]; p->t[1] = x[6]; p->t[2] = x[7];
    p->o0 = x[0]; p->scale = x[8]; p->o1 = x[9];
    p->f0 = kind == CALIB ? 1.0 : x[10];
    p->f1 = kind == CALIB ? 1.0 : (kind == TWO ? x[11] : x[10]);
}

// projection p = K Y (calibrated) or diag(f,f,1) Y, residual against (tx,ty), and the two
// rows D0, D1 of d(residual)/dY
template <int KIND> MP_HD void project_rows(const double *K, double f, const double *Y, double tx, double ty, double *r,
                                            double *D0, double *D1) {
    if (KIND == CALIB) {
        const double u = K[0] * Y[0] + K[1] * Y[1] + K[2] * Y[2], v = K[3] * Y[0] + K[4] * Y[1] + K[5] * Y[2];
        const double w = K[6] * Y[0] + K[7] * Y[1] + K[8] * Y[2], iw = 1.0 / w, uw = u * iw, vw = v * iw;
        r[0] = uw - tx; r[1] = vw - ty;
        for (int k = 0; k < 3; k++) { D0[k] = (K[k] - uw * K[6 + k]) * iw; D1[k] = (K[3 + k] - vw * K[6 + k]) * iw; }
    } else {
        const double iw = 1.0 / Y[2], uw = f * Y[0] * iw, vw = f * Y[1] * iw;
        r[0] = uw - tx; r[1] = vw - ty;
        D0[0] = f * iw; D0[1] = 0; D0[2] = -uw * iw;
        D1[0] = 0; D1[1] = f * iw; D1[2] = -vw * iw;
    }
}

// reprojection 0 -> 1 (cost_functions.h:27-43, 146-166, 273-293).  J[2][LM_NT] fully written.
template <int KIND> MP_HD void lm_row_reproj0(const PairConst &pc, const Match64 &m, const LmPoint &p, double *r,
                                              double (*J)[LM_NT]) {
    double a[3], da0 = 0, da1 = 0;
    if (KIND == CALIB) mat3_vec(pc.K0i, m.x0x, m.x0y, 1.0, a);
    else {
        const double i0 = 1.0 / p.f0;
        a[0] = m.x0x * i0; a[1] = m.x0y * i0; a[2] = 1.0;
        da0 = -a[0] * i0; da1 = -a[1] * i0;
    }
    const double s0 = m.d0 + p.o0;
    double Ra[3], RX[3], Y[3], D[2][3];
    mat3_vecp(p.R, a, Ra);
    for (int k = 0; k < 3; k++) { RX[k] = s0 * Ra[k]; Y[k] = RX[k] + p.t[k]; }
    project_rows<KIND>(pc.K1, p.f1, Y, m.x1x, m.x1y, r, D[0], D[1]);
    for (int e = 0; e < 2; e++) {
        const double *d = D[e];
        double *j = J[e], c[3];
        j[0] = dot3(d, Ra);
        cross3(RX, d, c);
        j[1] = 2 * c[0]; j[2] = 2 * c[1]; j[3] = 2 * c[2];
        j[4] = d[0]; j[5] = d[1]; j[6] = d[2];
        j[7] = 0; j[8] = 0; j[9] = 0; j[10] = 0;
        if (KIND != CALIB) {
            // f0 enters a; f1 enters the projection
            const double jf0 = s0 * (d[0] * (p.R[0] * da0 + p.R[1] * da1) + d[1] * (p.R[3] * da0 + p.R[4] * da1) +
                                     d[2] * (p.R[6] * da0 + p.R[7] * da1));
            const double jf1 = Y[e] / Y[2];
            if (KIND == SHARED) j[9] = jf0 + jf1;
            else { j[9] = jf0; j[10] = jf1; }
        }
    }
}

// reprojection 1 -> 0 (cost_functions.h:62-82, 182-205, 309-332)
template <int KIND> MP_HD void lm_row_reproj1(const PairConst &pc, const Match64 &m, const LmPoint &p, double *r,
                                              double (*J)[LM_NT]) {
    double b[3], db0 = 0, db1 = 0;
    if (KIND == CALIB) mat3_vec(pc.K1i, m.x1x, m.x1y, 1.0, b);
    else {
        const double i1 = 1.0 / p.f1;
        b[0] = m.x1x * i1; b[1] = m.x1y * i1; b[2] = 1.0;
        db0 = -b[0] * i1; db1 = -b[1] * i1;
    }
    const double dd = m.d1 + p.o1, s1 = dd * p.scale;
    double X[3], Y[3], Rtb[3], D[2][3];
    for (int k = 0; k < 3; k++) X[k] = s1 * b[k] - p.t[k];
    mat3T_vecp(p.R, X, Y);
    mat3T_vecp(p.R, b, Rtb);
    project_rows<KIND>(pc.K0, p.f0, Y, m.x0x, m.x0y, r, D[0], D[1]);
    for (int e = 0; e < 2; e++) {
        const double *d = D[e];
        double *j = J[e], Rd[3], c[3];
        mat3_vecp(p.R, d, Rd);
        const double dRtb = dot3(d, Rtb);
        j[0] = 0;
        cross3(Rd, X, c);
        j[1] = 2 * c[0]; j[2] = 2 * c[1]; j[3] = 2 * c[2];
        j[4] = -Rd[0]; j[5] = -Rd[1]; j[6] = -Rd[2];
        j[7] = dRtb * dd; j[8] = dRtb * p.scale; j[9] = 0; j[10] = 0;
        if (KIND != CALIB) {
            const double jf1 = s1 * (Rd[0] * db0 + Rd[1] * db1); // d . R^T (s1 db) = s1 (R d) . db
            const double jf0 = Y[e] / Y[2];
            if (KIND == SHARED) j[9] = jf0 + jf1;
            else { j[9] = jf0; j[10] = jf1; }
        }
    }
}

// weighted Sampson residual (cost_functions.h:100-124, 223-251, 350-382); one row
template <int KIND> MP_HD void lm_row_sampson(const PairConst &pc, const Match64 &m, const LmPoint &p, double W,
                                              double *r, double *J) {
    double a[3], b[3];
    const double i0 = 1.0 / p.f0, i1 = 1.0 / p.f1;
    if (KIND == CALIB) { mat3_vec(pc.K0i, m.x0x, m.x0y, 1.0, a); mat3_vec(pc.K1i, m.x1x, m.x1y, 1.0, b); }
    else { a[0] = m.x0x * i0; a[1] = m.x0y * i0; a[2] = 1.0; b[0] = m.x1x * i1; b[1] = m.x1y * i1; b[2] = 1.0; }
    double u[3], c[3], mm[3], g[3];
    mat3_vecp(p.R, a, u);
    cross3(p.t, u, c);   // E a
    cross3(b, p.t, mm);
    mat3T_vecp(p.R, mm, g); // E^T b
    const double e0 = c[0] * i1, e1 = c[1] * i1, g0 = g[0] * i0, g1 = g[1] * i0; // i0 = i1 = 1 when calibrated
    const double C = dot3(b, c), Dn = e0 * e0 + e1 * e1 + g0 * g0 + g1 * g1;
    const double isq = 1.0 / sqrt(Dn), k1 = W * isq, k2 = C / (2 * Dn);
    r[0] = C * k1;
    // directional derivative from (dc, dg[0..1], d(b).c, extra de, extra dg')
    auto dir = [&](const double *dc, double dg0, double dg1, double dbc, double xe0, double xe1, double xg0, double xg1) {
        const double dC = dbc + dot3(b, dc);
        const double de0 = dc[0] * i1 + xe0, de1 = dc[1] * i1 + xe1, dgp0 = dg0 * i0 + xg0, dgp1 = dg1 * i0 + xg1;
        const double dD = 2 * (e0 * de0 + e1 * de1 + g0 * dgp0 + g1 * dgp1);
        return k1 * (dC - k2 * dD);
    };
    J[0] = 0; J[7] = 0; J[8] = 0; J[9] = 0; J[10] = 0;
    for (int k = 0; k < 3; k++) {
        double ek[3] = {0, 0, 0}, dc[3], dm[3], w[3];
        ek[k] = 1.0;
        // translation t_k
        cross3(ek, u, dc);
        cross3(b, ek, dm);
        J[4 + k] = dir(dc, p.R[0] * dm[0] + p.R[3] * dm[1] + p.R[6] * dm[2], p.R[1] * dm[0] + p.R[4] * dm[1] + p.R[7] * dm[2],
                       0, 0, 0, 0, 0);
        // rotation delta_k: du = 2 e_k x u, dg = -2 R^T (e_k x m)
        cross3(ek, u, w);
        for (int i = 0; i < 3; i++) w[i] *= 2;
        cross3(p.t, w, dc);
        cross3(ek, mm, dm);
        J[1 + k] = dir(dc, -2 * (p.R[0] * dm[0] + p.R[3] * dm[1] + p.R[6] * dm[2]),
                       -2 * (p.R[1] * dm[0] + p.R[4] * dm[1] + p.R[7] * dm[2]), 0, 0, 0, 0, 0);
    }
    if (KIND != CALIB) {
        const double da[3] = {-a[0] * i0, -a[1] * i0, 0}, db[3] = {-b[0] * i1, -b[1] * i1, 0};
        double du[3], dc[3], dm[3], z3[3] = {0, 0, 0};
        mat3_vecp(p.R, da, du);
        cross3(p.t, du, dc);
        const double jf0 = dir(dc, 0, 0, 0, 0, 0, -g0 * i0, -g1 * i0);
        cross3(db, p.t, dm);
        const double jf1 = dir(z3, p.R[0] * dm[0] + p.R[3] * dm[1] + p.R[6] * dm[2],
                               p.R[1] * dm[0] + p.R[4] * dm[1] + p.R[7] * dm[2], dot3(db, c), -e0 * i1, -e1 * i1, 0, 0);
        if (KIND == SHARED) J[9] = jf0 + jf1;
        else { J[9] = jf0; J[10] = jf1; }
    }
}

// Partial sums of one evaluation: upper triangle of J^T J (full LM_NT slots), J^T r, cost.
struct LmAcc {
    double A[LM_NT * (LM_NT + 1) / 2], g[LM_NT], cost;
    MP_HD void clear() {
        for (int i = 0; i < LM_NT * (LM_NT + 1) / 2; i++) A[i] = 0;
        for (int i = 0; i < LM_NT; i++) g[i] = 0;
        cost = 0;
    }
    template <int NTK> MP_HD void add_row(double r, const double *j) {
        cost = fma(0.5 * r, r, cost);
        int k = 0;
#pragma unroll
        for (int a = 0; a < LM_NT; a++) {
#pragma unroll
            for (int b = a; b < LM_NT; b++, k++)
                if (a < NTK && b < NTK) A[k] = fma(j[a], j[b], A[k]);
            if (a < NTK) g[a] = fma(j[a], r, g[a]);
        }
    }
};
constexpr int LM_ACC_DOUBLES = LM_NT * (LM_NT + 1) / 2 + LM_NT + 1;

// one residual block (type t, match m) -> accumulators.  want_jac = false: cost only.
template <int KIND> MP_HD void lm_block(const PairConst &pc, const Match64 &m, int t, const LmPoint &p, double W,
                                        LmAcc *acc) {
    constexpr int NTK = KIND == CALIB ? 9 : (KIND == SHARED ? 10 : 11);
    if (t == 2) {
        double r, J[LM_NT];
        lm_row_sampson<KIND>(pc, m, p, W, &r, J);
        acc->add_row<NTK>(r, J);
    } else {
        double r[2], J[2][LM_NT];
        if (t == 0) lm_row_reproj0<KIND>(pc, m, p, r, J);
        else lm_row_reproj1<KIND>(pc, m, p, r, J);
        acc->add_row<NTK>(r[0], J[0]);
        acc->add_row<NTK>(r[1], J[1]);
    }
}

} // namespace mp
#include "mp_lm2.cuh"
#ifdef __CUDACC__
#include "mp_lm_warp.cuh"
#endif
