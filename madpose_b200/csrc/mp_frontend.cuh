// Front-end math of the three estimators (MP_HD): principal-point subtraction and the
// poselib::normalize_points-style isotropic scaling of the focal kinds
// (...shared_focal_estimator.cpp:14-25, ...two_focal_estimator.cpp:42-51), the "three data
// types" option rewrite (hybrid_pose_estimator.cpp:16-23), K^-1 and sampson_loss_scale
// (hybrid_pose_estimator.h:21,35-36), the fp64 match records and the fp32 staging records.
// The kernel that runs it is prep_pairs_kernel (mp_kernels.cuh).
#pragma once
#include "mp_common.cuh"

namespace mp {

// Host-filled description of one pair's raw input (already resident in HBM).
struct PairRaw {
    int n, kind;
    long long off;   // match offset into the packed raw arrays (x: 2*off)
    long long off32; // float4 offset of the fp32 staging block (multiple of 1)
    double K0[9], K1[9], pp0[2], pp1[2], min_depth[2];
    double thr[2], w[2]; // options.squared_inlier_thresholds / data_type_weights as given
    int scale_only, pad_;
};

// poselib::normalize_points(x1, x2, T1, T2, true, false, true) [PoseLib v2.0.4, not in the
// reference tree; SURVEY.md A.1]: scale = sum_k(|x1_k| + |x2_k|) / (sqrt(2) N) on the
// pp-subtracted points; this is one term of the sum.
MP_HD double norm_scale_term(const PairRaw &r, const double *x0, const double *x1) {
    const double ax = x0[0] - r.pp0[0], ay = x0[1] - r.pp0[1], bx = x1[0] - r.pp1[0], by = x1[1] - r.pp1[1];
    return sqrt(ax * ax + ay * ay) + sqrt(bx * bx + by * by);
}

// one match -> fp64 record + fp32 staging records P, Q (layout: mp_score.cuh TileF)
MP_HD void prep_match(const PairRaw &r, double ns, const double *K0i, const double *K1i, const double *x0,
                      const double *x1, double d0, double d1, Match64 *o, float *p, float *q) {
    o->d0 = d0; o->d1 = d1;
    if (r.kind == CALIB) {
        o->x0x = x0[0]; o->x0y = x0[1]; o->x1x = x1[0]; o->x1y = x1[1];
        double a[3], b[3];
        mat3_vec(K0i, o->x0x, o->x0y, 1.0, a);
        mat3_vec(K1i, o->x1x, o->x1y, 1.0, b);
        p[0] = (float)a[0]; p[1] = (float)a[1]; p[2] = (float)b[0]; p[3] = (float)b[1];
        q[0] = (float)d0; q[1] = (float)d1;
        if (r.scale_only) { // hybrid_pose_estimator.cpp:386,398: a prior depth below 1e-2 makes the reprojection residual
            // invalid whatever the model; a NaN depth in the fp32 record does that for free (fminf(NaN, thr) = thr)
            if (d0 < 1e-2) q[0] = nanf("");
            if (d1 < 1e-2) q[1] = nanf("");
        }
        q[2] = (float)(1.0 / sqrt(a[0] * a[0] + a[1] * a[1] + 1.0));
        q[3] = (float)(1.0 / sqrt(b[0] * b[0] + b[1] * b[1] + 1.0));
    } else {
        o->x0x = (x0[0] - r.pp0[0]) / ns; o->x0y = (x0[1] - r.pp0[1]) / ns;
        o->x1x = (x1[0] - r.pp1[0]) / ns; o->x1y = (x1[1] - r.pp1[1]) / ns;
        p[0] = (float)o->x0x; p[1] = (float)o->x0y; p[2] = (float)o->x1x; p[3] = (float)o->x1y;
        q[0] = (float)d0; q[1] = (float)d1; q[2] = 0.f; q[3] = 0.f;
    }
}

MP_HD void prep_constants(const PairRaw &r, double ns, PairConst *pcp) {
    PairConst &pc = *pcp;
    const int kind = r.kind;
    pc.n = r.n; pc.kind = kind; pc.off = r.off; pc.off32 = r.off32;
    double K0i[9], K1i[9];
    const double eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (kind == CALIB) { inv3(r.K0, K0i); inv3(r.K1, K1i); }
    for (int i = 0; i < 9; i++) {
        pc.K0[i] = kind == CALIB ? r.K0[i] : eye[i]; pc.K1[i] = kind == CALIB ? r.K1[i] : eye[i];
        pc.K0i[i] = kind == CALIB ? K0i[i] : eye[i]; pc.K1i[i] = kind == CALIB ? K1i[i] : eye[i];
    }
    double tr = r.thr[0], te = r.thr[1];
    if (kind != CALIB) { tr /= ns * ns; te /= ns * ns; }
    const double ws = r.w[1] * (2 * tr / te);
    pc.sampson_w = ws;
    pc.w[0] = r.w[0]; pc.w[1] = r.w[0]; pc.w[2] = ws;
    pc.thr[0] = tr; pc.thr[1] = tr; pc.thr[2] = te;
    pc.loss_scale = 1.0;
    if (kind == CALIB) {
        const double l = 1.0 / (r.K0[0] + r.K0[4]) + 1.0 / (r.K1[0] + r.K1[4]);
        pc.loss_scale = 1.0 / (l * l);
    }
    pc.norm_scale = ns;
    for (int k = 0; k < 2; k++) { pc.min_depth[k] = r.min_depth[k]; pc.pp0[k] = r.pp0[k]; pc.pp1[k] = r.pp1[k]; }
    pc.scale_only = r.scale_only; pc.pad_ = 0;
}

} // namespace mp
