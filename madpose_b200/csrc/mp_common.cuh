// madpose_b200 -- shared host/device definitions of the B200 hybrid LO-RANSAC path.
// Everything marked MP_HD compiles for the device (the product) and for the host
// (unit-test harness of the scalar math only; the product has no CPU path).
#pragma once
#include "../../include/madpose_b200.h"
#include <cfloat>
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define MP_HD __host__ __device__ __forceinline__
#define MP_D __device__ __forceinline__
#else
#define MP_HD inline
#define MP_D inline
#endif

namespace mp {

enum { CALIB = 0, SHARED = 1, TWO = 2 };
constexpr int kMaxModelsAffine[3] = {4, 8, 4};   // solver 0: solver.cpp:104,269,455
constexpr int kMaxModelsPoints[3] = {10, 15, 3}; // solver 1: PoseLib 5pt / 6pt shared focal / 7pt
constexpr int kMaxModelsAny = 15;

// A model in the estimator's internal units (R row-major).  Same layout as mp_model.
struct Model {
    double R[9], t[3], scale, o0, o1, f0, f1;
};
static_assert(sizeof(Model) == sizeof(mp_model), "Model must mirror mp_model");

// One match in internal units, fp64: pixels (calibrated) or pp-subtracted pixels /
// norm_scale (focal kinds).  48-byte records: a warp pass reads them fully coalesced
// and a minimal sample gathers one record per index.
struct Match64 {
    double x0x, x0y, x1x, x1y, d0, d1;
};

// Per-pair constants written by the front-end kernel (prep_pairs_kernel).
// Mirrors what the reference keeps in its Hybrid*PoseEstimator object plus the
// rewritten options (hybrid_pose_estimator.cpp:13-23, .h:14-37).
struct PairConst {
    int n;           // matches
    int kind;        // CALIB / SHARED / TWO
    long long off;   // first match of this pair in the packed arrays
    long long off32; // first float4 of this pair's fp32 staging block
    double K0[9], K1[9], K0i[9], K1i[9]; // calibrated intrinsics (identity for focal kinds)
    double min_depth[2];
    double thr[3], w[3]; // "three data types": [reproj 0->1, reproj 1->0, sampson]
    double sampson_w;    // sampson_squared_weight
    double loss_scale;   // sampson_loss_scale_ (calibrated), 1 otherwise
    double norm_scale;   // poselib::normalize_points scale (focal kinds), 1 otherwise
    double pp0[2], pp1[2];
    int scale_only;      // 1: HybridEstimatePoseAndScale (hybrid_pose_estimator.cpp:37-63,297-442): kind == CALIB,
                         // offsets absent, depth guards in the reprojection residuals (quirk Q16)
    int pad_;
};

MP_HD void mat3_mul(const double *A, const double *B, double *C) {
    double T[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) T[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    for (int i = 0; i < 9; i++) C[i] = T[i];
}
MP_HD void mat3_vec(const double *A, double x, double y, double z, double *o) {
    o[0] = A[0] * x + A[1] * y + A[2] * z;
    o[1] = A[3] * x + A[4] * y + A[5] * z;
    o[2] = A[6] * x + A[7] * y + A[8] * z;
}
MP_HD void mat3T_vec(const double *A, double x, double y, double z, double *o) {
    o[0] = A[0] * x + A[3] * y + A[6] * z;
    o[1] = A[1] * x + A[4] * y + A[7] * z;
    o[2] = A[2] * x + A[5] * y + A[8] * z;
}
MP_HD double det3(const double *A) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
}
MP_HD bool inv3(const double *A, double *B) {
    double d = det3(A);
    if (d == 0.0 || !(fabs(d) <= DBL_MAX)) return false;
    double id = 1.0 / d, T[9];
    T[0] = (A[4] * A[8] - A[5] * A[7]) * id; T[1] = (A[2] * A[7] - A[1] * A[8]) * id; T[2] = (A[1] * A[5] - A[2] * A[4]) * id;
    T[3] = (A[5] * A[6] - A[3] * A[8]) * id; T[4] = (A[0] * A[8] - A[2] * A[6]) * id; T[5] = (A[2] * A[3] - A[0] * A[5]) * id;
    T[6] = (A[3] * A[7] - A[4] * A[6]) * id; T[7] = (A[1] * A[6] - A[0] * A[7]) * id; T[8] = (A[0] * A[4] - A[1] * A[3]) * id;
    for (int i = 0; i < 9; i++) B[i] = T[i];
    return true;
}
MP_HD bool finite_d(double v) { return fabs(v) <= DBL_MAX; }
MP_HD bool model_finite(const Model &m) {
    const double *p = reinterpret_cast<const double *>(&m);
    bool ok = true;
    for (int i = 0; i < MP_MODEL_DOUBLES; i++) ok = ok && finite_d(p[i]);
    return ok;
}

} // namespace mp
