// K3 -- fused hypotheses x matches truncated-MSAC scoring (fp32), device side.
// Replaces ScoreModel + EvaluateModelOnPoint (hybrid_ransac.h:265-282 and the three
// estimator .cpp files) for the bulk of the hypotheses; the fp32 score is a FILTER:
// every decision the reference takes on a score (new best, argmin) is re-taken on the
// exact fp64 score of the few hypotheses inside a guard band (see mp_pair_kernel.cu).
//
// Mapping: lanes = matches (conflict-free LDS.128 from the staged tile), each warp
// holds MP_HT hypotheses in registers and walks a chunk of the tile; the three
// truncated sums are reduced with warp shuffles at the end of the chunk.
#pragma once
#include "mp_residuals.cuh"

namespace mp {

constexpr int MP_HT = 2;          // hypotheses held in registers per warp pass
constexpr int MP_CTA = 512;       // threads per CTA (16 warps)
constexpr int MP_WARPS = MP_CTA / 32;

// staged fp32 tile of one pair.  calibrated: P[i] = (ax, ay, bx, by) K^-1-normalised
// points, Q[i] = (d0, d1, 1/|a_h|, 1/|b_h|).  focal kinds: P[i] = (x0x, x0y, x1x, x1y),
// Q[i] = (d0, d1, 0, 0).
struct TileF {
    const float4 *P, *Q;
    int n;
};

struct HypR { // register copy of the HypF fields the evaluation uses
    float A[9], t[3], B[9], v1[3], o0, s, so1, f0, f1, if0, if1;
};

__device__ __forceinline__ void load_hyp(const HypF *h, HypR &r) {
    const float4 *p = reinterpret_cast<const float4 *>(h);
    float4 a = p[0], b = p[1], c = p[2], d = p[3], e = p[4], f = p[5], g = p[6], hh = p[7];
    r.A[0] = a.x; r.A[1] = a.y; r.A[2] = a.z; r.A[3] = a.w; r.A[4] = b.x; r.A[5] = b.y; r.A[6] = b.z; r.A[7] = b.w;
    r.A[8] = c.x; r.t[0] = c.y; r.t[1] = c.z; r.t[2] = c.w;
    r.B[0] = d.x; r.B[1] = d.y; r.B[2] = d.z; r.B[3] = d.w; r.B[4] = e.x; r.B[5] = e.y; r.B[6] = e.z; r.B[7] = e.w;
    r.B[8] = f.x; r.v1[0] = f.y; r.v1[1] = f.z; r.v1[2] = f.w;
    r.o0 = g.x; r.s = g.y; r.so1 = g.z; r.f0 = g.w; r.f1 = hh.x; r.if0 = hh.y; r.if1 = hh.z;
}

// one MUFU.RCP: __fdividef(1, x) guards denormal inputs with 2 selects, a compare and an extra multiply -- 24 of the 167
// instructions of the two-hypothesis loop body.  A denormal denominator flushes to zero -> inf -> the truncation
// returns the threshold (fminf drops a NaN), which is what the exact value would give as well.
__device__ __forceinline__ float rcpf(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// One hypothesis x one match: adds the three truncated residuals to s0,s1,s2.
template <int KIND>
__device__ __forceinline__ void eval32(const HypR &h, const PairF &pf, float4 P, float4 Q, float &s0, float &s1,
                                       float &s2) {
    // u = A [x0;1], w = B [x1;1]
    float ux = fmaf(h.A[0], P.x, fmaf(h.A[1], P.y, h.A[2]));
    float uy = fmaf(h.A[3], P.x, fmaf(h.A[4], P.y, h.A[5]));
    float uz = fmaf(h.A[6], P.x, fmaf(h.A[7], P.y, h.A[8]));
    float wx = fmaf(h.B[0], P.z, fmaf(h.B[1], P.w, h.B[2]));
    float wy = fmaf(h.B[3], P.z, fmaf(h.B[4], P.w, h.B[5]));
    float wz = fmaf(h.B[6], P.z, fmaf(h.B[7], P.w, h.B[8]));
    // type 0: q = (d0+o0) u + t, projected with K1, compared with x1
    float sa = Q.x + h.o0;
    float qx = fmaf(sa, ux, h.t[0]), qy = fmaf(sa, uy, h.t[1]), qz = fmaf(sa, uz, h.t[2]);
    float rz = rcpf(qz);
    float r0;
    if (KIND == CALIB) {
        float ex = fmaf(qx, rz, -P.z), ey = fmaf(qy, rz, -P.w);
        float rx = fmaf(pf.k1fx, ex, pf.k1sk * ey), ry = pf.k1fy * ey;
        r0 = fmaf(rx, rx, ry * ry);
    } else {
        float fr = h.f1 * rz;
        float rx = fmaf(qx, fr, -P.z), ry = fmaf(qy, fr, -P.w);
        r0 = fmaf(rx, rx, ry * ry);
    }
    r0 = (qz < 1e-2f) ? pf.thr0 : fminf(r0, pf.thr0);
    // type 1: q = scale (d1+o1) w - R^T t, projected with K0, compared with x0
    float sb = fmaf(h.s, Q.y, h.so1);
    float px = fmaf(sb, wx, h.v1[0]), py = fmaf(sb, wy, h.v1[1]), pz = fmaf(sb, wz, h.v1[2]);
    float rz1 = rcpf(pz);
    float r1;
    if (KIND == CALIB) {
        float ex = fmaf(px, rz1, -P.x), ey = fmaf(py, rz1, -P.y);
        float rx = fmaf(pf.k0fx, ex, pf.k0sk * ey), ry = pf.k0fy * ey;
        r1 = fmaf(rx, rx, ry * ry);
    } else {
        float fr = h.f0 * rz1;
        float rx = fmaf(px, fr, -P.x), ry = fmaf(py, fr, -P.y);
        r1 = fmaf(rx, rx, ry * ry);
    }
    r1 = (pz < 1e-2f) ? pf.thr1 : fminf(r1, pf.thr1);
    // type 2: Sampson.  E a = t x u, E^T b = v1 x w
    float c0 = fmaf(h.t[1], uz, -h.t[2] * uy), c1 = fmaf(h.t[2], ux, -h.t[0] * uz), c2 = fmaf(h.t[0], uy, -h.t[1] * ux);
    float g0 = fmaf(h.v1[1], wz, -h.v1[2] * wy), g1 = fmaf(h.v1[2], wx, -h.v1[0] * wz);
    float r2;
    if (KIND == CALIB) {
        float C = fmaf(P.z, c0, fmaf(P.w, c1, c2));
        float den = fmaf(c0, c0, fmaf(c1, c1, fmaf(g0, g0, g1 * g1)));
        r2 = C * C * pf.loss * rcpf(den);
        // cheirality on unit bearings (solver.cpp:1188-1206), Q.z = 1/|a_h|, Q.w = 1/|b_h|
        float cb = fmaf(ux, P.z, fmaf(uy, P.w, uz));
        float e1 = fmaf(ux, h.t[0], fmaf(uy, h.t[1], uz * h.t[2]));
        float e2 = fmaf(P.z, h.t[0], fmaf(P.w, h.t[1], h.t[2]));
        float al = -(Q.z * Q.w) * cb, b1 = -Q.z * e1, b2 = Q.w * e2;
        float l1 = fmaf(-al, b2, b1), l2 = fmaf(-al, b1, b2), md = fmaf(-1e-2f * al, al, 1e-2f);
        r2 = (l1 > md && l2 > md) ? fminf(r2, pf.thr2) : pf.thr2;
    } else {
        // F x0 = K1^-1 (t x u), F^T x1 = K0^-1 (v1 x w) on the (normalised) pixel coordinates
        float e0 = c0 * h.if1, e1 = c1 * h.if1, k0 = g0 * h.if0, k1 = g1 * h.if0;
        float C = fmaf(P.z, e0, fmaf(P.w, e1, c2));
        float den = fmaf(e0, e0, fmaf(e1, e1, fmaf(k0, k0, k1 * k1)));
        r2 = fminf(C * C * rcpf(den), pf.thr2);
    }
    s0 += r0; s1 += r1; s2 += r2;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One warp: MP_HT hypotheses (hyp[0..cnt)) over matches [lo, hi) of the staged tile.
// Returns (valid on every lane) the weighted truncated score of each hypothesis for the chunk.
template <int KIND>
__device__ __forceinline__ void warp_score_chunk(const TileF &tile, const PairF &pf, const float w0, const float w2,
                                                 const HypF *hyp, int cnt, int lo, int hi, float *out /*[MP_HT]*/) {
    const int lane = threadIdx.x & 31;
    HypR h[MP_HT];
#pragma unroll
    for (int j = 0; j < MP_HT; j++) load_hyp(hyp + (j < cnt ? j : 0), h[j]);
    float s0[MP_HT], s1[MP_HT], s2[MP_HT];
#pragma unroll
    for (int j = 0; j < MP_HT; j++) { s0[j] = 0.f; s1[j] = 0.f; s2[j] = 0.f; }
    for (int i = lo + lane; i < hi; i += 32) {
        float4 P = tile.P[i], Q = tile.Q[i];
#pragma unroll
        for (int j = 0; j < MP_HT; j++) eval32<KIND>(h[j], pf, P, Q, s0[j], s1[j], s2[j]);
    }
    const float m = (float)(hi - lo);
#pragma unroll
    for (int j = 0; j < MP_HT; j++) {
        float a = warp_sum(s0[j]), b = warp_sum(s1[j]), c = warp_sum(s2[j]);
        if (pf.gate == 1) { a = m * pf.thr0; b = m * pf.thr1; }  // EPI_ONLY: reprojection types masked
        if (pf.gate == 2) { c = m * pf.thr2; }                    // MD_ONLY: Sampson masked
        out[j] = fmaf(w0, a + b, w2 * c);
    }
}

// Same, for hypotheses addressed through an index list (the round's compacted hypotheses).
template <int KIND>
__device__ __forceinline__ void warp_score_list(const TileF &tile, const PairF &pf, const float w0, const float w2,
                                                const HypF *hyps, const unsigned short *list, int cnt, int lo, int hi,
                                                float *out /*[MP_HT]*/) {
    const int lane = threadIdx.x & 31;
    HypR h[MP_HT];
#pragma unroll
    for (int j = 0; j < MP_HT; j++) load_hyp(hyps + list[j < cnt ? j : 0], h[j]);
    float s0[MP_HT], s1[MP_HT], s2[MP_HT];
#pragma unroll
    for (int j = 0; j < MP_HT; j++) { s0[j] = 0.f; s1[j] = 0.f; s2[j] = 0.f; }
    for (int i = lo + lane; i < hi; i += 32) {
        float4 P = tile.P[i], Q = tile.Q[i];
#pragma unroll
        for (int j = 0; j < MP_HT; j++) eval32<KIND>(h[j], pf, P, Q, s0[j], s1[j], s2[j]);
    }
    const float m = (float)(hi - lo);
#pragma unroll
    for (int j = 0; j < MP_HT; j++) {
        float a = warp_sum(s0[j]), b = warp_sum(s1[j]), c = warp_sum(s2[j]);
        if (pf.gate == 1) { a = m * pf.thr0; b = m * pf.thr1; }
        if (pf.gate == 2) { c = m * pf.thr2; }
        out[j] = fmaf(w0, a + b, w2 * c);
    }
}

} // namespace mp
