// Stand-alone kernels of the path (front-end, scoring, inlier compaction, ALU peak
// probes).  The per-pair persistent kernel (mp_pair_kernel.cuh) reuses the same
// device functions; these entry points exist for mp_score_batch / mp_inliers and
// the roofline measurement.
#pragma once
#include "mp_block.cuh"

namespace mp {

// Front-end of the three estimators on the device (math in mp_frontend.cuh): one CTA per pair.
__global__ void __launch_bounds__(256) prep_pairs_kernel(const PairRaw *raw, const double *X0, const double *X1,
                                                         const double *D0, const double *D1, PairConst *pcs,
                                                         Match64 *M, float4 *S32) {
    __shared__ double red[32];
    const PairRaw &r = raw[blockIdx.x];
    const int n = r.n;
    const double *x0 = X0 + 2 * r.off, *x1 = X1 + 2 * r.off, *d0 = D0 + r.off, *d1 = D1 + r.off;
    Match64 *m = M + r.off;
    float4 *P = S32 + r.off32, *Q = P + n;
    double ns = 1.0, K0i[9], K1i[9];
    if (r.kind != CALIB) {
        double acc = 0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) acc += norm_scale_term(r, x0 + 2 * i, x1 + 2 * i);
        ns = block_sum(acc, red) / (sqrt(2.0) * n);
    } else {
        inv3(r.K0, K0i);
        inv3(r.K1, K1i);
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float p[4], q[4];
        prep_match(r, ns, K0i, K1i, x0 + 2 * i, x1 + 2 * i, d0[i], d1[i], &m[i], p, q);
        P[i] = make_float4(p[0], p[1], p[2], p[3]);
        Q[i] = make_float4(q[0], q[1], q[2], q[3]);
    }
    if (threadIdx.x == 0) prep_constants(r, ns, &pcs[blockIdx.x]);
}

// fp64 models -> fp32 hypothesis parameter blocks
__global__ void pack_hyps_kernel(int kind, int nm, const Model *models, HypF *hyps) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nm) pack_hyp(kind, models[i], hyps + i);
}

// K3 stand-alone: all hypotheses of one pair.  Persistent CTAs; every CTA stages the
// pair's fp32 tile once (bulk copy), then its warps walk hypothesis groups.
template <int KIND>
__global__ void __launch_bounds__(MP_CTA, 1) score_f32_kernel(const PairConst *pcp, const float4 *S32, int score_type,
                                                              int nm, const HypF *hyps, float *scores, int tile_cap) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    const PairConst &pc = *pcp;
    const PairF pf = make_pairf(pc, score_type);
    const float w0 = (float)pc.w[0], w2 = (float)pc.w[2];
    float4 *sP = reinterpret_cast<float4 *>(smem_raw), *sQ = sP + tile_cap;
    const float4 *gP = S32 + pc.off32, *gQ = gP + pc.n;
    if (threadIdx.x == 0) mbar_init(&bar, 1);
    __syncthreads();
    uint32_t phase = 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gw = blockIdx.x * MP_WARPS + warp, nw = gridDim.x * MP_WARPS;
    const int groups = (nm + MP_HT - 1) / MP_HT;
    for (int base = 0; base < pc.n; base += tile_cap) {
        const int len = min(tile_cap, pc.n - base);
        __syncthreads(); // previous tile fully consumed
        stage_two(sP, gP + base, len * 16u, sQ, gQ + base, len * 16u, &bar, phase);
        TileF tile{sP, sQ, len};
        for (int g = gw; g < groups; g += nw) {
            const int h0 = g * MP_HT, cnt = min(MP_HT, nm - h0);
            float out[MP_HT];
            warp_score_chunk<KIND>(tile, pf, w0, w2, hyps + h0, cnt, 0, len, out);
            if (lane == 0)
                for (int j = 0; j < cnt; j++) scores[h0 + j] = (base == 0 ? 0.f : scores[h0 + j]) + out[j];
        }
    }
}

// exact fp64 ScoreModel: one CTA per model, fixed summation tree
__global__ void __launch_bounds__(256) score_f64_kernel(const PairConst *pcp, const Match64 *M, int score_type, int nm,
                                                        const Model *models, double *scores) {
    __shared__ double red[32];
    const PairConst &pc = *pcp;
    const Match64 *m = M + pc.off;
    for (int h = blockIdx.x; h < nm; h += gridDim.x) {
        Eval64 e;
        make_eval64(pc, models[h], &e);
        double acc = 0;
        for (int i = threadIdx.x; i < pc.n; i += blockDim.x) {
            Match64 mm = m[i];
            double r0 = score_type == 1 ? DBL_MAX : eval64(e, mm, 0);
            double r1 = score_type == 1 ? DBL_MAX : eval64(e, mm, 1);
            double r2 = score_type == 2 ? DBL_MAX : eval64(e, mm, 2);
            acc += pc.w[0] * msac64(r0, pc.thr[0]) + pc.w[1] * msac64(r1, pc.thr[1]) + pc.w[2] * msac64(r2, pc.thr[2]);
        }
        double s = block_sum(acc, red);
        if (threadIdx.x == 0) scores[h] = s;
    }
}

// K4 stand-alone: GetInliers for one model -> three ascending index lists.
// One CTA; ballot + running offsets keep the lists ordered.
__global__ void __launch_bounds__(MP_CTA) inliers_kernel(const PairConst *pcp, const Match64 *M, const Model *model,
                                                         const double *thr3, int *out_idx, int *out_counts) {
    __shared__ int wcnt[3][MP_WARPS];
    __shared__ int base[3];
    const PairConst &pc = *pcp;
    const Match64 *m = M + pc.off;
    Eval64 e;
    make_eval64(pc, *model, &e);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 3) base[threadIdx.x] = 0;
    __syncthreads();
    for (int start = 0; start < pc.n; start += MP_CTA) {
        const int i = start + threadIdx.x;
        bool in[3] = {false, false, false};
        if (i < pc.n) {
            Match64 mm = m[i];
            for (int t = 0; t < 3; t++) in[t] = eval64(e, mm, t) < thr3[t];
        }
        unsigned bal[3];
        for (int t = 0; t < 3; t++) {
            bal[t] = __ballot_sync(0xffffffffu, in[t]);
            if (lane == 0) wcnt[t][warp] = __popc(bal[t]);
        }
        __syncthreads();
        for (int t = 0; t < 3; t++) {
            int off = base[t];
            for (int w = 0; w < warp; w++) off += wcnt[t][w];
            if (in[t]) out_idx[t * pc.n + off + __popc(bal[t] & ((1u << lane) - 1u))] = i;
        }
        __syncthreads();
        if (threadIdx.x < 3) {
            int s = 0;
            for (int w = 0; w < MP_WARPS; w++) s += wcnt[threadIdx.x][w];
            base[threadIdx.x] += s;
        }
        __syncthreads();
    }
    if (threadIdx.x < 3) out_counts[threadIdx.x] = base[threadIdx.x];
}

// register-only FMA chains: the measured ALU peaks the roofline fractions are quoted against
__global__ void __launch_bounds__(256) ffma_peak_kernel(float *out, int iters) {
    float a[8], b = 1.000001f, c = 1e-7f;
    for (int k = 0; k < 8; k++) a[k] = threadIdx.x * 1e-3f + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = fmaf(a[k], b, c);
    }
    float s = 0;
    for (int k = 0; k < 8; k++) s += a[k];
    if (s == 123.456f) out[0] = s;
}
__global__ void __launch_bounds__(256) dfma_peak_kernel(double *out, int iters) {
    double a[8], b = 1.000001, c = 1e-7;
    for (int k = 0; k < 8; k++) a[k] = threadIdx.x * 1e-3 + k;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = fma(a[k], b, c);
    }
    double s = 0;
    for (int k = 0; k < 8; k++) s += a[k];
    if (s == 123.456) out[0] = s;
}

} // namespace mp
