// K1/K2 translation unit: the per-sample fp64 minimal solvers (compiled once, shared by the
// stand-alone entry point and the loop's batch kernel) and mp_minimal_solve_batch.
#include <algorithm>
#include "mp_host.cuh"
#include "mp_loop.cuh"
#include "mp_minimal.cuh"
#include "mp_warp6.cuh"
#include "mp_estimate.cuh"

#ifndef MP_MIN_CTAS
#define MP_MIN_CTAS 12
#endif
namespace mp {

__device__ __noinline__ int minimal_solver_ni(const PairConst &pc, const Match64 *M, int solver_idx, const int *idx,
                                              unsigned mflags, Model *out, bool with_tail = true) {
    return minimal_solver(pc, M, solver_idx, idx, mflags, out, with_tail);
}
__device__ __noinline__ bool minimal_tail_ni(const PairConst &pc, const Match64 *M, const int *idx, unsigned mflags, Model *m) {
    if (pc.kind == CALIB) return minimal_points_tail<CALIB>(pc, M, idx, mflags, m);
    if (pc.kind == SHARED) return minimal_points_tail<SHARED>(pc, M, idx, mflags, m);
    return minimal_points_tail<TWO>(pc, M, idx, mflags, m);
}

// K1/K2 stand-alone: one thread per sample of one pair (mp_minimal_solve_batch)
__global__ void __launch_bounds__(64) minimal_kernel(const PairConst *pcp, const Match64 *M, int solver_idx, int k,
                                                     int n_samples, const int *idx, unsigned mflags, int max_models,
                                                     Model *out, int *counts) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_samples) return;
    const PairConst &pc = *pcp;
    Model tmp[kMaxModelsAny];
    int c = minimal_solver_ni(pc, M + pc.off, solver_idx, idx + (size_t)s * k, mflags, tmp);
    if (c > max_models) c = max_models;
    counts[s] = c;
    for (int j = 0; j < c; j++) out[(size_t)s * max_models + j] = tmp[j];
}

// ---- stand-alone solver bindings (bindings.cpp:156-166): one thread per explicit sample.
// xh / yh: per sample a 3 x NP row-major matrix whose COLUMNS are the homogeneous points (the layout numpy hands to the
// reference's Eigen::Matrix3d / Matrix3x4d arguments)
template <int KIND> __device__ void affine_sample_solve(const double *xh, const double *yh, const double *dx, const double *dy,
                                                        double *sols6, int *nsols, Model *models, int *nmodels) {
    constexpr int NPT = KIND == CALIB ? 3 : 4;
    double x0[NPT][3], x1[NPT][3], d0[NPT], d1[NPT];
    for (int k = 0; k < NPT; k++) {
        for (int c = 0; c < 3; c++) { x0[k][c] = xh[c * NPT + k]; x1[k][c] = yh[c * NPT + k]; }
        d0[k] = dx[k]; d1[k] = dy[k];
    }
    AffSol sols[8];
    int ns = 0;
    Model tmp[8];
    const int nm = affine_pose_models<KIND>(x0, x1, d0, d1, tmp, sols, &ns);
    for (int q = 0; q < ns; q++) { // solver.cpp:117-121, 285-289, 472-477: (1, b1, a2, b2 a2[, f | f1, f2])
        double *o = sols6 + 6 * q;
        o[0] = 1.0; o[1] = sols[q].b1; o[2] = sols[q].a2; o[3] = sols[q].b2 * sols[q].a2; o[4] = sols[q].f1; o[5] = sols[q].f2;
    }
    *nsols = ns;
    for (int q = 0; q < nm; q++) models[q] = tmp[q];
    *nmodels = nm;
}
__global__ void __launch_bounds__(64) affine_solve_kernel(int kind, int n_samples, const double *xh, const double *yh, const double *dx,
                                                          const double *dy, double *sols, int *nsols, Model *models, int *nmodels) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_samples) return;
    const int np_ = kind == CALIB ? 3 : 4;
    const double *a = xh + (size_t)s * 3 * np_, *b = yh + (size_t)s * 3 * np_, *c = dx + (size_t)s * np_, *d = dy + (size_t)s * np_;
    if (kind == CALIB) affine_sample_solve<CALIB>(a, b, c, d, sols + (size_t)s * 48, nsols + s, models + (size_t)s * 8, nmodels + s);
    else if (kind == SHARED) affine_sample_solve<SHARED>(a, b, c, d, sols + (size_t)s * 48, nsols + s, models + (size_t)s * 8, nmodels + s);
    else affine_sample_solve<TWO>(a, b, c, d, sols + (size_t)s * 48, nsols + s, models + (size_t)s * 8, nmodels + s);
}
// estimate_scale_and_pose (solver.cpp:5-33): X, Y are 3 x n row-major (columns = points), W n weights; one thread
__global__ void umeyama_kernel(int n, const double *X, const double *Y, const double *W, Model *out, int *ok) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    // two passes over the points straight from global memory (n is arbitrary)
    double cx[3] = {0, 0, 0}, cy[3] = {0, 0, 0}, S[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, ws = 0;
    for (int i = 0; i < n; i++) {
        ws += W[i];
        for (int k = 0; k < 3; k++) { cx[k] += X[k * n + i] * W[i]; cy[k] += Y[k * n + i] * W[i]; }
    }
    for (int k = 0; k < 3; k++) { cx[k] /= ws; cy[k] /= ws; }
    for (int i = 0; i < n; i++)
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) S[3 * a + b] += (Y[a * n + i] - cy[a]) * W[i] * (X[b * n + i] - cx[b]);
    Model m;
    m.o0 = 0; m.o1 = 0; m.f0 = 1; m.f1 = 1; m.scale = 1;
    for (int k = 0; k < 9; k++) m.R[k] = 0;
    for (int k = 0; k < 3; k++) m.t[k] = 0;
    *ok = 0;
    if (procrustes_rotation(S, m.R)) {
        double num = 0, den = 0;
        for (int i = 0; i < n; i++) {
            const double xc[3] = {X[i] - cx[0], X[n + i] - cx[1], X[2 * n + i] - cx[2]};
            double xr[3];
            mat3_vecp(m.R, xc, xr);
            for (int k = 0; k < 3; k++) { num += (Y[k * n + i] - cy[k]) * xr[k]; den += xr[k] * xr[k]; }
        }
        m.scale = num / den;
        double Rc[3];
        mat3_vecp(m.R, cx, Rc);
        for (int k = 0; k < 3; k++) m.t[k] = cy[k] - m.scale * Rc[k];
        *ok = 1;
    }
    *out = m;
}

// ---- K2, shared focal: ONE WARP per point-solver sample (mp_warp6.cuh).  Stand-alone form (mp_minimal_solve_batch):
// the candidates' tail (triangulation + depth fits) runs one candidate per lane, survivors are compacted in order.
constexpr int W6_WARPS = 4;
__device__ inline void warp6_load_sample(Warp6Work &W, const Match64 *M, const int *idx, int lane) {
    if (lane < 6) {
        const Match64 m = M[idx[lane]];
        W.x1[lane][0] = m.x0x; W.x1[lane][1] = m.x0y; W.x1[lane][2] = 1.0;
        W.x2[lane][0] = m.x1x; W.x2[lane][1] = m.x1y; W.x2[lane][2] = 1.0;
    }
    __syncwarp();
}
__global__ void __launch_bounds__(W6_WARPS * 32) minimal6_warp_kernel(const PairConst *pcp, const Match64 *M, int n_samples, const int *idx,
                                                                      unsigned mflags, int max_models, Model *out, int *counts, Model *scratch) {
    __shared__ Warp6Work work[W6_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PairConst &pc = *pcp;
    Warp6Work &W = work[warp];
    for (int s = blockIdx.x * W6_WARPS + warp; s < n_samples; s += gridDim.x * W6_WARPS) {
        warp6_load_sample(W, M + pc.off, idx + (size_t)s * 6, lane);
        Model *cand = scratch + (size_t)s * kMaxModelsAny;
        const int c = warp_relpose_6pt<false>(W, cand, lane);
        __syncwarp();
        bool ok = false;
        Model m;
        if (lane < c) { m = cand[lane]; ok = minimal_tail_ni(pc, M + pc.off, idx + (size_t)s * 6, mflags, &m); }
        const unsigned keep = __ballot_sync(0xffffffffu, ok);
        const int k = __popc(keep & ((1u << lane) - 1u));
        if (ok && k < max_models) out[(size_t)s * max_models + k] = m;
        if (lane == 0) counts[s] = min(__popc(keep), max_models);
        __syncwarp();
    }
}
// the loop's form: the point-solver samples of the round (list 1 of the sample kernel); candidates go to the model
// slots and to the candidate list of minimal_tail_kernel, exactly as minimal_batch_kernel leaves them
#ifndef W6B_WARPS
#define W6B_WARPS 12   // warps per CTA of the loop's kernel: one CTA per SM, stages in lockstep
#endif
#ifndef W6B_CTAS
#define W6B_CTAS 1      // resident CTAs per SM the kernel is compiled for
#endif
__global__ void __launch_bounds__(W6B_WARPS * 32, W6B_CTAS) minimal6_warp_batch_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    extern __shared__ __align__(16) unsigned char w6_smem[];
    Warp6Work *work = reinterpret_cast<Warp6Work *>(w6_smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = first / L.gsz;
    const int n1 = L.n_items_w[grp];
    Warp6Work &W = work[warp];
    for (int g0 = blockIdx.x * W6B_WARPS; g0 < n1; g0 += gridDim.x * W6B_WARPS) { // uniform trip count over the CTA
        const int g = g0 + warp;
        const bool valid = g < n1;
        const int item = valid ? L.items[(size_t)(2 * L.n_pairs + first) * LP_BMAX + g] : 0;
        const int p = first + item / LP_BMAX, b = item % LP_BMAX;
        PairWork &Wk = L.wk[p];
        const PairConst &pc = L.pc[p];
        if (valid) warp6_load_sample(W, L.M + pc.off, Wk.idx[b], lane);
        Model *out = L.models + ((size_t)p * LP_BMAX + b) * LP_MM;
        const int c = warp_relpose_6pt<true>(W, out, lane, valid);
        if (lane == 0 && valid) {
            Wk.count[b] = c;
            if (c > 0) {
                const int base = atomicAdd(L.n_cand + grp, c);
                int *cl = L.cand + (size_t)first * LP_BMAX * LP_MM;
                for (int j = 0; j < c; j++) cl[base + j] = item * LP_MM + j;
            }
        }
        __syncwarp();
    }
}

// ---- K1/K2 over the samples of the round, one thread per sample, ONE kernel per (estimator kind, solver): each launch
// carries only its own solver's code (all six in one kernel were 113 k instructions = 1.8 MB, thrashing the
// instruction cache, and every thread reserved the largest solver's 18 KB stack frame).  The samples come from the
// per-solver lists the sample kernel filled.
template <int KIND, int SOLVER>
__global__ void __launch_bounds__(64, MP_MIN_CTAS) minimal_batch_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const LoopParams &prm = *prmp;
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int grp = first / L.gsz;
    if (g >= L.n_items[2 * grp + SOLVER]) return;
    const int item = L.items[(size_t)(SOLVER * L.n_pairs + first) * LP_BMAX + g];
    const int p = first + item / LP_BMAX, b = item % LP_BMAX;
    PairWork &W = L.wk[p];
    const PairConst &pc = L.pc[p];
    constexpr int MM = SOLVER == 0 ? kMaxModelsAffine[KIND] : (KIND == TWO ? 8 : kMaxModelsPoints[KIND]); // TWO: + spare slots
    Model tmp[MM];
    const unsigned mflags = minimal_flags(prm.cfg, prm.opt);
    Model *out = L.models + ((size_t)p * LP_BMAX + b) * LP_MM;
    if (SOLVER == 0) {
        const int c = minimal_affine<KIND>(pc, L.M + pc.off, W.idx[b], mflags, tmp);
        W.count[b] = c;
        HypF *hy = L.hyps + ((size_t)p * LP_BMAX + b) * LP_MM;
        for (int j = 0; j < c; j++) { out[j] = tmp[j]; pack_hyp(KIND, tmp[j], hy + j); }
    } else {
        // point-only solver: candidates only; their tail (triangulation, depth fits) runs in minimal_tail_kernel with one
        // thread per candidate, so that it does not serialise over lanes holding different numbers of candidates
        // (the 6-point solver goes through the noinline entry: inlined into this kernel under the 80-register cap it lost
        // candidates on the GPU -- test_minimal_modes_whole_path[1] -- while the same source agrees with the oracle in
        // the stand-alone kernel and in the host build; not root-caused, the thread form is the non-default mode)
        const int c = KIND == SHARED ? minimal_solver_ni(pc, L.M + pc.off, 1, W.idx[b], mflags, tmp, false)
                                     : minimal_points<KIND>(pc, L.M + pc.off, W.idx[b], mflags, tmp, false, KIND == TWO);
        W.count[b] = c;
        for (int j = 0; j < c; j++) out[j] = tmp[j];
        if (KIND == TWO)
            for (int j = 0; j < c; j++)
                for (int k = 0; k < 9; k++) out[j + 4].R[k] = tmp[j + 4].R[k]; // second rotation of E's factorisation
        if (c > 0) {
            const int base = atomicAdd(L.n_cand + grp, c);
            int *cl = L.cand + (size_t)first * LP_BMAX * LP_MM;
            for (int j = 0; j < c; j++) cl[base + j] = item * LP_MM + j;
        }
    }
}

// cv::recoverPose (...two_focal_estimator.cpp:143) for the candidates of the 7-point solver, ONE WARP per candidate:
// lane = (pose hypothesis c, point i), 28 of 32 lanes triangulate one point each (one thread per sample walked the 84
// triangulations of its <= 3 candidates in sequence: 32 % of the GPU time of configs[2]).  The candidate arrives with the
// factors of E (minimal_points, defer_pose); the votes are one ballot.
__global__ void __launch_bounds__(128) recover_pose_warp_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const int grp = first / L.gsz;
    const int nc = L.n_cand[grp];
    for (int g = warp; g < nc; g += nwarps) {
        const int e = L.cand[(size_t)first * LP_BMAX * LP_MM + g];
        const int item = e / LP_MM, j = e % LP_MM;
        const int p = first + item / LP_BMAX, b = item % LP_BMAX;
        const PairConst &pc = L.pc[p];
        const PairWork &W = L.wk[p];
        Model *slot = L.models + ((size_t)p * LP_BMAX + b) * LP_MM + j;
        double Ra[9], Rb[9], tt[3]; // the factors of E, left by minimal_batch_kernel<TWO, 1> (Rb in the spare slot j + 4)
        for (int k = 0; k < 9; k++) { Ra[k] = slot->R[k]; Rb[k] = slot[4].R[k]; }
        for (int k = 0; k < 3; k++) tt[k] = slot->t[k];
        bool ok = false;
        if (lane < 28) {
            const Match64 m = L.M[pc.off + W.idx[b][lane % 7]];
            const double p0[2] = {m.x0x, m.x0y}, p1[2] = {m.x1x, m.x1y}; // un-calibrated points with K = I, as the reference does
            ok = recover_pose_point_ok(Ra, Rb, tt, lane / 7, p0, p1, 1e9);
        }
        const unsigned votes = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) { // (every lane read the factors before the ballot)
            int good[4];
            for (int c = 0; c < 4; c++) good[c] = __popc((votes >> (7 * c)) & 0x7fu);
            recover_pose_select(Ra, Rb, tt, good, slot->R, slot->t);
        }
        __syncwarp();
    }
}

// tail of the point-only solvers, one thread per candidate of the round (dense: the list was filled with atomics)
__global__ void __launch_bounds__(64, MP_MIN_CTAS) minimal_tail_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const LoopParams &prm = *prmp;
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int grp = first / L.gsz;
    if (g >= L.n_cand[grp]) return;
    const int e = L.cand[(size_t)first * LP_BMAX * LP_MM + g];
    const int item = e / LP_MM, j = e % LP_MM;
    const int p = first + item / LP_BMAX, b = item % LP_BMAX;
    const PairConst &pc = L.pc[p];
    const PairWork &W = L.wk[p];
    Model *slot = L.models + ((size_t)p * LP_BMAX + b) * LP_MM + j;
    Model m = *slot;
    const bool ok = minimal_tail_ni(pc, L.M + pc.off, W.idx[b], minimal_flags(prm.cfg, prm.opt), &m);
    if (ok) *slot = m;
    L.cand_ok[((size_t)p * LP_BMAX + b) * LP_MM + j] = ok ? 1 : 0;
}

// compaction of the surviving candidates of every point-solver sample, in order, + fp32 packs
__global__ void __launch_bounds__(64) minimal_pack_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int grp = first / L.gsz;
    const int n1 = L.n_items[2 * grp + 1], n2 = L.n_items_w[grp]; // both point-solver lists
    if (g >= n1 + n2) return;
    const int item = g < n1 ? L.items[(size_t)(L.n_pairs + first) * LP_BMAX + g] : L.items[(size_t)(2 * L.n_pairs + first) * LP_BMAX + (g - n1)];
    const int p = first + item / LP_BMAX, b = item % LP_BMAX;
    const PairConst &pc = L.pc[p];
    PairWork &W = L.wk[p];
    Model *out = L.models + ((size_t)p * LP_BMAX + b) * LP_MM;
    HypF *hy = L.hyps + ((size_t)p * LP_BMAX + b) * LP_MM;
    const unsigned char *ok = L.cand_ok + ((size_t)p * LP_BMAX + b) * LP_MM;
    const int c = W.count[b];
    int k = 0;
    for (int j = 0; j < c; j++) {
        if (!ok[j]) continue;
        const Model m = out[j];
        if (k != j) out[k] = m;
        pack_hyp(pc.kind, m, hy + k);
        k++;
    }
    W.count[b] = k;
}

// items_bound: upper bound of the samples of this round (pairs of the group that are not done x batch cap): the item
// and candidate lists are dense, so the grids only have to cover them (45000 empty CTAs cost 0.13 ms in a round that
// has one pair left)
cudaError_t launch_minimal_batch(LoopBuffers L, const LoopParams *dprm, int first, int count, long long items_bound, int kind,
                                 int sm_count, int warp_solvers, cudaStream_t s) {
    const long long ib = items_bound > 0 ? items_bound : 1;
    const int warp6 = (kind == SHARED && warp_solvers) ? 1 : 0; // lists routed to it are empty otherwise
    const unsigned grid1 = (unsigned)((ib + 63) / 64);
    if (kind == CALIB) {
        minimal_batch_kernel<CALIB, 0><<<grid1, 64, 0, s>>>(L, dprm, first, count);
        minimal_batch_kernel<CALIB, 1><<<grid1, 64, 0, s>>>(L, dprm, first, count);
    } else if (kind == SHARED) {
        minimal_batch_kernel<SHARED, 0><<<grid1, 64, 0, s>>>(L, dprm, first, count);
        minimal_batch_kernel<SHARED, 1><<<grid1, 64, 0, s>>>(L, dprm, first, count); // list 1 is empty when every sample takes the warp form
    } else {
        minimal_batch_kernel<TWO, 0><<<grid1, 64, 0, s>>>(L, dprm, first, count);
        minimal_batch_kernel<TWO, 1><<<grid1, 64, 0, s>>>(L, dprm, first, count);
    }
    if (warp6) { // grid-stride over the list: enough CTAs to fill the machine, no more than the list can use
        const long long want = (ib + W6B_WARPS - 1) / W6B_WARPS;
        const unsigned grid = (unsigned)std::max<long long>(1, std::min<long long>(want, (long long)sm_count * W6B_CTAS));
        // (function attributes are per device: set on every launch, the call is cheap)
        const cudaError_t attr = cudaFuncSetAttribute(minimal6_warp_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                      (int)(W6B_WARPS * sizeof(Warp6Work)));
        if (attr != cudaSuccess) return attr;
        minimal6_warp_batch_kernel<<<grid, W6B_WARPS * 32, W6B_WARPS * sizeof(Warp6Work), s>>>(L, dprm, first, count);
    }
    if (kind == TWO) recover_pose_warp_kernel<<<(unsigned)std::max<long long>(1, std::min<long long>((ib * 3 + 3) / 4, 8ll * sm_count)), 128, 0, s>>>(L, dprm, first, count);
    // worst case of the candidate list: every sample of the round is a point-solver sample with LP_MM candidates; the
    // threads beyond the actual count exit at once
    minimal_tail_kernel<<<(unsigned)((ib * LP_MM + 63) / 64), 64, 0, s>>>(L, dprm, first, count);
    minimal_pack_kernel<<<(unsigned)((ib + 63) / 64), 64, 0, s>>>(L, dprm, first, count);
    return cudaGetLastError();
}

} // namespace mp

extern "C" {
int mp_minimal_solve_batch(mp_ctx *c, int kind, int solver_idx, const mp_pair *pair, const mp_options *opt,
                           const mp_config *cfg, int n_samples, const int32_t *sample_idx, mp_model *out_models,
                           int32_t *out_counts) {
    using namespace mp;
    if (!c || !pair || !opt || !cfg || !sample_idx || !out_models || !out_counts || n_samples <= 0 || solver_idx < 0 ||
        solver_idx > 1)
        return fail(MP_ERR_ARG, "mp_minimal_solve_batch: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    int rc = upload_pairs(c, kind, 1, pair, opt);
    if (rc) return rc;
    const int kk = kind == MP_KIND_SCALE_ONLY ? CALIB : kind;
    const int k = solver_idx == 0 ? (kk == CALIB ? 3 : 4) : (kk == CALIB ? 5 : (kk == SHARED ? 6 : 7));
    const int mm = solver_idx == 0 ? kMaxModelsAffine[kk] : kMaxModelsPoints[kk];
    for (size_t i = 0; i < (size_t)n_samples * k; i++)
        if (sample_idx[i] < 0 || sample_idx[i] >= pair->n) return fail(MP_ERR_ARG, "sample index out of range");
    MP_CUDA(c->d_idx.reserve((size_t)n_samples * k * sizeof(int)));
    MP_CUDA(c->d_models.reserve((size_t)n_samples * mm * sizeof(Model)));
    MP_CUDA(c->d_misc.reserve((size_t)n_samples * sizeof(int)));
    MP_CUDA(cudaMemcpyAsync(c->d_idx.p, sample_idx, (size_t)n_samples * k * sizeof(int), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    if (kk == SHARED && solver_idx == 1 && c->minimal_mode != 1 && true) {
        // one warp per sample: the path small batches take (mp_set_minimal_mode)
        MP_CUDA(c->d_work.reserve((size_t)n_samples * kMaxModelsAny * sizeof(Model)));
        const unsigned grid = (unsigned)std::max(1, std::min((n_samples + W6_WARPS - 1) / W6_WARPS, c->sm_count * 8));
        minimal6_warp_kernel<<<grid, W6_WARPS * 32, 0, s>>>(c->d_pc.as<PairConst>(), c->d_m64.as<Match64>(), n_samples,
                                                             c->d_idx.as<int>(), minimal_flags(*cfg, *opt), mm, c->d_models.as<Model>(),
                                                             c->d_misc.as<int>(), c->d_work.as<Model>());
    } else
    minimal_kernel<<<(n_samples + 63) / 64, 64, 0, s>>>(c->d_pc.as<PairConst>(), c->d_m64.as<Match64>(), solver_idx, k,
                                                        n_samples, c->d_idx.as<int>(), minimal_flags(*cfg, *opt), mm,
                                                        c->d_models.as<Model>(), c->d_misc.as<int>());
    MP_CUDA(cudaGetLastError());
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    MP_CUDA(cudaMemcpyAsync(out_models, c->d_models.p, (size_t)n_samples * mm * sizeof(Model), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(out_counts, c->d_misc.p, (size_t)n_samples * sizeof(int), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    float ms = 0;
    MP_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->last_kernel_ms = ms; c->last_total_ms = ms; c->last_launches = 2;
    return MP_OK;
}

int mp_solve_affine(mp_ctx *c, int kind, int n_samples, const double *x_homo, const double *y_homo, const double *depth_x,
                    const double *depth_y, double *out_sols, int32_t *out_nsols, mp_model *out_models, int32_t *out_nmodels) {
    using namespace mp;
    if (!c || !x_homo || !y_homo || !depth_x || !depth_y || n_samples <= 0 || kind < 0 || kind > 2)
        return fail(MP_ERR_ARG, "mp_solve_affine: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    const int np_ = kind == CALIB ? 3 : 4;
    const size_t nx = (size_t)n_samples * 3 * np_, nd = (size_t)n_samples * np_;
    MP_CUDA(c->d_work.reserve((2 * nx + 2 * nd + (size_t)n_samples * 48) * sizeof(double)));
    MP_CUDA(c->d_models.reserve((size_t)n_samples * 8 * sizeof(Model)));
    MP_CUDA(c->d_misc.reserve(2 * (size_t)n_samples * sizeof(int)));
    double *dx = c->d_work.as<double>(), *dy = dx + nx, *dd0 = dy + nx, *dd1 = dd0 + nd, *dsol = dd1 + nd;
    int *dn = c->d_misc.as<int>();
    MP_CUDA(cudaMemcpyAsync(dx, x_homo, nx * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dy, y_homo, nx * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dd0, depth_x, nd * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dd1, depth_y, nd * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    affine_solve_kernel<<<(n_samples + 63) / 64, 64, 0, s>>>(kind, n_samples, dx, dy, dd0, dd1, dsol, dn, c->d_models.as<Model>(), dn + n_samples);
    MP_CUDA(cudaGetLastError());
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    if (out_sols) MP_CUDA(cudaMemcpyAsync(out_sols, dsol, (size_t)n_samples * 48 * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (out_nsols) MP_CUDA(cudaMemcpyAsync(out_nsols, dn, n_samples * sizeof(int), cudaMemcpyDeviceToHost, s));
    if (out_models) MP_CUDA(cudaMemcpyAsync(out_models, c->d_models.p, (size_t)n_samples * 8 * sizeof(Model), cudaMemcpyDeviceToHost, s));
    if (out_nmodels) MP_CUDA(cudaMemcpyAsync(out_nmodels, dn + n_samples, n_samples * sizeof(int), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    float ms = 0;
    MP_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->last_kernel_ms = ms; c->last_total_ms = ms; c->last_launches = 1;
    return MP_OK;
}

int mp_estimate_scale_and_pose(mp_ctx *c, int n, const double *X, const double *Y, const double *W, mp_model *out) {
    using namespace mp;
    if (!c || !X || !Y || !W || !out || n <= 0) return fail(MP_ERR_ARG, "mp_estimate_scale_and_pose: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    MP_CUDA(c->d_work.reserve((size_t)n * 7 * sizeof(double)));
    MP_CUDA(c->d_models.reserve(sizeof(Model)));
    MP_CUDA(c->d_misc.reserve(sizeof(int)));
    double *dX = c->d_work.as<double>(), *dY = dX + 3 * (size_t)n, *dW = dY + 3 * (size_t)n;
    MP_CUDA(cudaMemcpyAsync(dX, X, 3 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dY, Y, 3 * (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dW, W, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, s));
    umeyama_kernel<<<1, 32, 0, s>>>(n, dX, dY, dW, c->d_models.as<Model>(), c->d_misc.as<int>());
    MP_CUDA(cudaGetLastError());
    int ok = 0;
    MP_CUDA(cudaMemcpyAsync(out, c->d_models.p, sizeof(Model), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(&ok, c->d_misc.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    c->last_launches = 1;
    if (!ok) return fail(MP_ERR_ARG, "mp_estimate_scale_and_pose: degenerate point set (rank of the covariance below 2)");
    return MP_OK;
}
}
