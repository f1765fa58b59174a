// K1/K2 translation unit: the per-sample fp64 minimal solvers (compiled once, shared by the
// stand-alone entry point and the loop's batch kernel) and mp_minimal_solve_batch.
#include "mp_host.cuh"
#include "mp_loop.cuh"
#include "mp_minimal.cuh"

#ifndef MP_MIN_CTAS
#define MP_MIN_CTAS 12
#endif
namespace mp {

__device__ __noinline__ int minimal_solver_ni(const PairConst &pc, const Match64 *M, int solver_idx, const int *idx,
                                              bool min_depth_on, Model *out, bool with_tail = true) {
    return minimal_solver(pc, M, solver_idx, idx, min_depth_on, out, with_tail);
}
__device__ __noinline__ bool minimal_tail_ni(const PairConst &pc, const Match64 *M, const int *idx, bool min_depth_on, Model *m) {
    if (pc.kind == CALIB) return minimal_points_tail<CALIB>(pc, M, idx, min_depth_on, m);
    if (pc.kind == SHARED) return minimal_points_tail<SHARED>(pc, M, idx, min_depth_on, m);
    return minimal_points_tail<TWO>(pc, M, idx, min_depth_on, m);
}

// K1/K2 stand-alone: one thread per sample of one pair (mp_minimal_solve_batch)
__global__ void __launch_bounds__(64) minimal_kernel(const PairConst *pcp, const Match64 *M, int solver_idx, int k,
                                                     int n_samples, const int *idx, int min_depth_on, int max_models,
                                                     Model *out, int *counts) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_samples) return;
    const PairConst &pc = *pcp;
    Model tmp[kMaxModelsAny];
    int c = minimal_solver_ni(pc, M + pc.off, solver_idx, idx + (size_t)s * k, min_depth_on != 0, tmp);
    if (c > max_models) c = max_models;
    counts[s] = c;
    for (int j = 0; j < c; j++) out[(size_t)s * max_models + j] = tmp[j];
}

// ---- K1/K2 over all samples of the round: one thread per sample, taken from the two per-solver lists the
// sample kernel filled (solver 0 first, solver 1 from the next multiple of 32: a warp runs one solver only)
__global__ void __launch_bounds__(64, MP_MIN_CTAS) minimal_batch_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const LoopParams &prm = *prmp;
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int grp = first / L.gsz;
    const int n0 = L.n_items[2 * grp], n1 = L.n_items[2 * grp + 1], n0p = (n0 + 31) & ~31;
    int item;
    if (g < n0) item = L.items[(size_t)first * LP_BMAX + g];
    else if (g >= n0p && g < n0p + n1) item = L.items[(size_t)(L.n_pairs + first) * LP_BMAX + (g - n0p)];
    else return;
    const int p = first + item / LP_BMAX, b = item % LP_BMAX;
    PairWork &W = L.wk[p];
    const PairConst &pc = L.pc[p];
    Model tmp[LP_MM];
    const int solver = W.solver[b];
    // point-only solver: candidates only; their tail (triangulation, depth fits) runs in minimal_tail_kernel with one
    // thread per candidate, so that it does not serialise over lanes holding different numbers of candidates
    const int c = minimal_solver_ni(pc, L.M + pc.off, solver, W.idx[b], prm.cfg.min_depth_constraint != 0, tmp, solver == 0);
    W.count[b] = c;
    Model *out = L.models + ((size_t)p * LP_BMAX + b) * LP_MM;
    HypF *hy = L.hyps + ((size_t)p * LP_BMAX + b) * LP_MM;
    if (solver == 0) {
        for (int j = 0; j < c; j++) { out[j] = tmp[j]; pack_hyp(pc.kind, tmp[j], hy + j); }
    } else {
        for (int j = 0; j < c; j++) out[j] = tmp[j];
        if (c > 0) {
            const int base = atomicAdd(L.n_cand + grp, c);
            int *cl = L.cand + (size_t)first * LP_BMAX * LP_MM;
            for (int j = 0; j < c; j++) cl[base + j] = item * LP_MM + j;
        }
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
    const bool ok = minimal_tail_ni(pc, L.M + pc.off, W.idx[b], prm.cfg.min_depth_constraint != 0, &m);
    if (ok) *slot = m;
    L.cand_ok[((size_t)p * LP_BMAX + b) * LP_MM + j] = ok ? 1 : 0;
}

// compaction of the surviving candidates of every point-solver sample, in order, + fp32 packs
__global__ void __launch_bounds__(64) minimal_pack_kernel(LoopBuffers L, const LoopParams *prmp, int first, int count) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int grp = first / L.gsz;
    if (g >= L.n_items[2 * grp + 1]) return;
    const int item = L.items[(size_t)(L.n_pairs + first) * LP_BMAX + g];
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
cudaError_t launch_minimal_batch(LoopBuffers L, const LoopParams *dprm, int first, int count, long long items_bound, cudaStream_t s) {
    const long long ib = items_bound > 0 ? items_bound : 1;
    minimal_batch_kernel<<<(unsigned)((ib + 32 + 63) / 64), 64, 0, s>>>(L, dprm, first, count);
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
    const int k = solver_idx == 0 ? (kind == CALIB ? 3 : 4) : (kind == CALIB ? 5 : (kind == SHARED ? 6 : 7));
    const int mm = solver_idx == 0 ? kMaxModelsAffine[kind] : kMaxModelsPoints[kind];
    for (size_t i = 0; i < (size_t)n_samples * k; i++)
        if (sample_idx[i] < 0 || sample_idx[i] >= pair->n) return fail(MP_ERR_ARG, "sample index out of range");
    MP_CUDA(c->d_idx.reserve((size_t)n_samples * k * sizeof(int)));
    MP_CUDA(c->d_models.reserve((size_t)n_samples * mm * sizeof(Model)));
    MP_CUDA(c->d_misc.reserve((size_t)n_samples * sizeof(int)));
    MP_CUDA(cudaMemcpyAsync(c->d_idx.p, sample_idx, (size_t)n_samples * k * sizeof(int), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    minimal_kernel<<<(n_samples + 63) / 64, 64, 0, s>>>(c->d_pc.as<PairConst>(), c->d_m64.as<Match64>(), solver_idx, k,
                                                        n_samples, c->d_idx.as<int>(), cfg->min_depth_constraint, mm,
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
}
