// Launchers implemented in the other translation units + the per-context loop buffers.
#pragma once
#include "mp_host.cuh"
#include "mp_loop.cuh"

namespace mp {

cudaError_t launch_minimal_batch(LoopBuffers L, const LoopParams *dprm, int first, int count, long long items_bound, cudaStream_t s);
template <int KIND>
cudaError_t launch_score_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, int tile_cap, size_t smem, cudaStream_t s);
template <int KIND>
cudaError_t launch_scan_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, const long long *off3,
                             const int *caps, int round, cudaStream_t s);
template <int KIND>
cudaError_t launch_lo_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, const long long *off3,
                           const int *caps, int round, cudaStream_t s);
template <int KIND>
cudaError_t launch_lm_kind(const PairConst *pc, const Match64 *M, const LoopParams *dprm, int n_problems, const Model *start,
                           const int *idx, const int *off, Model *out, double *info, bool nonminimal, cudaStream_t s);

struct LoopAlloc {
    DevBuf trace, evals32, st, wk, rngA, rngB, models, hyps, inl_best, inl_cur, flat, sample, misc, off3, caps, prm, out_models, out_stats, lo_req, items, cand, cand_ok;
    PinBuf h_off3, h_caps, h_prm, h_active, h_out;
};

} // namespace mp
