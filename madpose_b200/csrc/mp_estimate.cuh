// Launchers implemented in the other translation units + the per-context loop buffers.
#pragma once
#include "mp_host.cuh"
#include "mp_loop.cuh"

namespace mp {

cudaError_t launch_minimal_batch(LoopBuffers L, const LoopParams *dprm, int first, int count, long long items_bound, int kind,
                                 int sm_count, int warp_solvers, cudaStream_t s);
// Batches of at most this many pairs run the point-only minimal solvers that have a warp-per-sample form (6-point shared
// focal) with ONE WARP per sample -- a round of a small batch holds too few samples to fill the machine with one thread
// each, and one thread needs ~1.1 ms for a 6-point solve against ~0.1 ms for a warp; larger batches keep one thread per
// sample (~1.4 x the throughput when every SM is full).  The choice depends on the size of the call only, never on the
// schedule, so results stay independent of timing (mp_set_minimal_mode overrides it).
constexpr int MP_WARP_SOLVER_MAX_PAIRS = 64;
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
