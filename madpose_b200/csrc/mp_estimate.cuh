// Launchers implemented in the other translation units + the per-context loop buffers.
#pragma once
#include "mp_host.cuh"
#include "mp_loop.cuh"

namespace mp {

cudaError_t launch_minimal_batch(LoopBuffers L, const LoopParams *dprm, int first, int count, long long items_bound, int kind,
                                 int sm_count, int warp_solvers, cudaStream_t s);
// The point-only minimal solvers that have a warp-per-sample form (6-point shared focal, mp_warp6.cuh) run with ONE WARP
// per sample: no thread-private arrays (the thread-per-sample form keeps 17 KB of local memory per thread) and ~0.1 ms
// instead of ~1.1 ms for one solve.  The loop's kernel runs 12 warps per CTA, one CTA per SM, with CTA barriers between
// the solver's stages so that all warps of an SM execute the same stage's code at the same time -- the solver is
// ~170 KB of instructions and free-running warps thrash the instruction cache.  Measured on configs[1]
// (tools/exp_warp_from.py): thread per sample 167.7 ms/step; warp per sample free-running 194.7; in lockstep 137.9
// (8 warps per CTA: 145.9, 16: 137.1).  mp_set_minimal_mode / mp_set_warp_from_iteration select the forms for
// experiments; for a given setting results never depend on the schedule.
constexpr int MP_WARP_SOLVER_MAX_PAIRS = 1 << 30; // (calls up to this size use the warp form for every iteration)
constexpr unsigned MP_WARP_FROM_IT = 0u;
template <int KIND>
cudaError_t launch_score_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, int tile_cap, size_t smem, cudaStream_t s);
template <int KIND>
cudaError_t launch_scan_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, const long long *off3,
                             const int *caps, int round, cudaStream_t s);
template <int KIND>
cudaError_t launch_lo_kind(LoopBuffers L, const LoopParams *dprm, int first, int count, const long long *off3,
                           const int *caps, int round, int widths, cudaStream_t s, cudaStream_t s_heavy, int heavy_from);
template <int KIND>
cudaError_t launch_lm_kind(const PairConst *pc, const Match64 *M, const LoopParams *dprm, int n_problems, const Model *start,
                           const int *idx, const int *off, Model *out, double *info, bool nonminimal, cudaStream_t s);

struct LoopAlloc {
    DevBuf trace, evals32, st, wk, rngA, rngB, models, hyps, inl_best, inl_cur, flat, sample, misc, off3, caps, prm, out_models, out_stats, lo_req, items, cand, cand_ok;
    PinBuf h_off3, h_caps, h_prm, h_active, h_out;
};

} // namespace mp
