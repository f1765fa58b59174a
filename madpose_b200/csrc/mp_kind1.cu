// explicit instantiation of the per-kind loop kernels (kind 1); see mp_kind.cuh
#include "mp_kind.cuh"
namespace mp {
template cudaError_t launch_score_kind<1>(LoopBuffers, const LoopParams *, int, int, int, size_t, cudaStream_t);
template cudaError_t launch_scan_kind<1>(LoopBuffers, const LoopParams *, int, int, const long long *, const int *, int, cudaStream_t);
template cudaError_t launch_lo_kind<1>(LoopBuffers, const LoopParams *, int, int, const long long *, const int *, int, int, cudaStream_t, cudaStream_t, int);
template cudaError_t launch_lm_kind<1>(const PairConst *, const Match64 *, const LoopParams *, int, const Model *, const int *, const int *, Model *, double *, bool, cudaStream_t);
}
