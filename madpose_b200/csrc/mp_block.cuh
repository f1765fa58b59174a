// Block-level device helpers shared by the stand-alone kernels and the loop kernels.
#pragma once
#include "mp_frontend.cuh"
#include "mp_score.cuh"
#include "mp_stage.cuh"

namespace mp {

// deterministic block sum (fixed tree), result valid on all threads
__device__ __forceinline__ double block_sum(double v, double *red /*[32]*/) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (lane == 0) red[wid] = v;
    __syncthreads();
    double s = 0;
    for (int i = 0; i < nw; i++) s += red[i];
    return s;
}

__device__ __forceinline__ PairF make_pairf(const PairConst &pc, int score_type) {
    PairF f;
    f.k0fx = (float)pc.K0[0]; f.k0sk = (float)pc.K0[1]; f.k0fy = (float)pc.K0[4];
    f.k1fx = (float)pc.K1[0]; f.k1sk = (float)pc.K1[1]; f.k1fy = (float)pc.K1[4];
    f.thr0 = (float)pc.thr[0]; f.thr1 = (float)pc.thr[1]; f.thr2 = (float)pc.thr[2];
    f.loss = (float)pc.loss_scale; f.n = pc.n; f.gate = score_type;
    return f;
}

} // namespace mp
