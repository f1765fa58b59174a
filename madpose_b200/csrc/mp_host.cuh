// Host-side plumbing of the C ABI: context, growable device / pinned buffers, the
// upload + front-end launch shared by every entry point.
#pragma once
#include "mp_block.cuh"
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include <string>
#include <vector>

namespace mp {

extern thread_local std::string g_err;
inline int fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}
#define MP_CUDA(x)                                                                                        \
    do {                                                                                                  \
        cudaError_t e_ = (x);                                                                             \
        if (e_ != cudaSuccess)                                                                            \
            return mp::fail(MP_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_));                \
    } while (0)

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
    ~DevBuf() { if (p) cudaFree(p); }
};
struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
    ~PinBuf() { if (p) cudaFreeHost(p); }
};

} // namespace mp

struct mp_ctx {
    int device = 0, sm_count = 0, clock_khz = 0, smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    // uploaded batch
    mp::PinBuf h_x0, h_x1, h_d0, h_d1, h_raw, h_out, h_misc;
    mp::DevBuf d_x0, d_x1, d_d0, d_d1, d_raw, d_pc, d_m64, d_s32;
    // scratch
    mp::DevBuf d_models, d_hyps, d_scores32, d_scores64, d_idx, d_misc, d_out, d_tape, d_work, d_opt;
    std::vector<long long> off, off32;
    long long total_n = 0, total32 = 0;
    double last_kernel_ms = 0, last_total_ms = 0;
    int last_launches = 0, last_rounds = 0;
    int resident_kind = -1, resident_pairs = 0;
    double up_thr[2] = {0, 0}, up_w[2] = {0, 0}; // thresholds / weights the resident PairConst records were built with
    std::vector<cudaEvent_t> tick_ev;
    std::vector<int> tick_cls;
    int n_tick = 0;
    unsigned long long phase_clk[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    double class_ms[4] = {0, 0, 0, 0};
    int class_launches[4] = {0, 0, 0, 0};
    int groups_override = 0, last_groups = 1; // 0 = automatic
    int warp_from_it = -1;                    // experiments: override of MP_WARP_FROM_IT (-1 = default)
    int minimal_mode = 0;                     // mp_set_minimal_mode: 0 automatic, 1 one thread per sample, 2 one warp per sample
    int trace_cap = 0;                        // mp_set_trace: events kept per pair (0 = tracing off)
    int batch_override = 0;                   // speculated iterations per pair per round, 0 = LP_BMAX
    std::vector<cudaStream_t> gstream, sstream; // per-group round pipelines; side streams of the lo_kernel launches
    mp::PinBuf h_inl;      // pinned landing buffer of the inlier lists (copied to the caller's pageable buffer by host threads)
    std::vector<cudaEvent_t> gev, sev, hev;
    std::vector<cudaStream_t> hstream; // high-priority twins of the side streams (heavy pairs' local optimisations)
    cudaEvent_t ev_init = nullptr;
    void *loop_alloc = nullptr; // mp::LoopAlloc, created on first mp_estimate_batch
    uint32_t tape_seed = 0;
    size_t tape_len = 0;
};

namespace mp {
// Packs the caller's (pageable, scattered) arrays into pinned staging buffers, copies them to
// HBM and runs the front-end kernel (defined in mp_capi.cu).
int upload_pairs(mp_ctx *c, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt);
} // namespace mp
