// Host-side plumbing of the C ABI: context, growable device / pinned buffers, the
// upload + front-end launch shared by every entry point.
#pragma once
#include "mp_kernels.cuh"
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
    int last_launches = 0;
    uint32_t tape_seed = 0;
    size_t tape_len = 0;
};

namespace mp {

// Packs the caller's (pageable, scattered) arrays into pinned staging buffers, copies
// them to HBM and runs the front-end kernel.  After this PairConst / Match64 / the fp32
// staging blocks of all pairs are resident.
inline int upload_pairs(mp_ctx *c, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt) {
    if (!c || !pairs || !opt || n_pairs <= 0) return fail(MP_ERR_ARG, "upload_pairs: bad arguments");
    if (kind < 0 || kind > 2) return fail(MP_ERR_ARG, "kind must be 0, 1 or 2");
    c->off.assign(n_pairs + 1, 0);
    c->off32.assign(n_pairs + 1, 0);
    for (int p = 0; p < n_pairs; p++) {
        if (pairs[p].n < 0 || (pairs[p].n > 0 && (!pairs[p].x0 || !pairs[p].x1 || !pairs[p].d0 || !pairs[p].d1)))
            return fail(MP_ERR_ARG, "pair with null data pointer or negative size");
        c->off[p + 1] = c->off[p] + pairs[p].n;
        c->off32[p + 1] = c->off32[p] + 2ll * pairs[p].n + 8; // P then Q, 16-byte records, padded
    }
    long long N = c->off[n_pairs];
    c->total_n = N; c->total32 = c->off32[n_pairs];
    size_t nb = (size_t)(N > 0 ? N : 1) * sizeof(double);
    MP_CUDA(c->h_x0.reserve(2 * nb)); MP_CUDA(c->h_x1.reserve(2 * nb));
    MP_CUDA(c->h_d0.reserve(nb)); MP_CUDA(c->h_d1.reserve(nb));
    MP_CUDA(c->h_raw.reserve(n_pairs * sizeof(PairRaw)));
    MP_CUDA(c->d_x0.reserve(2 * nb)); MP_CUDA(c->d_x1.reserve(2 * nb));
    MP_CUDA(c->d_d0.reserve(nb)); MP_CUDA(c->d_d1.reserve(nb));
    MP_CUDA(c->d_raw.reserve(n_pairs * sizeof(PairRaw)));
    MP_CUDA(c->d_pc.reserve(n_pairs * sizeof(PairConst)));
    MP_CUDA(c->d_m64.reserve((size_t)(N > 0 ? N : 1) * sizeof(Match64)));
    MP_CUDA(c->d_s32.reserve((size_t)c->total32 * sizeof(float4)));
    PairRaw *raw = c->h_raw.as<PairRaw>();
    for (int p = 0; p < n_pairs; p++) {
        const mp_pair &q = pairs[p];
        size_t n = q.n;
        std::memcpy(c->h_x0.as<double>() + 2 * c->off[p], q.x0, 2 * n * sizeof(double));
        std::memcpy(c->h_x1.as<double>() + 2 * c->off[p], q.x1, 2 * n * sizeof(double));
        std::memcpy(c->h_d0.as<double>() + c->off[p], q.d0, n * sizeof(double));
        std::memcpy(c->h_d1.as<double>() + c->off[p], q.d1, n * sizeof(double));
        PairRaw &r = raw[p];
        r.n = q.n; r.kind = kind; r.off = c->off[p]; r.off32 = c->off32[p];
        std::memcpy(r.K0, q.K0, sizeof(r.K0)); std::memcpy(r.K1, q.K1, sizeof(r.K1));
        for (int k = 0; k < 2; k++) {
            r.pp0[k] = q.pp0[k]; r.pp1[k] = q.pp1[k]; r.min_depth[k] = q.min_depth[k];
            r.thr[k] = opt->squared_inlier_thresholds[k]; r.w[k] = opt->data_type_weights[k];
        }
    }
    cudaStream_t s = c->stream;
    MP_CUDA(cudaMemcpyAsync(c->d_x0.p, c->h_x0.p, 2 * N * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(c->d_x1.p, c->h_x1.p, 2 * N * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(c->d_d0.p, c->h_d0.p, N * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(c->d_d1.p, c->h_d1.p, N * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(c->d_raw.p, c->h_raw.p, n_pairs * sizeof(PairRaw), cudaMemcpyHostToDevice, s));
    prep_pairs_kernel<<<n_pairs, 256, 0, s>>>(c->d_raw.as<PairRaw>(), c->d_x0.as<double>(), c->d_x1.as<double>(),
                                              c->d_d0.as<double>(), c->d_d1.as<double>(), c->d_pc.as<PairConst>(),
                                              c->d_m64.as<Match64>(), c->d_s32.as<float4>());
    MP_CUDA(cudaGetLastError());
    c->last_launches = 1;
    return MP_OK;
}

} // namespace mp
