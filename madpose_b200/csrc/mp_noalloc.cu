// mp_score_models_noalloc: fp64 ScoreModel of a list of models on ONE pair WITHOUT any allocation.  No counterpart in
// the reference; added for callers that cannot call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it).  It is the fp64 leg of mp_score_batch (front-end
// kernel + score_f64_kernel, mp_kernels.cuh) with the scratch of that path -- the pair's constants and its fp64 match
// records -- placed in ONE file-scope device array of fixed size instead of the context's cudaMalloc'ed buffers.
//
//  * inputs (raw matches, models) and the output (scores) are the CALLER's device buffers;
//  * the scratch is g_noalloc_scratch below: MP_NOALLOC_MAX_MATCHES records + a header, ~770 KB, held by every GPU
//    that loads the library until the process exits;
//  * every call on a device shares that array and overwrites it: calls that overlap on the GPU (different streams,
//    different host threads) RACE -- serialise them on one stream;
//  * no cudaMalloc / cudaFree / allocating library routine anywhere on this path: kernel launches, one
//    cudaMemsetAsync of the scratch header and cudaGetSymbolAddress / cudaDeviceGetAttribute at the first call per
//    device only.
#include <algorithm>
#include <mutex>
#include "mp_host.cuh"

namespace mp {

constexpr int NA_MAX_DEVICES = 64;
constexpr size_t NA_HEADER_BYTES = 1024; // PairConst (+ padding): cleared by every call before the front-end writes it
static_assert(sizeof(PairConst) <= NA_HEADER_BYTES, "scratch header too small for PairConst");
static_assert(NA_HEADER_BYTES % 256 == 0, "records start on a 256-byte boundary");
constexpr size_t NA_SCRATCH_BYTES = NA_HEADER_BYTES + (size_t)MP_NOALLOC_MAX_MATCHES * sizeof(Match64);

// The scratch.  In HOST code this name is only a handle for cudaGetSymbolAddress (a host-side shadow, not a device
// address): the kernels below receive the device address obtained once per device in noalloc_scratch_ptr().
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_SCRATCH_BYTES];

// front-end of ONE pair (= prep_pairs_kernel, mp_kernels.cuh, without the fp32 staging records): one CTA, the same
// summation tree for the normalisation scale, so the records are the bits mp_score_batch scores
__global__ void __launch_bounds__(256) noalloc_prep_kernel(PairRaw r, const double *x0, const double *x1, const double *d0,
                                                           const double *d1, unsigned char *scratch) {
    __shared__ double red[32];
    PairConst *pc = reinterpret_cast<PairConst *>(scratch);
    Match64 *m = reinterpret_cast<Match64 *>(scratch + NA_HEADER_BYTES);
    const int n = r.n;
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
        float p[4], q[4]; // the fp32 staging records are not kept here
        prep_match(r, ns, K0i, K1i, x0 + 2 * i, x1 + 2 * i, d0[i], d1[i], &m[i], p, q);
    }
    if (threadIdx.x == 0) prep_constants(r, ns, pc);
}

// = score_f64_kernel (mp_kernels.cuh) reading the pair from the static scratch: one CTA per model, fixed summation tree
__global__ void __launch_bounds__(256) noalloc_score_f64_kernel(const unsigned char *scratch, int score_type, int nm,
                                                                const Model *models, double *scores) {
    __shared__ double red[32];
    const PairConst &pc = *reinterpret_cast<const PairConst *>(scratch);
    const Match64 *m = reinterpret_cast<const Match64 *>(scratch + NA_HEADER_BYTES);
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

// device address of the scratch + SM count, looked up ONCE per device with that device current
struct NoAllocDevice { std::once_flag once; cudaError_t err = cudaSuccess; unsigned char *ptr = nullptr; int sm_count = 0; };
static NoAllocDevice g_noalloc_dev[NA_MAX_DEVICES];

static cudaError_t noalloc_scratch_ptr(int device, unsigned char **ptr, int *sm_count) {
    NoAllocDevice &d = g_noalloc_dev[device];
    std::call_once(d.once, [&] { // the caller made `device` current
        void *p = nullptr;
        d.err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (d.err == cudaSuccess) d.err = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device);
        if (d.err == cudaSuccess) d.ptr = static_cast<unsigned char *>(p);
    });
    *ptr = d.ptr; *sm_count = d.sm_count;
    return d.err;
}

} // namespace mp

extern "C" int mp_score_models_noalloc(int device, int kind, const mp_pair *pair_dev, const mp_options *opt, const mp_config *cfg,
                                       int n_models, const mp_model *d_models, double *d_scores64, void *stream) {
    using namespace mp;
    if (!pair_dev || !opt || !cfg || !d_models || !d_scores64 || n_models <= 0)
        return fail(MP_ERR_ARG, "mp_score_models_noalloc: bad arguments");
    if (device < 0 || device >= NA_MAX_DEVICES) return fail(MP_ERR_ARG, "mp_score_models_noalloc: device out of range");
    if (kind < 0 || kind > 3) return fail(MP_ERR_ARG, "kind must be 0 (calibrated), 1 (shared focal), 2 (two focal) or 3 (scale only)");
    if (pair_dev->n <= 0 || !pair_dev->x0 || !pair_dev->x1 || !pair_dev->d0 || !pair_dev->d1)
        return fail(MP_ERR_ARG, "mp_score_models_noalloc: pair with null data pointer or no matches");
    if (pair_dev->n > MP_NOALLOC_MAX_MATCHES)
        return fail(MP_ERR_ARG, "mp_score_models_noalloc: " + std::to_string(pair_dev->n) + " matches exceed the fixed scratch of MP_NOALLOC_MAX_MATCHES = " +
                                    std::to_string(MP_NOALLOC_MAX_MATCHES) + " (use mp_score_batch, which allocates)");
    MP_CUDA(cudaSetDevice(device));
    unsigned char *scratch = nullptr;
    int sm_count = 0;
    MP_CUDA(noalloc_scratch_ptr(device, &scratch, &sm_count));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int scale_only = kind == MP_KIND_SCALE_ONLY ? 1 : 0;
    PairRaw r; // as upload_pairs (mp_capi.cu) fills it
    std::memset(&r, 0, sizeof(r));
    r.n = pair_dev->n; r.kind = scale_only ? (int)CALIB : kind; r.off = 0; r.off32 = 0; r.scale_only = scale_only; r.pad_ = 0;
    std::memcpy(r.K0, pair_dev->K0, sizeof(r.K0)); std::memcpy(r.K1, pair_dev->K1, sizeof(r.K1));
    for (int k = 0; k < 2; k++) {
        r.pp0[k] = pair_dev->pp0[k]; r.pp1[k] = pair_dev->pp1[k]; r.min_depth[k] = scale_only ? 0.0 : pair_dev->min_depth[k];
        r.thr[k] = opt->squared_inlier_thresholds[k]; r.w[k] = opt->data_type_weights[k];
    }
    // the array keeps the previous call's contents: the header starts from zero, the records [0, n) are all rewritten
    MP_CUDA(cudaMemsetAsync(scratch, 0, NA_HEADER_BYTES, s));
    noalloc_prep_kernel<<<1, 256, 0, s>>>(r, pair_dev->x0, pair_dev->x1, pair_dev->d0, pair_dev->d1, scratch);
    MP_CUDA(cudaGetLastError());
    noalloc_score_f64_kernel<<<std::min(n_models, sm_count * 8), 256, 0, s>>>(scratch, cfg->score_type, n_models,
                                                                               reinterpret_cast<const Model *>(d_models), d_scores64);
    MP_CUDA(cudaGetLastError());
    return MP_OK; // asynchronous on `stream`: the caller synchronises
}
