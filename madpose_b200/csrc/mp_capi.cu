// C ABI of madpose_b200 (include/madpose_b200.h).  Every entry point runs CUDA
// kernels on the context's device; there is no CPU fallback.
#include <cstdlib>
#include <atomic>
#include <thread>
#include <vector>
#include <algorithm>
#include "mp_host.cuh"
#include "mp_kernels.cuh"
#include "mp_estimate.cuh"

namespace mp {
thread_local std::string g_err;
}
namespace mp {
// Packs the caller's (pageable, scattered) arrays into pinned staging buffers, copies
// them to HBM and runs the front-end kernel.  After this PairConst / Match64 / the fp32
// staging blocks of all pairs are resident.
int upload_pairs(mp_ctx *c, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt) {
    if (!c || !pairs || !opt || n_pairs <= 0) return fail(MP_ERR_ARG, "upload_pairs: bad arguments");
    if (kind < 0 || kind > 3) return fail(MP_ERR_ARG, "kind must be 0 (calibrated), 1 (shared focal), 2 (two focal) or 3 (scale only)");
    // whatever was resident is overwritten below: a failed or per-subsystem upload must not leave mp_estimate_resident
    // believing the previous batch is still there
    c->resident_pairs = 0; c->resident_kind = -1;
    const int scale_only = kind == MP_KIND_SCALE_ONLY ? 1 : 0;
    if (scale_only) kind = CALIB; // same kernels; PairConst::scale_only switches the solver 0 / guards / LM differences
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
    auto pack_range = [&](int p0, int p1) {
        for (int p = p0; p < p1; p++) {
            const mp_pair &q = pairs[p];
            size_t n = q.n;
            std::memcpy(c->h_x0.as<double>() + 2 * c->off[p], q.x0, 2 * n * sizeof(double));
            std::memcpy(c->h_x1.as<double>() + 2 * c->off[p], q.x1, 2 * n * sizeof(double));
            std::memcpy(c->h_d0.as<double>() + c->off[p], q.d0, n * sizeof(double));
            std::memcpy(c->h_d1.as<double>() + c->off[p], q.d1, n * sizeof(double));
            PairRaw &r = raw[p];
            r.n = q.n; r.kind = kind; r.off = c->off[p]; r.off32 = c->off32[p]; r.scale_only = scale_only; r.pad_ = 0;
            std::memcpy(r.K0, q.K0, sizeof(r.K0)); std::memcpy(r.K1, q.K1, sizeof(r.K1));
            for (int k = 0; k < 2; k++) {
                r.pp0[k] = q.pp0[k]; r.pp1[k] = q.pp1[k]; r.min_depth[k] = scale_only ? 0.0 : q.min_depth[k];
                r.thr[k] = opt->squared_inlier_thresholds[k]; r.w[k] = opt->data_type_weights[k];
            }
        }
    };
    // The caller's arrays are pageable and scattered: several threads pack them into the pinned staging buffers, and every
    // thread sends a slice to the device as soon as it is packed (~64 k matches = 3 MB at a time), so the copies over the
    // link run while the rest is still being packed.
    cudaStream_t s = c->stream;
    std::atomic<int> cuda_err{(int)cudaSuccess};
    auto send_range = [&](int pa, int pb) {
        const long long o = c->off[pa], m = c->off[pb] - o;
        if (m <= 0) return;
        cudaError_t e = cudaMemcpyAsync(c->d_x0.as<double>() + 2 * o, c->h_x0.as<double>() + 2 * o, 2 * m * sizeof(double), cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_x1.as<double>() + 2 * o, c->h_x1.as<double>() + 2 * o, 2 * m * sizeof(double), cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_d0.as<double>() + o, c->h_d0.as<double>() + o, m * sizeof(double), cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_d1.as<double>() + o, c->h_d1.as<double>() + o, m * sizeof(double), cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) cuda_err.store((int)e);
    };
    auto pack_and_send = [&](int p0, int p1, bool set_device) {
        if (set_device && cudaSetDevice(c->device) != cudaSuccess) { cuda_err.store((int)cudaErrorInvalidDevice); return; }
        int pa = p0;
        while (pa < p1) {
            int pb = pa;
            while (pb < p1 && c->off[pb] - c->off[pa] < 65536) pb++;
            pack_range(pa, pb);
            send_range(pa, pb);
            pa = pb;
        }
    };
    {
        static const int nt_cap = [] { const char *e = getenv("MP_PACK_THREADS"); const int v = e ? atoi(e) : 8; return v >= 1 && v <= 64 ? v : 8; }();
        int nt = (int)std::min<long long>(nt_cap, std::max<long long>(1, N / 200000));
        nt = std::min<int>(nt, std::max(1u, std::thread::hardware_concurrency()));
        if (nt <= 1) pack_and_send(0, n_pairs, false);
        else {
            std::vector<std::thread> th;
            int p0 = 0;
            for (int t = 0; t < nt; t++) { // equal shares of matches
                int p1 = p0;
                const long long want = N * (t + 1) / nt;
                while (p1 < n_pairs && c->off[p1 + 1] <= want) p1++;
                if (t == nt - 1) p1 = n_pairs;
                th.emplace_back(pack_and_send, p0, p1, true);
                p0 = p1;
            }
            for (auto &x : th) x.join();
        }
    }
    if (cuda_err.load() != (int)cudaSuccess)
        return fail(MP_ERR_CUDA, std::string("upload_pairs: host-to-device copy: ") + cudaGetErrorString((cudaError_t)cuda_err.load()));
    MP_CUDA(cudaMemcpyAsync(c->d_raw.p, c->h_raw.p, n_pairs * sizeof(PairRaw), cudaMemcpyHostToDevice, s));
    prep_pairs_kernel<<<n_pairs, 256, 0, s>>>(c->d_raw.as<PairRaw>(), c->d_x0.as<double>(), c->d_x1.as<double>(),
                                              c->d_d0.as<double>(), c->d_d1.as<double>(), c->d_pc.as<PairConst>(),
                                              c->d_m64.as<Match64>(), c->d_s32.as<float4>());
    MP_CUDA(cudaGetLastError());
    c->last_launches = 1;
    for (int k = 0; k < 2; k++) { c->up_thr[k] = opt->squared_inlier_thresholds[k]; c->up_w[k] = opt->data_type_weights[k]; }
    return MP_OK;
}

// madpose.utils.get_depths (madpose/utils.py:4-22) on the device: one thread per keypoint
template <class T>
__global__ void get_depths_kernel(const T *map, int map_h, int map_w, double rx, double ry, const double2 *kp, int n, double *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double2 p = kp[i];
    // numpy.round = round half to even = rint; astype(int) of an integral double; then numpy.clip
    double cx = rint(p.x * rx), cy = rint(p.y * ry);
    cx = fmin(fmax(cx, 0.0), (double)(map_w - 1));
    cy = fmin(fmax(cy, 0.0), (double)(map_h - 1));
    out[i] = (double)map[(size_t)(long long)cy * map_w + (long long)cx];
}

// madpose.utils.compute_pose_error (madpose/utils.py:46-78): one thread per pose
__global__ void pose_errors_kernel(int n, const double *Tgt, const double *R, const double *t, double t_thres, double *err_t,
                                   double *err_R) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *T = Tgt + 16 * (size_t)i, *Ri = R + 9 * (size_t)i, *ti = t + 3 * (size_t)i;
    const double RAD2DEG = 57.29577951308232;
    double tg[3] = {T[3], T[7], T[11]};
    double dot = 0, n1 = 0, n2 = 0, tr = 0;
    for (int k = 0; k < 3; ++k) {
        dot += ti[k] * tg[k]; n1 += ti[k] * ti[k]; n2 += tg[k] * tg[k];
        for (int j = 0; j < 3; ++j) tr += Ri[3 * k + j] * T[4 * k + j]; // trace(R^T R_gt)
    }
    double c = dot / (sqrt(n1) * sqrt(n2));
    c = c < -1.0 ? -1.0 : (c > 1.0 ? 1.0 : c); // numpy.clip keeps a NaN (zero-length translation)
    double et = RAD2DEG * acos(c);
    et = (et != et || 180.0 - et != 180.0 - et) ? et : fmin(et, 180.0 - et); // numpy.minimum propagates NaN
    double cr = (tr - 1.0) * 0.5;
    cr = cr < -1.0 ? -1.0 : (cr > 1.0 ? 1.0 : cr);
    if (t_thres >= 0.0 && sqrt(n2) < t_thres) et = 0.0;
    err_t[i] = et;
    err_R[i] = RAD2DEG * fabs(acos(cr));
}

} // namespace mp
using namespace mp;

extern "C" {

const char *mp_last_error(void) { return g_err.c_str(); }

int mp_create(int device, mp_ctx **out) {
    if (!out) return fail(MP_ERR_ARG, "mp_create: out is null");
    *out = nullptr;
    // The device loop keeps up to ~20 streams busy and the driver's default of 8 hardware queues makes them wait on each
    // other (measured: 646 -> 511 ms/step at 32 groups): CUDA_DEVICE_MAX_CONNECTIONS=32 is a CALLER-side setting that
    // must be in the environment before the process creates its CUDA context; the library does not
    // touch the environment.
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(MP_ERR_NO_DEVICE, "no CUDA device: madpose_b200 has no CPU path");
    if (device < 0 || device >= count) return fail(MP_ERR_ARG, "device index out of range");
    cudaDeviceProp prop;
    MP_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(MP_ERR_NO_DEVICE, "device is not sm_100 (B200) class");
    MP_CUDA(cudaSetDevice(device));
    mp_ctx *c = new mp_ctx();
    c->device = device; c->sm_count = prop.multiProcessorCount; c->smem_optin = (int)prop.sharedMemPerBlockOptin;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device);
    c->clock_khz = khz;
    MP_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (auto &e : c->ev) MP_CUDA(cudaEventCreate(&e));
    *out = c;
    return MP_OK;
}

void mp_destroy(mp_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (auto &e : c->ev) if (e) cudaEventDestroy(e);
    for (auto &e : c->tick_ev) cudaEventDestroy(e);
    for (auto &e : c->gev) cudaEventDestroy(e);
    for (auto &e : c->sev) cudaEventDestroy(e);
    for (auto &g : c->gstream) cudaStreamDestroy(g);
    for (auto &g : c->sstream) cudaStreamDestroy(g);
    for (auto &g : c->hstream) cudaStreamDestroy(g);
    for (auto &e : c->hev) cudaEventDestroy(e);
    if (c->ev_init) cudaEventDestroy(c->ev_init);
    if (c->loop_alloc) delete static_cast<mp::LoopAlloc *>(c->loop_alloc);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int mp_device_info(mp_ctx *c, int *sm_count, int *sm_clock_khz) {
    if (!c) return fail(MP_ERR_ARG, "null context");
    if (sm_count) *sm_count = c->sm_count;
    if (sm_clock_khz) *sm_clock_khz = c->clock_khz;
    return MP_OK;
}

int mp_measure_alu_peaks(mp_ctx *c, double *fp32_tflops, double *fp64_tflops) {
    if (!c) return fail(MP_ERR_ARG, "null context");
    MP_CUDA(cudaSetDevice(c->device));
    MP_CUDA(c->d_misc.reserve(64));
    const int blocks = c->sm_count * 8, threads = 256;
    float ms = 0;
    for (int pass = 0; pass < 2; pass++) { // pass 0 warms up
        const int it32 = 20000, it64 = 4000;
        MP_CUDA(cudaEventRecord(c->ev[0], c->stream));
        ffma_peak_kernel<<<blocks, threads, 0, c->stream>>>(c->d_misc.as<float>(), it32);
        MP_CUDA(cudaEventRecord(c->ev[1], c->stream));
        dfma_peak_kernel<<<blocks, threads, 0, c->stream>>>(c->d_misc.as<double>(), it64);
        MP_CUDA(cudaEventRecord(c->ev[2], c->stream));
        MP_CUDA(cudaStreamSynchronize(c->stream));
        MP_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
        if (fp32_tflops) *fp32_tflops = 2.0 * 8 * it32 * (double)blocks * threads / (ms * 1e-3) / 1e12;
        MP_CUDA(cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]));
        if (fp64_tflops) *fp64_tflops = 2.0 * 8 * it64 * (double)blocks * threads / (ms * 1e-3) / 1e12;
    }
    return MP_OK;
}

int mp_last_timing(mp_ctx *c, double *kernel_ms, double *total_ms, int *launches) {
    if (!c) return fail(MP_ERR_ARG, "null context");
    if (kernel_ms) *kernel_ms = c->last_kernel_ms;
    if (total_ms) *total_ms = c->last_total_ms;
    if (launches) *launches = c->last_launches;
    return MP_OK;
}

int mp_get_depths(mp_ctx *c, const void *depth_map, int elem_size, int map_h, int map_w, int img_h, int img_w,
                  const double *mkpts, int n, double *out) {
    if (!c || !depth_map || !mkpts || !out || n <= 0 || map_h <= 0 || map_w <= 0 || img_h <= 0 || img_w <= 0)
        return fail(MP_ERR_ARG, "mp_get_depths: bad arguments");
    if (elem_size != 4 && elem_size != 8) return fail(MP_ERR_ARG, "mp_get_depths: elem_size must be 4 (float) or 8 (double)");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    const size_t map_bytes = (size_t)map_h * map_w * elem_size;
    MP_CUDA(c->d_work.reserve(map_bytes));
    MP_CUDA(c->d_idx.reserve((size_t)n * 2 * sizeof(double)));
    MP_CUDA(c->d_scores64.reserve((size_t)n * sizeof(double)));
    MP_CUDA(cudaMemcpyAsync(c->d_work.p, depth_map, map_bytes, cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(c->d_idx.p, mkpts, (size_t)n * 2 * sizeof(double), cudaMemcpyHostToDevice, s));
    const double rx = (double)map_w / (double)img_w, ry = (double)map_h / (double)img_h; // utils.py:6-11
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    if (elem_size == 4)
        get_depths_kernel<float><<<(n + 255) / 256, 256, 0, s>>>(c->d_work.as<float>(), map_h, map_w, rx, ry, c->d_idx.as<double2>(), n,
                                                              c->d_scores64.as<double>());
    else
        get_depths_kernel<double><<<(n + 255) / 256, 256, 0, s>>>(c->d_work.as<double>(), map_h, map_w, rx, ry, c->d_idx.as<double2>(), n,
                                                               c->d_scores64.as<double>());
    MP_CUDA(cudaGetLastError());
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    MP_CUDA(cudaMemcpyAsync(out, c->d_scores64.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    float ms = 0;
    MP_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->last_kernel_ms = ms; c->last_total_ms = ms; c->last_launches = 1;
    return MP_OK;
}

int mp_pose_errors(mp_ctx *c, int n, const double *T_0to1, const double *R, const double *t, double t_thres, double *err_t,
                   double *err_R) {
    if (!c || n <= 0 || !T_0to1 || !R || !t || !err_t || !err_R) return fail(MP_ERR_ARG, "mp_pose_errors: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    MP_CUDA(c->d_work.reserve((size_t)n * 28 * sizeof(double)));
    MP_CUDA(c->d_scores64.reserve((size_t)n * 2 * sizeof(double)));
    double *dT = c->d_work.as<double>(), *dR = dT + 16 * (size_t)n, *dt = dR + 9 * (size_t)n;
    double *dout = c->d_scores64.as<double>();
    MP_CUDA(cudaMemcpyAsync(dT, T_0to1, (size_t)n * 16 * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dR, R, (size_t)n * 9 * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(dt, t, (size_t)n * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    pose_errors_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, dT, dR, dt, t_thres, dout, dout + n);
    MP_CUDA(cudaGetLastError());
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    MP_CUDA(cudaMemcpyAsync(err_t, dout, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(err_R, dout + n, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    float ms = 0;
    MP_CUDA(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
    c->last_kernel_ms = ms; c->last_total_ms = ms; c->last_launches = 1;
    return MP_OK;
}

int mp_score_batch(mp_ctx *c, int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, int nm,
                   const mp_model *models, float *s32, double *s64) {
    if (!c || !pair || !opt || !cfg || !models || nm <= 0) return fail(MP_ERR_ARG, "mp_score_batch: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    MP_CUDA(cudaEventRecord(c->ev[0], s));
    int rc = upload_pairs(c, kind, 1, pair, opt);
    if (rc) return rc;
    MP_CUDA(c->d_models.reserve(nm * sizeof(Model)));
    MP_CUDA(c->d_hyps.reserve(nm * sizeof(HypF)));
    MP_CUDA(c->d_scores32.reserve(nm * sizeof(float)));
    MP_CUDA(c->d_scores64.reserve(nm * sizeof(double)));
    MP_CUDA(cudaMemcpyAsync(c->d_models.p, models, nm * sizeof(Model), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaEventRecord(c->ev[1], s));
    int launches = 1;
    if (s32) {
        pack_hyps_kernel<<<(nm + 255) / 256, 256, 0, s>>>(kind == MP_KIND_SCALE_ONLY ? CALIB : kind, nm, c->d_models.as<Model>(), c->d_hyps.as<HypF>());
        int cap = (c->smem_optin - 2048) / 32;
        cap = std::min(cap, std::max(pair->n, 32));
        cap = (cap + 31) / 32 * 32;
        size_t smem = (size_t)cap * 32;
        int groups = (nm + MP_HT - 1) / MP_HT;
        int grid = std::max(1, std::min(c->sm_count, (groups + MP_WARPS - 1) / MP_WARPS));
        auto launch = [&](auto kern) -> cudaError_t {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
            kern<<<grid, MP_CTA, smem, s>>>(c->d_pc.as<PairConst>(), c->d_s32.as<float4>(), cfg->score_type, nm,
                                            c->d_hyps.as<HypF>(), c->d_scores32.as<float>(), cap);
            return cudaGetLastError();
        };
        if (kind == CALIB || kind == MP_KIND_SCALE_ONLY) MP_CUDA(launch(score_f32_kernel<CALIB>));
        else if (kind == SHARED) MP_CUDA(launch(score_f32_kernel<SHARED>));
        else MP_CUDA(launch(score_f32_kernel<TWO>));
        launches += 2;
    }
    MP_CUDA(cudaEventRecord(c->ev[2], s));
    if (s64) {
        score_f64_kernel<<<std::min(nm, c->sm_count * 8), 256, 0, s>>>(c->d_pc.as<PairConst>(), c->d_m64.as<Match64>(),
                                                                        cfg->score_type, nm, c->d_models.as<Model>(),
                                                                        c->d_scores64.as<double>());
        MP_CUDA(cudaGetLastError());
        launches += 1;
    }
    if (s32) MP_CUDA(cudaMemcpyAsync(s32, c->d_scores32.p, nm * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (s64) MP_CUDA(cudaMemcpyAsync(s64, c->d_scores64.p, nm * sizeof(double), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaEventRecord(c->ev[3], s));
    MP_CUDA(cudaStreamSynchronize(s));
    float k = 0, t = 0;
    MP_CUDA(cudaEventElapsedTime(&k, c->ev[1], c->ev[2])); // fp32 scoring (pack + fused kernel)
    MP_CUDA(cudaEventElapsedTime(&t, c->ev[0], c->ev[3]));
    c->last_kernel_ms = k; c->last_total_ms = t; c->last_launches = launches;
    return MP_OK;
}

int mp_inliers(mp_ctx *c, int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg,
               const mp_model *model, const double *thr3, int32_t *out_idx, int32_t *out_counts3) {
    if (!c || !pair || !opt || !cfg || !model || !thr3 || !out_idx || !out_counts3)
        return fail(MP_ERR_ARG, "mp_inliers: bad arguments");
    MP_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;
    int rc = upload_pairs(c, kind, 1, pair, opt);
    if (rc) return rc;
    const int n = pair->n;
    MP_CUDA(c->d_models.reserve(sizeof(Model)));
    MP_CUDA(c->d_misc.reserve(64));
    MP_CUDA(c->d_idx.reserve((3 * (size_t)n + 4) * sizeof(int)));
    MP_CUDA(cudaMemcpyAsync(c->d_models.p, model, sizeof(Model), cudaMemcpyHostToDevice, s));
    MP_CUDA(cudaMemcpyAsync(c->d_misc.p, thr3, 3 * sizeof(double), cudaMemcpyHostToDevice, s));
    inliers_kernel<<<1, MP_CTA, 0, s>>>(c->d_pc.as<PairConst>(), c->d_m64.as<Match64>(), c->d_models.as<Model>(),
                                        c->d_misc.as<double>(), c->d_idx.as<int>(), c->d_idx.as<int>() + 3 * n);
    MP_CUDA(cudaGetLastError());
    MP_CUDA(cudaMemcpyAsync(out_idx, c->d_idx.p, 3 * (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaMemcpyAsync(out_counts3, c->d_idx.as<int>() + 3 * n, 3 * sizeof(int), cudaMemcpyDeviceToHost, s));
    MP_CUDA(cudaStreamSynchronize(s));
    return MP_OK;
}

} // extern "C"
