// Batch sharding over the GPUs of one node INSIDE the library: pairs are independent units, so a batch
// is dealt to the devices (by decreasing estimated cost = matches, snake order), each device runs its share through
// mp_estimate_batch on its own context from its own host thread, and the results are scattered back into the caller's
// buffers in the caller's order.  No inter-GPU traffic, no collective.  Contexts are created on first use and kept
// (one per (device, slot); a slot is an estimator kind for the mixed entry so that the kinds of a mixed batch run
// concurrently on every device).
#include <algorithm>
#include <map>
#include <mutex>
#include <thread>
#include <vector>
#include "mp_host.cuh"

namespace mp {
namespace {
struct PoolEntry { mp_ctx *ctx = nullptr; std::mutex mu; };
std::mutex g_pool_mu;
std::map<std::pair<int, int>, PoolEntry *> g_pool;

PoolEntry *pool_get(int device, int slot) {
    std::lock_guard<std::mutex> lock(g_pool_mu);
    auto key = std::make_pair(device, slot);
    auto it = g_pool.find(key);
    if (it != g_pool.end()) return it->second;
    PoolEntry *e = new PoolEntry();
    g_pool[key] = e;
    return e;
}

struct Share { std::vector<int> idx; std::vector<mp_pair> pairs; std::vector<mp_model> models; std::vector<mp_stats> stats;
               std::vector<int32_t> inl; int rc = MP_OK; std::string err; };

int run_sharded(int n_devices, const int *devices, int slot, int kind, const std::vector<int> &which, const mp_pair *pairs,
                const mp_options *opt, const mp_config *cfg, mp_model *out_models, mp_stats *out_stats, int32_t *out_inliers,
                const std::vector<long long> &inl_off) {
    if (which.empty()) return MP_OK;
    std::vector<int> order(which);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return pairs[a].n > pairs[b].n; });
    std::vector<Share> sh(n_devices);
    for (size_t k = 0; k < order.size(); k++) { // snake deal: 0 1 2 2 1 0 0 1 2 ...
        const int r = (int)(k % (2 * (size_t)n_devices));
        const int d = r < n_devices ? r : 2 * n_devices - 1 - r;
        sh[d].idx.push_back(order[k]);
    }
    std::vector<std::thread> th;
    for (int d = 0; d < n_devices; d++) {
        Share &S = sh[d];
        if (S.idx.empty()) continue;
        std::sort(S.idx.begin(), S.idx.end()); // the caller's order inside a share
        th.emplace_back([&, d]() {
            Share &S2 = sh[d];
            const int m = (int)S2.idx.size();
            S2.pairs.resize(m); S2.models.resize(m); S2.stats.resize(m);
            size_t tot = 0;
            for (int k = 0; k < m; k++) { S2.pairs[k] = pairs[S2.idx[k]]; tot += (size_t)pairs[S2.idx[k]].n; }
            S2.inl.resize(3 * tot + 1);
            PoolEntry *e = pool_get(devices[d], slot);
            std::lock_guard<std::mutex> lock(e->mu);
            if (!e->ctx) {
                S2.rc = mp_create(devices[d], &e->ctx);
                if (S2.rc != MP_OK) { S2.err = mp_last_error(); return; }
            }
            S2.rc = mp_estimate_batch(e->ctx, kind, m, S2.pairs.data(), opt, cfg, S2.models.data(), S2.stats.data(),
                                      out_inliers ? S2.inl.data() : nullptr);
            if (S2.rc != MP_OK) S2.err = mp_last_error();
        });
    }
    for (auto &t : th) t.join();
    for (int d = 0; d < n_devices; d++)
        if (sh[d].rc != MP_OK) return fail(sh[d].rc, "device " + std::to_string(devices[d]) + ": " + sh[d].err);
    for (int d = 0; d < n_devices; d++) {
        const Share &S = sh[d];
        size_t o = 0;
        for (size_t k = 0; k < S.idx.size(); k++) {
            const int p = S.idx[k];
            out_models[p] = S.models[k]; out_stats[p] = S.stats[k];
            const size_t len = 3 * (size_t)pairs[p].n;
            if (out_inliers) std::copy(S.inl.begin() + o, S.inl.begin() + o + len, out_inliers + inl_off[p]);
            o += len;
        }
    }
    return MP_OK;
}
} // namespace
} // namespace mp

extern "C" {

int mp_estimate_batch_multi(int n_devices, const int *devices, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt,
                            const mp_config *cfg, mp_model *out_models, mp_stats *out_stats, int32_t *out_inliers) {
    using namespace mp;
    if (n_devices <= 0 || !devices || !pairs || !opt || !cfg || !out_models || !out_stats || n_pairs <= 0)
        return fail(MP_ERR_ARG, "mp_estimate_batch_multi: bad arguments");
    std::vector<long long> off(n_pairs + 1, 0);
    std::vector<int> all(n_pairs);
    for (int p = 0; p < n_pairs; p++) { off[p + 1] = off[p] + 3ll * pairs[p].n; all[p] = p; }
    return run_sharded(n_devices, devices, 0, kind, all, pairs, opt, cfg, out_models, out_stats, out_inliers, off);
}

int mp_estimate_mixed_multi(int n_devices, const int *devices, int n_pairs, const int *kinds, const mp_pair *pairs,
                            const mp_options *opt_by_kind, const mp_config *cfg, mp_model *out_models, mp_stats *out_stats,
                            int32_t *out_inliers) {
    using namespace mp;
    if (n_devices <= 0 || !devices || !kinds || !pairs || !opt_by_kind || !cfg || !out_models || !out_stats || n_pairs <= 0)
        return fail(MP_ERR_ARG, "mp_estimate_mixed_multi: bad arguments");
    std::vector<long long> off(n_pairs + 1, 0);
    std::vector<int> by_kind[4];
    for (int p = 0; p < n_pairs; p++) {
        if (kinds[p] < 0 || kinds[p] > 3) return fail(MP_ERR_ARG, "mp_estimate_mixed_multi: kind out of range");
        off[p + 1] = off[p] + 3ll * pairs[p].n;
        by_kind[kinds[p]].push_back(p);
    }
    int rc[4] = {MP_OK, MP_OK, MP_OK, MP_OK};
    std::string err[4];
    std::vector<std::thread> th;
    for (int k = 0; k < 4; k++) {
        if (by_kind[k].empty()) continue;
        th.emplace_back([&, k]() { // the kinds run concurrently: their own contexts (slot 1 + kind) on every device
            rc[k] = run_sharded(n_devices, devices, 1 + k, k, by_kind[k], pairs, opt_by_kind + k, cfg, out_models, out_stats,
                                out_inliers, off);
            if (rc[k] != MP_OK) err[k] = mp_last_error();
        });
    }
    for (auto &t : th) t.join();
    for (int k = 0; k < 4; k++)
        if (rc[k] != MP_OK) return fail(rc[k], "kind " + std::to_string(k) + ": " + err[k]);
    return MP_OK;
}

int mp_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

} // extern "C"
