// TEST INFRASTRUCTURE ONLY (never part of the product, never loaded by madpose_b200/).
// Compiles the MP_HD scalar math of madpose_b200/csrc/*.cuh for the HOST so that the
// per-sample solver / LM / RNG code can be checked against the oracle in the CPU-only
// container (the kernels that run this math on the GPU are tested by the -m gpu tests).
#include "../../madpose_b200/csrc/mp_frontend.cuh"
#include "../../madpose_b200/csrc/mp_minimal.cuh"
#include "../../madpose_b200/csrc/mp_lm.cuh"
#include "../../madpose_b200/csrc/mp_rng.cuh"
#include <random>
#include <cstring>
#include <vector>

using namespace mp;

namespace {
struct HostPair {
    PairConst pc;
    std::vector<Match64> M;
};
void build(int kind, const mp_pair *pr, const mp_options *opt, HostPair *hp) {
    PairRaw r;
    std::memset(&r, 0, sizeof(r));
    r.scale_only = kind == 3 ? 1 : 0;
    if (kind == 3) kind = CALIB;
    r.n = pr->n; r.kind = kind; r.off = 0; r.off32 = 0;
    std::memcpy(r.K0, pr->K0, 72); std::memcpy(r.K1, pr->K1, 72);
    for (int k = 0; k < 2; k++) {
        r.pp0[k] = pr->pp0[k]; r.pp1[k] = pr->pp1[k]; r.min_depth[k] = pr->min_depth[k];
        r.thr[k] = opt->squared_inlier_thresholds[k]; r.w[k] = opt->data_type_weights[k];
    }
    double ns = 1.0, K0i[9], K1i[9];
    if (kind != CALIB) {
        double acc = 0;
        for (int i = 0; i < pr->n; i++) acc += norm_scale_term(r, pr->x0 + 2 * i, pr->x1 + 2 * i);
        ns = acc / (std::sqrt(2.0) * pr->n);
    } else { inv3(r.K0, K0i); inv3(r.K1, K1i); }
    hp->M.resize(pr->n);
    for (int i = 0; i < pr->n; i++) {
        float p[4], q[4];
        prep_match(r, ns, K0i, K1i, pr->x0 + 2 * i, pr->x1 + 2 * i, pr->d0[i], pr->d1[i], &hp->M[i], p, q);
    }
    prep_constants(r, ns, &hp->pc);
}
} // namespace

extern "C" {

int hc_minimal_solve(int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg, int solver_idx,
                     int n_samples, const int *idx, int max_models, mp_model *out_models, int *out_counts) {
    HostPair hp;
    build(kind, pair, opt, &hp);
    if (kind == 3) kind = CALIB;
    const int k = solver_idx == 0 ? (kind == CALIB ? 3 : 4) : (kind == CALIB ? 5 : (kind == SHARED ? 6 : 7));
    Model tmp[kMaxModelsAny];
    for (int s = 0; s < n_samples; s++) {
        int c = minimal_solver(hp.pc, hp.M.data(), solver_idx, idx + (size_t)s * k, minimal_flags(*cfg, *opt), tmp);
        if (c > max_models) c = max_models;
        out_counts[s] = c;
        for (int j = 0; j < c; j++) std::memcpy(out_models + (size_t)s * max_models + j, &tmp[j], sizeof(mp_model));
    }
    return 0;
}

#include "mp_hostcheck_more.inc"

} // extern "C"
