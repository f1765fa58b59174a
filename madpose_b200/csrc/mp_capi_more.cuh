// Remaining C-ABI entry points (minimal solvers, LM, whole path).
#pragma once
#include "mp_host.cuh"

extern "C" {
int mp_minimal_solve_batch(mp_ctx *, int, int, const mp_pair *, const mp_options *, const mp_config *, int,
                           const int32_t *, mp_model *, int32_t *) {
    return mp::fail(MP_ERR_ARG, "mp_minimal_solve_batch: not implemented yet");
}
int mp_lm_batch(mp_ctx *, int, const mp_pair *, const mp_options *, const mp_config *, int, const mp_model *,
                const int32_t *, const int32_t *, mp_model *, double *) {
    return mp::fail(MP_ERR_ARG, "mp_lm_batch: not implemented yet");
}
int mp_estimate_batch(mp_ctx *, int, int, const mp_pair *, const mp_options *, const mp_config *, mp_model *,
                      mp_stats *, int32_t *) {
    return mp::fail(MP_ERR_ARG, "mp_estimate_batch: not implemented yet");
}
}
