#!/usr/bin/env python
"""Stand-alone K3: fused fp32 scoring of H hypotheses x N matches of one pair through mp_score_batch.
Prints one JSON line: residual evaluations/s, algorithmic TFLOP/s (99 / 161 flop per evaluation) and
the fraction of the FFMA probe measured in the same process."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madpose_b200 import capi  # noqa: E402
from tests import oracle_lib  # noqa: E402
from tests.helpers import models_around_gt, problem  # noqa: E402

F_PER_EVAL = {0: 161.0, 1: 99.0, 2: 99.0}


def main():
    config = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    nh = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
    d, pair, opt, cfg, keep = problem(config, p=0)
    ora = oracle_lib.load(build=False)
    ns = ora.frontend(d["kind"], pair, opt, cfg)["norm_scale"]
    models = models_around_gt(d, ns, nh, seed=0)
    ctx = capi.Context(0)
    peak32, peak64 = ctx.measure_alu_peaks()
    ms = []
    for _ in range(8):
        ctx.score_batch(d["kind"], pair, opt, cfg, models, want32=True, want64=False)
        ms.append(ctx.last_timing()[0])
    t = float(np.median(ms[3:])) * 1e-3
    evals = nh * pair.n
    tf = evals * F_PER_EVAL[d["kind"]] / t / 1e12
    print(json.dumps({"kernel": "score_f32_kernel (+ pack_hyps)", "config": config, "hypotheses": nh, "matches": pair.n,
                      "ms": t * 1e3, "residual_evals_per_s": evals / t, "algorithmic_tflops": tf, "ffma_probe_tflops": peak32,
                      "frac_of_ffma_probe": tf / peak32}))


if __name__ == "__main__":
    main()
