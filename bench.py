#!/usr/bin/env python
"""bench.py -- MADPose hybrid LO-RANSAC hot path on B200 (contract: task statement; design: DESIGN.md).

A step = one pass of the whole path (sampling, minimal solvers, scoring, local optimisation, termination) over
one batch of synthetic ScanNet-1500-shaped pairs: BASELINE.json configs[1] = shared-focal estimator, 1500 pairs,
640x480, ~2k matches each (SURVEY.md 8d recipe).  metric = image pairs/s.
  value : the batch is resident in HBM when the timed region starts (mp_upload_batch, then mp_estimate_resident)
  e2e   : mp_estimate_batch with HOST buffers: pack + H2D + front-end + loop + D2H inside the timed region
  --impl reference : the CPU path (reference's own control-loop headers around the ported solver math,
                     oracle/_ref; else the oracle port) on all host threads, bounded sample per step
N > 1 (torchrun): pairs shard over the ranks, no data-path collective, weak scaling (1500 pairs per GPU).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # before CUDA initialises (see mp_create)
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIG = 1                # index into BASELINE.json "configs"
PAIRS_PER_GPU = 1500
F_PER_EVAL = {0: 161.0, 1: 99.0, 2: 99.0}  # algorithmic flop per (model x match) evaluation
F_PER_LM_BLOCK = 500.0    # residual + analytic Jacobian + normal-equation update of one block


class ClockSampler(threading.Thread):
    def __init__(self, idx=0):
        super().__init__(daemon=True)
        self.idx, self.samples, self.reasons, self.stop_flag, self.max_mhz = idx, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.idx)], capture_output=True, text=True, timeout=5).stdout.strip()
                f = [v.strip() for v in out.split(",")]
                self.samples.append(float(f[0])); self.max_mhz = float(f[1])
                for nme, v in zip(names, f[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nme)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def make_batch(n_pairs, first):
    """Synthetic pairs first .. first+n_pairs-1 of configs[1]; returns (kind, pairs, opt, cfg, keepalive)."""
    from tests.helpers import problem
    ps = [problem(CONFIG, p=first + i) for i in range(n_pairs)]
    return ps[0][0]["kind"], [q[1] for q in ps], ps[0][2], ps[0][3], ps


def cpu_arm(kind, pairs, opt, cfg, threads):
    """One pass of the CPU path over `pairs`; returns (seconds, stats list, label)."""
    from tests import oracle_lib
    ora = oracle_lib.load(build=False)
    ref_path = os.path.join(oracle_lib.ORACLE_DIR, "_ref", "liboracle_ref.so")
    lib, label = None, "oracle port (own restatement of loop + solver math)"
    if os.path.exists(ref_path):
        lib = C.CDLL(ref_path)
        label = "oracle/_ref: the reference's own HybridLOMSAC / RansacLib headers around the ported solver math"
    t0 = time.perf_counter()
    _, stats, _ = ora.estimate_batch(kind, pairs, opt, cfg, threads=threads, lib=lib)
    return time.perf_counter() - t0, stats, label


def workload_name():
    return ("configs[1]: shared-focal HybridEstimatePoseScaleOffsetSharedFocal, %d synthetic ScanNet-1500-shaped pairs "
            "per GPU (640x480, N~U{1500..2500} matches, 30-50%% outliers), options of examples/shared_focal.py"
            % PAIRS_PER_GPU)


def reference_main(a, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_sample = max(threads * 4, 16)
    kind, pairs, opt, cfg, keep = make_batch(n_sample, 0)
    times = []
    for it in range(a.warmup + a.steps):
        dt, stats, label = cpu_arm(kind, pairs, opt, cfg, threads)
        if it >= a.warmup:
            times.append(dt)
    dt = float(np.mean(times))
    v = n_sample / dt
    hyps = sum(s.hypotheses_scored for s in stats); evals = sum(s.residual_evals for s in stats)
    cb = {"value": v, "unit": "pairs/s", "cores": threads, "kind": "port",
          "sample": f"first {n_sample} pairs of the same workload per step, one pair per thread; {label}; "
                    "the reference's Eigen/Ceres/PoseLib/OpenCV build is not possible offline (DESIGN.md)"}
    print(json.dumps({"impl": "reference", "metric": "image_pairs_per_s", "value": v, "unit": "pairs/s",
                      "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt * 1e3,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                      "data": "synthetic", "config": {"workload": workload_name(), "pairs_per_step": n_sample},
                      "hypotheses_scored_per_s": hyps / dt, "residual_evals_per_s": evals / dt,
                      "cpu_baseline": cb,
                      "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--pairs", type=int, default=PAIRS_PER_GPU, help="pairs per GPU (default = the named workload)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        reference_main(a, rank)
        return

    import torch
    import torch.distributed as dist
    from madpose_b200 import capi
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    n_pairs = a.pairs
    kind, pairs, opt, cfg, keep = make_batch(n_pairs, rank * n_pairs)  # rank r owns pairs [r*P, (r+1)*P): no exchange
    ctx = capi.Context(local)
    peak32, peak64 = ctx.measure_alu_peaks()
    total_n = sum(p.n for p in pairs)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: batch resident in HBM (inputs 1500 x ~2000 x 80 B = 240 MB + 190 MB of hypothesis slots > 126 MB L2)
    ctx.upload_batch(kind, pairs, opt)
    for _ in range(a.warmup):
        ctx.estimate_resident(opt, cfg)
    sampler = ClockSampler(local); sampler.start()
    sync_all()
    dev_ms, launches, rounds = [], 0, 0
    t0 = time.perf_counter()
    for _ in range(a.steps):
        models, stats, _ = ctx.estimate_resident(opt, cfg)
        k, t, l = ctx.last_timing()
        dev_ms.append(t); launches += l
        rounds += ctx.last_breakdown()[1]
    sync_all()
    wall_res = time.perf_counter() - t0
    # ---- e2e: host buffers through mp_estimate_batch (pack + H2D + front-end + loop + D2H of models, stats, inlier lists)
    pair_arr = (capi.Pair * n_pairs)(*pairs)  # the C ABI's argument: an array of mp_pair records pointing at the host arrays
    for _ in range(min(a.warmup, 1)):
        ctx.estimate_batch(kind, pair_arr, opt, cfg, split=False)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        ctx.estimate_batch(kind, pair_arr, opt, cfg, split=False)
    sync_all()
    wall_e2e = time.perf_counter() - t0
    # ---- one extra, untimed-for-the-metric pass on ONE stream: exact per-class kernel durations (CUDA events between
    # launches); in the timed passes the groups' kernels overlap on 16 streams, so only their union is observable
    ctx.set_groups(1)
    ctx.estimate_resident(opt, cfg)
    brk, _ = ctx.last_breakdown()
    single_stream_ms = ctx.last_timing()[1]
    phase_cycles = ctx.last_phase_cycles()
    ctx.set_groups(0)
    sampler.stop_flag = True; sampler.join()

    step_ms = float(np.sum(dev_ms)) / a.steps  # CUDA events on the launching stream, whole resident call
    tt = torch.tensor([step_ms, wall_res / a.steps * 1e3, wall_e2e / a.steps * 1e3], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    step_ms, wall_res_ms, wall_e2e_ms = tt.tolist()
    hyps = float(sum(s.hypotheses_scored for s in stats)); evals = float(sum(s.residual_evals for s in stats))
    e64 = float(sum(s.evals_fp64 for s in stats)); lmb = float(sum(s.lm_block_evals for s in stats))
    its = float(sum(s.num_iterations_total for s in stats))
    if rank == 0:
        share = {k2: v["ms"] for k2, v in brk.items()}
        dom = max(share, key=share.get)
        score_ms_launch = brk["score"]["ms"] / max(brk["score"]["launches"], 1)
        decide_ms_launch = brk["decide"]["ms"] / max(brk["decide"]["launches"], 1)
        # algorithmic flops of one step per kernel class (per-unit figures: DESIGN.md section 4)
        fl_score = evals * F_PER_EVAL[kind]                         # fp32: every scored hypothesis x match
        fl_decide = e64 * F_PER_EVAL[kind] + lmb * F_PER_LM_BLOCK   # fp64: re-score + inlier passes + LM blocks
        if dom == "score":
            ach, peak, bound = fl_score / (share["score"] * 1e-3) / 1e12, peak32, "alu_fp32"
        else:
            ach, peak, bound = fl_decide / (share["decide"] * 1e-3) / 1e12, peak64, "alu_fp64"
        h2d = total_n * 48 + n_pairs * 256
        d2h = n_pairs * (136 + C.sizeof(capi.Stats)) + 3 * total_n * 4
        out = {
            "metric": "image_pairs_per_s", "value": world * n_pairs / (step_ms * 1e-3), "unit": "pairs/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64 (minimal solvers, LM, decisions) + f32 (scoring filter)",
            "data": "synthetic",
            "config": {"workload": workload_name(), "pairs_per_gpu": n_pairs, "matches_total_per_gpu": total_n,
                       "l2": "inputs larger than L2: 80 B/match x %d matches + hypothesis slots, per step" % total_n,
                       "seed": 0, "timing": "CUDA events on the library's stream around the whole resident call, max over ranks"},
            "hypotheses_scored_per_s": world * hyps / (step_ms * 1e-3),
            "residual_evals_per_s": world * evals / (step_ms * 1e-3),
            "ransac_iterations_per_s": world * its / (step_ms * 1e-3),
            "e2e": {"value": world * n_pairs / (wall_e2e_ms * 1e-3), "unit": "pairs/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "note": "mp_estimate_batch on pageable host arrays, wall clock"},
            "resident_wall_pairs_per_s": world * n_pairs / (wall_res_ms * 1e-3),
            "gpu_launches": int(launches),
            "rounds_per_step": rounds / a.steps,
            "kernel_ms_per_step_single_stream": share, "single_stream_ms_per_step": single_stream_ms,
            "decide_phase_gcycles": {k2: round(v / 1e9, 3) for k2, v in phase_cycles.items()},
            "roofline": {"bound": bound, "kernel": {"decide": "scan_kernel + lo_kernel (fp64 re-score, inliers, local optimisation / LM)"}.get(dom, dom), "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                         "frac": ach / peak,
                         # dram__bytes_read.sum + dram__bytes_write.sum of ONE lo_kernel launch (750 CTAs, 7.4 ms: 22.06 MB read + 1.02 MB written), from the last
                         # `ncu --set full` capture summarised in profiles/r1_whole_path_v3.md; the class is ALU / latency
                         # bound (3.5 GB/s), the figure only shows that nothing is re-read from HBM
                         "traffic": 23.08e6 if dom == "decide" else None,
                         "peak_source": "register-only FFMA / DFMA probes measured in this run (fp32 %.1f, fp64 %.1f TFLOP/s); "
                                        "MEASURED_PEAKS.json holds HBM and bf16 tensor peaks only" % (peak32, peak64),
                         "avg_launch_ms": {"score": score_ms_launch, "decide": decide_ms_launch},
                         "scoring_in_loop": {"achieved": fl_score / max(share["score"] * 1e-3, 1e-12) / 1e12, "peak": peak32,
                                             "frac": fl_score / max(share["score"] * 1e-3, 1e-12) / 1e12 / peak32}},
            "clocks": sampler.summary(),
        }
        threads = os.cpu_count() or 1
        ns = min(n_pairs, max(threads * 4, 16))
        dt, cst, label = cpu_arm(kind, pairs[:ns], opt, cfg, threads)
        out["cpu_baseline"] = {"value": ns / dt, "unit": "pairs/s", "cores": threads, "kind": "port",
                               "sample": f"first {ns} pairs of rank 0's batch, one pair per thread; {label}"}
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
