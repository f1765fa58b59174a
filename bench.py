#!/usr/bin/env python
"""bench.py -- MADPose hybrid LO-RANSAC hot path on B200 (contract: task statement; design: DESIGN.md).

A step = one pass of the whole path (sampling, minimal solvers, scoring, local optimisation, termination) over one
batch of synthetic pairs of a BASELINE.json config (SURVEY.md 8d recipe, madpose_b200/workload.py).  metric = image pairs/s.
  --config 1 (default)  shared-focal estimator, 1500 ScanNet-1500-shaped pairs per GPU: the config the metric is quoted on
  --config 0 | 2 | 3    calibrated single pair | two-focal 1500 x 5k matches | calibrated dense 64 x 10k matches, 100k budget
  --config 4            15k mixed-estimator pairs, a FIXED batch split over the GPUs (strong scaling)
  value : the batch is resident in HBM when the timed region starts (mp_upload_batch, then mp_estimate_resident)
  e2e   : mp_estimate_batch with HOST buffers: pack + H2D + front-end + loop + D2H inside the timed region
  --impl reference : the CPU path (the reference's own control-loop headers around the ported solver math, oracle/_ref;
                     else the oracle port) on all host threads, a bounded sample of the same workload per step
The default line (config 1, N = 1) also carries short measurements of configs 0 / 2 / 3 under "configs", the stand-alone
fused-scoring roofline and the single-pair latency through the Python drop-in API (--no-extra skips them).
N > 1 (torchrun): pairs shard over the ranks, no data-path collective.
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

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # caller-side setting, before CUDA initialises
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

F_PER_EVAL = {0: 161.0, 1: 99.0, 2: 99.0, 3: 161.0}  # algorithmic flop per (model x match) evaluation
F_PER_LM_BLOCK = 500.0    # residual + analytic Jacobian + normal-equation update of one block
CPU_SAMPLE_PAIRS = {0: 1, 1: 256, 2: 96, 3: 16, 4: 192}  # pairs per step of the CPU arm (bounded sample of the workload)


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


# ------------------------------------------------------------------------------------------------ workload
def build_workload(config, n_pairs, rank, world):
    """The sub-batches this rank runs for one step: list of (kind, pairs, opt, cfg, keepalive).
    configs 0-3: rank r owns pairs [r*P, (r+1)*P) (weak scaling); config 4: the fixed 15k batch, pair i -> rank i % world."""
    from madpose_b200 import workload
    if config == 4:
        mixed = workload.make_mixed_batch(range(rank, n_pairs, world))
        return [(k, v[0], v[1], v[2], v[3]) for k, v in sorted(mixed.items())]
    return [workload.make_batch(config, n_pairs, first=rank * n_pairs)]


def workload_name(config, n_pairs):
    from madpose_b200 import workload
    base = workload.CONFIG_INFO[config]["name"]
    if n_pairs != workload.CONFIG_INFO[config]["pairs"]:
        base += f" [run with {n_pairs} pairs]"
    return base


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_arm(parts, threads):
    """One pass of the CPU path over the sub-batches; returns (seconds, stats list, label, kind)."""
    from oracle import oracle_lib
    ora = oracle_lib.load(build=False)
    ref_path = os.path.join(oracle_lib.ORACLE_DIR, "_ref", "liboracle_ref.so")
    lib, label = None, "oracle port (own restatement of loop + solver math)"
    if os.path.exists(ref_path):
        lib = C.CDLL(ref_path)
        label = "oracle/_ref: the reference's own HybridLOMSAC / RansacLib headers around the ported solver math"
    t0 = time.perf_counter()
    stats = []
    for kind, pairs, opt, cfg, _ in parts:
        _, st, _ = ora.estimate_batch(kind, pairs, opt, cfg, threads=threads, lib=lib)
        stats += st
    return time.perf_counter() - t0, stats, label


def cpu_sample(config, n_pairs):
    """The bounded sample both the reference arm and the cpu_baseline leg time: the first pairs of rank 0's workload."""
    ns = min(n_pairs, CPU_SAMPLE_PAIRS[config])
    return build_workload(config, ns, 0, 1), ns


def reference_main(a, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_pairs = a.pairs
    parts, ns = cpu_sample(a.config, n_pairs)
    times = []
    for it in range(a.warmup + a.steps):
        dt, stats, label = cpu_arm(parts, threads)
        if it >= a.warmup:
            times.append(dt)
    dt = float(np.mean(times))
    v = ns / dt
    hyps = sum(s.hypotheses_scored for s in stats); evals = sum(s.residual_evals for s in stats)
    sample = (f"first {ns} pairs of the same workload per step, one pair per thread on {threads} threads; {label}; "
              "the reference's Eigen/Ceres/PoseLib/OpenCV build is not possible offline (DESIGN.md section 5)")
    cb = {"value": v, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample}
    print(json.dumps({"impl": "reference", "metric": "image_pairs_per_s", "value": v, "unit": "pairs/s",
                      "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt * 1e3,
                      "higher_is_better": True, "scaling": "strong" if a.config == 4 else "weak", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic",
                      "config": {"workload": workload_name(a.config, n_pairs), "pairs_per_step": ns,
                                 "note": "bounded sample of the workload (rates are per pair)"},
                      "hypotheses_scored_per_s": hyps / dt, "residual_evals_per_s": evals / dt,
                      "cpu_baseline": cb,
                      "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------ GPU arm
class GpuRun:
    """One context per sub-batch (configs 0-3: one; config 4: one per estimator kind, run concurrently)."""

    def __init__(self, capi, device, parts):
        self.capi, self.parts = capi, parts
        self.ctxs = [capi.Context(device) for _ in parts]
        self.arrs = [(capi.Pair * len(p[1]))(*p[1]) for p in parts]
        self.n_pairs = sum(len(p[1]) for p in parts)
        self.total_n = sum(q.n for p in parts for q in p[1])

    def close(self):
        for c in self.ctxs:
            c.close()

    def _each(self, fn):
        if len(self.ctxs) == 1:
            return [fn(0)]
        out = [None] * len(self.ctxs)

        def work(i):
            out[i] = fn(i)
        th = [threading.Thread(target=work, args=(i,)) for i in range(len(self.ctxs))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        return out

    def upload(self):
        self._each(lambda i: self.ctxs[i].upload_batch(self.parts[i][0], self.parts[i][1], self.parts[i][2]))

    def resident_step(self):
        """Returns (device ms of the step, stats, launches, rounds): single sub-batch = CUDA events on the library's stream
        around the whole call; several concurrent sub-batches = the slowest of them (they start together)."""
        res = self._each(lambda i: (self.ctxs[i].estimate_resident(self.parts[i][2], self.parts[i][3], want_inliers=True)[1],
                                    self.ctxs[i].last_timing(), self.ctxs[i].last_breakdown()[1]))
        stats = [s for r in res for s in r[0]]
        return max(r[1][1] for r in res), stats, sum(r[1][2] for r in res), sum(r[2] for r in res)

    def e2e_step(self):
        self._each(lambda i: self.ctxs[i].estimate_batch(self.parts[i][0], self.arrs[i], self.parts[i][2], self.parts[i][3], split=False))


def measure_config(capi, torch, dist, device, config, n_pairs, rank, world, steps, warmup, breakdown=True):
    parts = build_workload(config, n_pairs, rank, world)
    run = GpuRun(capi, device, parts)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run.upload()
    for _ in range(warmup):
        run.resident_step()
    sync_all()
    dev_ms, launches, rounds = [], 0, 0
    t0 = time.perf_counter()
    for _ in range(steps):
        ms, stats, l, r = run.resident_step()
        dev_ms.append(ms); launches += l; rounds += r
    sync_all()
    wall_res = time.perf_counter() - t0
    for _ in range(min(warmup, 1)):
        run.e2e_step()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(steps):
        run.e2e_step()
    sync_all()
    wall_e2e = time.perf_counter() - t0
    out = dict(parts=parts, step_ms=float(np.sum(dev_ms)) / steps, wall_res_ms=wall_res / steps * 1e3,
               wall_e2e_ms=wall_e2e / steps * 1e3, stats=stats, launches=launches, rounds=rounds / steps,
               n_pairs=run.n_pairs, total_n=run.total_n)
    if breakdown and len(run.ctxs) == 1:
        # one extra, untimed-for-the-metric pass on ONE stream: exact per-class kernel durations (CUDA events between
        # launches); in the timed passes the groups' kernels overlap on many streams, so only their union is observable
        ctx = run.ctxs[0]
        ctx.set_groups(1)
        ctx.estimate_resident(parts[0][2], parts[0][3])
        out["brk"] = ctx.last_breakdown()[0]
        out["single_stream_ms"] = ctx.last_timing()[1]
        out["phase_cycles"] = ctx.last_phase_cycles()
        ctx.set_groups(0)
    run.close()
    return out


def scoring_standalone(capi, device, config, peak32, nh=65536):
    """K3 alone: nh hypotheses x the matches of pair 0 through mp_score_batch (pack + fused fp32 kernel, CUDA events)."""
    from madpose_b200 import workload
    d, pair, opt, cfg, keep = workload.problem(config, p=0)
    kind = d["kind"]
    ns = 1.0
    if kind != 0:  # poselib::normalize_points scale of the focal kinds (front-end), to put the models in internal units
        a = d["x0"] - d["pp0"]; b = d["x1"] - d["pp1"]
        ns = float((np.linalg.norm(a, axis=1).sum() + np.linalg.norm(b, axis=1).sum()) / (np.sqrt(2.0) * pair.n))
    models = workload.models_around_gt(d, ns, nh, seed=0)
    ctx = capi.Context(device)
    ms = []
    for _ in range(8):
        ctx.score_batch(kind, pair, opt, cfg, models, want32=True, want64=False)
        ms.append(ctx.last_timing()[0])
    ctx.close()
    t = float(np.median(ms[3:])) * 1e-3
    evals = nh * pair.n
    tf = evals * F_PER_EVAL[kind] / t / 1e12
    return {"kernel": "score_f32_kernel (+ pack_hyps_kernel)", "workload": f"{nh} hypotheses x {pair.n} matches, pair 0 of configs[{config}]",
            "ms": t * 1e3, "residual_evals_per_s": evals / t, "achieved": tf, "peak": peak32, "unit": "TFLOP/s", "frac": tf / peak32}


def single_pair_latency(device, config, n_calls=20):
    """One pair per call through the Python drop-in API (what the reference's API does), next to the CPU path's
    per-pair latency on one thread for the same pairs."""
    os.environ["MADPOSE_B200_DEVICE"] = str(device)
    import madpose_b200 as madpose
    from madpose_b200 import synth, workload
    from oracle import oracle_lib
    o = madpose.HybridLORansacOptions()
    for k, v in synth.options_for(config).items():
        setattr(o, k, list(v) if isinstance(v, tuple) else v)
    ds = [synth.make_pair(config, p) for p in range(4)]

    def call(d):
        if config == 0:
            return madpose.HybridEstimatePoseScaleOffset(d["x0"], d["x1"], d["d0"], d["d1"], d["min_depth"], d["K0"], d["K1"], o)
        return madpose.HybridEstimatePoseScaleOffsetSharedFocal(d["x0"], d["x1"], d["d0"], d["d1"], d["min_depth"], d["pp0"], d["pp1"], o)
    for d in ds[:2]:
        call(d)
    ms = []
    for k in range(n_calls):
        t0 = time.perf_counter(); call(ds[k % len(ds)]); ms.append((time.perf_counter() - t0) * 1e3)
    ora = oracle_lib.load(build=False)
    cms = []
    for p in range(len(ds)):
        d, pair, opt, cfg, keep = workload.problem(config, p=p)
        t0 = time.perf_counter(); ora.estimate_batch(d["kind"], [pair], opt, cfg, threads=1); cms.append((time.perf_counter() - t0) * 1e3)
    return {"api": "madpose_b200.HybridEstimatePoseScaleOffset" + ("" if config == 0 else "SharedFocal") + " (pybind11, one pair per call, host arrays)",
            "gpu_ms_median": float(np.median(ms)), "gpu_ms_min": float(np.min(ms)), "calls": n_calls,
            "cpu_one_thread_ms_median": float(np.median(cms)), "pairs": f"pairs 0-{len(ds) - 1} of configs[{config}]"}


def summarize(m, kind_for_flops, world, steps):
    """Whole-job numbers of one measured config (after the max over ranks has been applied to the times)."""
    stats = m["stats"]
    hyps = float(sum(s.hypotheses_scored for s in stats)); evals = float(sum(s.residual_evals for s in stats))
    its = float(sum(s.num_iterations_total for s in stats))
    n = m["n_pairs_global"]
    return {"value": n / (m["step_ms"] * 1e-3), "ms_per_step": m["step_ms"], "e2e": n / (m["wall_e2e_ms"] * 1e-3),
            "hypotheses_scored_per_s": world * hyps / (m["step_ms"] * 1e-3),
            "residual_evals_per_s": world * evals / (m["step_ms"] * 1e-3),
            "ransac_iterations_per_s": world * its / (m["step_ms"] * 1e-3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--config", type=int, default=1, choices=[0, 1, 2, 3, 4], help="index into BASELINE.json configs")
    ap.add_argument("--pairs", type=int, default=0, help="pairs per GPU (config 4: in total); default = the named workload")
    ap.add_argument("--no-extra", action="store_true", help="skip the short measurements of the other configs, the "
                    "stand-alone scoring roofline and the single-pair latency")
    a = ap.parse_args()
    from madpose_b200 import workload
    if a.pairs <= 0:
        a.pairs = workload.CONFIG_INFO[a.config]["pairs"]
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
    ctx0 = capi.Context(local)
    peak32, peak64 = ctx0.measure_alu_peaks()
    ctx0.close()

    sampler = ClockSampler(local); sampler.start()
    m = measure_config(capi, torch, dist, local, a.config, a.pairs, rank, world, a.steps, a.warmup)
    sampler.stop_flag = True; sampler.join()
    tt = torch.tensor([m["step_ms"], m["wall_res_ms"], m["wall_e2e_ms"]], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    m["step_ms"], m["wall_res_ms"], m["wall_e2e_ms"] = tt.tolist()
    strong = a.config == 4
    m["n_pairs_global"] = a.pairs if strong else world * m["n_pairs"]
    n_local_tt = torch.tensor([float(m["total_n"]), float(m["n_pairs"])], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(n_local_tt, op=dist.ReduceOp.SUM)

    if rank == 0:
        stats = m["stats"]
        kind0 = m["parts"][0][0]
        S = summarize(m, kind0, world, a.steps)
        fl = {k: 0.0 for k in ("score", "decide")}
        e32 = e64 = lmb = 0.0
        # per-kind flop figures (config 4 mixes kinds): walk the sub-batches in the order the stats were concatenated
        o = 0
        for kind, pairs, _, _, _ in m["parts"]:
            st = stats[o:o + len(pairs)]; o += len(pairs)
            k32 = float(sum(s.evals_fp32 for s in st)); k64 = float(sum(s.evals_fp64 for s in st)); kl = float(sum(s.lm_block_evals for s in st))
            e32 += k32; e64 += k64; lmb += kl
            fl["score"] += k32 * F_PER_EVAL[kind]
            fl["decide"] += k64 * F_PER_EVAL[kind] + kl * F_PER_LM_BLOCK
        total_n, n_pairs = m["total_n"], m["n_pairs"]
        h2d = total_n * 48 + n_pairs * 256
        d2h = n_pairs * (136 + C.sizeof(capi.Stats)) + 3 * total_n * 4
        out = {
            "metric": "image_pairs_per_s", "value": S["value"], "unit": "pairs/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": m["step_ms"], "higher_is_better": True,
            "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": "f64 (minimal solvers, LM, decisions) + f32 (scoring filter)", "data": "synthetic",
            "config": {"workload": workload_name(a.config, a.pairs), "pairs_per_gpu": n_pairs, "pairs_total": m["n_pairs_global"],
                       "matches_total_per_gpu": total_n,
                       "l2": "inputs larger than L2: 80 B/match x %d matches + hypothesis slots, per step" % total_n
                             if total_n * 80 > 126e6 else "inputs fit in L2 (single-pair / small config): latency, not bandwidth, is measured",
                       "seed": 0, "timing": "CUDA events on the library's stream around the whole resident call, max over ranks"},
            "hypotheses_scored_per_s": S["hypotheses_scored_per_s"], "residual_evals_per_s": S["residual_evals_per_s"],
            "ransac_iterations_per_s": S["ransac_iterations_per_s"],
            "e2e": {"value": S["e2e"], "unit": "pairs/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "note": "mp_estimate_batch on pageable host arrays, wall clock, max over ranks"},
            "resident_wall_pairs_per_s": m["n_pairs_global"] / (m["wall_res_ms"] * 1e-3),
            "gpu_launches": int(m["launches"]), "rounds_per_step": m["rounds"],
            "clocks": sampler.summary(),
        }
        if "brk" in m:
            brk = m["brk"]
            share = {k2: v["ms"] for k2, v in brk.items()}
            dom = max(share, key=share.get)
            if dom == "score":
                ach, peak, bound = fl["score"] / (share["score"] * 1e-3) / 1e12, peak32, "alu_fp32"
            elif dom == "decide":
                ach, peak, bound = fl["decide"] / (share["decide"] * 1e-3) / 1e12, peak64, "alu_fp64"
            else:  # sampling / minimal solvers: no per-unit flop figure is defined; report the class share only
                ach, peak, bound = None, peak64, "alu_fp64"
            traffic, traffic_src = None, None
            tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
            if os.path.exists(tpath):  # dram bytes per launch of the dominant kernel from the last `ncu --set full` capture
                tj = json.load(open(tpath))
                if tj.get("class") == dom:
                    traffic, traffic_src = tj.get("dram_bytes_per_launch"), "profiles/roofline_traffic.json: " + tj.get("source", "")
            names = {"decide": "scan_kernel + lo_kernel (fp64 re-score, inliers, local optimisation / LM)",
                     "minimal": "minimal solver kernels (fp64 K1/K2)", "score": "score_batch_kernel (fused fp32 scoring)",
                     "sample": "sample_kernel"}
            score_frac = fl["score"] / max(share["score"] * 1e-3, 1e-12) / 1e12
            out.update({
                "kernel_ms_per_step_single_stream": share, "single_stream_ms_per_step": m["single_stream_ms"],
                "decide_phase_gcycles": {k2: round(v / 1e9, 3) for k2, v in m["phase_cycles"].items()},
                "roofline": {"bound": bound, "kernel": names.get(dom, dom), "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                             "frac": (ach / peak) if ach is not None else None,
                             "traffic": traffic, "traffic_source": traffic_src or "not measured in this run (no profiler); the class is ALU / latency bound",
                             "peak_source": "register-only FFMA / DFMA probes measured in this run (fp32 %.1f, fp64 %.1f TFLOP/s); "
                                            "MEASURED_PEAKS.json holds HBM and bf16 tensor peaks only" % (peak32, peak64),
                             "avg_launch_ms": {k2: v["ms"] / max(v["launches"], 1) for k2, v in brk.items()},
                             "algorithmic_flop_per_step": {"score_fp32": fl["score"], "decide_fp64": fl["decide"]},
                             "scoring_in_loop": {"achieved": score_frac, "peak": peak32, "frac": score_frac / peak32,
                                                 "evals_fp32_per_step": e32,
                                                 "note": "evaluations counted by score_batch_kernel itself (speculated iterations included); "
                                                         "LocalOptimization's fp64 ScoreModel passes are in `decide`"}}})
        if not a.no_extra and world == 1:
            out["roofline"]["scoring_standalone"] = scoring_standalone(capi, local, 3, peak32) if "roofline" in out else None
            if "roofline" in out:
                out["roofline"]["scoring_standalone_focal"] = scoring_standalone(capi, local, 1, peak32)
            out["single_pair_latency_ms"] = {"configs[1]": single_pair_latency(local, 1), "configs[0]": single_pair_latency(local, 0)}
            if a.config == 1:
                extra = {}
                for c, st, wu in ((0, 10, 3), (2, 2, 1), (3, 2, 1)):
                    mm = measure_config(capi, torch, dist, local, c, workload.CONFIG_INFO[c]["pairs"], 0, 1, st, wu, breakdown=False)
                    mm["n_pairs_global"] = mm["n_pairs"]
                    s2 = summarize(mm, mm["parts"][0][0], 1, st)
                    extra[f"configs[{c}]"] = {"workload": workload_name(c, workload.CONFIG_INFO[c]["pairs"]), "steps": st, "warmup": wu,
                                              "value": s2["value"], "unit": "pairs/s", "ms_per_step": s2["ms_per_step"],
                                              "e2e": s2["e2e"], "residual_evals_per_s": s2["residual_evals_per_s"],
                                              "hypotheses_scored_per_s": s2["hypotheses_scored_per_s"],
                                              "ransac_iterations_per_s": s2["ransac_iterations_per_s"]}
                out["configs"] = extra
        threads = os.cpu_count() or 1
        parts, ns = cpu_sample(a.config, a.pairs)
        dt, cst, label = cpu_arm(parts, threads)
        out["cpu_baseline"] = {"value": ns / dt, "unit": "pairs/s", "cores": threads, "kind": "port",
                               "sample": f"first {ns} pairs of rank 0's workload, one pair per thread; {label}"}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
