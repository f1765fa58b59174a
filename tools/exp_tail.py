#!/usr/bin/env python
"""experiment: how the number of unfinished pairs falls at the end of a step (MP_TRACE timeline), per rank data range"""
import os, sys, re, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 2 and sys.argv[1] == "child":
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    from madpose_b200 import capi, workload
    r = int(sys.argv[2])
    kind, pairs, opt, cfg, keep = workload.make_batch(1, 1500, first=1500 * r)
    ctx = capi.Context(0)
    ctx.upload_batch(kind, pairs, opt)
    ctx.estimate_resident(opt, cfg); ctx.estimate_resident(opt, cfg)
    os.environ["MP_TRACE"] = "1"
    st = ctx.estimate_resident(opt, cfg)[1]
    lo = sorted(s.number_lo_iterations for s in st)
    lm = sorted(s.lm_iterations for s in st)
    print("LO per pair: max", lo[-5:], "LM iterations per pair: max", lm[-5:], "median", lm[len(lm) // 2])
    sys.exit(0)
for r in [int(v) for v in sys.argv[1:]] or [0, 7]:
    out = subprocess.run([sys.executable, __file__, "child", str(r)], capture_output=True, text=True)
    ev = []
    for l in out.stderr.splitlines():
        m = re.match(r"\[mp\]\s+([0-9.]+) ms\s+g(\d)\s+.*act (\d+)", l)
        if m: ev.append((float(m.group(1)), int(m.group(2)), int(m.group(3))))
    act = {0: 750, 1: 750}; tl = []
    for t, g, a in ev:
        act[g] = a; tl.append((t, act[0] + act[1]))
    end = tl[-1][0] if tl else 0
    marks = {}
    for t, a in tl:
        for thr in (1000, 500, 250, 100, 50, 20, 5, 0):
            if a <= thr and thr not in marks: marks[thr] = round(t, 1)
    print("rank", r, "end", end, "first time with <= N unfinished pairs:", marks)
    print(out.stdout.strip())
