"""N > 1 host logic of bench.py on CPU (gloo, world_size 2): every rank owns a disjoint contiguous slice of the
pair indices (no data-path collective), the per-rank step times are combined with an all-reduce MAX, and the
whole-job value is (pairs of all ranks) / max time -- the same arithmetic bench.py runs over NCCL."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, %r)
    import numpy as np, torch, torch.distributed as dist
    import bench
    from madpose_b200 import synth
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    P = 5
    (kind, pairs, opt, cfg, keep), = bench.build_workload(1, P, rank, world)   # weak: rank r owns pairs [r*P, (r+1)*P)
    first = synth.make_pair(1, rank * P)
    assert np.array_equal(keep[0][0]["x0"], first["x0"]) and kind == 1
    # strong (configs[4]): the fixed batch of 12 mixed pairs is dealt round-robin over the ranks, kinds grouped per rank
    parts = bench.build_workload(4, 12, rank, world)
    assert sum(len(part[1]) for part in parts) == 6 and sorted(part[0] for part in parts) == [0, 1, 2]
    sig = torch.tensor([float(sum(p.n for p in pairs))], dtype=torch.float64)
    gathered = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, sig)
    tt = torch.tensor([100.0 + 50.0 * rank], dtype=torch.float64)         # pretend step times (ms)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    value = world * P / (tt.item() * 1e-3)
    if rank == 0:
        print(json.dumps({"sizes": [g.item() for g in gathered], "max_ms": tt.item(), "value": value}))
    dist.destroy_process_group()
""") % ROOT


def test_two_rank_sharding_and_max_reduce(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29531", str(w)],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["max_ms"] == 150.0 and abs(r["value"] - 2 * 5 / 0.150) < 1e-9
    assert len(r["sizes"]) == 2 and r["sizes"][0] != r["sizes"][1]      # different pairs on the two ranks


def test_reference_arm_only_rank0_prints(tmp_path):
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""
