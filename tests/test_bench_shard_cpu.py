"""CPU tests of the multi-GPU plumbing (independent textures, no data-path collective).

* bench.py's N > 1 path -- rank -> shard, barrier, whole-job value = pixels of all ranks / max-over-ranks time -- run with
  world_size 2 over gloo (the CUDA arm runs the same Ranks / whole_job_value code over nccl);
* hkbcn_plan_batch, the host-only planning step of hkbcn_compress_batch (greedy by pixel-count-weighted cost)."""
import json
import os
import subprocess
import sys

import hkbcn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_sharding_world2_gloo():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29541", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--selftest-shard"],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, "exactly one JSON line, printed by rank 0"
    r = json.loads(lines[0])
    assert r["n_gpus"] == 2
    assert r["ms_per_step"] == 20.0                      # max over ranks of (10, 20)
    assert abs(r["value"] - 3000 / 20e-3 / 1e6) < 1e-12  # pixels of all ranks / max time
    assert r["seeds"][1] == r["seeds"][0] + 1            # one independent texture per rank


def test_plan_batch_balances_and_is_deterministic():
    BC1, BC5, BC7 = 54, 62, 66
    jobs = [(BC1, 2048, 2048)] * 16 + [(BC5, 2048, 2048)] * 8 + [(BC7, 2048, 2048)] * 8   # config 5's 50/25/25 mix
    for n in (1, 2, 4, 8):
        plan = hkbcn.plan_batch(jobs, n)
        assert plan == hkbcn.plan_batch(jobs, n)
        assert set(plan) == set(range(n))
        # the expensive format (BC7) is spread evenly; every job is assigned exactly once
        bc7 = [plan[i] for i in range(24, 32)]
        assert max(bc7.count(d) for d in range(n)) == 8 // n
    # unequal sizes: the largest job alone, the rest balanced
    plan = hkbcn.plan_batch([(BC7, 4096, 4096), (BC7, 2048, 2048), (BC7, 2048, 2048), (BC7, 2048, 2048), (BC7, 2048, 2048)], 2)
    assert plan[0] != plan[1] and plan[1:] == [plan[1]] * 4


def test_plan_batch_rejects_bad_input():
    import pytest
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.plan_batch([(99, 4, 4)], 2)
    with pytest.raises(hkbcn.HkbcnError):
        hkbcn.plan_batch([(54, 4, 4)], 0)
