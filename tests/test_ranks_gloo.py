"""CPU tier, world_size 2 over gloo: the N>1 plumbing (shard plan, max over ranks, AND of verify
bits, checksum of checksums).  Each rank builds its shard of the host oracle's triad and the
summed shard checksums must equal the whole vector's — the same property the GPU ranks rely on."""
import json
import os
import socket
import subprocess
import sys
import textwrap

import oracle_lib as O
from b200v_pkg import ROOT

WORKER = textwrap.dedent("""
    import importlib, json, os, sys
    sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
    import oracle_lib as O
    pkg = importlib.import_module("kubernetes-with-nvidia-gpu_b200")
    ranks = importlib.import_module("kubernetes-with-nvidia-gpu_b200.ranks")
    R = ranks.Ranks(backend="gloo")
    n, seed = {n}, 20261106
    begin, count = pkg.shard_range(n, R.rank, R.world)          # C ABI, host-only arithmetic
    x = O.fill("f32", begin, count, seed, 0)
    y = O.fill("f32", begin, count, seed, 1)
    out = O.op("triad", "f32", count, s=3.0, x=x, y=y)
    total = R.sum_checksums(O.checksum(out, begin))
    R.barrier()
    R.host_barrier()                                            # the CPU-only wait (rank 0 runs the launcher binaries meanwhile)
    slowest = R.max(10.0 + R.rank)                              # stands in for per-rank device ms
    ok_all = R.all_ok(True)
    ok_one_bad = R.all_ok(R.rank != 1)
    if R.rank == 0:
        print(json.dumps({{"world": R.world, "total": total, "slowest": slowest, "ok_all": ok_all,
                          "ok_one_bad": ok_one_bad, "shard": [begin, count]}}))
    R.close()
""")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_ranks_over_gloo(tmp_path):
    n = 1_000_003
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, n=n))
    port = free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, se[-2000:]
    res = json.loads(outs[0][0].strip().splitlines()[-1])
    x = O.fill("f32", 0, n, 20261106, 0)
    y = O.fill("f32", 0, n, 20261106, 1)
    whole = O.checksum(O.op("triad", "f32", n, s=3.0, x=x, y=y))
    assert tuple(res["total"]) == whole
    assert res["world"] == 2 and res["slowest"] == 11.0
    assert res["ok_all"] is True and res["ok_one_bad"] is False
    assert res["shard"] == [0, 500_002]
