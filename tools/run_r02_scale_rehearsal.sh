#!/bin/bash
# What the driver does at round end, rehearsed on one 8-GPU box: for N in 1 2 4 8 the reference arm, then our arm,
# with the driver's flags.  gpurun --gpus 8 -- bash tools/run_r02_scale_rehearsal.sh
set -u
O=gpurun_out
mkdir -p $O
for n in 1 2 4 8; do
  for impl in reference ours; do
    f=$O/r02_scale_${impl}_n$n.json
    t0=$(date +%s.%N)
    if [ $n -eq 1 ]; then
      python bench.py --impl $impl --gpus 1 --steps 20 --warmup 5 > $f 2> $f.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --impl $impl --gpus $n --steps 20 --warmup 5 > $f 2> $f.err
    fi
    rc=$?
    t1=$(date +%s.%N)
    python - <<PY
import json
rc, n, impl = $rc, $n, "$impl"
try:
    d = json.loads([l for l in open("$f") if l.startswith("{")][-1])
except Exception as e:
    print(f"N={n} {impl}: rc={rc} NO JSON LINE ({e})"); raise SystemExit
wall = $t1 - $t0
if impl == "reference":
    print(f"N={n} reference rc={rc} wall={wall:.1f}s value={d['value']} cores={d['cpu_baseline']['cores']}")
else:
    p2p = d.get("p2p")
    print(f"N={n} ours rc={rc} wall={wall:.1f}s value={d['value']} e2e={d['e2e']['value']} parity={d['parity']['device_matches_cpu_baseline']} "
          f"sharded={d['triad_sharded_2p32']['node_gbs']}/{d['triad_sharded_2p32']['matches_oracle']} native={d['native_node']['node_gbs']}/{d['native_node']['ok']} "
          f"burnin={d['burnin']['ok']} p2p={(p2p or {}).get('ok')} sustained={d['sustained']['add']['gbs']} clocks={d['clocks']['reasons']}")
PY
  done
done
