#!/bin/bash
# Final round evidence on an 8-GPU box: bench scaling with the final build, and the Job manifest's command line.
K=kubernetes-with-nvidia-gpu_b200
mkdir -p gpurun_out
for n in 1 2 4 8; do
  if [ $n -eq 1 ]; then
    timeout 300 python bench.py --steps 1000 --warmup 20 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 1000 --warmup 20 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  fi
  echo "bench n=$n rc=$?"; python -c "import json; d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['triad']['node_gbs'], d['e2e']['value'], d['parity'].get('device_matches_cpu_baseline'), d['clocks']['reasons'])"
done
timeout 600 ./$K/gpu-validate --mode full --seconds 20 --hbm-fraction 0.9 --fail-on-throttle --json gpurun_out/full8_job.json > gpurun_out/full8_job.log 2>&1; echo "job command rc=$?"; grep -E "rank|under load|ECC|\[node\]|\[p2p\]|Test" gpurun_out/full8_job.log | tail -40
