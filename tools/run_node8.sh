#!/bin/bash
# Round-end evidence run on an 8-GPU box (gpurun --gpus 8): tests, bench scaling, node/p2p/full modes.
K=kubernetes-with-nvidia-gpu_b200
mkdir -p gpurun_out
G=$(nvidia-smi -L | wc -l)
echo "GPUs: $G"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu${G}.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu${G}.log
timeout 200 ./$K/gpu-validate --mode bench --measured-copy-gbs 6444.1 --json gpurun_out/stream_sweep.json > gpurun_out/stream_sweep.log 2>&1; echo "sweep rc=$?"
for n in 1 2 4 8; do
  [ $n -gt $G ] && continue
  if [ $n -eq 1 ]; then
    timeout 300 python bench.py --steps 500 --warmup 10 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 500 --warmup 10 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  fi
  echo "bench n=$n rc=$?"; python -c "import json,sys; d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['triad']['node_gbs'], d['e2e']['value'], d['clocks'])"
done
timeout 120 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2>&1; echo "ref rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $G --steps 100 --warmup 5 --workload triad-sharded > gpurun_out/bench_sharded_n$G.json 2> gpurun_out/bench_sharded.err; echo "bench sharded rc=$?"
for n in 1 2 4 8; do
  [ $n -gt $G ] && continue
  timeout 300 ./$K/gpu-validate --mode node --gpus $n --json gpurun_out/node${n}_2p32.json > gpurun_out/node${n}_2p32.log 2>&1; echo "node $n rc=$?"; grep "^\[node\]" gpurun_out/node${n}_2p32.log | tail -2
done
timeout 400 ./$K/gpu-validate --mode p2p --json gpurun_out/p2p${G}.json > gpurun_out/p2p${G}.log 2>&1; echo "p2p rc=$?"; tail -4 gpurun_out/p2p${G}.log
timeout 400 ./$K/gpu-validate --mode full --seconds 20 --p2p-bytes 268435456 --json gpurun_out/full${G}.json > gpurun_out/full${G}.log 2>&1; echo "full rc=$?"; tail -3 gpurun_out/full${G}.log
