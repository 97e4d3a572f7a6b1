#!/bin/bash
# Round-2 multi-GPU evidence pass: run with `gpurun --gpus 8 -- bash tools/run_r02_8gpu.sh` (outputs in gpurun_out/).
set -u
O=gpurun_out
G=${1:-8}
mkdir -p $O
nvidia-smi topo -m > $O/r02_topo_${G}gpu.txt 2>&1
python -m pytest tests/test_gpu_launcher.py tests/test_zz_no_alloc.py -q -m gpu -k "sharded_over_all or p2p or peer or full_mode or per_device" 2>&1 | tail -5 | tee $O/r02_pytest_multi_${G}gpu.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $G --steps 1000 --warmup 20 > $O/r02_bench_n${G}.json 2> $O/r02_bench_n${G}.err; echo "bench N=$G rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29522 tools/e2e_scaling.py $O/r02_e2e_scaling_${G}gpubox.json > $O/r02_e2e_scaling_${G}gpubox.log 2>&1; echo "e2e_scaling rc=$?"; grep -E "^(table|solo|pairs|host|thp)" $O/r02_e2e_scaling_${G}gpubox.log
./kubernetes-with-nvidia-gpu_b200/gpu-validate --mode full --seconds 10 --hbm-fraction 0.9 --fail-on-throttle --json $O/r02_full_${G}gpu.json > $O/r02_full_${G}gpu_job_command.log 2>&1; echo "full rc=$?"; tail -14 $O/r02_full_${G}gpu_job_command.log
python - <<PY
import json
d = json.loads([l for l in open("$O/r02_bench_n${G}.json") if l.startswith("{")][-1])
print({k: d[k] for k in ("value", "n_gpus")}, d["e2e"]["value"], d["parity"], d["triad_sharded_2p32"]["node_gbs"], d["triad_sharded_2p32"]["matches_oracle"],
      d["native_node"]["node_gbs"], d["native_node"]["ok"], d["p2p"]["ok"], d["p2p"]["gbs"], d["sustained"]["add"]["gbs"], d["sustained"]["add"]["clocks"])
PY
