#!/bin/bash
# Round-2 single-GPU evidence pass (run through gpurun; outputs land in gpurun_out/).
# Every ncu pass comes after the same command has exited 0 without ncu.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi > $O/r02_nvidia-smi.txt 2>&1
python tools/time_fill.py > $O/r02_time_fill.txt 2>&1; echo "time_fill rc=$?"; cat $O/r02_time_fill.txt
python tools/profile_all.py > /dev/null 2>&1 && echo "profile_all ok"
python bench.py > $O/r02_bench_default.json 2> $O/r02_bench_default.err; echo "bench default rc=$?"
B="python bench.py --steps 20 --warmup 5 --no-cpu --no-native --no-sharded --e2e-steps 2 --sustained-seconds 0.02"
$B > $O/r02_bench_for_ncu.json 2>&1; echo "bench (ncu shape) rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02_ncu_launches_raw.csv $B > $O/r02_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 8 -c 3 -f -o $O/r02_add_full $B > $O/r02_ncu_add_full.log 2>&1; echo "ncu add full rc=$?"
ncu -i $O/r02_add_full.ncu-rep --page raw --csv --print-units base > $O/r02_ncu_add_full_raw.csv 2>/dev/null
ncu --set full --clock-control none -k regex:'stream_kernel|checksum_kernel|fill_kernel' -s 13 -c 13 -f -o $O/r02_all_kernels python tools/profile_all.py > $O/r02_ncu_all.log 2>&1; echo "ncu all rc=$?"
ncu -i $O/r02_all_kernels.ncu-rep --page raw --csv --print-units base > $O/r02_ncu_all_kernels_raw.csv 2>/dev/null
python tools/e2e_scaling.py $O/r02_e2e_scaling_1gpubox.json 2>&1 | tail -12
ls -la $O | tail -20
