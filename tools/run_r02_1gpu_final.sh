set -u
O=gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -4 | tee $O/r02_pytest_1gpu.txt
python bench.py > $O/r02_bench_default.json 2> $O/r02_bench_default.err; echo "bench default rc=$?"
python bench.py --impl reference --steps 20 --warmup 5 > $O/r02_bench_ref.json 2>&1; echo "ref rc=$?"
python tools/profile_all.py > /dev/null 2>&1 && echo "profile_all ok"
ncu --set full --clock-control none -k regex:'stream_kernel|checksum_kernel|fill_kernel' -s 13 -c 13 -f -o $O/r02_all_kernels python tools/profile_all.py > $O/r02_ncu_all.log 2>&1; echo "ncu all rc=$?"
ncu -i $O/r02_all_kernels.ncu-rep --page raw --csv --print-units base > $O/r02_ncu_all_kernels_raw.csv 2>/dev/null
./kubernetes-with-nvidia-gpu_b200/gpu-validate --mode bench --json $O/r02_stream_sweep.json > $O/r02_stream_sweep.log 2>&1; echo "sweep rc=$?"; tail -3 $O/r02_stream_sweep.log
