# ncu --set full captures behind profiles/traffic.json (run on one GPU after the same commands exited 0 without ncu).
# The 16-launch cfg4 report is ~80 MB: only its raw-page CSV comes back.
set -x
ncu --set full --clock-control none -k regex:k_validate --launch-skip 48 -c 16 -o /tmp/r02n_cfg4_validate -f python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02n_ncu_cfg4.log 2>&1; echo rc=$?
ncu -i /tmp/r02n_cfg4_validate.ncu-rep --page raw --csv > gpurun_out/r02n_cfg4_validate_raw.csv 2>/dev/null; rm -f /tmp/r02n_cfg4_validate.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:k_validate --launch-skip 3 -c 1 -o gpurun_out/r02n_cfg2_validate -f python bench.py --workload cfg2 --steps 2 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02n_ncu_cfg2.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:k_validate --launch-skip 3 -c 1 -o gpurun_out/r02n_cfg5_validate -f python bench.py --workload cfg5 --cfg5-views 40 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r02n_ncu_cfg5.log 2>&1; echo rc=$?
ls -la gpurun_out | tail -8; du -sh gpurun_out
