set -x
python bench.py > gpurun_out/r02k_n1_bench_line.json 2> gpurun_out/r02k_n1.err; echo rc=$?
python bench.py --impl reference > gpurun_out/r02k_reference_arm_line.json 2> gpurun_out/r02k_ref.err; echo rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r02k_launches.csv python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02k_ncu_launches.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:k_validate --launch-skip 20 -c 1 -o gpurun_out/r02k_validate_fused -f python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02k_ncu_val.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:"k_runs_cluster|k_runs_prepare|k_cluster_classes|k_form|k_cluster_coop|k_select" --launch-skip 120 -c 12 -o gpurun_out/r02k_form -f python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02k_ncu_form.log 2>&1; echo rc=$?
ncu --set full --clock-control none --import-source on -k regex:k_validate --launch-skip 3 -c 1 -o gpurun_out/r02k_cfg5 -f python bench.py --workload cfg5 --cfg5-views 40 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r02k_ncu_cfg5.log 2>&1; echo rc=$?
ls -la gpurun_out
