set -x
python -m pytest tests -q -m gpu 2>&1 | tail -3 > gpurun_out/r01_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_final.log
python bench.py > gpurun_out/bench_ds2_final.log 2>&1; cp gpurun_out/op_profile_ds2_n1.json gpurun_out/op_profile_ds2_final.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_final.log 2>&1
python bench.py --workload ds3 > gpurun_out/bench_ds3_final.log 2>&1; cp gpurun_out/op_profile_ds3_n1.json gpurun_out/op_profile_ds3_final.json
python bench.py --workload hgcal --no-cpu-baseline > gpurun_out/bench_hgcal_final.log 2>&1; cp gpurun_out/op_profile_hgcal_n1.json gpurun_out/op_profile_hgcal_final.json
python bench.py --workload ds1 --no-cpu-baseline > gpurun_out/bench_ds1_final.log 2>&1; cp gpurun_out/op_profile_ds1_n1.json gpurun_out/op_profile_ds1_final.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r01_launches_ds2_v10.csv python bench.py --steps 1 --warmup 3 --sampling-steps 2 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv_mma_kernel\|gn_apply --launch-skip 69 -c 12 -o gpurun_out/r01_ds2_L0_full_v10 python scripts/ncu_target.py --evals 2 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
