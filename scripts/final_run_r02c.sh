#!/bin/bash
# Round-2 closing run on one B200: GPU tests, smoke, the bench lines of every workload, the reference arm, then the ncu launch
# list and one `ncu --set full` capture of the dominant kernels (each only after its command ran clean without ncu).
set -x
O=gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -3 > $O/r02c_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02c_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02c_smoke.log
python bench.py > $O/r02c_bench_ds2.log 2>&1; cp $O/op_profile_ds2_n1.json $O/r02c_op_profile_ds2.json
python bench.py --impl reference --steps 1 --warmup 0 > $O/r02c_bench_ref.log 2>&1
python bench.py --workload ds3 > $O/r02c_bench_ds3.log 2>&1; cp $O/op_profile_ds3_n1.json $O/r02c_op_profile_ds3.json
python bench.py --workload hgcal --no-cpu-baseline > $O/r02c_bench_hgcal.log 2>&1; cp $O/op_profile_hgcal_n1.json $O/r02c_op_profile_hgcal.json
python bench.py --workload ds1 --no-cpu-baseline > $O/r02c_bench_ds1.log 2>&1; cp $O/op_profile_ds1_n1.json $O/r02c_op_profile_ds1.json
python bench.py --workload ds2_train > $O/r02c_bench_ds2_train.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/r02c_launches_ds2.csv python bench.py --steps 1 --warmup 3 --sampling-steps 2 --no-cpu-baseline > $O/r02c_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv_mma_kernel\|gn_apply\|fa_ --launch-skip 2 -c 14 -o $O/r02c_ds2_full python scripts/ncu_target.py --evals 1 > $O/r02c_ncu_full.log 2>&1
tail -2 $O/r02c_ncu_full.log
tail -1 $O/r02c_bench_*.log | cut -c1-400
