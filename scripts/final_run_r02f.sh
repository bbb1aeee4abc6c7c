#!/bin/bash
# Round-2 closing run on one B200 (final build): GPU tests, smoke, the bench lines of every workload with both reference legs,
# the reference arm, then the ncu launch list of the default bench command.
set -x
O=gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -4 > $O/r02f_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02f_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02f_smoke.log
python bench.py > $O/r02f_bench_ds2.log 2>&1; python scripts/show_profile.py $O/op_profile_ds2_n1.json > $O/r02f_op_profile_ds2_n1.txt
python bench.py --impl reference --steps 1 --warmup 0 > $O/r02f_bench_ref.log 2>&1
python bench.py --workload ds3 > $O/r02f_bench_ds3.log 2>&1; python scripts/show_profile.py $O/op_profile_ds3_n1.json > $O/r02f_op_profile_ds3_n1.txt
python bench.py --workload hgcal > $O/r02f_bench_hgcal.log 2>&1; python scripts/show_profile.py $O/op_profile_hgcal_n1.json > $O/r02f_op_profile_hgcal_n1.txt
python bench.py --workload ds1 > $O/r02f_bench_ds1.log 2>&1
python bench.py --workload ds2_train > $O/r02f_bench_ds2_train.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/r02f_launches_ds2.csv python bench.py --steps 1 --warmup 3 --sampling-steps 2 --no-cpu-baseline > $O/r02f_ncu_list.log 2>&1
cat $O/r02f_pytest_gpu_final.log $O/r02f_smoke.log
for w in ds2 ds3 hgcal ds1 ds2_train; do tail -1 $O/r02f_bench_$w.log | cut -c1-200; done
