#!/bin/bash
# Round-2 closing run after the Z-stacked conv: GPU tests, smoke, DS3 / HGCal / DS2 bench lines, then one ncu capture of the DS3 convs.
set -x
O=gpurun_out
python -m pytest tests -q -m gpu 2>&1 | tail -4 > $O/r02d_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02d_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02d_smoke.log
python bench.py --workload ds3 > $O/r02d_bench_ds3.log 2>&1; cp $O/op_profile_ds3_n1.txt $O/r02d_op_profile_ds3_n1.txt
python bench.py --workload hgcal --no-cpu-baseline > $O/r02d_bench_hgcal.log 2>&1
python bench.py --no-cpu-baseline > $O/r02d_bench_ds2.log 2>&1; cp $O/op_profile_ds2_n1.txt $O/r02d_op_profile_ds2_n1.txt
ncu --set full --clock-control none --import-source on -k regex:conv_zs_kernel\|conv_mma_kernel --launch-skip 0 -c 6 -o $O/r02d_ds3_convs python scripts/ncu_target.py --workload ds3 --evals 1 > $O/r02d_ncu_full.log 2>&1
tail -2 $O/r02d_ncu_full.log
cat $O/r02d_pytest_gpu_final.log $O/r02d_smoke.log
for f in $O/r02d_bench_ds3.log $O/r02d_bench_hgcal.log $O/r02d_bench_ds2.log; do tail -1 $f | cut -c1-300; done
