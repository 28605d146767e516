#!/bin/bash
# ncu evidence of one round, run on the GPU box through gpurun (one call): every ncu command follows a plain run of the
# same command line that exited 0 (B200_PROFILING.md).  Outputs land in gpurun_out/.
set -u
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 2 --warmup 3 --no-cpu"
$B > $O/ncu_bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $B > $O/ncu_bench.log 2>&1
echo "launch list rc=$?"
for k in mma res taco dm; do
  pat=mma_loop; [ $k = res ] && pat=res_loop; [ $k = taco ] && pat=taco_loop; [ $k = dm ] && pat=dm_loop
  python scripts/prof_target.py $k > $O/ncu_${k}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$pat -c 1 -f -o $O/prof_$k python scripts/prof_target.py $k > $O/ncu_$k.log 2>&1
  echo "$k capture rc=$?"
done
python scripts/prof_target.py mma > $O/ncu_cond_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:"cond_operand|resnet_batch" -c 2 -f -o $O/prof_cond python scripts/prof_target.py mma > $O/ncu_cond.log 2>&1
echo "cond capture rc=$?"
ls -la $O/*.ncu-rep $O/launches.csv
