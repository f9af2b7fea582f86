# launch list + ncu --set full of the loop / arg-max kernels (inst60000 and ML1M shapes), one GPU
OUT=${1:-ev}
mkdir -p gpurun_out/$OUT
C="python bench.py --steps 2 --warmup 3 --iters 16 --no-cpu-baseline --no-extra"
$C > gpurun_out/$OUT/plain.json 2>gpurun_out/$OUT/plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/$OUT/launches_inst60000.csv $C > gpurun_out/$OUT/ncu_launch.log 2>&1
for W in inst60000 ml1m; do
C="python bench.py --workload $W --steps 1 --warmup 3 --iters 3 --no-cpu-baseline --no-extra --option graph=0"
$C > gpurun_out/$OUT/plain_$W.json 2>gpurun_out/$OUT/plain_$W.err && ncu --set full --clock-control none --import-source on -k regex:"k_update_L|k_dot|k_chain|k_acc" -s 12 -c 8 -o gpurun_out/$OUT/full_$W $C > gpurun_out/$OUT/ncu_$W.log 2>&1; tail -1 gpurun_out/$OUT/ncu_$W.log
done
