# Evidence of a round on ONE GPU (run under gpurun; everything lands in gpurun_out/$OUT):
#   the bench line, the launch list of the same command (ncu gpu__time_duration: shares, not absolutes),
#   and one ncu --set full capture of the loop and arg-max kernels of the inst60000 and ML1M shapes.
OUT=${1:-ev}
mkdir -p gpurun_out/$OUT
python bench.py > gpurun_out/$OUT/bench_inst60000.json 2> gpurun_out/$OUT/bench_inst60000.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/$OUT/bench_reference_arm.json 2> gpurun_out/$OUT/bench_reference_arm.err
C="python bench.py --steps 2 --warmup 3 --iters 16 --no-cpu-baseline --no-extra"
$C > gpurun_out/$OUT/plain.json 2>gpurun_out/$OUT/plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/$OUT/launches_inst60000.csv $C > gpurun_out/$OUT/ncu_launch.log 2>&1
for W in inst60000 ml1m; do
C="python bench.py --workload $W --steps 1 --warmup 3 --iters 3 --no-cpu-baseline --no-extra --option graph=0"
$C > gpurun_out/$OUT/plain_$W.json 2>gpurun_out/$OUT/plain_$W.err && ncu --set full --clock-control none --import-source on -k regex:"k_update_L|k_dot|k_chain|k_acc|k_masked_argmax" -s 12 -c 10 -o gpurun_out/$OUT/full_$W $C > gpurun_out/$OUT/ncu_$W.log 2>&1; tail -2 gpurun_out/$OUT/ncu_$W.log
done
ls -la gpurun_out/$OUT
