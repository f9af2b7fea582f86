# one ncu --set full capture of the loop kernels of a bench workload (after the same command ran clean)
W=${1:-inst60000}; OUT=${2:-gpurun_out/ncu}; N=${3:-6}; shift 3 2>/dev/null
mkdir -p $OUT
CMD="python bench.py --workload $W --steps 1 --warmup 1 --iters 3 --no-cpu-baseline --option graph=0 $*"
$CMD > $OUT/plain_$W.json 2> $OUT/plain_$W.err && ncu --set full --clock-control none --import-source on -k regex:"k_update_L|k_dot|k_chain|k_item|k_acc" -c $N -o $OUT/full_$W $CMD > $OUT/ncu_$W.log 2>&1; tail -2 $OUT/ncu_$W.log
