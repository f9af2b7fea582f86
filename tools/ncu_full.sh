W=$1
mkdir -p gpurun_out/final
python bench.py --workload $W --steps 1 --warmup 1 --iters 3 --no-cpu-baseline --option graph=0 > gpurun_out/final/plain_$W.json 2>gpurun_out/final/plain_$W.err && ncu --set full --clock-control none --import-source on -k regex:"k_update_L|k_dot|k_chain|k_masked_argmax|k_argmax_final" -c 16 -o gpurun_out/final/full_$W python bench.py --workload $W --steps 1 --warmup 1 --iters 3 --no-cpu-baseline --option graph=0 > gpurun_out/final/ncu_$W.log 2>&1; tail -2 gpurun_out/final/ncu_$W.log
