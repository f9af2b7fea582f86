set -x
mkdir -p gpurun_out/f2
python bench.py > gpurun_out/f2/bench_inst60000.json 2> gpurun_out/f2/bench_inst60000.err
python bench.py --workload ml1m --no-cpu-baseline > gpurun_out/f2/bench_ml1m.json 2> gpurun_out/f2/bench_ml1m.err
python bench.py --workload ml100k --no-cpu-baseline > gpurun_out/f2/bench_ml100k.json 2> gpurun_out/f2/bench_ml100k.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/f2/bench_reference_arm.json 2> gpurun_out/f2/bench_reference_arm.err
python bench.py --steps 2 --warmup 3 --iters 16 --no-cpu-baseline > gpurun_out/f2/plain.json 2>gpurun_out/f2/plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f2/launches_inst60000.csv python bench.py --steps 2 --warmup 3 --iters 16 --no-cpu-baseline > gpurun_out/f2/ncu_launch.log 2>&1
for W in inst60000 ml1m; do
python bench.py --workload $W --steps 1 --warmup 1 --iters 3 --no-cpu-baseline --option graph=0 > gpurun_out/f2/plain_$W.json 2>gpurun_out/f2/plain_$W.err && ncu --set full --clock-control none --import-source on -k regex:"k_update_L|k_dot|k_chain|k_masked_argmax|k_argmax_final" -c 16 -o gpurun_out/f2/full_$W python bench.py --workload $W --steps 1 --warmup 1 --iters 3 --no-cpu-baseline --option graph=0 > gpurun_out/f2/ncu_$W.log 2>&1; tail -2 gpurun_out/f2/ncu_$W.log
done
ls -la gpurun_out/f2
