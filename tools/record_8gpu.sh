# Everything that needs the 8-GPU box, in one lease (run under `gpurun --gpus 8`); output in gpurun_out/g8.
mkdir -p gpurun_out/g8
python -m pytest tests -q -m gpu -rs > gpurun_out/g8/pytest_gpu_8gpu.txt 2>&1; tail -4 gpurun_out/g8/pytest_gpu_8gpu.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
    tests/rank_mode_worker.py gpurun_out/g8/soak_8ranks.txt soak 100000 1000 > gpurun_out/g8/soak_stdout.txt 2>&1; cat gpurun_out/g8/soak_8ranks.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/g8/bench_8gpu.json 2> gpurun_out/g8/bench_8gpu.err; tail -c 600 gpurun_out/g8/bench_8gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29523 \
    bench.py --gpus 4 --steps 10 --warmup 3 --no-extra > gpurun_out/g8/bench_4gpu.json 2> gpurun_out/g8/bench_4gpu.err
NGPUS="8" bash tools/cli_times.sh > gpurun_out/g8/cli_times_8gpu.txt 2>&1; grep "==" gpurun_out/g8/cli_times_8gpu.txt | head -12
