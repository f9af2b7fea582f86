# N-GPU check of the exchange inside k_acc (option xchg_fused): parity tests with it
# forced on, then the headline workload with it on and off.  usage: NGPUS=2 tools/overlap_check.sh out
OUT=${1:-ovl}; N=${NGPUS:-2}
mkdir -p gpurun_out/$OUT
MFGPU_COMM_TIMEOUT_MS=4000 MFGPU_XCHG_FUSED=1 timeout 900 python -m pytest tests -m gpu -x -q -k "several_gpus or multi_gpu or rank_mode or item_sharded" > gpurun_out/$OUT/pytest_on.txt 2>&1; tail -3 gpurun_out/$OUT/pytest_on.txt
for V in 1 0; do
MFGPU_COMM_TIMEOUT_MS=4000 MFGPU_XCHG_FUSED=$V timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/$OUT/bench_n${N}_ovl$V.json 2> gpurun_out/$OUT/bench_n${N}_ovl$V.err; tail -1 gpurun_out/$OUT/bench_n${N}_ovl$V.json | cut -c1-400
done
