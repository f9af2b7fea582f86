# the headline workload with the exchange inside k_acc on and off.  usage: NGPUS=8 tools/fused_xchg_bench.sh out
OUT=${1:-ovl}; N=${NGPUS:-2}
mkdir -p gpurun_out/$OUT
for V in ${VALS:-1 0}; do
MFGPU_COMM_TIMEOUT_MS=4000 MFGPU_XCHG_FUSED=$V timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps ${STEPS:-5} --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/$OUT/bench_n${N}_ovl$V.json 2> gpurun_out/$OUT/bench_n${N}_ovl$V.err; tail -1 gpurun_out/$OUT/bench_n${N}_ovl$V.json | cut -c1-330
done
