set -x
mkdir -p gpurun_out
N=${N:-2}
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
nvidia-smi -L
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --ndofs 50000000 > gpurun_out/bench_n${N}_small.json 2> gpurun_out/bench_n${N}_small.err; echo "rc=$?"; tail -5 gpurun_out/bench_n${N}_small.err; grep '^{' gpurun_out/bench_n${N}_small.json | cut -c1-1500
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err; echo "rc=$?"; tail -5 gpurun_out/bench_n${N}.err; grep '^{' gpurun_out/bench_n${N}.json | cut -c1-2500
