set -x
mkdir -p gpurun_out
N=${N:-8}
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
nvidia-smi -L | wc -l
python -m pytest tests/test_gpu_nccl.py -x -q -m gpu 2>&1 | tail -15
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/scale_q3_n$n.json 2> gpurun_out/scale_q3_n$n.err
    else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 10 --warmup 3 --no-e2e > gpurun_out/scale_q3_n$n.json 2> gpurun_out/scale_q3_n$n.err; fi
    echo "rc=$?"; tail -3 gpurun_out/scale_q3_n$n.err
    python - gpurun_out/scale_q3_n$n.json <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print(sys.argv[1],'n',d['n_gpus'],'value',round(d['value'],2),'frac',round(r['frac'],3),'ms',round(d['ms_per_step'],2),'kernel_ms',round(r['kernel_ms_per_apply'],2),'cg',d['cg'] and round(d['cg']['value'],2))
PY
  fi
done
for n in 8; do
  if [ $n -le $N ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 10 --warmup 3 --degree 6 --ndofs 500000000 --no-e2e > gpurun_out/scale_q6_n$n.json 2> gpurun_out/scale_q6_n$n.err
    echo "rc=$?"; tail -3 gpurun_out/scale_q6_n$n.err
    python - gpurun_out/scale_q6_n$n.json <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print(sys.argv[1],'n',d['n_gpus'],'value',round(d['value'],2),'frac',round(r['frac'],3),'ms',round(d['ms_per_step'],2),'kernel_ms',round(r['kernel_ms_per_apply'],2),'cg',d['cg'] and round(d['cg']['value'],2))
PY
  fi
done
