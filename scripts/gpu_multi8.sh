mkdir -p gpurun_out
N=${N:-8}
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
python -m pytest tests/test_gpu_nccl.py -x -q -m gpu 2>&1 | tail -2
summ() { python - "$1" <<'PY'
import json,sys
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        ok=True
        d=json.loads(l); r=d['roofline']; print(sys.argv[1].split('/')[-1],'n',d['n_gpus'],d['scaling'],'value',round(d['value'],2),'frac',round(r['frac'],3),'ms',round(d['ms_per_step'],2),'kernel_ms',round(r['kernel_ms_per_apply'],2),'cg',d['cg'] and round(d['cg']['value'],2))
if not ok: print(open(sys.argv[1]).read()[-800:])
PY
}
run() { # name n extra-args
  name=$1; n=$2; shift 2
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu --no-e2e "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $n --steps 10 --warmup 3 --no-e2e "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; fi
  summ gpurun_out/$name.json
}
for n in 1 2 4 8; do [ $n -le $N ] && run scale2_q3_n$n $n; done
for n in 1 2 4 8; do [ $n -le $N ] && run scale2_q6_n$n $n --degree 6 --ndofs 500000000; done
for n in 1 2 4 8; do [ $n -le $N ] && run strong_q3_f32_n$n $n --degree 3 --float 32 --ndofs_global 1000000000 --no-cg; done
for n in 1 2 4 8; do [ $n -le $N ] && run strong_q6_f32_n$n $n --degree 6 --float 32 --ndofs_global 1000000000 --no-cg; done
./benchmark_dolfinx_b200/bench_dolfinx --ngpus $N --ndofs 300000000 --degree 3 --cg --nreps 20 --json gpurun_out/cli_q3_cg_n$N.json 2>&1 | tail -8
./benchmark_dolfinx_b200/bench_dolfinx --ngpus $N --ndofs_global 100000 --degree 3 --nreps 1 --mat_comp --json gpurun_out/cli_matcomp_n$N.json 2>&1 | tail -8
cat gpurun_out/cli_matcomp_n$N.json
