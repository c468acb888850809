# N-GPU run (N = 2, 4, 8; gpurun --gpus N): NCCL parity test, then the default bench line (with its parity leg,
# both geometry variants and the halo timing sub-record) exactly as the driver launches it.
set -x
mkdir -p gpurun_out
N=${N:-2}
TAG=${TAG:-r2}
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_nccl.py tests/test_gpu_multirank.py -x -q 2>&1 | tail -5 | cut -c1-400 | tee gpurun_out/${TAG}_n${N}_tests.txt
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 $EXTRA > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err; echo "rc=$?"
tail -5 gpurun_out/${TAG}_bench_n${N}.err | cut -c1-400
python - gpurun_out/${TAG}_bench_n${N}.json <<'PY'
import json, sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d = json.loads(l)
        if 'error' in d:
            print(l[:2000]); continue
        print('N', d['n_gpus'], 'value', round(d['value'], 2), d['config']['geometry'], 'ms', round(d['ms_per_step'], 3),
              'kernel_ms', round(d['roofline']['kernel_ms_per_apply'], 3), 'halo', d.get('halo'))
        for k, v in (d.get('variants') or {}).items():
            print('  variant', k, round(v['value'], 2) if v.get('value') else v, 'halo', v.get('halo') if isinstance(v, dict) else None)
        q6 = d.get('q6')
        if q6:
            print('  q6', round(q6['value'], 2), q6['geometry'], 'halo', q6.get('halo'), 'cg', q6['cg'] and round(q6['cg']['value'], 2))
        print('  cg', d['cg'] and round(d['cg']['value'], 2), 'e2e', d['e2e'] and round(d['e2e']['value'], 2), d['e2e'] and d['e2e'].get('per_rank_copy_gbs_each_way'))
        print('  parity', d['parity'] and {k: d['parity'][k] for k in ('rel_err', 'cg_rel_err', 'ok')})
PY
