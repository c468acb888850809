# Full check on one GPU: GPU test suite, smoke, the default bench line (what the driver runs).
set -x
mkdir -p gpurun_out
TAG=${TAG:-r2}
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | cut -c1-400 | tee gpurun_out/${TAG}_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/${TAG}_smoke.txt
timeout 900 python bench.py > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err; echo "bench rc=$?"
tail -c 600 gpurun_out/${TAG}_bench_default.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc=$?"
python - gpurun_out/${TAG}_bench_default.json gpurun_out/${TAG}_bench_reference.json <<'PY'
import json, sys
for f in sys.argv[1:]:
    for l in open(f):
        if l.startswith('{'):
            d = json.loads(l)
            r = d.get('roofline') or {}
            print(f.split('/')[-1], 'value', round(d['value'], 3), d['config'].get('geometry'), 'bound', r.get('bound'), 'frac', r.get('frac') and round(r['frac'], 3),
                  'kernel_ms', r.get('kernel_ms_per_apply') and round(r['kernel_ms_per_apply'], 3))
            for k, v in (d.get('variants') or {}).items():
                print('  variant', k, round(v['value'], 2), v['roofline']['bound'], round(v['roofline']['frac'], 3), 'cg', v.get('cg') and round(v['cg'], 2))
            q6 = d.get('q6')
            if q6:
                print('  q6', round(q6['value'], 2), q6['geometry'], q6['roofline']['bound'], round(q6['roofline']['frac'], 3), 'cg', q6['cg'] and round(q6['cg']['value'], 2),
                      {k: round(v['value'], 2) for k, v in q6.get('variants', {}).items()})
            print('  cg', d.get('cg') and round(d['cg']['value'], 2), 'e2e', d.get('e2e') and round(d['e2e']['value'], 2), 'cpu', d.get('cpu_baseline') and d['cpu_baseline']['value'],
                  'gpu_q0', d.get('gpu_qmode0') and round(d['gpu_qmode0']['value'], 2))
            print('  refgpu', d.get('reference_gpu_kernel'))
            print('  parity', d.get('parity') and {k: d['parity'][k] for k in ('rel_err', 'cg_rel_err', 'ok')}, 'clocks', d.get('clocks'))
PY
