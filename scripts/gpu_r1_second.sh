set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -2
python -m pytest tests -x -q -m gpu 2>&1 | tail -8
python bench.py > gpurun_out/bench_q3_v1.json 2> gpurun_out/bench_q3_v1.err; echo "rc=$?"; tail -3 gpurun_out/bench_q3_v1.err
python bench.py --degree 6 --ndofs 500000000 --no-cpu > gpurun_out/bench_q6_v1.json 2> gpurun_out/bench_q6_v1.err; echo "rc=$?"; tail -3 gpurun_out/bench_q6_v1.err
python bench.py --degree 1 --ndofs 50000000 --no-cpu --no-e2e > gpurun_out/bench_q1_v1.json 2>&1; echo "rc=$?"
python bench.py --degree 2 --ndofs 50000000 --no-cpu --no-e2e > gpurun_out/bench_q2_v1.json 2>&1; echo "rc=$?"
python bench.py --degree 3 --ndofs 500000000 --float 32 --no-cpu --no-e2e > gpurun_out/bench_q3_f32_v1.json 2>&1; echo "rc=$?"
python bench.py --degree 6 --ndofs 500000000 --float 32 --no-cpu --no-e2e > gpurun_out/bench_q6_f32_v1.json 2>&1; echo "rc=$?"
python bench.py --degree 3 --qmode 0 --no-cpu --no-e2e > gpurun_out/bench_q3_q0_v1.json 2>&1; echo "rc=$?"
for f in gpurun_out/bench_*_v1.json; do python - "$f" <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print(sys.argv[1], 'value',round(d['value'],2),'frac',round(r['frac'],3),'kernel_ms',round(r['kernel_ms_per_apply'],2),'cg',d['cg'] and round(d['cg']['value'],2),'e2e',d['e2e'] and round(d['e2e']['value'],2),'cpu',d['cpu_baseline'] and d['cpu_baseline']['value'])
PY
done
