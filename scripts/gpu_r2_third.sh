# Round 2, third call: full suite, default bench line, ncu of the two on-the-fly kernels, one more A/B (nq = 7, Phi = I).
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | cut -c1-400 | tee gpurun_out/r2_third_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 | tee gpurun_out/r2_third_smoke.txt
timeout 900 python bench.py > gpurun_out/r2_third_bench_default.json 2> gpurun_out/r2_third_bench_default.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2_third_bench_default.err
for g in precomputed onthefly; do
timeout 200 python bench.py --degree 6 --qmode 0 --ndofs 300000000 --geometry $g --no-cpu --no-e2e --no-cg --no-variants --no-q6 --no-parity --no-reference-gpu-kernel --steps 10 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT Q6 q0 $g: %.2f GDoF/s kernel_ms %.3f' % (d['value'], r['kernel_ms_per_apply']))
" | tee -a gpurun_out/r2_third_q6q0.txt
done
ARGS="--no-cpu --no-e2e --no-cg --no-variants --no-q6 --no-parity --no-reference-gpu-kernel --steps 2 --warmup 3 --geometry onthefly"
ncu --set full --clock-control none --import-source on -k regex:otf -s 3 -c 1 -o gpurun_out/r2_prof_q3_otf -f python bench.py $ARGS > gpurun_out/r2_ncu_q3_otf.log 2>&1; echo "ncu q3 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:otf -s 3 -c 1 -o gpurun_out/r2_prof_q6_otf -f python bench.py --degree 6 --ndofs 500000000 $ARGS > gpurun_out/r2_ncu_q6_otf.log 2>&1; echo "ncu q6 rc=$?"
ls -la gpurun_out/*.ncu-rep
