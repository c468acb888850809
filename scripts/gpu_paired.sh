# FFMA2 variant of the fp32 pencil kernel: parity, then A/B (action only).
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
timeout 400 python -m pytest tests/test_gpu_pencil.py -q -k "0-1" 2>&1 | tail -8 | cut -c1-600
one() { # paired degree ndofs
  BD_PENCIL_PAIRED=$1 timeout 150 python bench.py --degree $2 --ndofs $3 --float 32 --qmode 1 --no-cpu --no-e2e --no-cg --steps 8 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT paired=$1 P=$2 n=$3 f32: %.2f GDoF/s frac %.3f kernel_ms %.3f | %s' % (d['value'], r['frac'], r['kernel_ms_per_apply'], r['kernel'][:40]))
"
}
for cfg in "6 500000000" "7 300000000" "5 300000000"; do
  set -- $cfg
  one 0 $1 $2
  one 1 $1 $2
done 2>&1 | grep RESULT | tee gpurun_out/pencil_paired_ab.txt
