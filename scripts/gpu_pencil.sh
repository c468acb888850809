# Parity of the pencil kernel, then A/B against the default kernels (action only).
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_pencil.py -q 2>&1 | tail -25 | cut -c1-600
one() { # variant cpb ring degree ndofs float
  BD_STIFFNESS_VARIANT=$1 BD_PENCIL_CPB=$2 BD_PENCIL_RING=$3 timeout 150 python bench.py --degree $4 --ndofs $5 --float $6 --qmode 1 --no-cpu --no-e2e --no-cg --steps 8 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT variant=$1 cpb=$2 ring=$3 P=$4 n=$5 f=$6: %.2f GDoF/s frac %.3f kernel_ms %.3f | %s' % (d['value'], r['frac'], r['kernel_ms_per_apply'], r['kernel'][:40]))
"
}
for cfg in "7 200000000 64" "6 300000000 32" "4 200000000 64" "5 200000000 64" "7 300000000 32" "5 300000000 32"; do
  set -- $cfg
  one default 0 0 $1 $2 $3
  one pencil 0 0 $1 $2 $3
  one pencil 1 0 $1 $2 $3
  one pencil 2 0 $1 $2 $3
  one pencil 0 1 $1 $2 $3
done 2>&1 | grep RESULT | tee gpurun_out/pencil_ab.txt
