# A/B of the FFMA2 slab kernel (BD_SLAB_F2=1) against the default fp32 slab kernel: Q3, Q2, Q1 qmode 1.
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
timeout 200 python -m pytest tests/test_gpu_slab_f2.py -q 2>&1 | tail -3
one() { # f2 degree ndofs
  BD_SLAB_F2=$1 timeout 150 python bench.py --degree $2 --ndofs $3 --float 32 --qmode 1 --no-cpu --no-e2e --no-cg --steps 10 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT f2=$1 P=$2 n=$3 f32: %.2f GDoF/s frac %.3f kernel_ms %.3f' % (d['value'], r['frac'], r['kernel_ms_per_apply']))
"
}
for cfg in "3 500000000" "3 1000000000" "2 100000000" "1 100000000"; do
  set -- $cfg
  one 0 $1 $2
  one 1 $1 $2
done 2>&1 | grep RESULT | tee gpurun_out/slab_f2_ab.txt
