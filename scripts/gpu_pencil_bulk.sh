# First run of the experimental bulk-copy pencil kernel (BD_PENCIL_BULK=1): parity under a timeout, then A/B
# at Q6 fp32 against the shipped pencil kernel, with and without FFMA2.
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
BD_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_pencil_bulk.py -x -q 2>&1 | tail -8 | cut -c1-600 || exit 1
one() { # bulk paired ndofs
  BD_PENCIL_BULK=$1 BD_PENCIL_PAIRED=$2 timeout 150 python bench.py --degree 6 --ndofs $3 --float 32 --qmode 1 --no-cpu --no-e2e --no-cg --steps 10 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT bulk=$1 paired=$2 n=$3 Q6 f32: %.2f GDoF/s frac %.3f kernel_ms %.3f' % (d['value'], r['frac'], r['kernel_ms_per_apply']))
"
}
for n in 500000000 1000000000; do
  one 0 0 $n
  one 1 0 $n
  one 1 1 $n
done 2>&1 | grep RESULT | tee gpurun_out/pencil_bulk_ab.txt
