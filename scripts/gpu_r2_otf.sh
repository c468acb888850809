# Round 2: first run of the on-the-fly geometry kernels: parity, then timing with the register-cap variants.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_onthefly.py -x -q 2>&1 | tail -15 | cut -c1-600 | tee gpurun_out/r2_otf_tests.txt
one() { # label env degree qmode float ndofs geometry
  env $2 timeout 200 python bench.py --degree $3 --qmode $4 --float $5 --ndofs $6 --geometry $7 --no-cpu --no-e2e --no-cg --no-variants --no-q6 --no-parity --no-reference-gpu-kernel --steps 10 --warmup 3 2>gpurun_out/err_$1.txt | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT %-22s P=$3 q=$4 f$5 n=$6 %-12s: %.2f GDoF/s hbm_frac %.3f kernel_ms %.3f fp64_frac %s ynorm %.15e' % ('$1', r['geometry'], d['value'], r['frac'], r['kernel_ms_per_apply'], r.get('fp64_pipe',{}).get('frac'), d['y_norm']))
"
}
{
  one pre        "BD_X=0"          3 1 64 300000000 precomputed
  one otf_cap0   "BD_OTF_MINB=0"   3 1 64 300000000 onthefly
  one otf_cap10  "BD_OTF_MINB=10"  3 1 64 300000000 onthefly
  one otf_cap12  "BD_OTF_MINB=12"  3 1 64 300000000 onthefly
  one otf_cap14  "BD_OTF_MINB=14"  3 1 64 300000000 onthefly
  one pre        "BD_X=0"          6 1 64 500000000 precomputed
  one otf_cap0   "BD_OTF_MINB=0"   6 1 64 500000000 onthefly
  one otf_cap3   "BD_OTF_MINB=3"   6 1 64 500000000 onthefly
  one otf_cap30  "BD_OTF_MINB=30"  6 1 64 500000000 onthefly
  one pre        "BD_X=0"          3 1 32 500000000 precomputed
  one otf        "BD_X=0"          3 1 32 500000000 onthefly
  one pre        "BD_X=0"          6 1 32 500000000 precomputed
  one otf        "BD_X=0"          6 1 32 500000000 onthefly
  one pre        "BD_X=0"          1 1 64 50000000 precomputed
  one otf        "BD_X=0"          1 1 64 50000000 onthefly
  one pre        "BD_X=0"          2 1 64 50000000 precomputed
  one otf        "BD_X=0"          2 1 64 50000000 onthefly
  one pre        "BD_X=0"          3 0 64 300000000 precomputed
  one otf        "BD_X=0"          3 0 64 300000000 onthefly
} 2>&1 | grep RESULT | tee gpurun_out/r2_otf_ab.txt
tail -3 gpurun_out/err_*.txt | tail -30
