# Round 2: on-the-fly kernels, second run: parity of everything new, register-cap A/B over the degrees, general path.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_onthefly.py tests/test_gpu_jacobi.py -x -q 2>&1 | tail -15 | cut -c1-600 | tee gpurun_out/r2_otf2_tests.txt
one() { # label env degree qmode float ndofs geometry [extra]
  env $2 timeout 200 python bench.py --degree $3 --qmode $4 --float $5 --ndofs $6 --geometry $7 $8 --no-cpu --no-e2e --no-cg --no-variants --no-q6 --no-parity --no-reference-gpu-kernel --steps 10 --warmup 3 2>gpurun_out/err_$1.txt | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT %-14s P=$3 q=$4 f$5 n=$6 $8 %-16s: %.2f GDoF/s hbm_frac %.3f kernel_ms %.3f fp64_frac %s ynorm %.15e' % ('$1', r['geometry'], d['value'], r['frac'], r['kernel_ms_per_apply'], r.get('fp64_pipe',{}).get('frac'), d['y_norm']))
"
}
{
  for cfg in "1 1 64 50000000" "2 1 64 50000000" "2 0 64 100000000" "3 0 64 300000000" "4 0 64 300000000" "4 1 64 300000000" "5 0 64 300000000" "3 1 32 500000000" "4 1 32 500000000"; do
    one cap1  "BD_OTF_MINB=0"  $cfg onthefly
    one cap12 "BD_OTF_MINB=12" $cfg onthefly
  done
  one pre   "BD_X=0"          6 1 64 500000000 precomputed
  one cap1  "BD_OTF_MINB=0"   6 1 64 500000000 onthefly
  one cap3  "BD_OTF_MINB=3"   6 1 64 500000000 onthefly
  one cap30 "BD_OTF_MINB=30"  6 1 64 500000000 onthefly
  one pre   "BD_X=0"          5 1 64 300000000 precomputed
  one cap3  "BD_OTF_MINB=3"   5 1 64 300000000 onthefly
  one pre   "BD_X=0"          7 0 64 300000000 precomputed
  one cap3  "BD_OTF_MINB=3"   7 0 64 300000000 onthefly
  one pre   "BD_X=0"          6 1 32 500000000 precomputed
  one otf   "BD_X=0"          6 1 32 500000000 onthefly
  # general (perturbed) meshes: stored G against G rebuilt per point
  one pre   "BD_X=0"          3 1 64 300000000 precomputed "--geom_perturb_fact 0.2"
  one gen1  "BD_OTF_MINB=0"   3 1 64 300000000 onthefly    "--geom_perturb_fact 0.2"
  one gen12 "BD_OTF_MINB=12"  3 1 64 300000000 onthefly    "--geom_perturb_fact 0.2"
  one pre   "BD_X=0"          6 1 64 500000000 precomputed "--geom_perturb_fact 0.2"
  one gen1  "BD_OTF_MINB=0"   6 1 64 500000000 onthefly    "--geom_perturb_fact 0.2"
  one gen3  "BD_OTF_MINB=3"   6 1 64 500000000 onthefly    "--geom_perturb_fact 0.2"
  one pre   "BD_X=0"          3 1 32 500000000 precomputed "--geom_perturb_fact 0.2"
  one gen1  "BD_OTF_MINB=0"   3 1 32 500000000 onthefly    "--geom_perturb_fact 0.2"
} 2>&1 | grep RESULT | tee gpurun_out/r2_otf2_ab.txt
for f in gpurun_out/err_*.txt; do echo $f; tail -n 3 $f | cut -c1-300; done
