timeout 900 python -m pytest tests/test_gpu_onthefly.py -x -q 2>&1 | tail -5 | cut -c1-500
for f in 0 1; do BD_TMP_AFFINE=$f TAG=affine$f CFGS="3:1:64:300000000:onthefly 3:0:64:300000000:onthefly 2:1:64:50000000:onthefly 1:1:64:50000000:onthefly 3:1:32:500000000:onthefly 2:1:32:100000000:onthefly" bash scripts/gpu_quick.sh; done
