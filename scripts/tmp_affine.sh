timeout 900 python -m pytest tests/test_gpu_onthefly.py -x -q 2>&1 | tail -3 | cut -c1-500
BD_TMP_AFFINE=1 TAG=affine1 CFGS="3:1:64:300000000:onthefly 2:1:64:50000000:onthefly 1:1:64:50000000:onthefly 3:1:32:500000000:onthefly" bash scripts/gpu_quick.sh
BD_TMP_AFFINE=0 TAG=affine0 CFGS="3:1:64:300000000:onthefly 3:0:64:300000000:onthefly 4:1:64:300000000:onthefly 6:1:64:500000000:onthefly 3:1:32:500000000:onthefly" bash scripts/gpu_quick.sh
