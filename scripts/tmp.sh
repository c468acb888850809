for f in 3 4; do BD_TMP_COL=$f TAG=col$f CFGS="7:1:64:300000000:onthefly 7:1:32:300000000:onthefly 5:1:32:300000000:onthefly" bash scripts/gpu_quick.sh; done
BD_TMP_COL=0 TAG=col0 CFGS="7:1:64:300000000:auto 7:1:32:300000000:auto" bash scripts/gpu_quick.sh
