# DWT tuning sweep (run on the GPU box): parity tests per kernel variant, then level-1 times
for v in "4 0" "4 2" "8 1" "8 3" "4 3" "8 2"; do set -- $v
  echo "TEST ncol=$1 pf=$2"; J2KCUDA_DWT_NCOL=$1 J2KCUDA_DWT_PF=$2 python -m pytest tests/test_gpu_stages.py tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -1
done
for nc in 4 8; do for pf in 0 1 2 3; do
  for items in 7104 14208 28416; do
    echo "VARIANT ncol=$nc pf=$pf items=$items"
    J2KCUDA_DWT_NCOL=$nc J2KCUDA_DWT_PF=$pf J2KCUDA_DWT_ITEMS=$items python tools/stage_times.py dwt 2>&1 | grep -E "mct=" | sed 's/t1ctx.*host_t2=[0-9.]*//'
    J2KCUDA_DWT_NCOL=$nc J2KCUDA_DWT_PF=$pf J2KCUDA_DWT_ITEMS=$items python tools/stage_times.py decode 2>&1 | grep -E "idwt" | head -1 | sed 's/host_parse.*t1=[0-9.]*//'
  done
done; done
