# DWT tuning sweep (run on the GPU box): parity tests per strip geometry, then level-1 times
for v in "24 8" "28 4" "16 8"; do set -- $v
  echo "TEST owners=$1 gran=$2"; J2KCUDA_DWT_OWNERS=$1 J2KCUDA_DWT_GRAN=$2 python -m pytest tests/test_gpu_stages.py tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -1
done
for v in "30 1" "28 4" "24 8" "16 8" "30 2"; do set -- $v
  for items in 14208 28416; do
    echo "VARIANT owners=$1 gran=$2 items=$items"
    J2KCUDA_DWT_OWNERS=$1 J2KCUDA_DWT_GRAN=$2 J2KCUDA_DWT_ITEMS=$items python tools/stage_times.py dwt 2>&1 | grep -E "mct=" | sed 's/t1ctx.*host_t2=[0-9.]*//'
    J2KCUDA_DWT_OWNERS=$1 J2KCUDA_DWT_GRAN=$2 J2KCUDA_DWT_ITEMS=$items python tools/stage_times.py decode 2>&1 | grep -E "idwt" | head -1 | sed 's/host_parse.*t1=[0-9.]*//'
  done
done
