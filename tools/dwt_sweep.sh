# DWT tuning sweep (run on the GPU box): segment-length policy of the small levels
for v in "2 1" "8 2368" "8 1184" "16 2368" "4 2368" "8 4736"; do set -- $v
  echo "VARIANT minsp=$1 floor=$2"
  J2KCUDA_DWT_MINSP=$1 J2KCUDA_DWT_FLOOR=$2 python tools/stage_times.py dwt 2>&1 | grep -E "mct=" | sed 's/t1ctx.*host_t2=[0-9.]*//'
  J2KCUDA_DWT_MINSP=$1 J2KCUDA_DWT_FLOOR=$2 python tools/stage_times.py decode 2>&1 | grep -E "idwt" | head -1 | sed 's/host_parse.*t1=[0-9.]*//'
done
