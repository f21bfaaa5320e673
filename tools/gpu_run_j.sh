# bench sweep: MQ lanes per warp / long-block count with 16 frames in flight
mkdir -p gpurun_out
for nl in 0 592; do for l in 4 8 16 32; do
  echo "NLONG $nl LANES $l"
  J2KCUDA_MQ_NLONG=$nl J2KCUDA_MQ_LANES=$l python bench.py --no-cpu-baseline --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(' value %.1f e2e %.1f enc %.1f dec %.1f pipe %.1f' % (d['value'], d['e2e']['value'], d['encode_mpix_s'], d['decode_mpix_s'], d['e2e_pipelined']['value']))"
done; done > gpurun_out/j_sweep.log 2>&1
cat gpurun_out/j_sweep.log
