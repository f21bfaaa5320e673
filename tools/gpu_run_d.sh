# bench sweeps: decoder register variants and frames in flight
mkdir -p gpurun_out
for mb in 4 5 6; do for w in 8 16; do
  echo "MINB $mb WORKERS $w"
  J2KCUDA_DEC_MINB=$mb python bench.py --no-cpu-baseline --steps 3 --warmup 3 --workers $w --frames 16 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(' value %.1f e2e %.1f enc %.1f dec %.1f' % (d['value'], d['e2e']['value'], d['encode_mpix_s'], d['decode_mpix_s']))"
done; done > gpurun_out/d_sweep.log 2>&1
cat gpurun_out/d_sweep.log
