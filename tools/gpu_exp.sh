#!/bin/bash
# One parametrised GPU experiment script (run through gpurun): tools/gpu_exp.sh <what> ...; results under gpurun_out/.
mkdir -p gpurun_out
what=${1:-tests}
case $what in
tests)   python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/tests.log ;;
decsweep)
  : > gpurun_out/decsweep.log
  [ -z "$NOV1" ] && J2KCUDA_DEC_IMPL=0 python tools/dec_bench.py --tag v1 >> gpurun_out/decsweep.log 2>&1
  for cfg in ${CFGS:-4_6 8_6 16_8 32_8}; do
    a=${cfg%_*}; b=${cfg#*_}
    J2KCUDA_DEC_LPB=$a J2KCUDA_DEC_WARPS=$b timeout 300 python tools/dec_bench.py --tag multi >> gpurun_out/decsweep.log 2>&1
  done
  cat gpurun_out/decsweep.log ;;
decprof)   # ncu full capture of the Tier-1 decode kernel of one 4K frame: tools/gpu_exp.sh decprof <name> [env assignments...]
  name=${2:-dec}; shift 2
  env "$@" python tools/dec_bench.py --frames 1 --alone-only --tag $name > gpurun_out/${name}_plain.log 2>&1 &&
  env "$@" ncu --set full --clock-control none --import-source on -k regex:k_t1_decode -s 2 -c 1 -f -o gpurun_out/$name \
      python tools/dec_bench.py --frames 1 --alone-only --tag $name > gpurun_out/${name}_ncu.log 2>&1
  python tools/ncu_summary.py gpurun_out/$name.ncu-rep > gpurun_out/${name}_summary.txt 2>&1
  python tools/ncu_lines.py gpurun_out/$name.ncu-rep k_t1_decode 70 > gpurun_out/${name}_lines.txt 2>&1
  cat gpurun_out/${name}_plain.log; tail -3 gpurun_out/${name}_ncu.log; head -60 gpurun_out/${name}_lines.txt ;;
dwtsweep)  # per-level tunables of the DWT kernels (J2KCUDA_DWT_*), forward and inverse totals of one 4K RGBA16 frame
  : > gpurun_out/dwtsweep.log
  for cfg in "" "J2KCUDA_DWT_MINSP=4" "J2KCUDA_DWT_MINSP=2" "J2KCUDA_DWT_MINSP=4 J2KCUDA_DWT_FLOOR=4736" "J2KCUDA_DWT_MINSP=4 J2KCUDA_DWT_FLOOR=9472" \
             "J2KCUDA_DWT_PF=2" "J2KCUDA_DWT_PF=3" "J2KCUDA_DWT_PF_INV=3" "J2KCUDA_DWT_PF_INV=1" "J2KCUDA_DWT_NCOL=8" "J2KCUDA_DWT_NCOL=8 J2KCUDA_DWT_PF=2" \
             "J2KCUDA_DWT_ITEMS=9472" "J2KCUDA_DWT_ITEMS=28416" "J2KCUDA_DWT_MINSP=4 J2KCUDA_DWT_PF=2 J2KCUDA_DWT_PF_INV=3"; do
    echo "== $cfg" >> gpurun_out/dwtsweep.log
    env $cfg python tools/stage_times.py dwt 2>&1 | grep -E "mct=|dwt_l1" >> gpurun_out/dwtsweep.log
    env $cfg python tools/stage_times.py decode1 2>&1 | grep -E "idwt" >> gpurun_out/dwtsweep.log
  done
  cat gpurun_out/dwtsweep.log ;;
bench)     # bench.py on this box: tools/gpu_exp.sh bench <name> [bench.py args...]
  name=${2:-bench}; shift 2
  python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; tail -2 gpurun_out/$name.err
  python tools/bench_summary.py gpurun_out/$name.json ;;
cxprof)    # ncu full capture of k_t1_context of one 4K frame: tools/gpu_exp.sh cxprof <name> [env assignments...]
  name=${2:-cx}; shift 2
  env "$@" ncu --set full --clock-control none --import-source on -k regex:k_t1_context -s 2 -c 1 -f -o gpurun_out/$name \
      python tools/stage_times.py dwt > gpurun_out/${name}_ncu.log 2>&1
  python tools/ncu_summary.py gpurun_out/$name.ncu-rep > gpurun_out/${name}_summary.txt 2>&1
  python tools/ncu_lines.py gpurun_out/$name.ncu-rep k_t1_context 60 > gpurun_out/${name}_lines.txt 2>&1
  cat gpurun_out/${name}_summary.txt; head -40 gpurun_out/${name}_lines.txt ;;
cxsweep)   # resident CTAs of k_t1_context (J2KCUDA_CX_MINB): per-stage times alone and the batch throughput
  : > gpurun_out/cxsweep.log
  for mb in 4 5 6; do
    echo "== J2KCUDA_CX_MINB=$mb" >> gpurun_out/cxsweep.log
    J2KCUDA_CX_MINB=$mb python tools/stage_times.py dwt 2>&1 | grep -E "t1ctx" >> gpurun_out/cxsweep.log
    J2KCUDA_CX_MINB=$mb python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python tools/bench_summary.py /dev/stdin | head -1 >> gpurun_out/cxsweep.log
  done
  cat gpurun_out/cxsweep.log ;;
cpuexp)    # host cores against frames in flight (profiles/r2_cpuexp.log): taskset to 4 and 2 cores, 6 and 16 workers, 32 frames per step
  : > gpurun_out/cpuexp.log; nproc >> gpurun_out/cpuexp.log
  run() { echo "== $*" >> gpurun_out/cpuexp.log; "$@" 2>/dev/null | python tools/bench_summary.py /dev/stdin 2>/dev/null | head -5 | cut -c1-330 >> gpurun_out/cpuexp.log; }
  B="python bench.py --steps 3 --warmup 2 --no-cpu-baseline --frames 16"
  run $B --workers 16; run $B --workers 6; run taskset -c 0-3 $B --workers 6; run taskset -c 0-3 $B --workers 16
  run taskset -c 0-1 $B --workers 16; run python bench.py --steps 3 --warmup 2 --no-cpu-baseline --workers 16 --frames 32
  cat gpurun_out/cpuexp.log ;;
mqexp)     # one-block warps of the MQ encoder (J2KCUDA_MQ_NLONG, profiles/r2_mqexp.log)
  : > gpurun_out/mqexp.log
  for nl in ${NLONGS:-0 64}; do
    echo "== J2KCUDA_MQ_NLONG=$nl" >> gpurun_out/mqexp.log
    J2KCUDA_MQ_NLONG=$nl python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python tools/bench_summary.py /dev/stdin 2>/dev/null | head -4 | cut -c1-300 >> gpurun_out/mqexp.log
  done
  cat gpurun_out/mqexp.log ;;
check)     # the two things the round's end runs: GPU tests and the bench line: tools/gpu_exp.sh check <tag>
  tag=${2:-check}
  python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/${tag}_gpu_tests.log
  python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
  cat gpurun_out/${tag}_gpu_tests.log; python tools/bench_summary.py gpurun_out/${tag}_bench.json ;;
esac
