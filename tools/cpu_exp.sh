mkdir -p gpurun_out; : > gpurun_out/cpuexp.log
nproc >> gpurun_out/cpuexp.log
run() { echo "== $*" >> gpurun_out/cpuexp.log; "$@" 2>/dev/null | python tools/bench_summary.py /dev/stdin 2>/dev/null | head -5 | cut -c1-330 >> gpurun_out/cpuexp.log; }
B="python bench.py --steps 3 --warmup 2 --no-cpu-baseline"
run $B --workers 16
run $B --workers 6
run taskset -c 0-3 $B --workers 6
run taskset -c 0-3 $B --workers 16
run taskset -c 0-1 $B --workers 16
run $B --workers 16 --frames 32
cat gpurun_out/cpuexp.log
