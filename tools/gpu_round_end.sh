#!/bin/bash
# Round-end evidence on one B200 (run through gpurun): GPU tests, smoke, both bench arms, ncu launch list; optionally (FULL=1) a
# full capture of one frame.  Everything lands under gpurun_out/ (kept below the 64 MiB that travel back).
mkdir -p gpurun_out
tag=${1:-r2}
python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/${tag}_gpu_tests.log
if ! grep -q " passed" gpurun_out/${tag}_gpu_tests.log || grep -q "failed" gpurun_out/${tag}_gpu_tests.log; then
  J2KCUDA_SPEC_D2H=0 python -m pytest tests -m gpu -q -x 2>&1 | tail -8 > gpurun_out/${tag}_gpu_tests_nospec.log
fi
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${tag}_smoke.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
if [ $? -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${tag}_launches.csv \
      python bench.py --steps 1 --warmup 1 --frames 16 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1
  if [ -n "$FULL" ]; then
    timeout 900 ncu --set full --clock-control none --import-source on -f -o gpurun_out/${tag}_frame_full \
        python tools/stage_times.py roundtrip1 > gpurun_out/${tag}_ncu_full.log 2>&1
    ncu -i gpurun_out/${tag}_frame_full.ncu-rep --page raw --csv > gpurun_out/${tag}_frame_full_raw.csv 2>/dev/null
    [ $(stat -c %s gpurun_out/${tag}_frame_full.ncu-rep) -gt 40000000 ] && rm -f gpurun_out/${tag}_frame_full.ncu-rep
  fi
fi
cat gpurun_out/${tag}_gpu_tests.log gpurun_out/${tag}_gpu_tests_nospec.log gpurun_out/${tag}_smoke.log 2>/dev/null
python tools/bench_summary.py gpurun_out/${tag}_bench.json
cat gpurun_out/${tag}_bench_reference.json
