set -x; mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/a_gputests.log
python tools/stage_times.py decode > gpurun_out/a_stage_dec.log 2>&1
python tools/stage_times.py > gpurun_out/a_stage_enc.log 2>&1
python bench.py > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/a_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/a_ncu.log 2>&1
tail -3 gpurun_out/a_gputests.log; cat gpurun_out/a_stage_dec.log | tail -12
