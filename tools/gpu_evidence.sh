# round evidence run: GPU tests, compute-sanitizer (memcheck + racecheck) over the one-launch pipeline tests, bench,
# launch list and one `ncu --set full` capture of the bench command (each profiler pass only after its command exited 0)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -q -x \
      -k "single_launch_pipeline_equals_per_frame_oracle or tma_warpaffine_fast_path" > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?" >> gpurun_out/sanitizer_$tool.log; tail -4 gpurun_out/sanitizer_$tool.log
done
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ev.json 2> gpurun_out/bench_ev.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-ops > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pipe_box_warp -c 3 -o gpurun_out/r2_headline python bench.py --steps 2 --warmup 3 --no-ops > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
