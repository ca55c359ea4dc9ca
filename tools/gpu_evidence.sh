# round evidence run: GPU tests, bench,
# launch list and one `ncu --set full` capture of the bench command (each profiler pass only after its command exited 0)
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
# (compute-sanitizer is closed on this pool; the pipeline's protocol is covered by the soak test in tests/test_gpu_parity.py)
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ev.json 2> gpurun_out/bench_ev.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-ops > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pipe_box_warp -c 3 -o gpurun_out/r2_headline python bench.py --steps 2 --warmup 3 --no-ops > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
python tools/exp_pyr.py > gpurun_out/exp_pyr.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:pyrdown -c 2 -o gpurun_out/r2_pyrdown python tools/exp_pyr.py > gpurun_out/ncu_pyr.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pyrup -c 2 -o gpurun_out/r2_pyrup python tools/exp_pyr.py > gpurun_out/ncu_pyrup.log 2>&1
