# 8-GPU evidence: host-feed control, the multi-GPU tests, bench at N=8 (short)
set -x
for n in 1 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n tools/exp_h2d_multi.py 2>&1 | grep "GPU(s)"
done > gpurun_out/h2d_control.log 2>&1
python -m pytest tests -m gpu -q -k "two_gpus or not_current or all_visible" 2>&1 | tail -3 > gpurun_out/pytest_multi.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err
cat gpurun_out/h2d_control.log gpurun_out/pytest_multi.log; tail -c 400 gpurun_out/bench_8gpu.err
