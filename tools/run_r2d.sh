# 2 GPUs: multi-GPU parity test, then the bench line with the strong-scaling + sharded-parity blocks
set -x
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_commit.py -m gpu -x -q -rs -k "sharded or materialised" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log; tail -5 gpurun_out/r2d_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2d_bench2.log 2>&1; echo "bench rc=$?" >> gpurun_out/r2d_bench2.log
tail -c 5000 gpurun_out/r2d_bench2.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2d_bench1.log 2>&1; tail -c 1500 gpurun_out/r2d_bench1.log
