set -x
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log; tail -8 gpurun_out/r2c_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2c_bench.log 2>&1; tail -c 6000 gpurun_out/r2c_bench.log
