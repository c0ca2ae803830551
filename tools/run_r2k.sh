set -x
N=$1
bash tools/run_r2j.sh $N
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
