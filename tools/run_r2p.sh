set -x
python -m pytest tests/test_gpu_air_eval.py tests/test_gpu_plonk.py tests/test_gpu_cpp_mirror.py tests/test_gpu_quotient.py -m gpu -x -q 2>&1 | tail -4
python tools/proof_phases.py 16 2>&1 | tail -18
