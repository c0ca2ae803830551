set -x
python -m pytest tests/test_gpu_plonk.py tests/test_gpu_air_eval.py -m gpu -x -q 2>&1 | tail -15
python bench.py --workload plonk_proof_demo --plonk-degree-bits 12 --steps 2 2>&1 | tail -c 2500
