set -x
python -m pytest tests/test_gpu_plonk.py tests/test_gpu_air_eval.py tests/test_gpu_cpp_mirror.py -m gpu -x -q 2>&1 | tail -4
python tools/proof_phases.py 16 2>&1 | head -9
python bench.py --workload plonk_proof_demo --steps 3 > gpurun_out/r2r_plonk.log 2>&1; tail -c 900 gpurun_out/r2r_plonk.log
