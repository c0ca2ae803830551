set -x
python -m pytest tests/test_gpu_plonk.py -m gpu -x -q 2>&1 | tail -4
python bench.py --workload plonk_proof_demo --steps 3 > gpurun_out/r2q_plonk.log 2>&1; tail -c 1500 gpurun_out/r2q_plonk.log
python tools/proof_phases.py 16 2>&1 | head -9
