#!/bin/bash
# tools/gpu_validate.sh -- what a round's final check runs on a B200 box (under gpurun): the GPU test suite, smoke(),
# the contract bench line, the launch list and the two `ncu --set full` captures that profiles/traffic.json is rebuilt
# from (tools/update_traffic.py), and the whole-proof workloads.  Outputs under gpurun_out/<tag>_*.
set -x
TAG=${1:-val}
python -m pytest tests -m gpu -x -q -rs > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log; tail -6 gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()"
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.log 2>&1; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/${TAG}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_ncu1.log 2>&1
$B > gpurun_out/${TAG}_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_ntt -c 4 -o gpurun_out/${TAG}_ntt $B > gpurun_out/${TAG}_ncu2.log 2>&1
$B > gpurun_out/${TAG}_plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_leaf_hash -c 1 -o gpurun_out/${TAG}_leaf $B > gpurun_out/${TAG}_ncu3.log 2>&1
python bench.py --workload stark_proof_fpmul --steps 5 > gpurun_out/${TAG}_stark.log 2>&1
python bench.py --workload stark_proof_ed_add --steps 3 > gpurun_out/${TAG}_stark_ed_add.log 2>&1
python bench.py --workload stark_proof_ed_scalar_mul --steps 3 > gpurun_out/${TAG}_stark_ed_scalar_mul.log 2>&1
python bench.py --workload plonk_proof_demo --steps 5 > gpurun_out/${TAG}_plonk.log 2>&1
python bench.py --workload commit_2^20x400_r1_cap4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_config4.log 2>&1
