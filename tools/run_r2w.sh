set -x
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2w_pytest.log; tail -6 gpurun_out/r2w_pytest.log
python -c "import __graft_entry__ as g; g.smoke()"
python bench.py --steps 20 --warmup 3 > gpurun_out/r2w_bench.log 2>&1; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r2w_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2w_launches.csv $B > gpurun_out/r2w_ncu1.log 2>&1
$B > gpurun_out/r2w_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_ntt -c 4 -o gpurun_out/r2w_ntt $B > gpurun_out/r2w_ncu2.log 2>&1
$B > gpurun_out/r2w_plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_leaf_hash -c 1 -o gpurun_out/r2w_leaf $B > gpurun_out/r2w_ncu3.log 2>&1
python bench.py --workload stark_proof_fpmul --steps 5 > gpurun_out/r2w_stark.log 2>&1
python bench.py --workload plonk_proof_demo --steps 5 > gpurun_out/r2w_plonk.log 2>&1
python bench.py --workload commit_2^20x400_r1_cap4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2w_config4.log 2>&1
ls gpurun_out | wc -l
