set -x
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log; tail -8 gpurun_out/r2b_pytest.log
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r2b_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2b_launches.csv $B > gpurun_out/r2b_ncu1.log 2>&1
$B > gpurun_out/r2b_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_ntt -s 2 -c 2 -o gpurun_out/r2b_ntt $B > gpurun_out/r2b_ncu2.log 2>&1
$B > gpurun_out/r2b_plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_leaf_hash -c 1 -o gpurun_out/r2b_leaf $B > gpurun_out/r2b_ncu3.log 2>&1
ls -la gpurun_out | tail
