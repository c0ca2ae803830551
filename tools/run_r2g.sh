set -x
python bench.py --workload stark_proof_fpmul --steps 3 > gpurun_out/r2g_stark_proof.log 2>&1; tail -c 2500 gpurun_out/r2g_stark_proof.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29533 tools/sweep.py multi --min-log 14 --max-log 15 > gpurun_out/r2g_sweep1.log 2>&1; tail -5 gpurun_out/r2g_sweep1.log
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r2g_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2g_launches.csv $B > gpurun_out/r2g_ncu1.log 2>&1
$B > gpurun_out/r2g_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_ntt -s 2 -c 2 -o gpurun_out/r2g_ntt $B > gpurun_out/r2g_ncu2.log 2>&1
$B > gpurun_out/r2g_plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_leaf_hash -c 1 -o gpurun_out/r2g_leaf $B > gpurun_out/r2g_ncu3.log 2>&1
ls -la gpurun_out | tail
