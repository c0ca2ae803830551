set -x
NLC_COMMIT_LIB=$PWD/tools/ubench/libnlc_nttminb3.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2l_bench_minb3.log 2>&1
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2l_bench.log 2>&1
python - <<PY
import json
for f in ("r2l_bench_minb3","r2l_bench"):
    d=json.loads([l for l in open("gpurun_out/%s.log"%f) if l.startswith("{")][-1])
    print(f, d["ms_per_step"], d["phases_ms"])
PY
python bench.py --workload plonk_proof_demo --steps 3 > gpurun_out/r2l_plonk_proof.log 2>&1; tail -c 1800 gpurun_out/r2l_plonk_proof.log
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r2l_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2l_launches.csv $B > gpurun_out/r2l_ncu1.log 2>&1
$B > gpurun_out/r2l_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_ntt -c 4 -o gpurun_out/r2l_ntt $B > gpurun_out/r2l_ncu2.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29533 tools/sweep.py multi > gpurun_out/r2l_sweep_1gpu.jsonl 2> gpurun_out/r2l_sweep_1gpu.err; tail -2 gpurun_out/r2l_sweep_1gpu.jsonl | cut -c1-300
