set -x
tools/ubench/ubench_perm2 > gpurun_out/r2e_ubench_perm2.txt 2>&1; cat gpurun_out/r2e_ubench_perm2.txt
for v in coop8192 coop4096; do
  NLC_COMMIT_LIB=$PWD/tools/ubench/libnlc_$v.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_bench_$v.log 2>&1
  python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2e_bench_$v.log") if l.startswith("{")][-1])
print("$v", d["ms_per_step"], d["phases_ms"])
PY
done
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2e_bench.log 2>&1
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2e_bench.log") if l.startswith("{")][-1])
print("coop2048", d["ms_per_step"], d["phases_ms"]); print(json.dumps(d["e2e_materialised"]))
PY
python -m pytest tests/test_gpu_merkle.py tests/test_gpu_commit.py tests/test_gpu_fri.py -m gpu -x -q 2>&1 | tail -3
