set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench.log 2>&1
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2h_bench.log") if l.startswith("{")][-1])
print(d["ms_per_step"], d["phases_ms"], d["gpu_launches"], d["e2e"]["ms_per_step"], d["e2e"]["pageable_host_ms_per_step"])
PY
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload commit_2^16x20_r3_cap4 > gpurun_out/r2h_bench20.log 2>&1
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload commit_2^20x400_r1_cap4 > gpurun_out/r2h_bench_c4.log 2>&1
python - <<PY
import json
for f in ("r2h_bench20","r2h_bench_c4"):
    d=json.loads([l for l in open("gpurun_out/%s.log"%f) if l.startswith("{")][-1])
    print(f, d["ms_per_step"], d["phases_ms"], d["e2e"]["ms_per_step"])
PY
