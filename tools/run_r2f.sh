set -x
python -m pytest tests/test_gpu_commit.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_bench.log 2>&1
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2f_bench.log") if l.startswith("{")][-1])
print(d["ms_per_step"], d["phases_ms"]); print(json.dumps(d["e2e_materialised"]))
PY
