set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/proof_phases.py 16 2>&1 | grep "proof,\|trace commit\|wires commit" 
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2o_bench.log 2>&1
python bench.py --workload stark_proof_fpmul --steps 5 > gpurun_out/r2o_stark.log 2>&1
python bench.py --workload plonk_proof_demo --steps 5 > gpurun_out/r2o_plonk.log 2>&1
python bench.py --workload replay_plonk --steps 5 > gpurun_out/r2o_replay_plonk.log 2>&1
python bench.py --workload replay_stark --steps 5 > gpurun_out/r2o_replay_stark.log 2>&1
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2o_bench.log") if l.startswith("{")][-1])
print(d["ms_per_step"], d["phases_ms"], d["e2e"]["ms_per_step"], d["e2e"]["pageable_host_ms_per_step"], d["e2e_materialised"])
for f in ("stark","plonk","replay_plonk","replay_stark"):
    d=json.loads([l for l in open("gpurun_out/r2o_%s.log"%f) if l.startswith("{")][-1])
    print(f, d["ms_per_step"], d["cpu_baseline"]["value"], d["cpu_baseline"].get("bit_exact", d["cpu_baseline"].get("proof_bit_exact")))
PY
