set -x
N=$1
nvidia-smi -L | wc -l
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tools/sweep.py multi --steps 3 --warmup 2 > gpurun_out/r2j_sweep_${N}gpu.jsonl 2> gpurun_out/r2j_sweep_${N}gpu.err; echo "sweep rc=$?"; tail -3 gpurun_out/r2j_sweep_${N}gpu.jsonl | cut -c1-600; tail -5 gpurun_out/r2j_sweep_${N}gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2j_bench_${N}gpu.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/r2j_bench_${N}gpu.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])
for k,v in d.get('strong',{}).items(): print(k, v['ms_per_commit'], v['speedup_vs_1gpu'], v['e2e_ms'], v['phases_ms_max_over_ranks'], v['cap_equals_1gpu_commit_on_every_rank'])
print(d.get('sharded_parity'))
"
