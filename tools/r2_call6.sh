cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err ) 2>&1 | tail -3; echo "bench n8 rc=$?"
tail -4 gpurun_out/r2_bench_n8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n8.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","e2e","view_sharded"):
    print(k, json.dumps(d.get(k))[:3500])
PY
( time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err ) 2>&1 | tail -3; echo "bench n4 rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n4.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","e2e","view_sharded"):
    print(k, json.dumps(d.get(k))[:3500])
PY
