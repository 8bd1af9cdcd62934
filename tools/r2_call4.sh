cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo.txt 2>&1
timeout 300 python -m pytest tests/test_sharding_nccl_gpu.py -x -q -s > gpurun_out/r2_sharding_nccl.log 2>&1; echo "sharding test rc=$?"; tail -5 gpurun_out/r2_sharding_nccl.log
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err ) 2>&1 | tail -3; echo "bench n2 rc=$?"
tail -8 gpurun_out/r2_bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n2.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","e2e","view_sharded"):
    print(k, json.dumps(d.get(k))[:3000])
PY
