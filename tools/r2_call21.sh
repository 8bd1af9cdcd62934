cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_sharding_nccl_gpu.py -x -q -s > gpurun_out/r2_sharding_nccl.log 2>&1; echo "sharding test rc=$?"; tail -3 gpurun_out/r2_sharding_nccl.log
( time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2c.json 2> gpurun_out/r2_bench_n2c.err ) 2>&1 | tail -3; echo "bench n2 rc=$?"
tail -3 gpurun_out/r2_bench_n2c.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n2c.json').read().strip().splitlines()[-1])
for r in d['view_sharded']:
    print(r['workload'], r['mode'][:12], r['scenes_per_step'], 'ms', round(r['ms_per_step'],3), {k:round(v,3) for k,v in r['stages'].items()}, 'pipelined', round(r.get('pipelined',{}).get('ms_per_step',0),3), r.get('self_check',{}).get('within_1e-5'))
PY
