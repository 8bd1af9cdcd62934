cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2b.json 2> gpurun_out/r2_bench_n2b.err ) 2>&1 | tail -3; echo "bench n2 rc=$?"
tail -8 gpurun_out/r2_bench_n2b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n2b.json').read().strip().splitlines()[-1])
for r in d['view_sharded']:
    print(r['workload'], r['mode'], r['scenes_per_step'], 'ms', r['ms_per_step'], r['stages'], 'pipelined', r.get('pipelined',{}).get('ms_per_step'), r['timing'], r['scenes_per_s'])
PY
