cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for n in 8 4 2; do
( time timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_bench_n${n}d.json 2> gpurun_out/r2_bench_n${n}d.err ) 2>&1 | grep real; echo "bench n$n rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_bench_n${n}d.json').read().strip().splitlines()[-1])
print('headline ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), 'ceiling', round(d['e2e']['copy_ceiling']['ms_per_step'],3))
for r in d['view_sharded']:
    print(' ', r['workload'][10:], r['mode'][:10], r['scenes_per_step'], 'ms', round(r['ms_per_step'],3), {k:round(v,3) for k,v in r['stages'].items()}, 'pipelined', round(r.get('pipelined',{}).get('ms_per_step',0),3), r.get('self_check',{}).get('within_1e-5'), 'scenes/s', round(r['scenes_per_s']), 'pip', round(r.get('pipelined',{}).get('scenes_per_s',0)))
PY
done
