cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err ) 2>&1 | tail -3; echo "bench rc=$?"
tail -5 gpurun_out/r2_bench_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n1.json').read().strip().splitlines()[-1])
for k in ("value","ms_per_step","roofline","roofline_backproject","stages","e2e","neck_fp32","gpu_reference","view_sharded","cpu_baseline","clocks"):
    print(k, json.dumps(d.get(k))[:1500])
PY
( time timeout 600 python bench.py --steps 10 --warmup 3 --neck-precision fp32 --no-extras > gpurun_out/r2_bench_n1_fp32.json 2> gpurun_out/r2_bench_n1_fp32.err ) 2>&1 | tail -3
tail -3 gpurun_out/r2_bench_n1_fp32.err; python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_n1_fp32.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], json.dumps(d['roofline']), json.dumps(d['e2e'])[:400])"
