N=8
python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/r2_multi_n$N.log 2>&1; tail -3 gpurun_out/r2_multi_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --config c5 > gpurun_out/r2_bench_c5_n$N.json 2> gpurun_out/r2_bench_c5_n$N.err
for f in r2_bench_n$N r2_bench_c5_n$N; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1])
    r=d['roofline']
    print('$f value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
    print('step_frac',r['step_frac'],'phase',r['phase_ms'])
    print('exchange',r.get('exchange'))
    print('parity',d['parity']['ok'], d['parity']['max_err_over_tol'], d['parity']['exchange'])
    e=d['eval'] or {}; print('eval',e.get('value'),e.get('sharding')); print('item_sharded',e.get('item_sharded'))
    print('strong',d['fixed_global_work'])
except Exception as ex:
    print('$f FAILED', repr(ex))
PY
done
