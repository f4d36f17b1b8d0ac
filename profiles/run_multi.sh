N=$1
python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/r2_multi_n$N.log 2>&1; tail -3 gpurun_out/r2_multi_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_bench_n$N.json').read().strip().splitlines()[-1])
r=d['roofline']
print('N=$N value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])
print('step_frac',r['step_frac'],'phase',r['phase_ms'])
print('exchange',r.get('exchange'))
print('parity',d['parity']['ok'], d['parity']['max_err_over_tol'], d['parity']['exchange'])
e=d['eval']; print('eval',e['value'],e['sharding']); print('item_sharded',e.get('item_sharded'))
print('strong',d['fixed_global_work'])
PY
