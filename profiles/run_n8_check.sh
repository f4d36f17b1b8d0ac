N=8
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/r2_multi_n$N.log 2>&1; tail -2 gpurun_out/r2_multi_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_bench_n8.json').read().strip().splitlines()[-1])
print('N=8 value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'parity',d['parity']['ok'],'eval',d['eval']['value'], 'item_sharded', d['eval']['item_sharded']['value'])
PY
