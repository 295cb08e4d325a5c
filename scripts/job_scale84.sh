set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02z_bench_n8.json 2> gpurun_out/r02z_bench_n8.err; echo rc8=$?
timeout 600 $TR --nproc-per-node 4 --master-port 29512 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02z_bench_n4.json 2> gpurun_out/r02z_bench_n4.err; echo rc4=$?
grep -h '^{' gpurun_out/r02z_bench_n8.json gpurun_out/r02z_bench_n4.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['n_gpus'], d['ms_per_step'], d['value'], d.get('state_hash'), d['partition']['collective_ms'], d['partition']['pack_ms'], d['flood']['ms_per_step'])
"
