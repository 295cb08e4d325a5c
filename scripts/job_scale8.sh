set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02w_bench_n8.json 2> gpurun_out/r02w_bench_n8.err; echo rc8=$?
timeout 600 $TR --nproc-per-node 4 --master-port 29512 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02w_bench_n4.json 2> gpurun_out/r02w_bench_n4.err; echo rc4=$?
timeout 600 $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02w_bench_n2.json 2> gpurun_out/r02w_bench_n2.err; echo rc2=$?
timeout 900 $TR --nproc-per-node 8 --master-port 29514 bench.py --gpus 8 --steps 10 --warmup 3 --config x4 --flood-steps 0 > gpurun_out/r02w_bench_x4_n8.json 2> gpurun_out/r02w_bench_x4_n8.err; echo rcx4=$?
grep -h '^{' gpurun_out/r02w_bench_n8.json gpurun_out/r02w_bench_n4.json gpurun_out/r02w_bench_n2.json gpurun_out/r02w_bench_x4_n8.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['n_gpus'], d['config']['workload'][:30], d['ms_per_step'], d['value'], d.get('state_hash'), d.get('partition', {}).get('collective_ms'), d.get('partition', {}).get('pack_ms'))
"
