# one GPU, end of round: smoke, the GPU tests, the default bench line, the ensemble with 16 members, the reference arm, ncu launch list + captures
set -x
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02x_pytest.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/r02x_pytest.log
timeout 900 python bench.py > gpurun_out/r02x_bench_n1.json 2> gpurun_out/r02x_bench_n1.err; echo bench rc=$?
timeout 600 python bench.py --members 16 --no-cpu-baseline --steps 5 > gpurun_out/r02x_bench_m16.json 2> gpurun_out/r02x_bench_m16.err; echo bench16 rc=$?
timeout 900 python bench.py --impl reference --steps 10 --warmup 1 > gpurun_out/r02x_bench_reference.json 2> gpurun_out/r02x_bench_reference.err; echo ref rc=$?
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --flood-steps 0 --members 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02x_launches.csv $B > gpurun_out/r02x_ncu0.log 2>&1; echo ncu0 rc=$?
for K in Purchase SellerProduction StorageDelivery; do
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:$K -s 3 -c 1 -f -o gpurun_out/r02x_$K $B > gpurun_out/r02x_ncu_$K.log 2>&1; echo ncu $K rc=$?
done
python - <<'PY'
import json
for f in ("r02x_bench_n1", "r02x_bench_m16", "r02x_bench_reference"):
    for l in open("gpurun_out/%s.json" % f):
        if l.startswith("{"):
            d = json.loads(l)
            print(f, d.get("ms_per_step"), d.get("value"), d.get("state_hash"), (d.get("flood") or {}).get("ms_per_step"), d.get("ensemble"), d.get("e2e_outputs"))
PY
