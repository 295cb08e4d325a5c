# one GPU: the default bench line, then the ncu launch list and one --set full capture per dominant kernel (numbers printed under ncu are not bench values)
set -x
timeout 900 python bench.py > gpurun_out/r02n_bench_n1.json 2> gpurun_out/r02n_bench_n1.err; echo bench rc=$?
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --flood-steps 0 --members 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02n_launches.csv $B > gpurun_out/r02n_ncu0.log 2>&1; echo ncu0 rc=$?
for K in Purchase SellerProduction StorageDelivery; do
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:$K -s 3 -c 1 -f -o gpurun_out/r02n_$K $B > gpurun_out/r02n_ncu_$K.log 2>&1; echo ncu $K rc=$?
done
ls -la gpurun_out/*.ncu-rep
