mkdir -p gpurun_out
timeout 900 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"
tail -c 3500 gpurun_out/bench_c5.json; tail -5 gpurun_out/bench_c5.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'K_BasinAccum|K_CclUnion|K_K3Union' -c 6 -o gpurun_out/top_c3 -f python bench.py --workload c3 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_c3.log 2>&1; echo "ncu full rc=$?"
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
