timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/r2_bench_v4.json 2> gpurun_out/bench_v4.err; tail -c 300 gpurun_out/r2_bench_v4.json
timeout 300 python bench.py --steps 1 --warmup 1 --playouts 100 --no-cpu-baseline --generation-plies 0 > /dev/null 2>&1 && \
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -s 2400 -c 400 --csv --log-file gpurun_out/r2_launches_v4.csv python bench.py --steps 1 --warmup 1 --playouts 100 --no-cpu-baseline --generation-plies 0 > gpurun_out/ncu_l.log 2>&1
tail -2 gpurun_out/ncu_l.log | cut -c1-200
