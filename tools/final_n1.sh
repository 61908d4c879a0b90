# single-GPU round-end evidence: the bench line, one ncu --set full capture of the walk kernel, the launch list
python bench.py --steps 3 --warmup 3 > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err; echo "bench rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:lf_walk_kernel -c 1 -f -o gpurun_out/r2n_lf_walk python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-single-block > gpurun_out/r2n_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/r2n_launches.csv python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-single-block > gpurun_out/r2n_ncu_list.log 2>&1; echo "ncu list rc=$?"
ls -la gpurun_out/ | tail -8
