# 2-GPU development check: dist_check (parity of the sharded schedule) + a traced and an untraced bench run
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/dist_check.py > gpurun_out/n2_distcheck.log 2>&1; echo "dist_check n2 rc=$?"; tail -3 gpurun_out/n2_distcheck.log
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 2 > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/n2_bench.json'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','verified','comm_per_step')}); print(d.get('e2e')); print(d.get('phases_ms'))
PY
PSC_TRACE=1 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 1 --warmup 2 2>&1 | grep "psc trace" | tail -40 > gpurun_out/n2_trace.txt; tail -40 gpurun_out/n2_trace.txt | cut -c1-160
