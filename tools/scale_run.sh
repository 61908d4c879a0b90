# usage: bash tools/scale_run.sh N [check]   -- bench line at N GPUs (+ dist_check parity run when "check" is given)
N=$1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 3 --warmup 2 > gpurun_out/scale_bench_n$N.json 2> gpurun_out/scale_bench_n$N.err; echo "bench n$N rc=$?"
python - <<PY
import json
for l in open('gpurun_out/scale_bench_n$N.json'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d[k] for k in ('value','ms_per_step','verified','n_gpus')}); print(d['e2e']); print(d['phases_ms'])
PY
if [ "$2" = "check" ]; then
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N tools/dist_check.py > gpurun_out/scale_distcheck_n$N.log 2>&1; echo "dist_check n$N rc=$?"; grep -c "ok=True" gpurun_out/scale_distcheck_n$N.log; grep -c "ok=False" gpurun_out/scale_distcheck_n$N.log
fi
