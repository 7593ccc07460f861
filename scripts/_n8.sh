export PYTHONUNBUFFERED=1
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"
DP_CHECK_ONLY=dense timeout 600 $TR scripts/dp_check.py > gpurun_out/dp_check_n$N.log 2>&1; echo "dp_check rc=$?"; grep -E "FAIL|PASS|us/step|exchange==|unavailable|batch_norm" gpurun_out/dp_check_n$N.log | head -30
for e in "DFM_DP_EXCHANGE=dense" "DFM_DP_EXCHANGE=sparse"; do env $e timeout 300 $TR bench.py --gpus $N --steps 300 --warmup 30 --no_baseline > "gpurun_out/bench_n${N}_$e.log" 2>&1; python -c "
import sys,json
l=[x for x in open('gpurun_out/bench_n${N}_$e.log') if x.startswith('{')]
if l:
    d=json.loads(l[-1]); print('$e', d['ms_per_step'], d['e2e']['ms_per_step'], d['config']['comm'], d['kernels_per_step'], d['final_loss'], d['e2e_last_loss'], d['self_check'])
else:
    print('$e FAILED'); print(open('gpurun_out/bench_n${N}_$e.log').read()[-2000:])
"; done
timeout 300 $TR scripts/step_trace.py --steps 1 > gpurun_out/trace_n$N.md 2>&1; grep -A22 "rank 0, traced" gpurun_out/trace_n$N.md | head -30
