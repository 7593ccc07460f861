# N-GPU evidence session (default 8): dense dp check, headline bench with the graph-proxy arm and the in-bench self-check,
# step timeline (trace build), the bucketed variant, row-sharded benches (a2a roofline block). Raw logs: gpurun_out/n$N/.
export PYTHONUNBUFFERED=1
N=${1:-8}
O=gpurun_out/n$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"
J='import sys,json
l=[x for x in sys.stdin if x.startswith("{")]
d=json.loads(l[-1]) if l else {}
'
timeout 300 $TR bench.py --gpus $N --steps 1000 --warmup 50 > $O/bench_dp.log 2>&1; grep "^{" $O/bench_dp.log | python -c "$J
print('dp', d.get('ms_per_step'), d.get('e2e',{}).get('ms_per_step'), d.get('value'), d.get('vs_baseline'), (d.get('baseline') or {}).get('ms_per_step'), d.get('self_check'), d.get('config',{}).get('comm'))"
if [ "${SESSION:-full}" = "bench" ]; then exit 0; fi
DP_CHECK_ONLY=dense timeout 600 $TR scripts/dp_check.py > $O/dp_check.log 2>&1; echo "dp_check rc=$?"; grep -E "FAIL|PASS|us/step" $O/dp_check.log | head -12
DFM_KTRACE=1 timeout 300 $TR scripts/step_trace.py --steps 1 --out $O/trace.md > $O/trace.log 2>&1; grep -A24 "rank 0, traced" $O/trace.md | head -30
DFM_DP_BUCKETS=1 timeout 300 $TR bench.py --gpus $N --steps 500 --warmup 50 --no_baseline --no_selfcheck > $O/bench_dp_buckets.log 2>&1; grep "^{" $O/bench_dp_buckets.log | python -c "$J
print('dp buckets', d.get('ms_per_step'), d.get('e2e',{}).get('ms_per_step'))"
timeout 300 $TR bench.py --gpus $N --strategy sharded --batch_size 26624 --steps 50 --warmup 10 > $O/bench_sharded_1m.log 2>&1; grep "^{" $O/bench_sharded_1m.log | python -c "$J
print('sharded 1M ids/GPU', d.get('value'), d.get('ms_per_step'), d.get('a2a'))"
timeout 400 $TR bench.py --gpus $N --strategy sharded --feature_size 33554432 --embedding_size 64 --batch_size 1024 --steps 100 --warmup 10 > $O/bench_sharded_cfg3.log 2>&1; grep "^{" $O/bench_sharded_cfg3.log | python -c "$J
print('sharded cfg3', d.get('value'), d.get('ms_per_step'), d.get('a2a'))"
timeout 300 $TR scripts/sharded_check.py > $O/sharded_check.log 2>&1; grep -E "FAIL|PASS|us/step" $O/sharded_check.log | head -8
