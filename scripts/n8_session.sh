export PYTHONUNBUFFERED=1
N=${1:-8}
O=gpurun_out/n$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"
J='import sys,json
l=[x for x in sys.stdin if x.startswith("{")]
d=json.loads(l[-1]) if l else {}
'
DP_CHECK_ONLY=dense timeout 600 $TR scripts/dp_check.py > $O/dp_check.log 2>&1; echo "dp_check rc=$?"; grep -E "FAIL|PASS|us/step" $O/dp_check.log | head -12
timeout 300 $TR bench.py --gpus $N --steps 300 --warmup 30 > $O/bench_dp.log 2>&1; grep "^{" $O/bench_dp.log | python -c "$J
print('dp routed', d.get('ms_per_step'), d.get('e2e',{}).get('ms_per_step'), d.get('value'), d.get('vs_baseline'), (d.get('baseline') or {}).get('ms_per_step'), d.get('self_check'))"
DFM_DP_ROUTED=0 timeout 300 $TR bench.py --gpus $N --steps 300 --warmup 30 --no_baseline --no_selfcheck > $O/bench_dp_unrouted.log 2>&1; grep "^{" $O/bench_dp_unrouted.log | python -c "$J
print('dp unrouted', d.get('ms_per_step'), d.get('e2e',{}).get('ms_per_step'))"
timeout 300 $TR scripts/step_trace.py --steps 1 --out $O/trace.md > $O/trace.log 2>&1; grep -A20 "rank 0, traced" $O/trace.md | head -24
timeout 300 $TR scripts/flag_stress.py --steps 3000 --check_every 1500 --modes dense/p2p,dense/nvls > $O/flag_stress.log 2>&1; grep -E "FAIL|PASS|world=" $O/flag_stress.log
timeout 300 $TR bench.py --gpus $N --from_tfrecords /tmp/dfm_tfr > $O/bench_tfrecords.log 2>&1; grep "^{" $O/bench_tfrecords.log | python -c "$J
print('tfrecords', {k:(round(v['samples_per_s']/1e6,2)) for k,v in d.get('modes',{}).items()}, d.get('host_cpus'))"
timeout 300 $TR bench.py --gpus $N --strategy ps --batch_size 8192 --steps 100 --warmup 10 > $O/bench_ps.log 2>&1; grep "^{" $O/bench_ps.log | python -c "$J
print('ps B=8192', d.get('value'), d.get('ms_per_step'), d.get('ps_profile_us_per_step'), d.get('ps_max_observed_staleness'), d.get('final_loss'))"
timeout 300 $TR bench.py --gpus $N --strategy sharded --batch_size 26624 --steps 50 --warmup 10 > $O/bench_sharded_1m.log 2>&1; grep "^{" $O/bench_sharded_1m.log | python -c "$J
print('sharded 1M ids/GPU', d.get('value'), d.get('ms_per_step'), d.get('a2a'))"
timeout 400 $TR bench.py --gpus $N --strategy sharded --feature_size 33554432 --embedding_size 64 --batch_size 1024 --steps 100 --warmup 10 > $O/bench_sharded_cfg3.log 2>&1; grep "^{" $O/bench_sharded_cfg3.log | python -c "$J
print('sharded cfg3', d.get('value'), d.get('ms_per_step'), d.get('a2a'))"
tail -3 $O/bench_sharded_cfg3.log | cut -c1-300
