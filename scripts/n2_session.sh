# Two-GPU evidence session: multi-GPU test files, dp / sharded checks, headline bench with baseline arm and self-check.
export PYTHONUNBUFFERED=1
N=${1:-2}
O=gpurun_out/n$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"
J='import sys,json
l=[x for x in sys.stdin if x.startswith("{")]
d=json.loads(l[-1]) if l else {}
'
timeout 900 $TR scripts/dp_check.py > $O/dp_check.log 2>&1; echo "dp_check rc=$?"; grep -E "FAIL|PASS|us/step" $O/dp_check.log | head -16
timeout 600 $TR scripts/sharded_check.py > $O/sharded_check.log 2>&1; echo "sharded_check rc=$?"; grep -E "FAIL|PASS|us/step" $O/sharded_check.log | head -8
timeout 300 $TR scripts/flag_stress.py --steps 2000 --check_every 1000 > $O/flag_stress.log 2>&1; grep -E "FAIL|PASS|world=" $O/flag_stress.log | head
timeout 300 $TR bench.py --gpus $N --steps 1000 --warmup 50 > $O/bench_dp.log 2>&1; grep "^{" $O/bench_dp.log | python -c "$J
print('dp', d.get('ms_per_step'), d.get('e2e',{}).get('ms_per_step'), d.get('value'), d.get('vs_baseline'), (d.get('baseline') or {}).get('ms_per_step'), d.get('self_check'), d.get('config',{}).get('comm'))"
