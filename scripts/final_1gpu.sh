# One-GPU evidence session: full GPU suite, the headline bench with the graph-proxy baseline arm, the step timeline.
export PYTHONUNBUFFERED=1
O=gpurun_out/final; mkdir -p $O
python -m pytest tests -m gpu -q -x > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log
python bench.py --steps 2000 --warmup 50 > $O/bench_n1.log 2>&1; echo "bench rc=$?"; grep "^{" $O/bench_n1.log | tail -1 | cut -c1-1500
DFM_KTRACE=1 python scripts/step_trace.py --steps 1 --out $O/trace_n1.md > $O/trace_n1.log 2>&1; echo "trace rc=$?"; grep -A24 "traced" $O/trace_n1.md | head -40
