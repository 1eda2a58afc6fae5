timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_full.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_full.log; grep -n "^FAILED" gpurun_out/pytest_full.log | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_full.err | cut -c1-200
