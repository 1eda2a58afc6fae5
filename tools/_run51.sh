timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_full.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_full.log; grep -n "^FAILED" gpurun_out/pytest_full.log | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --workload head_miou > gpurun_out/bench_head_miou_r02.json 2>/dev/null; echo "head rc=$?"
timeout 600 python bench.py --workload enet --no-library-baseline > gpurun_out/bench_enet_r02.json 2>/dev/null; echo "enet rc=$?"
