set -x
python -m pytest tests -m gpu -q -rA -p no:cacheprovider > gpurun_out/pytest3.log 2>&1; echo "pytest rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed" gpurun_out/pytest3.log | tail -20
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke3.log 2>&1; echo "smoke rc=$?"; tail -6 gpurun_out/smoke3.log
timeout 900 python bench.py > gpurun_out/bench3.json 2> gpurun_out/bench3.err; echo "bench rc=$?"; cat gpurun_out/bench3.json; tail -5 gpurun_out/bench3.err
