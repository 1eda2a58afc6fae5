python -m pytest tests -m gpu -q -rA -p no:cacheprovider > gpurun_out/pytest11.log 2>&1; echo "pytest rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed" gpurun_out/pytest11.log | tail -30
