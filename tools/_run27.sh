for w in ddrnet bisenetv2; do
timeout 900 python bench.py --workload $w --no-library-baseline > gpurun_out/bench_${w}_r02.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; tail -2 gpurun_out/bench_$w.err | cut -c1-300
done
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_full.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_full.log
