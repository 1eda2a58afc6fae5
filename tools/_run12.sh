python -m pytest tests -m gpu -q -rA -p no:cacheprovider > gpurun_out/pytest12.log 2>&1; echo "pytest rc=$?"
grep -n "^FAILED\|^ERROR\| passed\| failed" gpurun_out/pytest12.log | tail -30
timeout 600 python bench.py --workload enet > gpurun_out/bench12_enet.json 2> gpurun_out/bench12_enet.err; echo "bench enet rc=$?"; cut -c1-900 gpurun_out/bench12_enet.json; tail -4 gpurun_out/bench12_enet.err
timeout 600 python bench.py --workload enet --impl reference --steps 20 --warmup 10 | cut -c1-600
timeout 600 python bench.py --workload erfnet_train > gpurun_out/bench12_train.json 2> gpurun_out/bench12_train.err; echo "bench train rc=$?"; cut -c1-700 gpurun_out/bench12_train.json; tail -4 gpurun_out/bench12_train.err
RTSEG_TORCH_SGD=1 timeout 600 python bench.py --workload erfnet_train --no-cpu-baseline 2>/dev/null | cut -c1-300
