timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest_full.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_full.log; grep -n "^FAILED" gpurun_out/pytest_full.log | head
for w in ddrnet bisenetv2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $w > gpurun_out/bench_${w}_n2_r02.json 2> gpurun_out/bench_${w}_n2.err; echo "$w n2 rc=$?"; tail -2 gpurun_out/bench_${w}_n2.err | cut -c1-200
done
