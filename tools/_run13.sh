nvidia-smi -L | head -3
timeout 900 python -m pytest tests/test_gpu_train.py -q -rA -p no:cacheprovider -k "ddp" > gpurun_out/ddp_test_n2.log 2>&1; echo "ddp test rc=$?"; tail -5 gpurun_out/ddp_test_n2.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $T bench.py --gpus 2 > gpurun_out/bench13_n2.json 2> gpurun_out/bench13_n2.err; echo "bench n2 rc=$?"; cut -c1-300 gpurun_out/bench13_n2.json; tail -3 gpurun_out/bench13_n2.err
timeout 900 $T bench.py --gpus 2 --workload head_miou > gpurun_out/bench13_head_n2.json 2> gpurun_out/bench13_head_n2.err; echo "head n2 rc=$?"; cut -c1-1500 gpurun_out/bench13_head_n2.json; tail -3 gpurun_out/bench13_head_n2.err
timeout 900 python bench.py --workload head_miou > gpurun_out/bench13_head_n1.json 2> gpurun_out/bench13_head_n1.err; echo "head n1 rc=$?"; cut -c1-400 gpurun_out/bench13_head_n1.json; tail -3 gpurun_out/bench13_head_n1.err
timeout 900 $T bench.py --gpus 2 --workload erfnet_train > gpurun_out/bench13_train_n2.json 2> gpurun_out/bench13_train_n2.err; echo "train n2 rc=$?"; cut -c1-300 gpurun_out/bench13_train_n2.json; tail -3 gpurun_out/bench13_train_n2.err
