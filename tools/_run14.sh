T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
timeout 420 $T bench.py --gpus 2 --workload erfnet_train --no-cpu-baseline > gpurun_out/bench14_train_n2.json 2> gpurun_out/bench14_train_n2.err; echo "train n2 graph rc=$?"; cut -c1-260 gpurun_out/bench14_train_n2.json; grep -v "OMP_NUM\|^\*\*\*" gpurun_out/bench14_train_n2.err | tail -8
