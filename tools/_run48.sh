run() { # name args...
name=$1; shift
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; grep -c . gpurun_out/$name.json; }
run bench_erfnet_fastscnn_n2_r02 --no-cpu-baseline
run bench_head_miou_n2_r02 --workload head_miou --no-cpu-baseline
