set -x
mkdir -p gpurun_out
nvidia-smi -L; nproc
python -m pytest tests -m gpu -q -rA -p no:cacheprovider > gpurun_out/pytest1.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest1.log
for w in dabnet fastscnn erfnet; do timeout 600 python tools/stress_engine.py --workload $w --iters 300 --host > gpurun_out/stress_$w.json 2> gpurun_out/stress_$w.err; echo "stress $w rc=$?"; cat gpurun_out/stress_$w.json; tail -3 gpurun_out/stress_$w.err; done
timeout 900 python bench.py > gpurun_out/bench1_erfnet.json 2> gpurun_out/bench1_erfnet.err; echo "bench rc=$?"; cat gpurun_out/bench1_erfnet.json; tail -3 gpurun_out/bench1_erfnet.err
for w in fastscnn dabnet lednet; do timeout 900 python bench.py --workload $w > gpurun_out/bench1_$w.json 2> gpurun_out/bench1_$w.err; echo "bench $w rc=$?"; cat gpurun_out/bench1_$w.json; tail -3 gpurun_out/bench1_$w.err; done
