set -x
python -m pytest tests -m gpu -q -rA -p no:cacheprovider > gpurun_out/pytest2.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest2.log
python tools/ab_short_ring.py --tree gpurun_scratch/r01 > gpurun_out/ab_r01.json 2> gpurun_out/ab_r01.err; echo "ab r01 rc=$?"; cat gpurun_out/ab_r01.json; tail -2 gpurun_out/ab_r01.err
python tools/ab_short_ring.py > gpurun_out/ab_r02.json 2> gpurun_out/ab_r02.err; echo "ab r02 rc=$?"; cat gpurun_out/ab_r02.json; tail -2 gpurun_out/ab_r02.err
for w in dabnet fastscnn erfnet lednet; do timeout 600 python tools/stress_engine.py --workload $w --iters 300 --host > gpurun_out/stress_$w.json 2> gpurun_out/stress_$w.err; echo "stress $w rc=$?"; cat gpurun_out/stress_$w.json; tail -3 gpurun_out/stress_$w.err; done
timeout 900 python bench.py > gpurun_out/bench2_erfnet.json 2> gpurun_out/bench2_erfnet.err; echo "bench rc=$?"; cat gpurun_out/bench2_erfnet.json; tail -5 gpurun_out/bench2_erfnet.err
timeout 900 python bench.py --input f32 --single --no-library-baseline --no-cpu-baseline > gpurun_out/bench2_erfnet_f32.json 2> gpurun_out/bench2_erfnet_f32.err; echo "bench f32 rc=$?"; cat gpurun_out/bench2_erfnet_f32.json; tail -5 gpurun_out/bench2_erfnet_f32.err
