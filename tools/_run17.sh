python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/pytest17.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest17.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke17.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke17.log
timeout 900 python bench.py > gpurun_out/bench17.json 2> gpurun_out/bench17.err; echo "bench rc=$?"; tail -3 gpurun_out/bench17.err
timeout 900 ncu --kernel-name-base demangled -k regex:rtseg --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches17.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-library-baseline > gpurun_out/ncu17.log 2>&1; echo "ncu rc=$?"
