timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_nets.py -q -p no:cacheprovider -x -k "argmax or head or upsample or confmat or inf or nan" 2>&1 | tail -2
python tools/layer_bench.py --set head --batch 32 --reps 5 --only upsample 2>&1 | cut -c1-160



