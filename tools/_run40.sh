(python tools/layer_bench.py --set erfnet --reps 10; python tools/layer_bench.py --set fastscnn --batch 32 --reps 10; python tools/layer_bench.py --set head --batch 32 --reps 10; python tools/layer_bench.py --set net:ddrnet --reps 5; python tools/layer_bench.py --set net:bisenetv2 --reps 5; python tools/_dwpw_bench.py) > gpurun_out/layers_r02.txt 2>&1
wc -l gpurun_out/layers_r02.txt
