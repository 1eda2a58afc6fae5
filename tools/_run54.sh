timeout 900 python -m pytest tests/test_gpu_convert.py -q -p no:cacheprovider -x 2>&1 | tail -3
