timeout 900 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_parity.py -q -m gpu -x 2>&1 | tail -2
python profiles/microbench/ab.py build_ab/libcur.so util_b200/csrc/libb200approx.so 1000000 37888 64 16 2
python profiles/microbench/ab.py build_ab/libcur.so util_b200/csrc/libb200approx.so 100000 100000 16 18 1
python profiles/microbench/ab.py build_ab/libcur.so util_b200/csrc/libb200approx.so 1000000 500000 64 16 1
