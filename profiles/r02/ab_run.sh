timeout 900 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_configs.py -q -m gpu -x 2>&1 | tail -2
for sh in "50000 50000 32 64" "100000 100000 16 64" "50000 50000 32 48" "100000 100000 16 40" "200000 200000 64 64"; do
python profiles/microbench/ab.py build_ab/libcur.so util_b200/csrc/libb200approx.so $sh 1
done
timeout 600 python profiles/microbench/fuzz_knn.py 30 13 2>&1 | tail -2
