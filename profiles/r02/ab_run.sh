timeout 900 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_configs.py -q -m gpu -x -s 2>&1 | grep -E "uncertified.*k=(33|40|100|24|64)|passed|failed"
for sh in "50000 50000 32 48" "100000 100000 16 40" "200000 200000 64 33" "200000 200000 64 24" "50000 50000 32 36" "20000 20000 100 40"; do
python profiles/microbench/ab.py build_ab/libcur.so util_b200/csrc/libb200approx.so $sh 1
done
timeout 600 python profiles/microbench/fuzz_knn.py 40 17 2>&1 | tail -2
