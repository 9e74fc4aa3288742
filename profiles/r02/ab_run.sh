for sh in "50000 50000 32 48" "50000 50000 32 40" "100000 100000 16 52" "100000 100000 16 40" "50000 50000 32 36"; do
python profiles/microbench/ab.py build_ab/libcur.so util_b200/csrc/libb200approx.so $sh 1
done
timeout 900 python -m pytest tests/test_gpu_tensor.py -q -m gpu -x 2>&1 | tail -2
