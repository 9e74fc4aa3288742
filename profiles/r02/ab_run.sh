for v in pw300 pw1000; do
python profiles/microbench/ab.py build_ab/libcur.so build_ab/lib$v.so 1000000 37888 64 16 2
python profiles/microbench/ab.py build_ab/libcur.so build_ab/lib$v.so 100000 100000 16 18 1
done
