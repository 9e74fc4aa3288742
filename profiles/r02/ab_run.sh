python profiles/microbench/ab.py build_ab/libcur.so build_ab/libnew.so 1000000 37888 64 16 1
python profiles/microbench/ab.py build_ab/libnew.so build_ab/liballhint.so 1000000 37888 64 16 2
python profiles/microbench/ab.py build_ab/libnew.so build_ab/liballhint.so 100000 100000 16 18 1
python profiles/microbench/ab.py build_ab/libcur.so build_ab/libnew.so 1000000 1000000 64 16 1
python profiles/microbench/ab.py build_ab/libnew.so build_ab/liballhint.so 1000000 1000000 64 16 1
