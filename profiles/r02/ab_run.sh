python profiles/microbench/ab.py build_ab/libpair.so build_ab/libpairpre.so 1000000 37888 64 16 2
python profiles/microbench/ab.py build_ab/libpair.so build_ab/libpairpre.so 1000000 1000000 64 16 2
python profiles/microbench/ab.py build_ab/libpair.so build_ab/libpairpre.so 100000 100000 16 18 1
