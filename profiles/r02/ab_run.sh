python profiles/microbench/ab.py build_ab/libpair.so build_ab/libpairpre.so 1000000 37888 64 16 2
python profiles/microbench/ab.py build_ab/libpair.so build_ab/libpairpre.so 1000000 1000000 64 16 2
python profiles/microbench/ab.py build_ab/libpair.so build_ab/libpairpre.so 100000 100000 16 18 1
for sl in 0 2 0 2; do
timeout 300 python bench.py --config c5q --no-cpu --no-e2e --steps 5 --tc-slots $sl 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('slots $sl', j['stages_ms']['ms_stage1'], j['tc'])"
done
