# Round-2 measurement batch (one B200): GPU tests, headline bench + reference arm, C1-C4, ncu launch list and
# full captures of the two kernels of the C5 step.  Every ncu pass runs after its command exited 0 without ncu.
mkdir -p gpurun_out/final
O=gpurun_out/final
timeout 2400 python -m pytest tests -q -m gpu 2>&1 | tail -15 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee $O/smoke.txt
timeout 900 python bench.py > $O/bench_c5.json 2> $O/bench_c5.err || tail -5 $O/bench_c5.err
timeout 900 python bench.py --impl reference > $O/bench_reference_arm.json 2> $O/bench_ref.err || tail -5 $O/bench_ref.err
for c in c1 c2 c3 c4; do
  timeout 900 python bench.py --config $c > $O/bench_$c.json 2> $O/bench_$c.err || tail -5 $O/bench_$c.err
done
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu > $O/b_plain.json 2> $O/b_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c5.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > $O/ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:knn_tc_kernel -c 1 -o $O/knn_tc_c5 -f \
    python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > $O/ncu_full_tc.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:rerank_kernel -c 1 -o $O/rerank_c5 -f \
    python bench.py --no-cpu --no-e2e --steps 1 --warmup 3 > $O/ncu_full_rr.log 2>&1
for f in $O/bench_c5.json $O/bench_reference_arm.json; do head -c 1500 $f; echo; done
ls -la $O
