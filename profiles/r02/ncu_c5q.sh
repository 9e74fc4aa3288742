mkdir -p gpurun_out
ncu --set full --import-source on --clock-control none -k regex:knn_tc_kernel -c 1 -o gpurun_out/knn_tc_c5q_r02 -f python bench.py --config c5q --no-cpu --no-e2e --steps 1 --warmup 3 > gpurun_out/ncu_c5q.log 2>&1
tail -5 gpurun_out/ncu_c5q.log
ls -la gpurun_out/*.ncu-rep
