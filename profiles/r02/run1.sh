set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensor.py -x -q -m gpu 2>&1 | tail -30 > gpurun_out/t_tensor.log
cat gpurun_out/t_tensor.log | tail -15
timeout 600 python bench.py --config c5q --no-cpu --no-e2e --steps 3 > gpurun_out/b_c5q.json 2> gpurun_out/b_c5q.err; tail -3 gpurun_out/b_c5q.err; cat gpurun_out/b_c5q.json
timeout 600 python bench.py --config c5q --no-cpu --no-e2e --steps 3 --tc-mode 1 > gpurun_out/b_c5q_m1.json 2>>gpurun_out/b_c5q.err; cat gpurun_out/b_c5q_m1.json
timeout 600 python bench.py --config c5q --no-cpu --no-e2e --steps 3 --tc-mode 2 > gpurun_out/b_c5q_m2.json 2>>gpurun_out/b_c5q.err; cat gpurun_out/b_c5q_m2.json
timeout 900 python bench.py --steps 3 --cpu-seconds 5 > gpurun_out/b_c5.json 2> gpurun_out/b_c5.err; tail -3 gpurun_out/b_c5.err; cat gpurun_out/b_c5.json
