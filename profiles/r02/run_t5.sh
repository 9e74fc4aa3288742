mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "ripple or robust" 2>&1 | tail -25 | tee gpurun_out/t_ripple.log
