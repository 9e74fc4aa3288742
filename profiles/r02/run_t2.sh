mkdir -p gpurun_out
timeout 2400 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py -x -q -m gpu -k "c2 or c3 or c4 or lshep" 2>&1 | tail -25 > gpurun_out/t_cfg.log; cat gpurun_out/t_cfg.log
