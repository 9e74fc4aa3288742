mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/t_multi_2gpu.log
timeout 900 python profiles/microbench/multi_handle.py 2 2>&1 | tail -4 | tee gpurun_out/multi_handle_2gpu.jsonl
for sh in query data; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --shard $sh --no-cpu --no-ref-class > gpurun_out/bench_c5_2gpu_$sh.json 2> gpurun_out/bench_2gpu_$sh.err || tail -5 gpurun_out/bench_2gpu_$sh.err
python -c "
import json
j=json.load(open('gpurun_out/bench_c5_2gpu_$sh.json')); print('$sh', j['value'], j['e2e']['value'], j['stages_ms'], j['roofline']['frac'])"
done
