# bash profiles/r02/run_ngpu.sh N [multi]: C5 on N GPUs of one box, query- and data-sharded, under torchrun;
# with "multi" also the one-process handle (b200_multi_*) and, at N = 2, tests/test_gpu_multi.py
N=$1
mkdir -p gpurun_out
if [ "$2" = "multi" ]; then
  [ "$N" = "2" ] && timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/pytest_gpu_multi_2gpu.txt
  timeout 900 python profiles/microbench/multi_handle.py $N 2>&1 | grep "^{" | tee gpurun_out/multi_handle_${N}gpu.jsonl
fi
for sh in query data; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --shard $sh --no-cpu --no-ref-class > gpurun_out/bench_c5_${N}gpu_${sh}shard.json 2> gpurun_out/bench_${N}gpu_$sh.err || tail -5 gpurun_out/bench_${N}gpu_$sh.err
sleep 1
python -c "
import json
j=json.loads(open('gpurun_out/bench_c5_${N}gpu_${sh}shard.json').read().strip().splitlines()[-1]); print('$sh', round(j['value']), round(j['e2e']['value']), {k:round(v,2) for k,v in j['stages_ms'].items()}, round(j['roofline']['frac'],3))"
done
