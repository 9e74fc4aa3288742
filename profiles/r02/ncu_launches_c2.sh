mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python profiles/microbench/c2_breakdown.py > gpurun_out/ncu_c2.log 2>&1
python profiles/summarize.py --launches gpurun_out/launches_c2.csv gpurun_out/launches_c2.txt | head -30
