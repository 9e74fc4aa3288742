run() { # sub flags mode halves margin
timeout 120 python bench.py --config c5q --steps 3 --no-cpu --no-e2e --tc-sub $1 --tc-flags $2 --tc-mode $3 --tc-halves $4 --margin ${5:--1} > gpurun_out/x.log 2>&1
python - <<PY
import json
l=[x for x in open("gpurun_out/x.log") if x.startswith("{")]
if not l: print("FAIL sub=$1 fl=$2 m=$3 h=$4", open("gpurun_out/x.log").read()[-400:])
else:
    j=json.loads(l[-1]); print("sub=$1 flags=$2 mode=$3 halves=$4 margin=${5:--1}", "stage1_ms", round(j["stages_ms"]["ms_stage1"],2), "frac", round(j["roofline"]["frac"],3), "cap", j["tc"]["cap"], "keep", j["tc"]["keep"], "slots", j["tc"]["slots"], "H", j["tc"]["halves"], "uncert", j["uncertified_per_step"])
PY
}
for m in 2 1 0; do run 128 0 $m 2; done
run 128 0 0 2 4; run 128 0 0 2 12; run 128 0 0 2 16
