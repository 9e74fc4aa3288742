mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "voronoi" 2>&1 | tail -15
python - <<'PY'
import time, numpy as np, util_b200 as ub
for n,q in ((5000,2000),(50000,64)):
    x=np.random.default_rng(0).random((n,32)); z=np.random.default_rng(1).random((q,32)); y=x[:,0]
    m=ub.Voronoi(); m.fit(x,y); m(z[:2]); t=time.perf_counter(); p=m(z); dt=time.perf_counter()-t
    print(f"Voronoi() n={n} d=32: {q} queries in {dt:.3f} s = {q/dt:.1f} q/s")
PY
