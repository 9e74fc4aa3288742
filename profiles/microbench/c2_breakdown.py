import numpy as np, time, sys
sys.path.insert(0,'.')
import util_b200 as ub
n=q=100000; d=16
x=np.random.default_rng(0).random((n,d)); z=np.random.default_rng(1).random((q,d)); y=x[:,0]
m=ub.ShepMod(); m.fit(x,y); m.predict(z); m.model.set_option("reset_stats",1)
t=time.perf_counter(); p=m.predict(z); dt=time.perf_counter()-t
print("predict", dt*1e3, "ms; redo", m.model.stat("uncertified"), "stage1", m.model.stat("ms_stage1"), "weights", m.model.stat("ms_weights"), "ierrors", m.ierrors)
ids,wts=m._predict(z[:2000]); c=np.array([len(i) for i in ids]); print("support mean", c.mean(), "max", c.max())
