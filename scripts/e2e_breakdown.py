import time, numpy as np, torch
from augcl_b200 import utils as U
B=128
buf = U.ReplayBuffer((9,84,84),(6,),100_000,B); buf.fill_synthetic(seed=0)
gs = U.GraphSampler(buf, kind="drq", pad=4, batch=B, host_reward=True)
rng = np.random.RandomState(0)
idx = rng.randint(0,100000,(3000,B)); off = rng.randint(0,9,(3000,2,B,2)).astype(np.int32)
st = torch.cuda.current_stream()
T = [0.0]*6
for t in range(3000):
    a=time.perf_counter()
    gs._done.synchronize()
    b=time.perf_counter()
    gs._h_idx[:] = idx[t]; gs._h_off[0] = off[t,0]; gs._h_off[1] = off[t,1]
    c=time.perf_counter()
    ok = int(gs._h_idx_u.max()) < 100000 and int(gs._h_off_u.max()) <= 8
    d=time.perf_counter()
    gs.graph.replay()
    e=time.perf_counter()
    gs._done.record()
    f=time.perf_counter()
    st.synchronize()
    g=time.perf_counter()
    if t>=1000:
        for i,(x,y) in enumerate(((a,b),(b,c),(c,d),(d,e),(e,f),(f,g))): T[i]+=y-x
print("us per step: evsync %.2f  writes %.2f  validate %.2f  replay %.2f  record %.2f  sync %.2f  total %.2f" % tuple([x/2000*1e6 for x in T]+[sum(T)/2000*1e6]))
# kernel-only graph for comparison
g2 = U.GraphSampler(buf, kind="drq", pad=4, batch=B)
t0=time.perf_counter()
for t in range(2000):
    g2.graph.replay(); st.synchronize()
print("replay+sync only (H2D + kernel): %.2f us" % ((time.perf_counter()-t0)/2000*1e6))
