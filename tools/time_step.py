"""Dev tool: time mg_step (the one-call device step) at bench size; MG_STEP_PARTS=<n> selects the overlapped form.
usage: [MG_STEP_PARTS=8] python tools/time_step.py [env id] [B]"""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
from magical_b200 import _lib as L
eid = sys.argv[1] if len(sys.argv) > 1 else 'MoveToCorner-Demo-LoRes4E-v0'
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
env = m.make(eid, num_envs=B, seed=0, obs_keys=('past_obs',))
env.auto_reset = False
env.reset()
lib, h = env._lib, env._h
acts = torch.randint(0, 18, (70, B), dtype=torch.int32, device='cuda', generator=torch.Generator('cuda').manual_seed(1))
st = torch.cuda.current_stream().cuda_stream
def step(t):
    L.check(h, lib.mg_step(h, acts[t].data_ptr(), env._obs.data_ptr(), env._reward.data_ptr(), env._done.data_ptr(), env._score.data_ptr(), st))
for t in range(10):
    step(t)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for t in range(10, 70):
    step(t)
e1.record(); torch.cuda.synchronize()
chk = int(env.get_bodies().view(np.int64).sum() & 0xffffffff)
ochk = int(env._obs.to(torch.int64).sum().item())
print('%s B=%d  mg_step %.3f ms  state-hash %08x obs-sum %d' % (eid, B, e0.elapsed_time(e1) / 60, chk, ochk))
