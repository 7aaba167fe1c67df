"""Dev tool: time the physics and render kernels of one build of the library at bench size
(MG_LIB_PATH selects the .so).  usage: MG_LIB_PATH=... python tools/time_variants.py [env id] [B]"""
import sys
import numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
eid = sys.argv[1] if len(sys.argv) > 1 else 'MoveToCorner-Demo-LoRes4E-v0'
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
env = m.make(eid, num_envs=B, seed=0, obs_keys=('past_obs',))
env.auto_reset = False
env.reset()
acts = torch.randint(0, 18, (60, B), dtype=torch.int32, device='cuda', generator=torch.Generator('cuda').manual_seed(1))
for t in range(10):
    env.step(acts[t])
env.set_timing(True)
env.kernel_ms(0); env.kernel_ms(1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for t in range(10, 60):
    env.step(acts[t])
e1.record(); torch.cuda.synchronize()
p, pn = env.kernel_ms(0); r, rn = env.kernel_ms(1)
chk = int(env.get_bodies().view(np.int64).sum() & 0xffffffff)
print('%s B=%d  physics %.3f ms  render %.3f ms  step %.3f ms  state-hash %08x' % (eid, B, p / pn, r / rn, e0.elapsed_time(e1) / 50, chk))
