"""Dev tool: clock64 section timers of mg_physics_kernel (build with -DMG_SECTION_TIMERS, MG_LIB_PATH selects the .so):
per-CTA cycles of the phases of one env-step, mean and max over the CTAs, averaged over the timed steps.
usage: MG_LIB_PATH=variants/libtimers.so MG_STEP_PARTS=1 python tools/section_timers.py [env id] [B] [first timed step]"""
import ctypes as C
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
lib = env._lib
acts = torch.randint(0, 18, (80, B), dtype=torch.int32, device='cuda', generator=torch.Generator('cuda').manual_seed(1))
NCTA = 8192
buf = np.zeros((NCTA, 16), np.int64)
W0 = int(sys.argv[3]) if len(sys.argv) > 3 else 20        # steps into the episode at which the timed window starts
for t in range(W0):
    env.step(acts[t % 80])
nt = lib.mg_debug_read_timers(buf.ctypes.data, NCTA)      # also clears
steps = 40
for t in range(W0, W0 + steps):
    env.step(acts[t % 80])
lib.mg_debug_read_timers(buf.ctypes.data, NCTA)
names = ['parallel half', 'barrier wait', 'serial half', '  spring pre-steps', '  cached impulses', '  contacts (arbiters)', '  block constraints',
         '  robot constraints', 'CTA total', 'staging in', '  robot regs load/store',
         '  (parallel) integrate', '  (parallel) boxes', '  (parallel) pairs + narrow phase', '  (parallel) pre-steps']
used = buf[:, 8] > 0
b = buf[used] / steps
print(f'{eid} B={B}: {used.sum()} CTAs, cycles per CTA and env-step (mean / max over CTAs)')
for k, n in enumerate(names):
    print(f'  {n:26s} {b[:, k].mean():10.0f} {b[:, k].max():10.0f}')
order = np.argsort(b[:, 8])
print('  slowest CTA:', ', '.join(f'{names[k].strip()} {b[order[-1], k]:.0f}' for k in (0, 1, 2, 4, 5, 6, 7, 8)))
