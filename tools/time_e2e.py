"""Dev tool: end-to-end env-steps/s of mg_step_host at bench size with host assembly off / fixed / auto.
usage: python tools/time_e2e.py [B] [threads]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
TH = int(sys.argv[2]) if len(sys.argv) > 2 else 0
env = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=B, seed=0, obs_keys=('past_obs',))
env.reset()
env._episode_steps[:] = np.arange(B) % 80
acts = torch.randint(0, 18, (64, B), dtype=torch.int32).pin_memory()
obs = torch.empty((B, 96, 96, 12), dtype=torch.uint8).pin_memory()
rew = torch.empty(B, dtype=torch.float32).pin_memory()
done = torch.empty(B, dtype=torch.uint8).pin_memory()
score = torch.empty(B, dtype=torch.float32).pin_memory()
t_all = [0]
def run(n):
    for _ in range(n):
        env.step_host(acts[t_all[0] % 64], obs, rew, done, score)
        t_all[0] += 1
def timed(label, warm=6, n=20):
    run(warm)
    t0 = time.perf_counter(); run(n); dt = (time.perf_counter() - t0) / n
    a = env.host_assembly()
    print('%-22s %.3f ms per step  %.3f M env-steps/s   parts %d/%d threads %d  copy %.2f ms host %.2f ms' % (label, dt * 1e3, B / dt / 1e6, a[0], a[1], a[2], a[3], a[4]), flush=True)
import os
for mode in os.environ.get('E2E_MODES', 'off,4,8,10,11,12,14,16,auto,off').split(','):
    if mode == 'off':
        env.set_host_assembly('off')
        timed('off')
    elif mode == 'auto':
        env.set_host_assembly('auto', 0, TH)
        timed('auto', warm=40)
    else:
        env.set_host_assembly('fixed', int(mode), TH)
        timed('fixed ' + mode)
