"""Dev tool: where the HOST time of one device-resident step goes (the bench's roll-out loop through
BatchedMagicalEnv.step_into with auto-reset and the reset feeder).  Runs the cfg1 workload, then:
 (a) issues N steps without waiting for the GPU and reports host ms per step (how far the host could run ahead),
 (b) the same under cProfile, top entries by cumulative time.
usage: python tools/profile_host_step.py [N]"""
import cProfile, pstats, sys, time
import numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
N = int(sys.argv[1]) if len(sys.argv) > 1 else 200
B = 4096
env = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=B, seed=0, obs_keys=('past_obs',))
env.reset()
env._episode_steps[:] = np.arange(B) % 80
acts = torch.randint(0, 18, (64, B), dtype=torch.int32, device='cuda')
rew = torch.zeros(B, device='cuda'); done = torch.zeros(B, dtype=torch.uint8, device='cuda'); score = torch.zeros(B, device='cuda')
t = [0]
def run(n):
    for _ in range(n):
        env.step_into(acts[t[0] % 64], rew, done, score)
        t[0] += 1
run(100); torch.cuda.synchronize()
for rep in range(3):
    t0 = time.perf_counter(); run(N); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print('issue %.3f ms per step (host), %.3f ms per step until the GPU is done' % ((t1 - t0) * 1e3 / N, (t2 - t0) * 1e3 / N), flush=True)
pr = cProfile.Profile(); pr.enable(); run(N); pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('cumulative').print_stats(14)
