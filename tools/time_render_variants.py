"""Time the render kernel's output variants at bench size (dev tool): lo-res out (no ring, no stack) vs the
fused ring push + 4-frame stack, CUDA events around 50 launches each."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
B = 4096
env = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=B, seed=0, obs_keys=('past_obs',))
env.auto_reset = False
env.reset()
a = torch.randint(0, 18, (B,), dtype=torch.int32, device='cuda')
for _ in range(5):
    env.step(a)
def timed(fn, n=50):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
L = env._lib
import magical_b200._lib as ML
st = env._stream()
print('lo-res ego out  (geometry + raster + downsample)      %.3f ms' % timed(lambda: env.render_view(1, False)))
print('full-res ego out (geometry + raster, 1.8 GB written)  %.3f ms' % timed(lambda: env.render_view(1, True)))
print('step_render     (geometry + raster + ring + stack)    %.3f ms' % timed(lambda: L.mg_step_render(env._h, ML.ptr(env._obs), st)))
