"""Dev tool: two handles with the same seed stepped side by side at full batch size (auto-reset on): every observation
byte and every body state must agree -- a statistical check for races in the kernels (two warps per rendered band, the
overlapped step).  usage: python tools/soak_determinism.py"""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
for eid, B, T in (('MoveToCorner-Demo-LoRes4E-v0', 4096, 400), ('FindDupe-TestLayout-LoRes4E-v0', 8192, 200), ('ClusterShape-TestAll-LoRes4E-v0', 4096, 200)):
    a = m.make(eid, num_envs=B, seed=1, obs_keys=('past_obs',)); b = m.make(eid, num_envs=B, seed=1, obs_keys=('past_obs',))
    a.reset(); b.reset()
    acts = torch.randint(0, 18, (T, B), dtype=torch.int32, device='cuda', generator=torch.Generator('cuda').manual_seed(3))
    bad = 0
    for t in range(T):
        oa = a.step(acts[t])[0]['past_obs']; ob = b.step(acts[t])[0]['past_obs']
        if not torch.equal(oa, ob): bad += 1
    same_bodies = np.array_equal(a.get_bodies(), b.get_bodies())
    print(eid, B, T, 'steps with differing observations:', bad, 'bodies equal:', same_bodies, 'status flags:', int((a.status() != 0).sum()))
    a.close(); b.close()
