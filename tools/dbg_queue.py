"""debug: the queued-resets test, one env id at a time, reporting the first failing call"""
import ctypes as C, sys
import numpy as np, torch
sys.path.insert(0, '.')
import magical_b200 as m
from magical_b200 import _lib as L
env_id, n = sys.argv[1], int(sys.argv[2])
a = m.make(env_id, num_envs=n, seed=9)
b = m.make(env_id, num_envs=n, seed=9)
a.auto_reset = b.auto_reset = False
a.reset()
b.reset([a.scenes[i] for i in range(n)])
dev = a.device
rew = [torch.zeros(n, device=dev) for _ in range(2)]
done = [torch.zeros(n, dtype=torch.uint8, device=dev) for _ in range(2)]
score = [torch.zeros(n, device=dev) for _ in range(2)]
st = torch.cuda.current_stream(dev).cuda_stream
rs = np.random.RandomState(3)
for t in range(5):
    acts = torch.from_numpy(rs.randint(0, 18, size=n).astype(np.int32)).to(dev)
    sub = np.arange(t, n, 5).astype(np.int32)
    new = L.scenes_array([a.scenes[int((i * 13 + t) % n)] for i in sub])
    for name, fn in (('queue', lambda: a._lib.mg_queue_resets(a._h, sub.ctypes.data, len(sub), C.addressof(new), st)),
                     ('step', lambda: a._lib.mg_step(a._h, acts.data_ptr(), a._obs.data_ptr(), rew[0].data_ptr(), done[0].data_ptr(), score[0].data_ptr(), st)),
                     ('b.physics', lambda: b._lib.mg_step_physics(b._h, acts.data_ptr(), rew[1].data_ptr(), done[1].data_ptr(), score[1].data_ptr(), st)),
                     ('b.set_scenes', lambda: b._lib.mg_set_scenes(b._h, sub.ctypes.data, len(sub), C.addressof(new), st)),
                     ('b.render', lambda: b._lib.mg_step_render(b._h, b._obs.data_ptr(), st))):
        rc = fn()
        h = a._h if not name.startswith('b.') else b._h
        if rc != 0:
            print('FAIL', env_id, 't', t, name, rc, a._lib.mg_last_error(h).decode())
            sys.exit(1)
    torch.cuda.synchronize()
    print(env_id, 't', t, 'ok, obs equal:', bool((a._obs == b._obs).all().item()))
