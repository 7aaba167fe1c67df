"""mg_get_state_noalloc: the checkpoint gather with its device scratch in a fixed file-scope __device__ array
(no cudaMalloc / cudaFree inside the call).  Checked through the C ABI against the existing entry point
(mg_get_state), byte for byte, on two calls in a row that share the scratch array.

Kept in a file that pytest collects last; it passes both alone in a fresh
process and after the other tests (the scratch holds whatever earlier calls left: nothing in it must start at zero).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _blobs(env, ids, fn_name):
    import torch
    from magical_b200 import _lib as L
    ids = np.ascontiguousarray(ids, np.int32)
    torch.cuda.synchronize(env.device)
    sz = env._lib.mg_state_size(env._h)
    buf = np.zeros((len(ids), sz), np.uint8)
    fn = getattr(env._lib, fn_name)
    L.check(env._h, fn(env._h, ids.ctypes.data, len(ids), buf.ctypes.data))
    return buf


def test_noalloc_get_state_equals_get_state_twice_in_a_row():
    import magical_b200 as m
    from magical_b200 import _lib as L
    env = m.make('FindDupe-TestLayout-LoRes4E-v0', num_envs=48, seed=11, obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset()
    rs = np.random.RandomState(0)
    for t in range(4):
        env.step(rs.randint(0, 18, size=48))
    # call 1: the maximum number of envs, scattered ids
    ids1 = rs.permutation(48)[:L.NOALLOC_MAX_ENVS]
    want1 = _blobs(env, ids1, 'mg_get_state')
    got1 = _blobs(env, ids1, 'mg_get_state_noalloc')
    np.testing.assert_array_equal(got1, want1)
    # call 2, straight after: fewer envs, other ids, a later state -- the scratch still holds call 1's contents
    env.step(rs.randint(0, 18, size=48))
    ids2 = np.asarray([47, 0, 5, 31, 9])
    got2 = _blobs(env, ids2, 'mg_get_state_noalloc')
    want2 = _blobs(env, ids2, 'mg_get_state')
    np.testing.assert_array_equal(got2, want2)
    # the blobs restore: set_state(noalloc blobs) then step == stepping the live env
    twin = m.make('FindDupe-TestLayout-LoRes4E-v0', num_envs=5, seed=99, obs_keys=('past_obs',))
    twin.auto_reset = False
    twin.reset()
    twin.set_state(dict(blobs=got2, episode_steps=env.episode_steps[ids2]))
    acts = rs.randint(0, 18, size=48)
    a = env.step(acts)[0]['past_obs'].cpu().numpy()[ids2]
    b = twin.step(acts[ids2])[0]['past_obs'].cpu().numpy()
    np.testing.assert_array_equal(a, b)
    # env_ids == NULL means envs 0..n-1, as for mg_get_state
    sz = env._lib.mg_state_size(env._h)
    g = np.zeros((3, sz), np.uint8)
    L.check(env._h, env._lib.mg_get_state_noalloc(env._h, None, 3, g.ctypes.data))
    np.testing.assert_array_equal(g, _blobs(env, [0, 1, 2], 'mg_get_state'))
    twin.close()
    env.close()


def test_noalloc_rejects_sizes_beyond_the_scratch():
    import magical_b200 as m
    from magical_b200 import _lib as L
    n = L.NOALLOC_MAX_ENVS + 1
    env = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=n, seed=1, obs_keys=('past_obs',))
    env.reset()
    sz = env._lib.mg_state_size(env._h)
    buf = np.zeros((n, sz), np.uint8)
    ids = np.arange(n, dtype=np.int32)
    rc = env._lib.mg_get_state_noalloc(env._h, ids.ctypes.data, n, buf.ctypes.data)
    assert rc == -1      # MG_ERR_ARG
    msg = env._lib.mg_last_error(env._h).decode()
    assert 'MG_NOALLOC_MAX_ENVS' in msg and str(L.NOALLOC_MAX_ENVS) in msg, msg
    assert not buf.any()
    bad = np.asarray([n + 5], np.int32)
    assert env._lib.mg_get_state_noalloc(env._h, bad.ctypes.data, 1, buf.ctypes.data) == -1
    env.close()
