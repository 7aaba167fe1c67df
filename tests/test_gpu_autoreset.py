"""Auto-reset through mg_queue_resets (resets applied between a step's physics and its render, inside the one-call
step) and the asynchronous reset feeder (the next episodes are sampled on a host thread while the GPU steps)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

IDS = ['MoveToRegion-Demo-v0', 'FixColour-TestAll-v0', 'MoveToCorner-Demo-v0', 'FindDupe-TestLayout-v0']   # horizons 40 / 60 / 80 / 100


def _mixed(m, per_id, seed, async_reset):
    env = m.make_mixed([i for i in IDS for _ in range(per_id)], seed=seed, preproc='LoRes4E', obs_keys=('past_obs',))
    env.async_reset = async_reset
    return env


def test_async_feeder_equals_synchronous_resets():
    """Same observations, rewards, scores, scenes and RandomState streams whether the next episodes are sampled ahead
    by the feeder thread or synchronously on the step that needs them; mixed horizons, so some env finishes on many
    steps and the feeder is consumed out of phase with any single horizon."""
    import magical_b200 as m
    a, b = _mixed(m, 6, 5, True), _mixed(m, 6, 5, False)
    oa, ob = a.reset()['past_obs'], b.reset()['past_obs']
    assert bool((oa == ob).all().item())
    B = a.num_envs
    actions = np.random.RandomState(1).randint(0, 18, size=(215, B))
    n_done = 0
    for t in range(215):
        oa, ra, da, ia = a.step(actions[t])
        ob, rb, db, ib = b.step(actions[t])
        assert bool((oa['past_obs'] == ob['past_obs']).all().item()), t
        np.testing.assert_array_equal(ra.cpu().numpy(), rb.cpu().numpy())
        np.testing.assert_array_equal(da.cpu().numpy(), db.cpu().numpy())
        np.testing.assert_array_equal(ia['eval_score'].cpu().numpy(), ib['eval_score'].cpu().numpy())
        np.testing.assert_array_equal(a.episode_steps, b.episode_steps)
        n_done += int(da.sum().item())
    assert n_done == 6 * (5 + 3 + 2 + 2)
    assert a._feeder is not None and b._feeder is None
    np.testing.assert_array_equal(a._scene_raw, b._scene_raw)
    # the feeder sampled episodes that have not started: looking at the streams rolls those draws back
    for r1, r2 in zip(a.rngs, b.rngs):
        assert r1.randint(1 << 30) == r2.randint(1 << 30)
    assert a._feeder is None
    # ... and stepping on restarts it with the same results
    for t in range(45):
        acts = np.random.RandomState(100 + t).randint(0, 18, size=B)
        oa, ra, da, ia = a.step(acts)
        ob, rb, db, ib = b.step(acts)
        assert bool((oa['past_obs'] == ob['past_obs']).all().item()), t
    np.testing.assert_array_equal(a._scene_raw, b._scene_raw)
    a.close(); b.close()


def test_queued_resets_equal_the_split_step(orc):
    """mg_queue_resets + mg_step == mg_step_physics; mg_set_scenes; mg_step_render (the round-1 reset path), also when
    the step runs in parts on internal streams and for a two-view stack (LoResStack, geometry pre-pass per view)."""
    import torch
    import magical_b200 as m
    from magical_b200 import _lib as L
    for env_id, n in (('FindDupe-TestLayout-LoRes4E-v0', 1024), ('MatchRegions-Demo-LoResStack-v0', 96)):
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
            L.check(a._h, a._lib.mg_queue_resets(a._h, sub.ctypes.data, len(sub), C.addressof(new), st))
            L.check(a._h, a._lib.mg_step(a._h, acts.data_ptr(), a._obs.data_ptr(), rew[0].data_ptr(), done[0].data_ptr(), score[0].data_ptr(), st))
            L.check(b._h, b._lib.mg_step_physics(b._h, acts.data_ptr(), rew[1].data_ptr(), done[1].data_ptr(), score[1].data_ptr(), st))
            L.check(b._h, b._lib.mg_set_scenes(b._h, sub.ctypes.data, len(sub), C.addressof(new), st))
            L.check(b._h, b._lib.mg_step_render(b._h, b._obs.data_ptr(), st))
            assert bool((a._obs == b._obs).all().item()), (env_id, t)
            assert bool((rew[0] == rew[1]).all().item())
            np.testing.assert_array_equal(a.get_bodies(), b.get_bodies())
        a.close(); b.close()


def test_queue_resets_is_one_call_per_step():
    import magical_b200 as m
    from magical_b200 import _lib as L
    env = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=4, seed=0, obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset()
    sc = L.scenes_array([env.scenes[0]])
    ids = np.asarray([2], np.int32)
    st = env._stream()
    assert env._lib.mg_queue_resets(env._h, ids.ctypes.data, 1, C.addressof(sc), st) == 0
    assert env._lib.mg_queue_resets(env._h, ids.ctypes.data, 1, C.addressof(sc), st) == -1
    assert b'already queued' in env._lib.mg_last_error(env._h)
    assert env._lib.mg_set_scenes(env._h, ids.ctypes.data, 1, C.addressof(sc), st) == -1
    env.step(np.zeros(4, int))                      # consumes the queue
    assert env._lib.mg_queue_resets(env._h, ids.ctypes.data, 1, C.addressof(sc), st) == 0
    env.step(np.zeros(4, int))
    bad = np.asarray([7], np.int32)
    assert env._lib.mg_queue_resets(env._h, bad.ctypes.data, 1, C.addressof(sc), st) == -1
    env.close()


def test_step_host_auto_reset_equals_device_step():
    """step_host with auto_reset: the finishing envs start their next episode inside the call, like step()."""
    import torch
    import magical_b200 as m
    n = 64
    a = m.make('MoveToRegion-Demo-LoRes4E-v0', num_envs=n, seed=4, obs_keys=('past_obs',))      # horizon 40
    b = m.make('MoveToRegion-Demo-LoRes4E-v0', num_envs=n, seed=4, obs_keys=('past_obs',))
    a.reset(); b.reset()
    obs_h = torch.empty((n, 96, 96, 12), dtype=torch.uint8).pin_memory()
    rew_h = torch.empty(n, dtype=torch.float32).pin_memory()
    done_h = torch.empty(n, dtype=torch.uint8).pin_memory()
    score_h = torch.empty(n, dtype=torch.float32).pin_memory()
    act_h = torch.empty(n, dtype=torch.int32).pin_memory()
    rs = np.random.RandomState(6)
    for t in range(85):
        acts = rs.randint(0, 18, size=n)
        act_h.copy_(torch.from_numpy(acts.astype(np.int32)))
        a.step_host(act_h, obs_h, rew_h, done_h, score_h)
        obs, rew, done, info = b.step(acts)
        np.testing.assert_array_equal(obs_h.numpy(), obs['past_obs'].cpu().numpy())
        np.testing.assert_array_equal(score_h.numpy(), info['eval_score'].cpu().numpy())
        np.testing.assert_array_equal(done_h.numpy().astype(bool), done.cpu().numpy())
    a.close(); b.close()
