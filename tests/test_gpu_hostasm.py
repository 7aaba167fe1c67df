"""mg_step_host with host assembly (mg_set_host_assembly): part of the batch sends only its NEW frame over PCIe and
host threads build the stacked observations; the result must be byte-identical to the device-side stack of mg_step —
across episode ends inside the step, first steps (ring not yet mirrored), a device-path step in between (mirror stale)
and a boundary that moves (auto balance)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pinned(n):
    import torch
    return (torch.empty((n, 96, 96, 12), dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory(),
            torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory(),
            torch.empty(n, dtype=torch.int32).pin_memory())


@pytest.mark.parametrize('mode,chunks,threads', [('fixed', 16, 4), ('fixed', 5, 2), ('auto', 0, 3), ('fixed', 0, 1)])
def test_host_assembled_step_equals_device_step(mode, chunks, threads):
    import torch
    import magical_b200 as m
    n = 256
    a = m.make('MoveToRegion-Demo-LoRes4E-v0', num_envs=n, seed=11, obs_keys=('past_obs',))      # horizon 40
    b = m.make('MoveToRegion-Demo-LoRes4E-v0', num_envs=n, seed=11, obs_keys=('past_obs',))
    a.reset(); b.reset()
    a.set_host_assembly(mode, chunks, threads)
    assert a.host_assembly()[1] == 16 and a.host_assembly()[2] == threads
    # episodes out of phase: a few envs finish (and restart inside the step) on every step
    for env in (a, b):
        env._episode_steps[:] = np.arange(n) % 40
    obs_h, rew_h, done_h, score_h, act_h = _pinned(n)
    rs = np.random.RandomState(2)
    seen_chunks = set()
    for t in range(70):
        acts = rs.randint(0, 18, size=n)
        act_h.copy_(torch.from_numpy(acts.astype(np.int32)))
        if t in (30, 31):
            # the device path in between: the rings move without the host mirror
            o_a = a.step(acts)[0]['past_obs'].cpu().numpy()
            o_b = b.step(acts)[0]['past_obs'].cpu().numpy()
            np.testing.assert_array_equal(o_a, o_b)
            continue
        obs_h.fill_(7)
        a.step_host(act_h, obs_h, rew_h, done_h, score_h)
        obs, rew, done, info = b.step(acts)
        want = obs['past_obs'].cpu().numpy()
        diff = np.nonzero((obs_h.numpy() != want).reshape(n, -1).any(1))[0]
        assert len(diff) == 0, (t, diff[:8], a.host_assembly())
        np.testing.assert_array_equal(score_h.numpy(), info['eval_score'].cpu().numpy())
        np.testing.assert_array_equal(done_h.numpy().astype(bool), done.cpu().numpy())
        seen_chunks.add(a.host_assembly()[0])
    if mode == 'fixed':
        assert seen_chunks == {chunks}
    a.set_host_assembly('off')
    act_h.zero_()
    a.step_host(act_h, obs_h, rew_h, done_h, score_h)
    np.testing.assert_array_equal(obs_h.numpy(), b.step(np.zeros(n, int))[0]['past_obs'].cpu().numpy())
    a.close(); b.close()


def test_host_assembly_needs_a_single_view_stack():
    import magical_b200 as m
    from magical_b200 import _lib as L
    env = m.make('MoveToCorner-Demo-LoRes3EA-v0', num_envs=64, seed=0, obs_keys=('past_obs',))
    env.reset()
    with pytest.raises(L.MgError):
        env.set_host_assembly('auto')
    env.close()
    small = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=8, seed=0, obs_keys=('past_obs',))
    small.reset()
    with pytest.raises(L.MgError):
        small.set_host_assembly('auto')
    small.close()


def test_host_assembly_ragged_batch_and_restored_state():
    """A batch that does not divide into the 16 parts evenly, and mg_set_state between two host-assembled steps (the
    restored rings are not what the host mirror holds: the handle refetches them)."""
    import torch
    import magical_b200 as m
    n = 100
    a = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=n, seed=3, obs_keys=('past_obs',))
    b = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=n, seed=3, obs_keys=('past_obs',))
    a.auto_reset = b.auto_reset = False
    a.reset(); b.reset()
    a.set_host_assembly('fixed', 9, 2)
    obs_h, rew_h, done_h, score_h, act_h = _pinned(n)
    rs = np.random.RandomState(8)
    saved = None
    for t in range(12):
        acts = rs.randint(0, 18, size=n)
        act_h.copy_(torch.from_numpy(acts.astype(np.int32)))
        if t == 4:
            saved = (a.get_state(), b.get_state())
        if t == 8:
            a.set_state(saved[0]); b.set_state(saved[1])       # back to step 4: rings and poses of that time
        a.step_host(act_h, obs_h, rew_h, done_h, score_h)
        want = b.step(acts)[0]['past_obs'].cpu().numpy()
        np.testing.assert_array_equal(obs_h.numpy(), want)
    a.close(); b.close()
