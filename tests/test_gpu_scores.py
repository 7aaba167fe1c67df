"""GPU parity of the score kernels and of long / crowded episodes.

* the CUDA score kernels (mg_score / mg_debug_reward) against the REFERENCE-generated vectors in
  tests/golden/reference_scores.json (tools/gen_goldens.py executed the reference's own numpy code):
  MakeLine, ClusterColour, ClusterShape, MoveToCorner, PickAndPlace -- <= 1e-5 (float32 output);
* MatchRegions / MakeLine / ClusterColour / ClusterShape rolled to their horizons (120 / 180 / 240 steps)
  against the oracle, with env.score() compared mid-episode every 20 steps;
* a worst-case contact load (ten stars packed around the robot) against the oracle with the health flags
  (mg_get_status) checked, and the NaN guard;
* full-resolution frames at steps 0 / 1 / 10 / 50 / 100.
"""
import json
import os

import numpy as np
import pytest

from conftest import to_oracle_scene

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'reference_scores.json')))


def _product_scene(orc_scene):
    """oracle Scene -> product Scene (both mirror include/mg_scene.h byte for byte)."""
    import ctypes as C
    from magical_b200 import _lib as L
    out = L.Scene()
    assert C.sizeof(out) == C.sizeof(orc_scene)
    C.memmove(C.byref(out), C.byref(orc_scene), C.sizeof(out))
    return out


def _score_env(m, scenes):
    env = m.make('MoveToCorner-Demo-v0', num_envs=len(scenes), seed=0)
    env.auto_reset = False
    env.set_scenes(np.arange(len(scenes)), [_product_scene(s) for s in scenes])
    return env


def _scene(orc, task, blocks, robot=(0.0, 0.0), **kw):
    return orc.make_scene(task, 80, robot_pose=(robot, 0.3), blocks=blocks, **kw)


def test_cuda_scores_match_reference_vectors(orc):
    """mg_score / mg_debug_reward against scores the REFERENCE's own code produced (not the oracle's)."""
    import magical_b200 as m
    scenes, want, want_dbg = [], [], []
    for case in GOLD['make_line']:
        blocks = [(orc.SQUARE, orc.RED, x, y, 0.0) for x, y in case['points']]
        scenes.append(_scene(orc, orc.TASK_MAKE_LINE, blocks, robot=(5.0, 5.0)))
        want.append(case['score']); want_dbg.append(None)
    for case in GOLD['cluster']:
        blocks = [(t, c, x, y, 0.0) for (x, y), c, t in zip(case['points'], case['colours'], case['types'])]
        task = orc.TASK_CLUSTER_SHAPE if case['by_type'] else orc.TASK_CLUSTER_COLOUR
        scenes.append(_scene(orc, task, blocks, robot=(5.0, 5.0)))
        want.append(case['score']); want_dbg.append(None)
    for case in GOLD['move_to_corner']:
        blocks = [(orc.SQUARE, orc.RED, case['block'][0], case['block'][1], 0.0)]
        scenes.append(_scene(orc, orc.TASK_MOVE_TO_CORNER, blocks, robot=case['robot'], robot_first=True, debug_reward=True))
        want.append(case['score']); want_dbg.append(case['debug_reward'])
    for case in GOLD['pick_and_place']:
        blocks = [(orc.SQUARE, orc.RED, case['block'][0], case['block'][1], 0.0)] * 3
        scenes.append(_scene(orc, orc.TASK_PICK_AND_PLACE, blocks, robot=case['robot'], robot_first=True, target_block=0,
                             n_valid_targets=case['n_valid'], target_pos=case['target'], debug_reward=True))
        want.append(case['score']); want_dbg.append(case['debug_reward'])
    env = _score_env(m, scenes)
    got = env.score().cpu().numpy()
    got_dbg = env.debug_reward().cpu().numpy()
    assert len(want) == 200
    for i, (w, wd) in enumerate(zip(want, want_dbg)):
        assert abs(got[i] - w) <= 1e-5, (i, got[i], w)
        if wd is not None:
            assert abs(got_dbg[i] - wd) <= 1e-5, (i, got_dbg[i], wd)
    assert (env.status() == 0).all()
    env.close()


LONG_IDS = ['MatchRegions-Demo-v0', 'MatchRegions-TestAll-v0', 'MakeLine-Demo-v0', 'MakeLine-TestAll-v0',
            'ClusterColour-Demo-v0', 'ClusterColour-TestAll-v0', 'ClusterShape-Demo-v0', 'ClusterShape-TestAll-v0']


def test_long_horizon_tasks_reach_their_scores(orc):
    """MatchRegions (120), MakeLine (180), ClusterColour / ClusterShape (240) rolled to the horizon: poses 1e-3,
    done exactly at the horizon, eval_score / reward 1e-5, and mg_score mid-episode every 20 steps."""
    import magical_b200 as m
    env_ids = [i for i in LONG_IDS for _ in range(3)]
    env = m.make_mixed(env_ids, seed=17, preproc='LoRes4E', obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset()
    B = env.num_envs
    oracles = [orc.Env(to_oracle_scene(orc, s)) for s in env.scenes]
    horizons = np.asarray([o.scene.horizon for o in oracles])
    assert sorted(set(horizons.tolist())) == [120, 180, 240]
    T = int(horizons.max())
    # a bias towards pushing forward so that blocks actually get moved into / out of regions and lines
    rs = np.random.RandomState(99)
    actions = rs.randint(0, 18, size=(T, B))
    n_done = np.zeros(B, int)
    max_err = 0.0
    for t in range(T):
        _, rew, done, info = env.step(actions[t])
        rew, done, score = rew.cpu().numpy(), done.cpu().numpy(), info['eval_score'].cpu().numpy()
        live = (t + 1) <= horizons          # past its horizon an env keeps stepping but is no longer compared
        for i, o in enumerate(oracles):
            if live[i]:
                o.step(int(actions[t, i]))
        if (t + 1) % 20 == 0 or (t + 1) in (120, 180, 240):
            got = env.get_bodies()
            mid = env.score().cpu().numpy()
            for i, o in enumerate(oracles):
                if not live[i]:
                    continue
                max_err = max(max_err, np.abs(got[i, :o.n_bodies, :3] - o.bodies()[:, :3]).max())
                assert abs(mid[i] - o.score()) <= 1e-5, (env_ids[i], t, mid[i], o.score())
        for i, o in enumerate(oracles):
            if not live[i]:
                continue
            at_h = (t + 1) == horizons[i]
            assert bool(done[i]) == at_h, (env_ids[i], t)
            want = o.score() if at_h else 0.0
            assert abs(score[i] - want) <= 1e-5, (env_ids[i], t, score[i], want)
            assert abs(rew[i] - want) <= 1e-5, (env_ids[i], t, rew[i], want)
            n_done[i] += int(at_h)
    assert (n_done == 1).all()
    assert max_err <= 1e-3, max_err
    assert (env.status() == 0).all()
    print('long-horizon max pose error:', max_err)
    env.close()


def _crowd_scene(orc, kind, seed):
    """Ten blocks packed on two rings around the robot, touching or overlapping: every sub-step starts with dozens of
    simultaneous contacts (stars are six convex pieces each)."""
    rs = np.random.RandomState(seed)
    blocks = []
    for k in range(10):
        ring = 0.36 if k < 6 else 0.58
        ang = 2 * np.pi * (k / 6.0 if k < 6 else (k - 6) / 4.0 + 0.3) + rs.uniform(-0.1, 0.1)
        blocks.append((kind, k % 4, ring * np.cos(ang), ring * np.sin(ang), rs.uniform(-3, 3)))
    return orc.make_scene(orc.TASK_CLUSTER_COLOUR, 240, robot_pose=((0.0, 0.0), 0.4), blocks=blocks, robot_first=False)


def test_worst_case_contact_load_vs_oracle(orc):
    """Ten stars (60 convex pieces) / squares / pentagons / circles crowded around the robot for 50 steps: the
    arbiter cache must hold every contact (no overflow flag) and the poses follow the oracle."""
    import magical_b200 as m
    scenes = [_crowd_scene(orc, kind, seed) for kind in (orc.STAR, orc.SQUARE, orc.PENTAGON, orc.CIRCLE) for seed in range(3)]
    env = m.make('ClusterColour-Demo-LoRes4E-v0', num_envs=len(scenes), seed=0, obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset([_product_scene(s) for s in scenes])
    oracles = [orc.Env(s) for s in scenes]
    # drive into the crowd, turn, squeeze: keeps many pieces in contact
    script = [1] * 12 + [4] * 8 + [10] * 10 + [7] * 10 + [1] * 10
    max_err = 0.0
    for t, a in enumerate(script):
        env.step(np.full(len(scenes), a))
        got = env.get_bodies()
        for i, o in enumerate(oracles):
            o.step(a)
            max_err = max(max_err, np.abs(got[i, :o.n_bodies, :3] - o.bodies()[:, :3]).max())
    status = env.status()
    assert (status == 0).all(), status
    assert max_err <= 1e-3, max_err
    print('crowded scenes: max pose error', max_err)
    env.close()


def test_nan_guard_flags_the_env_only():
    """A non-finite body state raises MG_STATUS_NONFINITE for that env alone, and the flag is sticky until cleared."""
    import magical_b200 as m
    from magical_b200 import _lib as L
    env = m.make('MoveToCorner-Demo-LoRes4E-v0', num_envs=4, seed=2, obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset()
    bad = L.copy_scene(env.scenes[2])
    bad.blocks[0].x = float('nan')
    env.set_scenes([2], [bad])
    env.step(np.zeros(4, int))
    st = env.status()
    assert st[2] & L.STATUS_NONFINITE and not st[[0, 1, 3]].any(), st
    env.set_scenes([2], [env.scenes[1]])
    env.step(np.zeros(4, int))
    assert env.status(clear=True)[2] & L.STATUS_NONFINITE      # sticky across the reset
    env.step(np.zeros(4, int))
    assert not env.status().any()
    env.close()


@pytest.mark.parametrize('view', [0, 1])
def test_full_res_frames_along_a_long_rollout(orc, view):
    """Full-resolution frames at steps 0 / 1 / 10 / 50 / 100 of a random roll-out (after long contact sequences and,
    for the ego view, large rotations next to the walls): <= 1 % of pixels off by > 8 levels vs the oracle."""
    import magical_b200 as m
    ids = ['MoveToCorner-Demo-v0', 'MatchRegions-TestAll-v0', 'MakeLine-TestAll-v0', 'FindDupe-TestLayout-v0',
           'FixColour-TestAll-v0', 'ClusterColour-TestAll-v0', 'ClusterShape-TestAll-v0', 'PickAndPlace-Demo-v0']
    env_ids = [i for i in ids for _ in range(2)]
    env = m.make_mixed(env_ids, seed=41, preproc='LoRes4E', obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset()
    oracles = [orc.Env(to_oracle_scene(orc, s)) for s in env.scenes]
    rs = np.random.RandomState(5)
    # long runs of one action: the robot reaches walls and corners and spins there
    actions = np.repeat(rs.randint(0, 18, size=(20, env.num_envs)), 5, axis=0)
    t = 0
    n_diff = 0
    for stop in (0, 1, 10, 50, 100):
        while t < stop:
            env.step(actions[t])
            for i, o in enumerate(oracles):
                o.step(int(actions[t, i]))
            t += 1
        img = env.render_view(view, True).cpu().numpy()
        for i, o in enumerate(oracles):
            want = o.render(view)
            frac = (np.abs(img[i].astype(int) - want.astype(int)).max(axis=-1) > 8).mean()
            n_diff += int((img[i] != want).any(axis=-1).sum())
            assert frac <= 0.01, (env_ids[i], stop, frac)
    print(f'view {view}: pixels differing at all over 5 x {len(oracles)} frames: {n_diff}')
    env.close()


def test_step_host_after_reset_on_the_callers_stream():
    """mg_step_host runs on the handle's own stream: it must order itself behind resets (mg_set_scenes + mg_reset_obs)
    the caller queued on ITS stream without a synchronise in between."""
    import torch
    import magical_b200 as m
    n = 512
    a = m.make('FindDupe-TestLayout-LoRes4E-v0', num_envs=n, seed=3, obs_keys=('past_obs',))
    b = m.make('FindDupe-TestLayout-LoRes4E-v0', num_envs=n, seed=3, obs_keys=('past_obs',))
    a.auto_reset = b.auto_reset = False
    a.reset()
    b.reset([a.scenes[i] for i in range(n)])
    obs_h = torch.empty((n, 96, 96, 12), dtype=torch.uint8).pin_memory()
    rew_h = torch.empty(n, dtype=torch.float32).pin_memory()
    done_h = torch.empty(n, dtype=torch.uint8).pin_memory()
    score_h = torch.empty(n, dtype=torch.float32).pin_memory()
    act_h = torch.empty(n, dtype=torch.int32).pin_memory()
    rs = np.random.RandomState(8)
    side = torch.cuda.Stream()
    for t in range(6):
        if t in (2, 4):
            # reset a subset on a side stream that is kept busy, then step at once
            sub = np.arange(t, n, 3)
            with torch.cuda.stream(side):
                junk = torch.empty(64 << 20, dtype=torch.uint8, device='cuda')
                for _ in range(8):
                    junk.fill_(1)
                a.set_scenes(sub, [a.scenes[(i * 7 + 1) % n] for i in sub])
            b.set_scenes(sub, [a.scenes[i] for i in sub])
        acts = rs.randint(0, 18, size=n)
        act_h.copy_(torch.from_numpy(acts.astype(np.int32)))
        a.step_host(act_h, obs_h, rew_h, done_h, score_h)
        obs, rew, done, info = b.step(acts)
        np.testing.assert_array_equal(obs_h.numpy(), obs['past_obs'].cpu().numpy())
        np.testing.assert_array_equal(rew_h.numpy(), rew.cpu().numpy())
    a.close(); b.close()


def test_two_devices_in_one_process():
    """The shared-memory opt-in of the kernels is tracked per device."""
    import torch
    import magical_b200 as m
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    outs = []
    for dev in (0, 1):
        env = m.make('ClusterColour-TestAll-LoRes4E-v0', num_envs=64, device=dev, seed=1, obs_keys=('past_obs',))
        env.auto_reset = False
        env.reset()
        with torch.cuda.device(dev):
            for t in range(3):
                obs = env.step(np.full(64, t + 1))[0]['past_obs']
            outs.append(obs.cpu().numpy())
        env.close()
    np.testing.assert_array_equal(outs[0], outs[1])
