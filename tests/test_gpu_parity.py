"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on identical initial
states and scripted random actions.  Criteria from BASELINE.json north_star: body poses within
1e-3 after 100 steps, reward/score within 1e-5, <= 1 % of pixels differing by more than 8 levels.
The oracle restates Chipmunk/pygame (parity vs the third-party libraries themselves is unpinned,
see oracle/orc.h)."""
import numpy as np
import pytest

from conftest import to_oracle_scene

pytestmark = pytest.mark.gpu

IDS = ['MoveToCorner-Demo-v0', 'MoveToCorner-TestShape-v0', 'MoveToCorner-TestDynamics-v0', 'MoveToRegion-TestAll-v0', 'MatchRegions-Demo-v0',
       'MatchRegions-TestAll-v0', 'MakeLine-Demo-v0', 'MakeLine-TestAll-v0', 'FindDupe-Demo-v0',
       'FindDupe-TestLayout-v0', 'FixColour-Demo-v0', 'FixColour-TestAll-v0', 'ClusterColour-Demo-v0',
       'ClusterColour-TestAll-v0', 'ClusterShape-TestAll-v0', 'PickAndPlace-Demo-v0']


def _mixed_env(m, ids, per_id, preproc='LoRes4E', seed=0):
    env_ids = [i for i in ids for _ in range(per_id)]
    return m.make_mixed(env_ids, seed=seed, preproc=preproc, obs_keys=('past_obs',)), env_ids


def test_physics_poses_and_scores_vs_oracle(orc):
    """100 steps of scripted random actions on 16 env ids x 4 seeds: poses 1e-3, rewards/scores 1e-5."""
    import magical_b200 as m
    env, env_ids = _mixed_env(m, IDS, 4)
    env.auto_reset = False
    env.reset()
    B = env.num_envs
    oracles = [orc.Env(to_oracle_scene(orc, s)) for s in env.scenes]
    np.testing.assert_allclose(env.get_bodies()[0, :oracles[0].n_bodies], oracles[0].bodies(), atol=1e-12)
    actions = np.random.RandomState(1234).randint(0, 18, size=(100, B))
    max_err = np.zeros(B)
    for t in range(100):
        _, rew, done, info = env.step(actions[t])
        rew, done, score = rew.cpu().numpy(), done.cpu().numpy(), info['eval_score'].cpu().numpy()
        got = env.get_bodies()
        for i, o in enumerate(oracles):
            o.step(actions[t, i])
            want = o.bodies()
            max_err[i] = max(max_err[i], np.abs(got[i, :o.n_bodies, :3] - want[:, :3]).max())
            horizon = o.scene.horizon
            exp_done = o.steps >= horizon
            assert bool(done[i]) == exp_done, (env_ids[i], t)
            exp_score = o.score() if exp_done else 0.0
            exp_rew = o.debug_reward() if o.scene.debug_reward else exp_score
            assert abs(score[i] - exp_score) <= 1e-5, (env_ids[i], t, score[i], exp_score)
            assert abs(rew[i] - exp_rew) <= 1e-5, (env_ids[i], t, rew[i], exp_rew)
    print('max pose error over 100 steps:', max_err.max())
    bad = [(env_ids[i], float(max_err[i])) for i in range(B) if max_err[i] > 1e-3]
    assert not bad, bad
    env.close()


@pytest.mark.parametrize('view', [0, 1])
def test_full_res_render_vs_oracle(orc, view):
    """Full-resolution allo/ego frames after 0 and 7 steps: <= 1 % of pixels differ by > 8 levels."""
    import magical_b200 as m
    env, env_ids = _mixed_env(m, IDS, 2, seed=3)
    env.auto_reset = False
    env.reset()
    oracles = [orc.Env(to_oracle_scene(orc, s)) for s in env.scenes]
    actions = np.random.RandomState(7).randint(0, 18, size=(7, env.num_envs))
    for phase in range(2):
        img = env.render_view(view, True).cpu().numpy()
        worst, n_any = 0.0, 0
        for i, o in enumerate(oracles):
            want = o.render(view)
            frac = (np.abs(img[i].astype(int) - want.astype(int)).max(axis=-1) > 8).mean()
            n_any += int((img[i] != want).any(axis=-1).sum())
            worst = max(worst, frac)
            assert frac <= 0.01, (env_ids[i], phase, frac)
        print(f'view {view} phase {phase}: worst >8-level fraction {worst:.6f}, pixels differing at all {n_any}')
        if phase == 0:
            for t in range(7):
                env.step(actions[t])
                for i, o in enumerate(oracles):
                    o.step(actions[t, i])
    env.close()


def test_lores_stack_vs_oracle_and_cv2_rule(orc):
    """LoRes4E past_obs = INTER_AREA(4x) of the last four ego frames, oldest first; first frame repeated
    after reset (FlattenFrameStack.reset)."""
    import magical_b200 as m
    env, env_ids = _mixed_env(m, ['MoveToCorner-Demo-v0', 'FindDupe-TestLayout-v0', 'ClusterColour-TestAll-v0'], 2, seed=5)
    env.auto_reset = False
    obs = env.reset()['past_obs'].cpu().numpy()
    oracles = [orc.Env(to_oracle_scene(orc, s)) for s in env.scenes]
    frames = [[orc.downsample4(o.render(1))] * 4 for o in oracles]
    actions = np.random.RandomState(11).randint(0, 18, size=(6, env.num_envs))
    for t in range(7):
        for i in range(env.num_envs):
            want = np.concatenate(frames[i][-4:], axis=-1)
            frac = (np.abs(obs[i].astype(int) - want.astype(int)).max(axis=-1) > 8).mean()
            assert frac <= 0.01, (env_ids[i], t, frac)
        # the downsample itself is bit-exact given the kernel's own full-res frame
        full = env.render_view(1, True).cpu().numpy()
        lo = env.render_view(1, False).cpu().numpy()
        for i in range(env.num_envs):
            np.testing.assert_array_equal(lo[i], orc.downsample4(full[i]))
            np.testing.assert_array_equal(obs[i][..., 9:], lo[i])
        if t == 6:
            break
        prev = obs
        obs = env.step(actions[t])[0]['past_obs'].cpu().numpy()
        np.testing.assert_array_equal(obs[..., :9], prev[..., 3:])   # the stack shifts by one frame
        for i, o in enumerate(oracles):
            o.step(actions[t, i])
            frames[i].append(orc.downsample4(o.render(1)))
    env.close()
