"""Host-side logic (no GPU): env-id registry, name grammar, and the reset samplers' RNG draw order /
placement invariants (reference: magical/benchmarks/__init__.py:308-432, 1017-1102, task on_reset()s,
magical/geom.py:116-341)."""
import json
import os

import numpy as np
import pytest

import magical_b200 as m
from magical_b200 import _lib as L
from magical_b200 import tasks

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
m.register_envs()


def sample(env_id, seed):
    sp = m.spec(env_id)
    kw = dict(sp['kwargs'])
    rd = kw.pop('rand_dynamics', False)
    for k in ('res_hw', 'fps', 'phys_steps', 'phys_iter', 'max_episode_steps'):
        kw.pop(k)
    return tasks.sample_scene(sp['entry_point'], np.random.RandomState(seed), sp['max_episode_steps'], rand_dynamics=rd, **kw)


def test_registry_matches_reference_counts():
    assert m.register_envs() is False            # idempotent, like the reference (:427-432)
    assert len(m.ALL_REGISTERED_ENVS) == 62 * 7 + 7 == 441
    assert len(set(m.ALL_REGISTERED_ENVS)) == 441
    assert m.AVAILABLE_PREPROCESSORS == ['LoRes3EA', 'LoRes4E', 'LoRes4A', 'LoResStack', 'LoResCHW4E', 'LoResCHW4A']
    assert len(m.DEMO_ENVS_TO_TEST_ENVS_MAP) == 9 * 7
    assert m.DEMO_ENVS_TO_TEST_ENVS_MAP['MoveToCorner-Demo-v0'] == (
        'MoveToCorner-TestColour-v0', 'MoveToCorner-TestShape-v0', 'MoveToCorner-TestJitter-v0',
        'MoveToCorner-TestDynamics-v0', 'MoveToCorner-TestAll-v0')
    horizons = {'PickAndPlace': 80, 'MoveToCorner': 80, 'MoveToRegion': 40, 'MatchRegions': 120, 'MakeLine': 180,
                'FindDupe': 100, 'FixColour': 60, 'ClusterColour': 240, 'ClusterShape': 240}
    for name in m.ALL_REGISTERED_ENVS:
        assert m.spec(name)['max_episode_steps'] == horizons[m.EnvName(name).task]


def test_env_name_grammar():
    n = m.EnvName('FindDupe-TestLayout-LoRes4E-v0')
    assert (n.task, n.variant, n.preproc, n.version, n.is_test) == ('FindDupe', 'TestLayout', 'LoRes4E', 'v0', True)
    assert n.demo_env_name == 'FindDupe-Demo-LoRes4E-v0'
    assert m.update_magical_env_name('MoveToCorner-Demo-v0', preproc='LoResStack') == 'MoveToCorner-Demo-LoResStack-v0'
    assert m.splice_in_preproc_name('ClusterShape-TestAll-v0', 'LoRes4A') == 'ClusterShape-TestAll-LoRes4A-v0'
    with pytest.raises(ValueError):
        m.EnvName('NotAnEnv')
    with pytest.raises(KeyError):
        m.spec('MoveToCorner-TestNope-v0')
    # DebugReward ids are registered un-wrapped by the reference (:1075-1100)
    assert m.spec('MoveToCorner-Demo-DebugReward-LoRes4E-v0')['preproc'] is None
    assert m.spec('MoveToCorner-Demo-DebugReward-v0')['kwargs']['debug_reward'] is True


def test_move_to_corner_draw_order():
    """rand(2) for the robot first (move_to_corner.py:34), then choice(colours), choice(types)."""
    for seed in range(5):
        s = sample('MoveToCorner-Demo-v0', seed)
        ref = np.random.RandomState(seed)
        pos = ref.rand(2)
        assert (s.robot[0][0], s.robot[0][1]) == (pos[0], pos[1])
        assert s.robot_first == 1 and s.n_blocks == 1 and s.horizon == 80
        assert (s.blocks[0].type, s.blocks[0].colour) == (L.SQUARE, L.RED)
        assert (s.blocks[0].x, s.blocks[0].y, s.blocks[0].angle) == (0.1, -0.65, 0.13 * np.pi)
        s2 = sample('MoveToCorner-TestColour-v0', seed)
        assert s2.blocks[0].colour == int(ref.choice(np.arange(4)))


def test_rand_dynamics_draws_five_uniforms_first():
    g = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'constants.json')))
    s = sample('MoveToCorner-TestDynamics-v0', 7)
    assert [s.phys[i] for i in range(5)] == g['phys_sample_seed7']      # PhysVar.sample of the reference
    ref = np.random.RandomState(7)
    for _, default, (lo, hi) in g['phys_vars']:
        ref.uniform(lo, hi)
    pos = ref.rand(2)
    assert (s.robot[0][0], s.robot[0][1]) == (pos[0], pos[1])
    d = sample('MoveToCorner-Demo-v0', 7)
    assert [d.phys[i] for i in range(5)] == [v[1] for v in g['phys_vars']]


def test_robot_group_poses():
    """Robot.setup: fingers at (+-0.09, 0.02) rotated by the robot angle, opened by pi/8 (entities.py:307-331)."""
    poses = tasks.robot_body_poses((0.3, -0.2), 0.7)
    c, s = np.cos(0.7), np.sin(0.7)
    for k, side in enumerate((-1, 1)):
        x, y, a = poses[4 + k]
        assert abs(x - (0.3 + side * 0.09 * c - 0.02 * s)) < 1e-15 and abs(y - (-0.2 + side * 0.09 * s + 0.02 * c)) < 1e-15
        assert a == 0.7 + (np.pi / 8 if side < 0 else -(np.pi / 8))
    assert poses[2] == [0.0, 0.0, 0.7]


@pytest.mark.parametrize('env_id', ['MatchRegions-TestAll-v0', 'FindDupe-TestLayout-v0', 'FixColour-TestAll-v0',
                                    'ClusterColour-TestAll-v0', 'MakeLine-TestLayout-v0', 'MoveToRegion-TestLayout-v0',
                                    'PickAndPlace-Test-v0', 'ClusterShape-TestJitter-v0'])
def test_layout_randomisation_leaves_no_overlaps(env_id):
    """After pm_randomise_all_poses nothing that was placed touches a wall, a goal or another entity,
    except the pairs the reference ignores (query block in its goal, FixColour blocks in their goals)."""
    lib = L.lib()
    for seed in range(6):
        s = sample(env_id, seed)
        task = m.EnvName(env_id).task
        all_b = (1 << s.n_blocks) - 1
        all_g = (1 << s.n_goals) - 1
        assert not lib.mg_host_entity_overlaps(s, L.ENT_ROBOT, 0, 0, all_b, all_g)
        for b in range(s.n_blocks):
            gmask = all_g
            if task == 'FindDupe' and b == s.n_blocks - 1:
                gmask = 0            # the query block sits inside the goal region
            if task == 'FixColour':
                gmask &= ~(1 << b)   # block b sits inside region b
            assert not lib.mg_host_entity_overlaps(s, L.ENT_BLOCK, b, 1, all_b, gmask), (env_id, seed, b)
        if task == 'FindDupe':
            q = s.blocks[s.n_blocks - 1]
            g = s.goals[0]
            assert abs(q.x - g.cx) <= g.w / 2 and abs(q.y - g.cy) <= g.h / 2
        if task == 'FixColour':
            for b in range(s.n_blocks):
                g = s.goals[b]
                assert abs(s.blocks[b].x - g.cx) <= g.w / 2 and abs(s.blocks[b].y - g.cy) <= g.h / 2


def test_count_ranges_and_roles():
    seen = set()
    for seed in range(40):
        s = sample('ClusterColour-TestAll-v0', seed)
        assert 7 <= s.n_blocks <= 10
        seen.add(s.n_blocks)
        assert sorted(set(s.blocks[i].colour for i in range(s.n_blocks))) == [0, 1, 2, 3]   # every colour occurs
        f = sample('FindDupe-TestAll-v0', seed)
        assert 3 <= f.n_blocks <= 7 and f.n_goals == 1
        q = f.blocks[f.n_blocks - 1]
        dup = f.blocks[f.n_blocks - 2]
        assert (q.type, q.colour, q.role) == (dup.type, dup.colour, L.ROLE_TARGET) and dup.role == L.ROLE_TARGET
        x = sample('FixColour-TestAll-v0', seed)
        assert 2 <= x.n_goals <= 3 and x.n_blocks == x.n_goals
        odd = [i for i in range(x.n_goals) if x.goals[i].want_block < 0]
        assert len(odd) == 1 and x.blocks[odd[0]].colour != x.goals[odd[0]].colour
    assert seen == {7, 8, 9, 10}


def test_jitter_stays_within_bounds():
    base = sample('MatchRegions-Demo-v0', 0)
    for seed in range(5):
        s = sample('MatchRegions-TestJitter-v0', seed)
        for b in range(s.n_blocks):
            assert abs(s.blocks[b].x - base.blocks[b].x) <= tasks.JITTER_POS_BOUND + 1e-12
            assert abs(s.blocks[b].angle - base.blocks[b].angle) <= tasks.JITTER_ROT_BOUND + 1e-12
        assert abs(s.goals[0].h - base.goals[0].h) <= tasks.JITTER_TARGET_BOUND + 1e-12


def test_placement_error_like_reference():
    """MoveToCorner-TestJitter with the robot spawned against a wall cannot be placed: the reference raises
    PlacementError after 10 x 10000 tries (geom.py:300-341); so do we."""
    found = False
    for seed in range(20):
        pos = np.random.RandomState(seed).rand(2)
        if pos.max() > 0.9:
            with pytest.raises(tasks.PlacementError):
                sample('MoveToCorner-TestJitter-v0', seed)
            found = True
            break
    assert found
