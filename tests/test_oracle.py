"""Pin the CPU oracle (oracle/, test infrastructure) against what the reference itself computes:
  * the 4x INTER_AREA downsample against cv2 (the reference's own call, benchmarks/__init__.py:181-183)
  * the numpy score functions against golden vectors produced by executing the reference's source
    (tools/gen_goldens.py -> tests/golden/reference_scores.json)
The Chipmunk / pygame halves have no reference-side golden vectors (parity unpinned, see oracle/orc.h)."""
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'reference_scores.json')))


def test_downsample_matches_cv2(orc):
    cv2 = pytest.importorskip('cv2')
    rng = np.random.RandomState(0)
    for c in (3, 12):
        img = rng.randint(0, 256, size=(384, 384, c)).astype(np.uint8)
        np.testing.assert_array_equal(orc.downsample4(img), cv2.resize(img, (96, 96), interpolation=cv2.INTER_AREA))
    # exact ties (sum % 16 == 8) round half to even
    img = np.zeros((4, 4, 1), np.uint8)
    img.flat[:8] = 1
    assert orc.downsample4(img)[0, 0, 0] == 0
    img.flat[:] = 1
    img.flat[:8] = 2
    assert orc.downsample4(img)[0, 0, 0] == 2
    img = np.full((4, 4, 3), 255, np.uint8)
    assert (orc.downsample4(img) == 255).all()


def _scene(orc, task, blocks, robot=(0.0, 0.0), **kw):
    return orc.make_scene(task, 80, robot_pose=(robot, 0.3), blocks=blocks, **kw)


def test_make_line_score_matches_reference(orc):
    for case in GOLD['make_line']:
        blocks = [(orc.SQUARE, orc.RED, x, y, 0.0) for x, y in case['points']]
        env = orc.Env(_scene(orc, orc.TASK_MAKE_LINE, blocks, robot=(5.0, 5.0)))
        assert abs(env.score() - case['score']) <= 1e-12, case


def test_cluster_score_matches_reference(orc):
    for case in GOLD['cluster']:
        blocks = [(t, c, x, y, 0.0) for (x, y), c, t in zip(case['points'], case['colours'], case['types'])]
        task = orc.TASK_CLUSTER_SHAPE if case['by_type'] else orc.TASK_CLUSTER_COLOUR
        env = orc.Env(_scene(orc, task, blocks, robot=(5.0, 5.0)))
        assert abs(env.score() - case['score']) <= 1e-12, case


def test_move_to_corner_score_matches_reference(orc):
    for case in GOLD['move_to_corner']:
        blocks = [(orc.SQUARE, orc.RED, case['block'][0], case['block'][1], 0.0)]
        env = orc.Env(_scene(orc, orc.TASK_MOVE_TO_CORNER, blocks, robot=case['robot'], robot_first=True,
                             debug_reward=True))
        assert abs(env.score() - case['score']) <= 1e-12
        assert abs(env.debug_reward() - case['debug_reward']) <= 1e-12


def test_pick_and_place_score_matches_reference(orc):
    for case in GOLD['pick_and_place']:
        blocks = [(orc.SQUARE, orc.RED, case['block'][0], case['block'][1], 0.0)] * 3
        env = orc.Env(_scene(orc, orc.TASK_PICK_AND_PLACE, blocks, robot=case['robot'], robot_first=True,
                             target_block=0, n_valid_targets=case['n_valid'], target_pos=case['target'],
                             debug_reward=True))
        assert abs(env.score() - case['score']) <= 1e-12
        assert abs(env.debug_reward() - case['debug_reward']) <= 1e-12


def test_oracle_free_motion_and_walls(orc):
    """Closed-form checks of the restated step: with no action the robot stays put; driving forward into the
    top wall stops at the wall (inner face y = 1, robot radius 0.2)."""
    env = orc.Env(_scene(orc, orc.TASK_MOVE_TO_CORNER, [(orc.SQUARE, orc.RED, -0.6, -0.6, 0.0)],
                         robot=(0.0, 0.0), robot_first=True))
    p0 = env.bodies()[0, :3].copy()
    for _ in range(10):
        env.step(0)
    assert np.abs(env.bodies()[0, :3] - p0).max() < 1e-9
    env2 = orc.Env(orc.make_scene(orc.TASK_MOVE_TO_CORNER, 80, robot_pose=((0.0, 0.5), 0.0),
                                  blocks=[(orc.SQUARE, orc.RED, -0.6, -0.6, 0.0)], robot_first=True))
    for _ in range(40):
        env2.step(1)   # UP
    y = env2.bodies()[0, 1]
    assert 0.55 < y < 0.81, y
