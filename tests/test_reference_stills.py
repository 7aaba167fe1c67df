"""Reference-held pin of the rasteriser: the eight stills the reference ships (images/static-<task>-demo-v0.png,
192x192, rendered by the upstream renderer from the deterministic <Task>-Demo-v0 layouts; copied unchanged to
tests/golden/reference_stills/) against the allocentric frame of the same layout from (a) the CPU oracle and
(b) the CUDA rasteriser, area-averaged 384 -> 192.

The stills come from upstream's GL renderer (antialiased, slightly different outline widths), so this is a pin of
layout, palette, draw order and geometry, not of single pixels.  Thresholds (measured: 3.5-5.4 % / 0.4-1.2 %):
<= 6 % of pixels off by more than 24 levels and <= 1.5 % by more than 64.  MoveToCorner: this fork draws the robot
position at random (magical/benchmarks/move_to_corner.py:33-34); the still shows the upstream default
(0.4, -0.0) kept there as a comment, so the scene is built with that pose.
"""
import ctypes as C
import math
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STILLS = os.path.join(ROOT, 'tests', 'golden', 'reference_stills')
TASKS = {'clustercolour': 'ClusterColour', 'clustershape': 'ClusterShape', 'finddupe': 'FindDupe', 'fixcolour': 'FixColour',
         'makeline': 'MakeLine', 'matchregions': 'MatchRegions', 'movetocorner': 'MoveToCorner', 'movetoregion': 'MoveToRegion'}


def _demo_scene(orc, task):
    import magical_b200 as m
    from magical_b200 import tasks
    m.register_envs()
    sp = m.spec(f'{task}-Demo-v0')
    kw = {a: b for a, b in sp['kwargs'].items() if a not in ('res_hw', 'fps', 'phys_steps', 'phys_iter', 'max_episode_steps')}
    rd = kw.pop('rand_dynamics', False)
    sc = tasks.sample_scene(sp['entry_point'], np.random.RandomState(0), sp['max_episode_steps'], rand_dynamics=rd, **kw)
    if task == 'MoveToCorner':
        rb = orc.robot_default_poses(0.4, -0.0, 0.55 * math.pi)
        for i in range(6):
            for k in range(3):
                sc.robot[i][k] = rb[i][k]
    return sc


def _still(key):
    cv2 = pytest.importorskip('cv2')
    im = cv2.imread(os.path.join(STILLS, f'static-{key}-demo-v0.png'), cv2.IMREAD_COLOR)
    assert im is not None and im.shape == (192, 192, 3)
    return im[..., ::-1]


def _check(frame384, key):
    import cv2
    lo = cv2.resize(np.ascontiguousarray(frame384), (192, 192), interpolation=cv2.INTER_AREA)
    d = np.abs(lo.astype(int) - _still(key).astype(int)).max(-1)
    f24, f64 = float((d > 24).mean()), float((d > 64).mean())
    assert f24 <= 0.06 and f64 <= 0.015, (key, f24, f64)
    return f24, f64


@pytest.mark.parametrize('key', sorted(TASKS))
def test_oracle_allo_frame_matches_reference_still(orc, key):
    from conftest import to_oracle_scene
    sc = _demo_scene(orc, TASKS[key])
    frame = orc.Env(to_oracle_scene(orc, sc)).render(0)
    print(key, _check(frame, key))


@pytest.mark.gpu
def test_cuda_allo_frames_match_reference_stills(orc):
    import magical_b200 as m
    keys = sorted(TASKS)
    scenes = [_demo_scene(orc, TASKS[k]) for k in keys]
    env = m.make_mixed([f'{TASKS[k]}-Demo-v0' for k in keys], seed=0, preproc='LoRes4E', obs_keys=('past_obs',))
    env.auto_reset = False
    env.reset(scenes)
    frames = env.render_view(0, True).cpu().numpy()
    for k, fr in zip(keys, frames):
        print(k, _check(fr, k))
    env.close()
