#!/usr/bin/env python
"""Record golden roll-outs from the REAL reference (pymunk + pygame + gym), BASELINE.md section 3(i).

Needs an environment where `import magical` works (this container and the GPU box do not: pymunk, pygame and gym
are not installed and there is no network -- the probe below prints what is missing and exits 2, and the loader
test tests/test_reference_rollouts.py then skips with "PARITY UNPINNED").  Run it wherever the reference installs:

    PYTHONPATH=/path/to/magical-1 python tools/record_reference_goldens.py [--out tests/golden/reference_rollouts]

For each env id (the BASELINE.json configs' ids plus one id per remaining task) and each seed 0..7 it writes
<id>_seed<k>.npz holding, with the seeding protocol of the reference's own roll-out test
(tests/test_rollout_preproc.py:17-36: env.seed(k) before reset):
  bodies0      f64 [n_bodies, 3]   x, y, angle after reset(), canonical order: robot, control, pupil L/R, finger L/R, blocks
  traj         f64 [100, n_bodies, 3]  the same after each of 100 steps
  actions      i64 [100]           RandomState(1234 + k).randint(0, 18, 100)
  rewards      f64 [100], scores f64 [100] (info['eval_score']), dones bool [100]
  allo_<t>, ego_<t>  u8 [384,384,3]  render('rgb_array') at steps t = 0, 1, 10, 50, 100
  past_obs_<t> u8 [96,96,12]       only for LoRes4E ids
  versions     str                 pymunk / pygame / gym / cv2 versions
tests/test_reference_rollouts.py replays them through the oracle and the CUDA path: initial state bit-exact (reset
sampler), poses within 1e-3 after 100 steps, rewards / scores within 1e-5, <= 1 % of pixels off by more than 8 levels.
"""
import argparse
import os
import sys

import numpy as np

IDS = ['MoveToCorner-Demo-v0', 'MoveToCorner-Demo-LoRes4E-v0', 'FindDupe-TestLayout-LoRes4E-v0',
       'ClusterColour-TestAll-LoRes4E-v0', 'MoveToRegion-TestAll-v0', 'MatchRegions-TestAll-v0', 'MakeLine-TestAll-v0',
       'FixColour-TestAll-v0', 'ClusterShape-TestAll-v0', 'PickAndPlace-Demo-v0']
FRAME_STEPS = (0, 1, 10, 50, 100)


def probe():
    missing = []
    vers = {}
    for name in ('pymunk', 'pygame', 'gym', 'cv2', 'magical'):
        try:
            mod = __import__(name)
            v = getattr(mod, '__version__', None) or getattr(getattr(mod, 'version', None), 'ver', None) or getattr(mod, 'version', '?')
            vers[name] = str(v)
        except Exception as exc:      # noqa: BLE001
            missing.append(f'{name} ({type(exc).__name__}: {exc})')
    return vers, missing


def bodies_of(base):
    """x, y, angle of every dynamic / kinematic body in the canonical order of mg_get_bodies."""
    robot = base._robot
    bodies = [robot.robot_body, robot.control_body, *robot.pupil_bodies, *robot.finger_bodies]
    for ent in base._entities:
        if hasattr(ent, 'shape_body'):
            bodies.append(ent.shape_body)
    return np.asarray([[b.position.x, b.position.y, b.angle] for b in bodies])


def record(env_id, seed, out_dir, versions):
    import gym
    env = gym.make(env_id)
    base = env.unwrapped
    env.seed(seed)
    obs = env.reset()
    data = {'versions': np.asarray(versions), 'env_id': np.asarray(env_id), 'seed': np.asarray(seed)}
    actions = np.random.RandomState(1234 + seed).randint(0, 18, size=100)

    def frames(t, obs):
        rgb = base.render(mode='rgb_array')
        data[f'allo_{t}'], data[f'ego_{t}'] = np.asarray(rgb['allo']), np.asarray(rgb['ego'])
        if isinstance(obs, dict) and 'past_obs' in obs:
            data[f'past_obs_{t}'] = np.asarray(obs['past_obs'])

    data['bodies0'] = bodies_of(base)
    frames(0, obs)
    traj, rewards, scores, dones = [], [], [], []
    for t in range(100):
        obs, rew, done, info = env.step(int(actions[t]))
        traj.append(bodies_of(base)); rewards.append(float(rew)); scores.append(float(info['eval_score'])); dones.append(bool(done))
        if t + 1 in FRAME_STEPS:
            frames(t + 1, obs)
        if done:
            break       # episodes shorter than 100 steps (MoveToRegion: 40) stop at their horizon
    data.update(traj=np.asarray(traj), actions=actions, rewards=np.asarray(rewards), scores=np.asarray(scores),
                dones=np.asarray(dones))
    np.savez_compressed(os.path.join(out_dir, f'{env_id}_seed{seed}.npz'), **data)
    env.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                  'tests', 'golden', 'reference_rollouts'))
    ap.add_argument('--seeds', type=int, default=8)
    ap.add_argument('--probe', action='store_true', help='only print the import probe')
    args = ap.parse_args()
    vers, missing = probe()
    print('import probe:', vers, 'missing:', missing)
    if missing:
        print('PARITY UNPINNED: the reference cannot be imported here; nothing recorded')
        return 2
    if args.probe:
        return 0
    import magical
    magical.register_envs()
    os.makedirs(args.out, exist_ok=True)
    versions = ' '.join(f'{k}={v}' for k, v in vers.items())
    for env_id in IDS:
        for seed in range(args.seeds):
            record(env_id, seed, args.out, versions)
            print('recorded', env_id, seed)
    return 0


if __name__ == '__main__':
    sys.exit(main())
