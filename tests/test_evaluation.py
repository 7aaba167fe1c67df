"""EvaluationProtocol / latexify_results against fixtures produced by the reference's own evaluation.py source
(tools/gen_goldens_eval.py -> tests/golden/evaluation.json); reference: magical/evaluation.py:13-154."""
import json
import os
import warnings

import numpy as np
import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'evaluation.json')))


def _proto(case):
    from magical_b200.evaluation import EvaluationProtocol

    class P(EvaluationProtocol):
        run_id = case['run_id']

        def obtain_scores(self, env_name):
            return case['scores'][env_name]
    return P(case['demo_env'], case['n_rollouts'])


def test_t_interval_known_answers():
    from magical_b200.evaluation import t_confint_mean
    lo, hi = t_confint_mean([0.0, 1.0])           # mean .5, std/sqrt(n) = .5, t(0.975, 1) = 12.7062047
    assert abs(lo - (0.5 - 6.35310237)) < 1e-6 and abs(hi - (0.5 + 6.35310237)) < 1e-6
    lo, hi = t_confint_mean([1, 2, 3, 4, 5])      # t(0.975, 4) = 2.77644511, std_mean = sqrt(2.5 / 5)
    assert abs(hi - (3 + 2.77644511 * np.sqrt(0.5))) < 1e-6 and abs(lo + hi - 6) < 1e-12


@pytest.mark.parametrize('k', range(len(GOLD['cases'])))
def test_do_eval_matches_reference_frame(k):
    case = GOLD['cases'][k]
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        frame = _proto(case).do_eval()
    assert list(frame.columns) == case['columns']
    extra = any(len(v) > case['n_rollouts'] for v in case['scores'].values())
    assert bool(w) == extra                          # the truncation warning, and only then
    got = json.loads(frame.to_json(orient='records'))
    assert len(got) == len(case['records'])
    for g, r in zip(got, case['records']):
        for key in case['columns']:
            if isinstance(r[key], float):
                assert abs(g[key] - r[key]) < 1e-9, (key, g[key], r[key])
            else:
                assert g[key] == r[key]


def test_test_env_list_is_demo_then_registered_variants():
    import magical_b200 as m
    m.register_envs()
    p = _proto(GOLD['cases'][0])
    assert p.test_env_names[0] == 'MoveToCorner-Demo-v0'
    assert tuple(p.test_env_names[1:]) == tuple(m.DEMO_ENVS_TO_TEST_ENVS_MAP['MoveToCorner-Demo-v0'])


def test_too_few_scores_raise():
    case = dict(GOLD['cases'][0])
    case['n_rollouts'] = 100
    with pytest.raises(ValueError):
        _proto(case).do_eval()


def test_latexify_is_the_reference_text():
    import pandas as pd
    from magical_b200.evaluation import latexify_results
    frames = []
    for k in GOLD['latex_input_cases']:
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            frames.append(_proto(GOLD['cases'][k]).do_eval())
    both = pd.concat(frames, ignore_index=True)
    assert latexify_results(both) == GOLD['latex']
    dup = pd.concat([both, both], ignore_index=True)
    with pytest.raises(ValueError):
        latexify_results(dup)


@pytest.mark.gpu
def test_batched_protocol_random_policy():
    """One batch per test variant; scores equal a plain roll-out of the same seeds and actions."""
    import torch
    import magical_b200 as m
    from magical_b200.evaluation import BatchedEvaluationProtocol
    n = 12
    acts = np.random.RandomState(5).randint(0, 18, size=(80, n))
    t = {'i': 0}

    def policy(obs, env):
        a = acts[t['i'] % 80]
        t['i'] += 1
        return a
    p = BatchedEvaluationProtocol('MoveToCorner-Demo-v0', n, policy=policy, seed=3, run_id='scripted')
    frame = p.do_eval()
    assert list(frame['test_env']) == p.test_env_names and (frame['run_id'] == 'scripted').all()
    assert ((frame['mean_score'] >= 0) & (frame['mean_score'] <= 1)).all()
    assert (frame['ci95_lower'] <= frame['mean_score']).all() and (frame['mean_score'] <= frame['ci95_upper']).all()
    env = m.make('MoveToCorner-Demo-v0', num_envs=n, seed=3, obs_keys=())
    env.auto_reset = False
    env.reset()
    for s in range(80):
        _, _, done, info = env.step(acts[s])
    want = info['eval_score'].cpu().numpy().astype(np.float64)
    assert bool(done.all())
    assert abs(frame['mean_score'][0] - want.mean()) < 1e-9
    assert abs(frame['std_score'][0] - want.std(ddof=1)) < 1e-9
