"""Dev-time generator of tests/golden/evaluation.json from the REFERENCE's magical/evaluation.py source text
(run in the build container only; the GPU box never reads /root/reference).

`import magical.evaluation` fails here (gym / statsmodels are not installed), so the module's source is executed
with two stand-ins: DEMO_ENVS_TO_TEST_ENVS_MAP (the registry's map) and statsmodels' DescrStatsW, replaced
by the textbook Student-t interval it implements (mean +- t(1 - alpha/2, n-1) * std(ddof=1)/sqrt(n)) -- so the
interval columns are pinned to that formula, everything else (row order, truncation, columns, std, the LaTeX
text) to the reference's own code."""
import json
import os
import sys
import types
import warnings

import numpy as np
from scipy import stats

REF = '/root/reference/magical/evaluation.py'
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden', 'evaluation.json')


class DescrStatsW:
    def __init__(self, x):
        self.x = np.asarray(x, dtype=np.float64)

    def tconfint_mean(self, alpha, alternative):
        assert alternative == 'two-sided'
        n = len(self.x)
        half = stats.t.ppf(1 - alpha / 2, n - 1) * self.x.std(ddof=1) / np.sqrt(n)
        return self.x.mean() - half, self.x.mean() + half


def main():
    src = open(REF).read()
    src = src.replace('from statsmodels.stats.weightstats import DescrStatsW', '')
    src = src.replace('from magical.benchmarks import DEMO_ENVS_TO_TEST_ENVS_MAP', '')
    # the demo -> test-variant map of the registry (checked against register_envs() by tools/gen_registry.py)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import magical_b200
    magical_b200.register_envs()
    test_map = {k: list(v) for k, v in magical_b200.DEMO_ENVS_TO_TEST_ENVS_MAP.items()}
    ns = {'DescrStatsW': DescrStatsW, 'DEMO_ENVS_TO_TEST_ENVS_MAP': test_map, '__name__': 'ref_evaluation'}
    exec(compile(src, REF, 'exec'), ns)
    rng = np.random.RandomState(11)
    cases = []
    frames = []
    for demo, n_rollouts, extra, run_id in [('MoveToCorner-Demo-v0', 10, 0, 'alg-a'), ('MoveToCorner-Demo-v0', 10, 3, 'alg-b'),
                                            ('MatchRegions-Demo-v0', 5, 0, 'alg-a')]:
        scores = {name: rng.uniform(0, 1, n_rollouts + extra).round(4).tolist() for name in [demo] + test_map[demo]}

        class Proto(ns['EvaluationProtocol']):
            run_id = None

            def obtain_scores(self, env_name):
                return scores[env_name]
        Proto.run_id = run_id
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            frame = Proto(demo, n_rollouts).do_eval()
        frames.append(frame)
        cases.append({'demo_env': demo, 'n_rollouts': n_rollouts, 'run_id': run_id, 'scores': scores,
                      'columns': list(frame.columns), 'records': json.loads(frame.to_json(orient='records'))})
    import pandas as pd
    both = pd.concat(frames[:2], ignore_index=True)
    latex = ns['latexify_results'](both)
    json.dump({'cases': cases, 'latex_input_cases': [0, 1], 'latex': latex}, open(OUT, 'w'), indent=1)
    print(latex)


if __name__ == '__main__':
    sys.exit(main())
