"""Multi-task evaluation protocol, the caller one level above the step path.

Mirrors magical/evaluation.py:
  EvaluationProtocol (:13-98)   same constructor, abstract `run_id` / `obtain_scores`, `do_eval()` returning a
                                DataFrame with the columns demo_env, test_env, mean_score, ci95_lower,
                                ci95_upper, std_score, run_id (one row per test variant of the demo env);
  latexify_results (:101-154)   same table text.
The reference takes the 95 % interval from statsmodels' DescrStatsW(scores).tconfint_mean(0.05): the
two-sided Student-t interval mean +- t(0.975, n-1) * std(ddof=1) / sqrt(n).  statsmodels is not a dependency
here; `t_confint_mean` computes the same formula with scipy.stats.t.

BatchedEvaluationProtocol is the batched form: the `n_rollouts` roll-outs of a test env run as ONE batch of
envs on the GPU (env i seeded seed + i, one episode each), sharded over ranks by env index when
torch.distributed is initialised, and the per-env `eval_score`s are all-gathered (shard.all_gather_scores).
"""
import abc
import collections
import warnings

import numpy as np


def t_confint_mean(scores, alpha=0.05):
    """DescrStatsW(scores).tconfint_mean(alpha, 'two-sided') (evaluation.py:71)."""
    from scipy import stats
    x = np.asarray(scores, dtype=np.float64)
    n = x.shape[0]
    mean = x.mean()
    std_mean = x.std(ddof=1) / np.sqrt(n)          # DescrStatsW.std_mean = std(ddof=0) / sqrt(n - 1)
    half = stats.t.ppf(1.0 - alpha / 2.0, n - 1) * std_mean
    return mean - half, mean + half


class EvaluationProtocol(abc.ABC):
    """Common elements of the multi-task evaluation protocol (evaluation.py:13-98): subclasses say how to
    roll out `n_rollouts` trajectories on a named env; this class aggregates the statistics."""
    _called_init = False

    def __init__(self, demo_env_name, n_rollouts):
        from . import DEMO_ENVS_TO_TEST_ENVS_MAP, register_envs
        register_envs()
        self.n_rollouts = n_rollouts
        self.demo_env_name = demo_env_name
        self.test_env_names = [demo_env_name, *DEMO_ENVS_TO_TEST_ENVS_MAP[demo_env_name]]
        self._called_init = True

    @property
    @abc.abstractmethod
    def run_id(self):
        """String identifying the model / algorithm; becomes the `run_id` column."""

    @abc.abstractmethod
    def obtain_scores(self, env_name):
        """`self.n_rollouts` (or more) `eval_score`s from roll-outs on `env_name`."""

    _COLUMNS = ('demo_env', 'test_env', 'mean_score', 'ci95_lower', 'ci95_upper', 'std_score', 'run_id')

    def _row(self, env_name):
        """Statistics of the first `n_rollouts` scores on one test env."""
        scores = np.asarray(self.obtain_scores(env_name), dtype=np.float64)
        want = self.n_rollouts
        if scores.shape[0] < want:
            raise ValueError(f".obtain_scores() returned only {scores.shape[0]} scores, but we asked for {want} scores")
        if scores.shape[0] > want:
            # extra roll-outs would narrow this row's interval relative to the others: drop them
            warnings.warn(f"Asked for {want} scores but got {scores.shape[0]} scores instead. Will truncate to "
                          f"only consider the first {want} scores.")
            scores = scores[:want]
        ci_lo, ci_hi = t_confint_mean(scores, 0.05)
        values = (self.demo_env_name, env_name, scores.mean(), ci_lo, ci_hi, scores.std(ddof=1), self.run_id)
        return collections.OrderedDict(zip(self._COLUMNS, values))

    def do_eval(self, verbose=False):
        import pandas as pd
        if not self._called_init:
            raise ValueError("EvaluationProtocol.__init__() was not called. Did you include a "
                             "super().__init__(…) call in your subclass?")
        frame = pd.DataFrame.from_records([self._row(name) for name in self.test_env_names], columns=list(self._COLUMNS))
        if verbose:
            print(f"Final mean scores for '{self.run_id}':")
            print(frame[['test_env', 'mean_score', 'ci95_lower', 'ci95_upper']])
        return frame


class BatchedEvaluationProtocol(EvaluationProtocol):
    """`obtain_scores` as one batched roll-out on the GPU.

    policy(obs_dict, env) -> int actions [B] (numpy, list, or an int32 CUDA tensor); the default policy
    draws uniform random actions from RandomState(seed) (the reference's smoke test samples the action
    space the same way, tests/test_rollout_preproc.py:30).  With torch.distributed initialised every rank
    runs its contiguous share of the roll-outs and the scores are all-gathered in roll-out order.
    """

    def __init__(self, demo_env_name, n_rollouts, policy=None, seed=0, run_id='magical_b200', device=0,
                 make_kwargs=None):
        super().__init__(demo_env_name, n_rollouts)
        self._policy = policy
        self._seed = int(seed)
        self._run_id = run_id
        self._device = device
        self._make_kwargs = dict(make_kwargs or {})
        # some variants cannot place every seed (the fork's MoveToCorner draws the robot uniformly, then TestJitter
        # may find it inside a wall: geom.PlacementError); a roll-out worker would call reset() again
        self.placement_retries = 50

    @property
    def run_id(self):
        return self._run_id

    def obtain_scores(self, env_name):
        import torch
        import torch.distributed as dist
        from . import make, shard
        rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_available() and dist.is_initialized() else (0, 1)
        lo, hi = shard.shard_range(self.n_rollouts, rank, world)
        n_local = hi - lo
        # every rank needs at least one env to own a handle; a surplus env's score is simply not gathered
        env = make(env_name, num_envs=max(n_local, 1), seed=self._seed + lo, device=self._device, **self._make_kwargs)
        try:
            env.auto_reset = False
            env.placement_retries = self.placement_retries
            obs = env.reset()
            rng = np.random.RandomState(self._seed + 7919 * (rank + 1))
            scores = None
            for _ in range(env.max_episode_steps):
                if self._policy is None:
                    actions = rng.randint(0, 18, size=env.num_envs)
                else:
                    actions = self._policy(obs, env)
                obs, _, done, info = env.step(actions)
                scores = info['eval_score']
            assert bool(done.all()), 'episodes end by horizon only (base_env.py:284-292)'
            local = scores[:n_local].to(torch.float32)
            if world > 1:
                local = shard.all_gather_scores(local, self.n_rollouts, rank, world)
            return local.cpu().numpy().astype(np.float64)
        finally:
            env.close()


def latexify_results(eval_data, id_column='run_id'):
    """LaTeX table of a frame produced by `do_eval()` (evaluation.py:101-154): one column per test env, one
    row per algorithm, cells 'mean ($\\pm$ std)'.  The text is the reference's, byte for byte (including
    its habit of closing the table after every algorithm row)."""
    envs = list(eval_data['test_env'].unique())
    lines = [r"\centering",
             r"\begin{tabular}{l@{\hspace{1em}}%s}" % ("c" * len(envs)),
             r"\toprule",
             r"\textbf{Randomisation} & " + ' & '.join(r"\textbf{%s}" % e for e in envs) + "\\\\",
             r"\midrule"]
    for alg in eval_data[id_column].unique():
        cells = []
        for env_name in envs:
            rows = eval_data[(eval_data[id_column] == alg) & (eval_data['test_env'] == env_name)]
            if len(rows) != 1:
                raise ValueError(f"got {len(rows)} rows corresponding to {id_column}={alg} and test_env={env_name}, "
                                 f"but expected one (maybe IDs in column {id_column} aren't unique?)")
            row = rows.iloc[0]
            cells.append(f'{row["mean_score"]:.2f} ($\\pm$ {row["std_score"]:.2f})')
        lines += [r"\textbf{%s} & " % alg + ' & '.join(cells) + "\\\\", r"\bottomrule", r"\end{tabular}"]
    return "\n".join(lines) + "\n"
