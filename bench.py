"""bench.py -- env-steps/sec of the MAGICAL hot path with 96x96 LoRes4E observations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg1|cfg2|cfg3|cfg5] [--impl reference]

Workloads (BASELINE.json configs; the default, cfg1, is the one the metric is quoted on):
  cfg1  MoveToCorner-Demo-LoRes4E-v0          4096 envs per GPU   (configs[1])
  cfg2  FindDupe-TestLayout-LoRes4E-v0       16384 envs per GPU   (configs[2])
  cfg3  ClusterColour-TestAll-LoRes4E-v0     32768 envs per GPU, block count forced to 10 (configs[3])
  cfg5  the 60 benchmark ids (8 tasks x all variants) mixed, 4096 envs per GPU, sharded by env index over the
        ranks (shard.shard_env_ids), per-step NCCL all-gather of the episode scores (configs[4])
One "step" = one env-step of every env of the batch: action decode, 10 physics sub-steps, score at the horizon, ego
render at 384^2, 4x area downsample, 4-frame stack.  Episodes are PHASED: env i starts its first episode i mod horizon
steps late (an untimed pre-roll), so on EVERY step about B / horizon envs reach their horizon, report their score and
start their next episode inside the step (mg_queue_resets: reset between the step's physics and its render, scenes
sampled ahead by the reset feeder thread with the reference RNG) -- any timed window, the driver's 20 steps included,
holds its proportional share of horizon steps.  With N > 1 ranks every step also all-gathers the ranks' eval scores over
NCCL (asynchronously, double-buffered).

--impl reference times the CPU oracle port (oracle/, the restated pymunk+pygame path; the reference itself
cannot be imported here: pymunk/pygame/gym are not installed) on all host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    'cfg1': dict(ids=['MoveToCorner-Demo-LoRes4E-v0'], envs=4096, label='BASELINE.json configs[1]'),
    'cfg2': dict(ids=['FindDupe-TestLayout-LoRes4E-v0'], envs=16384, label='BASELINE.json configs[2]'),
    'cfg3': dict(ids=['ClusterColour-TestAll-LoRes4E-v0'], envs=32768, force_blocks=10,
                 label='BASELINE.json configs[3] (CountPlus, block count forced to the maximum, 10)'),
    'cfg5': dict(ids='bench60', envs=4096, reset_pool=3, label='BASELINE.json configs[4]: 8 tasks x all variants (60 ids) mixed'),
}
BENCH_TASKS = ('MoveToCorner', 'MoveToRegion', 'MatchRegions', 'MakeLine', 'FindDupe', 'FixColour', 'ClusterColour', 'ClusterShape')
OBS_BYTES_PER_ENV = 96 * 96 * 12       # algorithmic bytes written per env-step
# dram__bytes_read.sum + dram__bytes_write.sum of one mg_render_kernel launch at 4096 MoveToCorner envs (ncu --set full);
# per env it does not depend on the batch size
NCU_DRAM_BYTES_PER_ENV = (356.310e6 + 537.062e6) / 4096
NCU_TRAFFIC_SOURCE = 'profiles/r2c_ncu_summary.txt (ncu --set full, one launch, 4096 MoveToCorner envs), scaled per env'
METRIC = 'env_steps_per_sec_lores4e'
UNIT = 'env-steps/s'
SCENE_NBLOCKS_WORD = 2                  # int32 index of mg_scene.n_blocks (include/mg_scene.h)


def workload_ids(m, wl):
    """The registered ids of a workload (cfg5: the 60 base ids of the eight benchmark tasks, LoRes4E stacks)."""
    if wl['ids'] != 'bench60':
        return list(wl['ids'])
    ids = [i for i in m.ALL_REGISTERED_ENVS
           if m.spec(i)['preproc'] is None and i.split('-')[0] in BENCH_TASKS and 'DebugReward' not in i]
    assert len(ids) == 60, len(ids)
    return ids


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def sample_scenes(m, env_ids, seed0, force_blocks=None):
    """One host-sampled initial state per entry of `env_ids` (entry i from RandomState(seed0 + i), re-drawn from the
    same stream after a PlacementError / until the block count is `force_blocks`), packed as a ctypes array."""
    from magical_b200 import _lib as L
    from magical_b200 import tasks
    scenes = []
    for i, env_id in enumerate(env_ids):
        sp = m.spec(env_id)
        kw = dict(sp['kwargs'])
        rand_dynamics = kw.pop('rand_dynamics', False)
        for k in ('res_hw', 'fps', 'phys_steps', 'phys_iter', 'max_episode_steps'):
            kw.pop(k)
        rng = np.random.RandomState(seed0 + i)
        while True:
            try:
                sc = tasks.sample_scene(sp['entry_point'], rng, sp['max_episode_steps'], rand_dynamics=rand_dynamics, **kw)
            except tasks.PlacementError:
                continue
            if force_blocks is None or sc.n_blocks == force_blocks:
                break
        scenes.append(sc)
    return L.scenes_array(scenes)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.lines = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                          '-lms', '100', '-i', str(gpu_index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.lines:
            parts = [p.strip() for p in line.split(',')]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); smax.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(smax) if smax else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def cpu_baseline(n_threads, wl_name, seconds_target=12.0):
    """The oracle port timed on the host cores on a bounded sample of the same workload (full reference
    step: 10 sub-steps, allo+ego 384^2 renders, three INTER_AREA resizes, score at the horizon)."""
    import magical_b200 as m
    from oracle import orc
    lib = orc.lib()
    m.register_envs()
    wl = WORKLOADS[wl_name]
    ids = workload_ids(m, wl)
    n_envs = max(max(n_threads, 1) * 2, len(ids))
    env_ids = [ids[i % len(ids)] for i in range(n_envs)]
    packed = sample_scenes(m, env_ids, 0, wl.get('force_blocks'))
    scenes = (orc.Scene * n_envs)()
    C.memmove(scenes, packed, C.sizeof(scenes))
    steps = 20
    actions = np.ascontiguousarray(np.random.RandomState(1234).randint(0, 18, size=(steps, n_envs)), np.int32)
    ck = C.c_uint64(0)
    t0 = time.perf_counter()
    lib.orc_bench_rollout(scenes, n_envs, steps, actions.ctypes.data, n_threads, 1, C.byref(ck))
    dt = time.perf_counter() - t0
    # scale the sample so that the measured run lasts about `seconds_target`
    steps2 = int(max(20, min(6000, steps * seconds_target / max(dt, 1e-3))))
    actions = np.ascontiguousarray(np.random.RandomState(1234).randint(0, 18, size=(steps2, n_envs)), np.int32)
    t0 = time.perf_counter()
    lib.orc_bench_rollout(scenes, n_envs, steps2, actions.ctypes.data, n_threads, 1, C.byref(ck))
    dt = time.perf_counter() - t0
    rate = n_envs * steps2 / dt
    what = ids[0] if len(ids) == 1 else f'{len(ids)} benchmark ids'
    return {'value': rate, 'unit': UNIT, 'cores': n_threads, 'kind': 'port',
            'sample': f'{n_envs} envs x {steps2} steps of {what} (oracle C port of the pymunk+pygame+cv2 path, '
                      f'allo+ego 384^2 render, 3 resizes), {dt:.1f} s'}, n_envs * steps2, dt


def run_reference(args, rank, world):
    """--impl reference: the CPU path on all host cores, rank 0 only."""
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    n_threads = host_cores()
    per_step_s = max(2.0, min(20.0, 150.0 / max(args.steps + args.warmup, 1)))
    times, units = [], []
    for i in range(args.warmup + args.steps):
        cb, n_units, dt = cpu_baseline(n_threads, args.workload, seconds_target=per_step_s)
        if i >= args.warmup:
            times.append(dt); units.append(n_units)
    value = sum(units) / sum(times)
    cb['value'] = value
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': 1e3 * sum(times) / len(times), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': f'{args.workload}: {wl["label"]}; each step = a bounded CPU sample: {cb["sample"]}',
                   'note': 'reference python path (pymunk/pygame/gym) not installable here; oracle C port timed'},
        'cpu_baseline': cb,
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def make_workload_env(m, wl, B, rank, world, local_rank):
    """The rank's env batch of a workload, reset, phased (pre-roll) and ready to step.  Returns (env, description)."""
    from magical_b200 import shard
    ids = workload_ids(m, wl)
    if len(ids) == 1:
        env = m.make(ids[0], num_envs=B, device=local_rank, seed=rank * B, obs_keys=('past_obs',))
        local_ids = None
    else:
        # the global batch is the id list repeated round-robin; every rank owns a contiguous block of it, sorted by id
        # inside the rank so that neighbouring physics lanes see the same task layout (shard.shard_env_ids)
        global_ids = [ids[i % len(ids)] for i in range(world * B)]
        gidx, local_ids = shard.shard_env_ids(global_ids, rank, world)
        env = m.make_mixed([i.replace('-LoRes4E', '') for i in local_ids], device=local_rank, seed=rank * B,
                           preproc='LoRes4E', obs_keys=('past_obs',))
    env.placement_retries = 64         # what a reference roll-out loop does after a PlacementError: reset() again
    force = wl.get('force_blocks')
    n_pool = wl.get('reset_pool', 0)
    ids_all = np.arange(B)
    if force is None and not n_pool:
        env.reset()                    # resets sampled live by the env's feeder thread (reference RNG streams)
        return env, local_ids
    if force is not None:
        # CountPlus at its maximum: the registered id draws 7..10 blocks; keep the draws with `force` blocks as a pool
        # (reference RNG and placement sampler) and restart finished envs from it in turn
        raw = np.frombuffer(env._sample_many(ids_all, packed=True), np.uint8).reshape(B, -1)
        pool = raw[raw.view(np.int32)[:, SCENE_NBLOCKS_WORD] == force].copy()
        assert len(pool) >= 64, len(pool)
        cursor = {'k': 0}

        def from_pool(finished):
            k = cursor['k']
            cursor['k'] = (k + len(finished)) % len(pool)
            return pool[(k + np.arange(len(finished))) % len(pool)]

        first = np.ascontiguousarray(pool[ids_all % len(pool)])
    else:
        # every env restarts from `n_pool` initial states of its own id drawn before the timed region from its own
        # RandomState stream, in turn (cfg5: two of the 60 ids, MoveToCorner-TestJitter / -TestAll, fail ~30 % of their
        # placement draws after ~10 ms of rejection sampling each -- the fork's random robot position -- and would make
        # the host sampler, not the GPU, the thing measured)
        per_env = np.stack([np.frombuffer(env._sample_many(ids_all, packed=True), np.uint8).reshape(B, -1).copy()
                            for _ in range(n_pool + 1)], axis=1)
        turn = np.zeros(B, np.int64)

        def from_pool(finished):
            out = per_env[finished, 1 + turn[finished] % n_pool]
            turn[finished] += 1
            return out

        first = np.ascontiguousarray(per_env[:, 0])
    env.reset_source = from_pool
    env.set_scenes_packed(ids_all, (C.c_uint8 * first.size).from_buffer(first), B)
    from magical_b200 import _lib as L
    L.check(env._h, env._lib.mg_reset_obs(env._h, env._obs.data_ptr(), env._stream()))
    env._needs_reset = False
    return env, local_ids


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import magical_b200 as m
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        # stdout carries the one JSON line: NCCL prints its version banner to stdout when the communicator comes
        # up, so file descriptor 1 points at stderr until the first collective is through
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group('nccl', device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            torch.cuda.synchronize(dev)
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    m.register_envs()
    wl = WORKLOADS[args.workload]
    B = args.envs or wl['envs']
    K, W = args.steps, args.warmup
    env, local_ids = make_workload_env(m, wl, B, rank, world, local_rank)
    horizons = env._horizons
    Hmax = int(horizons.max())
    Kk = max(3, min(K, 40))       # per-kernel timing pass
    Ka = max(3, min(K, 40))       # public-API arm
    Ke = max(3, min(K, 20))       # end-to-end arm
    total_steps = Hmax + W + K + Kk + Ka + Ke + 16
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    # scripted random actions, resident in HBM; re-used cyclically past 512 steps (the pre-roll of long horizons)
    T_act = min(total_steps, 512)
    actions = torch.randint(0, 18, (T_act, B), dtype=torch.int32, device=dev, generator=gen)
    rew = torch.zeros(B, dtype=torch.float32, device=dev)
    done = torch.zeros(B, dtype=torch.uint8, device=dev)
    score = [torch.zeros(B, dtype=torch.float32, device=dev) for _ in range(2)]
    scores_all = [torch.zeros(world * B, dtype=torch.float32, device=dev) for _ in range(2)]
    works = [None, None]
    state = {'t': 0}
    stream = torch.cuda.current_stream(dev)

    def dev_step(restart=None):
        """one pass of the hot path, inputs resident in HBM: resets of the finishing envs queued (host), one mg_step,
        and the ranks' eval scores of this step all-gathered over NCCL while the next step runs"""
        t = state['t']
        p = t & 1
        if works[p] is not None:
            works[p].wait()                 # stream-level: score[p] / scores_all[p] are free again
        env.step_into(actions[t % T_act], rew, done, score[p], restart=restart)
        if world > 1:
            works[p] = dist.all_gather_into_tensor(scores_all[p], score[p], async_op=True)
        state['t'] = t + 1

    def sync_all():
        for p in range(2):
            if works[p] is not None:
                works[p].wait()
                works[p] = None
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # pre-roll (untimed): env i restarts i mod horizon steps into its first episode, which spreads the episode phases
    # evenly; from then on about B / horizon envs finish on every step
    offs = np.arange(B) % horizons
    for t in range(1, Hmax):
        dev_step(restart=np.nonzero(offs == t)[0])
    for _ in range(W):
        dev_step()
    sync_all()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    launches0 = env.launch_count()
    resets0 = int(env._abs_step)
    n_fin = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    host_t0 = time.perf_counter()
    for _ in range(K):
        n_fin += int((env._episode_steps + 1 >= horizons).sum())
        dev_step()
    host_ms = (time.perf_counter() - host_t0) * 1e3 / K      # host time to ISSUE one step (it may block on the device when it runs ahead)
    e1.record(stream)
    sync_all()
    ms_total = e0.elapsed_time(e1)
    launches = env.launch_count() - launches0
    # per-kernel durations for the roofline: the same steps again with the library's CUDA-event timers on, which
    # also serialises the step's kernels (in the timed region above mg_step renders finished parts of the batch
    # under the physics tail of the others, so a render launch there has no duration of its own)
    env.set_timing(True)
    env.kernel_ms(0); env.kernel_ms(1)
    for _ in range(Kk):
        dev_step()
    sync_all()
    env.set_timing(False)
    phys_ms, phys_n = env.kernel_ms(0)
    rend_ms, rend_n = env.kernel_ms(1)
    clock_info = clocks.stop() if clocks else None
    t_ms = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_max = float(t_ms.item())
    value = world * B * K / (ms_max / 1e3)

    # ---- the same loop through the public API: BatchedMagicalEnv.step (auto-reset, tensors cloned per step)
    sync_all()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(stream)
    for _ in range(Ka):
        env.step(actions[state['t'] % T_act])
        state['t'] += 1
    a1.record(stream)
    sync_all()
    api_ms = torch.tensor([a0.elapsed_time(a1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(api_ms, op=dist.ReduceOp.MAX)
    api_value = world * B * Ka / (float(api_ms.item()) / 1e3)

    # ---- e2e: the C-ABI call with HOST buffers (pinned): H2D actions, D2H obs/reward/done/score every step,
    # finishing envs restarted inside the call as above
    obs = env._obs
    a_host = actions.cpu().pin_memory()
    obs_host = torch.empty(obs.shape, dtype=torch.uint8).pin_memory()
    rew_host = torch.empty(B, dtype=torch.float32).pin_memory()
    score_host = torch.empty(B, dtype=torch.float32).pin_memory()
    done_host = torch.empty(B, dtype=torch.uint8).pin_memory()

    def host_step():
        t = state['t']
        env.step_host(a_host[t % T_act], obs_host, rew_host, done_host, score_host)
        state['t'] = t + 1

    def time_host_steps(warm):
        for _ in range(warm):
            host_step()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(Ke):
            host_step()
            if os.environ.get('MG_BENCH_DEBUG'):
                print('e2e step', '%.3f' % ((time.perf_counter() - t0) * 1e3), env.host_assembly(), file=sys.stderr)
        torch.cuda.synchronize(dev)
        s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(s, op=dist.ReduceOp.MAX)
        return world * B * Ke / float(s.item())

    # (a) every stacked observation crosses the link whole (110 592 B per env-step)
    e2e_direct = time_host_steps(min(W, 3))
    obs_bytes_env = obs.numel() // B
    e2e_value, e2e_d2h, e2e_mode = e2e_direct, world * (obs.numel() + B * 9), 'every stacked observation copied whole'
    asm_info = None
    if args.e2e_assembly != 'off':
        # (b) mg_set_host_assembly: part of the batch sends only its new frame and host threads build those stacks
        # (same bytes in obs_host); the handle balances link against host over the warm-up steps
        local_world = int(os.environ.get('LOCAL_WORLD_SIZE', world))
        threads = int(os.environ.get('MG_BENCH_E2E_THREADS', 0)) or max(1, host_cores() // max(local_world, 1))
        env.set_host_assembly('auto', 0, threads)
        e2e_asm = time_host_steps(36)
        parts, total, threads, copy_ms, asm_ms = env.host_assembly()
        env.set_host_assembly('off')
        asm_info = {'value': e2e_asm, 'parts_built_on_host': parts, 'parts': total, 'host_threads': threads,
                    'last_step_copy_ms': copy_ms, 'last_step_host_ms': asm_ms}
        if e2e_asm > e2e_direct:
            n_asm_envs = B * parts // total
            e2e_value = e2e_asm
            e2e_d2h = world * (n_asm_envs * (obs_bytes_env // 4) + (B - n_asm_envs) * obs_bytes_env + B * 9)
            e2e_mode = (f'mg_set_host_assembly(auto): {parts} of {total} parts of the batch send only their new 96x96x3 frame and '
                        f'{threads} host threads build those stacked observations in the pinned output buffer')
    status = env.status()
    if rank == 0:
        peaks_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))['hbm_gbs'], 'MEASURED_PEAKS.json hbm_gbs'
        else:
            peak, peak_src = 6650.0, 'fallback (B200_PROFILING.md)'
        rend_avg_ms = rend_ms / max(rend_n, 1)
        achieved = B * OBS_BYTES_PER_ENV / (rend_avg_ms * 1e-3) / 1e9
        step_gbs = value / world * OBS_BYTES_PER_ENV / 1e9
        cb = None
        if world == 1 and not args.no_cpu_baseline:
            cb, _, _ = cpu_baseline(host_cores(), args.workload)
        ids = workload_ids(m, wl)
        what = ids[0] if len(ids) == 1 else f'{len(ids)} benchmark ids (8 tasks x all variants) mixed, sharded by env index'
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W,
            'ms_per_step': ms_max / K, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': f'{args.workload}: {what} x {B} envs per GPU ({wl["label"]}), scripted random actions, '
                                   f'horizons {sorted(set(horizons.tolist()))}, episodes phased so that ~B/horizon envs finish on every step: '
                                   f'{n_fin} episode ends + restarts inside the {K} timed steps on rank 0 (mg_queue_resets; scenes ' + ('from the reset feeder thread, reference RNG streams sampled live' if env.reset_source is None else 'from a pool sampled before the timed region with the reference RNG') + ')'
                                   + (f'; per-step NCCL all-gather of {world} x {B} eval scores' if world > 1 else ''),
                       'envs_per_gpu': B, 'obs': 'u8[B,96,96,12] past_obs (4 ego frames)',
                       'l2': f'per-step working set (obs {B * OBS_BYTES_PER_ENV / 1e6:.0f} MB + rings + state) exceeds the 126 MB L2',
                       'parallelism': f'env-sharded x{world}, no data-path collective; score all-gather only',
                       'episode_ends_in_timed_region': n_fin, 'status_flags_raised': int((status != 0).sum())},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'frac_step': step_gbs / peak, 'achieved_step': step_gbs,
                         'traffic': B * NCU_DRAM_BYTES_PER_ENV, 'traffic_source': NCU_TRAFFIC_SOURCE,
                         'kernel': 'mg_render_kernel', 'peak_source': peak_src,
                         'kernel_ms_avg': rend_avg_ms, 'physics_kernel_ms_avg': phys_ms / max(phys_n, 1),
                         'kernel_timing': f'CUDA events around every launch over {Kk} further steps of the same roll-out with the step\'s kernels serialised (mg_set_timing); in the timed region the render launches of finished parts overlap the physics tail',
                         'note': 'frac: algorithmic bytes (110592 B x envs per launch) / render-kernel time; frac_step: SURVEY 8(d) / BASELINE.md 4 definition, env-steps/s per GPU x 110592 B / peak (whole step). The kernels are issue/latency-bound, not HBM-bound'},
            'cpu_baseline': cb,
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': world * B * 4,
                    'd2h_bytes_per_step': e2e_d2h, 'steps': Ke, 'how': e2e_mode,
                    'direct': {'value': e2e_direct, 'd2h_bytes_per_step': world * (obs.numel() + B * 9)},
                    'host_assembly': asm_info},
            'api': {'value': api_value, 'unit': UNIT, 'steps': Ka, 'call': 'BatchedMagicalEnv.step (auto_reset=True, device tensors)'},
            'gpu_launches': launches, 'host_issue_ms_per_step': host_ms,
            'clocks': clock_info,
        }
        print(json.dumps(line), flush=True)
    env.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=160)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='cfg1', choices=sorted(WORKLOADS))
    ap.add_argument('--envs', type=int, default=0, help='envs per GPU (default: the workload\'s)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--e2e-assembly', default='auto', choices=['auto', 'off'],
                    help='end-to-end arm: also time mg_step_host with mg_set_host_assembly (new frame over the link, stacks built by host threads)')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        if args.steps > 20:
            args.steps = 20
        run_reference(args, rank, world)
        return
    run_gpu(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
