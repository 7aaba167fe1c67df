#!/usr/bin/env python
"""bench.py -- env-steps/sec of the MAGICAL hot path with 96x96 LoRes4E observations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): MoveToCorner-Demo-LoRes4E-v0, 4096 envs per GPU, scripted random
actions, envs sharded by index across ranks with no data-path collective (weak scaling); one NCCL
all-gather of episode scores when episodes end.  One "step" = one env-step of every env of the batch:
action decode, 10 physics sub-steps, score at the horizon, ego render at 384^2, 4x area downsample, 4-frame
stack.  Finished episodes are reset inside the timed region from initial states the native reset engine sampled
on host threads while the previous episode was running (reference RNG, one RandomState stream per env).

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

ENV_ID = 'MoveToCorner-Demo-LoRes4E-v0'
ENVS_PER_GPU = 4096
OBS_BYTES_PER_ENV = 96 * 96 * 12       # algorithmic bytes written per env-step
# dram__bytes_read.sum + dram__bytes_write.sum of one mg_render_kernel launch at 4096 envs (ncu --set full,
# profiles/r1_ncu_summary.txt): 361.9 MB + 546.3 MB; per env it does not depend on the batch size
NCU_DRAM_BYTES_PER_ENV = (359.468e6 + 513.831e6) / 4096
NCU_TRAFFIC_SOURCE = 'profiles/r1e_ncu_summary.txt (ncu --set full, one launch, 4096 envs), scaled per env'
METRIC = 'env_steps_per_sec_lores4e'
UNIT = 'env-steps/s'


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def sample_pool(m, num_envs, n_episodes, seed0):
    """Initial states for `n_episodes` consecutive episodes of every env: env i draws from its own
    RandomState(seed0 + i), episode after episode, exactly as successive reset() calls would."""
    from magical_b200 import _lib as L
    from magical_b200 import tasks
    sp = m.spec(ENV_ID)
    kw = dict(sp['kwargs'])
    rand_dynamics = kw.pop('rand_dynamics', False)
    for k in ('res_hw', 'fps', 'phys_steps', 'phys_iter', 'max_episode_steps'):
        kw.pop(k)
    rngs = [np.random.RandomState(seed0 + i) for i in range(num_envs)]
    pool = []
    for _ in range(n_episodes):
        scenes = [tasks.sample_scene(sp['entry_point'], rngs[i], sp['max_episode_steps'], rand_dynamics=rand_dynamics, **kw)
                  for i in range(num_envs)]
        pool.append(L.scenes_array(scenes))
    return pool, sp['max_episode_steps']


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


def cpu_baseline(n_threads, seconds_target=12.0):
    """The oracle port timed on the host cores on a bounded sample of the same workload (full reference
    step: 10 sub-steps, allo+ego 384^2 renders, three INTER_AREA resizes, score at the horizon)."""
    import magical_b200 as m
    from oracle import orc
    lib = orc.lib()
    m.register_envs()
    n_envs = max(n_threads, 1) * 2
    pool, horizon = sample_pool(m, n_envs, 1, 0)
    scenes = (orc.Scene * n_envs)()
    C.memmove(scenes, pool[0], C.sizeof(scenes))
    steps = 20
    actions = np.ascontiguousarray(np.random.RandomState(1234).randint(0, 18, size=(steps, n_envs)), np.int32)
    ck = C.c_uint64(0)
    t0 = time.perf_counter()
    lib.orc_bench_rollout(scenes, n_envs, steps, actions.ctypes.data, n_threads, 1, C.byref(ck))
    dt = time.perf_counter() - t0
    rate = n_envs * steps / dt
    # scale the sample so that the measured run lasts about `seconds_target`
    steps2 = int(max(20, min(6000, steps * seconds_target / max(dt, 1e-3))))
    actions = np.ascontiguousarray(np.random.RandomState(1234).randint(0, 18, size=(steps2, n_envs)), np.int32)
    t0 = time.perf_counter()
    lib.orc_bench_rollout(scenes, n_envs, steps2, actions.ctypes.data, n_threads, 1, C.byref(ck))
    dt = time.perf_counter() - t0
    rate = n_envs * steps2 / dt
    return {'value': rate, 'unit': UNIT, 'cores': n_threads, 'kind': 'port',
            'sample': f'{n_envs} envs x {steps2} steps of {ENV_ID} (oracle C port of the pymunk+pygame+cv2 path, '
                      f'allo+ego 384^2 render, 3 resizes), {dt:.1f} s'}, n_envs * steps2, dt


def run_reference(args, rank, world):
    """--impl reference: the CPU path on all host cores, rank 0 only."""
    if rank != 0:
        return
    n_threads = host_cores()
    per_step_s = max(2.0, min(20.0, 150.0 / max(args.steps + args.warmup, 1)))
    times, units = [], []
    for i in range(args.warmup + args.steps):
        cb, n_units, dt = cpu_baseline(n_threads, seconds_target=per_step_s)
        if i >= args.warmup:
            times.append(dt); units.append(n_units)
    value = sum(units) / sum(times)
    cb['value'] = value
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': 1e3 * sum(times) / len(times), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': f'{ENV_ID} (BASELINE.json configs[1]); each step = a bounded CPU sample: {cb["sample"]}',
                   'note': 'reference python path (pymunk/pygame/gym) not installable here; oracle C port timed'},
        'cpu_baseline': cb,
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


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
    B = args.envs
    K, W = args.steps, args.warmup
    total_steps = 3 * (W + K) + 8
    horizon = m.spec(ENV_ID)['max_episode_steps']
    env = m.make(ENV_ID, num_envs=B, device=local_rank, seed=rank * B, obs_keys=('past_obs',))
    env.auto_reset = False
    ids = np.arange(B, dtype=np.int32)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    actions = torch.randint(0, 18, (total_steps, B), dtype=torch.int32, device=dev, generator=gen)
    scores_all = torch.zeros(world * B, dtype=torch.float32, device=dev)
    state = {'episode': 0, 't': 0}
    lib, h = env._lib, env._h
    from magical_b200 import _lib as L
    obs, rew, done, score = env._obs, env._reward, env._done, env._score
    stream = torch.cuda.current_stream(dev)

    # initial states of the next episode: the native reset engine (every env on its own RandomState stream)
    # on host threads, started as soon as the previous reset is installed so that it runs under the episode
    import threading
    feeder = {'thread': None, 'scenes': None}

    def feed():
        feeder['scenes'] = env._sample_many(ids, packed=True)

    def start_feed():
        feeder['thread'] = threading.Thread(target=feed)
        feeder['thread'].start()

    def reset_all():
        feeder['thread'].join()
        env.set_scenes_packed(ids, feeder['scenes'], B)
        state['episode'] += 1
        start_feed()

    start_feed()

    def dev_step():
        """one pass of the hot path, inputs resident in HBM"""
        t = state['t']
        st = stream.cuda_stream
        if env._episode_steps[0] + 1 < horizon:
            # no episode ends on this step: the one-call step of the C ABI
            L.check(h, lib.mg_step(h, actions[t].data_ptr(), obs.data_ptr(), rew.data_ptr(), done.data_ptr(), score.data_ptr(), st))
            env._episode_steps += 1
            state['t'] = t + 1
            return
        L.check(h, lib.mg_step_physics(h, actions[t].data_ptr(), rew.data_ptr(), done.data_ptr(), score.data_ptr(), st))
        env._episode_steps += 1
        if env._episode_steps[0] >= horizon:
            # episode scores leave the device path here: all-gather over ranks (north_star), then reset
            if world > 1:
                dist.all_gather_into_tensor(scores_all, score)
            reset_all()
        L.check(h, lib.mg_step_render(h, obs.data_ptr(), st))
        state['t'] = t + 1

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    reset_all()
    L.check(h, lib.mg_reset_obs(h, obs.data_ptr(), stream.cuda_stream))
    for _ in range(W):
        dev_step()
    sync_all()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    launches0 = env.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(K):
        dev_step()
    e1.record(stream)
    sync_all()
    ms_total = e0.elapsed_time(e1)
    launches = env.launch_count() - launches0
    # per-kernel durations for the roofline: the same steps again with the library's CUDA-event timers on, which
    # also serialises the step's kernels (in the timed region above mg_step renders finished parts of the batch
    # under the physics tail of the others, so a render launch there has no duration of its own)
    Kk = max(3, min(K, 40))
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

    # ---- e2e: the C-ABI call with HOST buffers (pinned): H2D actions, D2H obs/reward/done/score every step
    a_host = actions.cpu().pin_memory()
    obs_host = torch.empty(obs.shape, dtype=torch.uint8).pin_memory()
    rew_host = torch.empty(B, dtype=torch.float32).pin_memory()
    score_host = torch.empty(B, dtype=torch.float32).pin_memory()
    done_host = torch.empty(B, dtype=torch.uint8).pin_memory()

    def host_step():
        t = state['t']
        env.step_host(a_host[t], obs_host, rew_host, done_host, score_host)
        if env._episode_steps[0] >= horizon:
            reset_all()
        state['t'] = t + 1

    for _ in range(min(W, 3)):
        host_step()
    sync_all()
    Ke = max(3, min(K, 20))
    t0 = time.perf_counter()
    for _ in range(Ke):
        host_step()
    torch.cuda.synchronize(dev)
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = world * B * Ke / float(e2e_s.item())
    if rank == 0:
        peaks_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))['hbm_gbs'], 'MEASURED_PEAKS.json hbm_gbs'
        else:
            peak, peak_src = 6650.0, 'fallback (B200_PROFILING.md)'
        rend_avg_ms = rend_ms / max(rend_n, 1)
        achieved = B * OBS_BYTES_PER_ENV / (rend_avg_ms * 1e-3) / 1e9
        cb = None
        if world == 1 and not args.no_cpu_baseline:
            cb, _, _ = cpu_baseline(host_cores())
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': K, 'warmup': W,
            'ms_per_step': ms_max / K, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': f'{ENV_ID} x {B} envs per GPU (BASELINE.json configs[1]), scripted random actions, '
                                   f'horizon {horizon}, resets inside the timed region (native reset engine on host threads, sampled while the previous episode runs)',
                       'envs_per_gpu': B, 'obs': 'u8[B,96,96,12] past_obs (4 ego frames)',
                       'l2': f'per-step working set (obs {B * OBS_BYTES_PER_ENV / 1e6:.0f} MB + rings + state) exceeds the 126 MB L2',
                       'parallelism': f'env-sharded x{world}, no data-path collective'},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': B * NCU_DRAM_BYTES_PER_ENV, 'traffic_source': NCU_TRAFFIC_SOURCE,
                         'kernel': 'mg_render_kernel', 'peak_source': peak_src,
                         'kernel_ms_avg': rend_avg_ms, 'physics_kernel_ms_avg': phys_ms / max(phys_n, 1),
                         'kernel_timing': f'CUDA events around every launch over {Kk} further steps of the same roll-out with the step\'s kernels serialised (mg_set_timing); in the timed region the render launches of finished parts overlap the physics tail',
                         'note': 'algorithmic bytes = 110592 B x envs per launch; the kernel is issue/latency-bound, not HBM-bound'},
            'cpu_baseline': cb,
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': world * B * 4,
                    'd2h_bytes_per_step': world * (obs.numel() + B * 9), 'steps': Ke},
            'gpu_launches': launches,
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
    ap.add_argument('--envs', type=int, default=ENVS_PER_GPU)
    ap.add_argument('--no-cpu-baseline', action='store_true')
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
