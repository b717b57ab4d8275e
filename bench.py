#!/usr/bin/env python
"""Headline benchmark of the CIDEr-D reward path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the CPU scorer, all host cores

A "step" is one pass of the hot path over one batch: the CST_MS_SCB* workload of
BASELINE.json configs[4] -- 65,536 synthetic videos x 20 sampled captions, each
scored against the video's 20 references, SCB* baseline (scb_baseline=2,
scb_captions=20, use_eos=1, expand_feat=1: the README recipe of the reference).
One process per GPU; every rank owns its own 65,536 videos (weak scaling), the df
table is replicated, and the only exchange is an all-reduce of the reward
statistics.  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# stdout carries the ONE JSON line and nothing else: library chatter (NCCL prints its
# version to stdout) is sent to stderr by pointing fd 1 there for the whole run.
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_JSON_FD, (json.dumps(line) + '\n').encode())


METRIC = 'cider_d_reward_captions_per_sec_20refs'
UNIT = 'captions/s'
SEQ_PER_IMG = 20
N_REFS = 20
DF_VIDEOS = 6513
CAND_W, REF_W = 28, 30
# int32 candidate rows + int32 ref rows read once + fp32 reward written
BYTES_PER_CAPTION = (4 * CAND_W * SEQ_PER_IMG + 4 * REF_W * N_REFS + 4 * SEQ_PER_IMG) / SEQ_PER_IMG


def workload_config(n_videos, n_gpus):
    return {
        'workload': 'CST_MS_SCB* (BASELINE.json configs[4]): %d videos x %d sampled captions vs %d refs per GPU, '
                    'SCB* baseline k=20, use_eos=1' % (n_videos, SEQ_PER_IMG, N_REFS),
        'videos_per_gpu': n_videos, 'captions_per_step_per_gpu': n_videos * SEQ_PER_IMG,
        'seq_per_img': SEQ_PER_IMG, 'refs_per_video': N_REFS, 'scb_baseline': 2, 'scb_captions': 20,
        'df_corpus': '%dx%d synthetic captions, table built on device' % (DF_VIDEOS, N_REFS),
        'parallelism': 'videos sharded, dp%d, df table replicated' % n_gpus,
        'l2_policy': 'inputs (%.0f MB/step) exceed the 126 MB L2; no flush needed'
                     % (n_videos * SEQ_PER_IMG * BYTES_PER_CAPTION / 1e6),
    }


# ---------------------------------------------------------------------------
# clocks during the timed region
# ---------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: 'hw_slowdown', 0x40: 'hw_thermal_slowdown', 0x20: 'sw_thermal_slowdown',
               0x4: 'sw_power_cap', 0x80: 'hw_power_brake_slowdown', 0x2: 'applications_clocks_setting'}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                for bit, name in self.REASONS.items():
                    if bits & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.01)

    def __enter__(self):
        if self._nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': ['unavailable']}
        return {'sm_mhz': float(np.median(self.samples)), 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(self.samples)}


def bind_to_gpu_numa_node(index):
    """Pin this rank to the CPUs NVML reports as local to its GPU, so the pinned host buffers
    of the end-to-end leg are first-touched on that socket.  Best effort; a no-op on the
    measured pool, whose boxes expose one NUMA node (profiles/r1k_h2d_scaling_n8.jsonl)."""
    try:
        import math
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, math.ceil((os.cpu_count() or 64) / 64))
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        allowed = set(os.sched_getaffinity(0))
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def load_synth():
    """The synthetic-corpus module WITHOUT importing the package: its __init__ dlopens
    libcst_reward.so, and the reference arm must not map any of this repo's native code."""
    import importlib.util
    spec = importlib.util.spec_from_file_location('cst_synth', os.path.join(ROOT, 'cst_captioning_b200', 'synth.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def measured_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        return float(json.load(open(path))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


def ncu_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture."""
    try:
        return json.load(open(os.path.join(ROOT, 'profiles', 'roofline_traffic.json')))
    except Exception:
        return None


# ---------------------------------------------------------------------------
# CPU arms (oracle port; the reference itself is Python whose scorer is an absent
# dependency, so there is no oracle/_ref to compile)
# ---------------------------------------------------------------------------
_CPU = {}


def _cpu_init(df, n_doc):
    from oracle import ciderd_oracle
    _CPU['scorer'] = ciderd_oracle.CiderD(df=(df, n_doc))


def _cpu_shard(args):
    """Oracle port of utils.get_cst_reward on a shard -> (scores, rewards) per caption, float64.
    The scorer runs once (score_rows is what utils.py:249-284 does); the baseline block of
    utils.py:292-322 is then applied to those scores through the function's own WXE entry
    (bcmrscores given, use_mixer=0 -> scores = bcmrscores.copy()), so both come out of one pass."""
    from oracle import reward_oracle
    sampled, refs = args
    rows = sampled.astype(np.int64)
    gts = [refs[v].astype(np.int64) for v in range(refs.shape[0])]
    scores = reward_oracle.score_rows(list(rows), gts, _CPU['scorer'], 1, SEQ_PER_IMG, 1)
    rew, m, b = reward_oracle.get_cst_reward(rows, gts, _CPU['scorer'], bcmrscores=scores.reshape(-1, SEQ_PER_IMG),
                                             expand_feat=1, seq_per_img=SEQ_PER_IMG, scb_captions=20, scb_baseline=2,
                                             use_eos=1, use_mixer=0)
    return np.asarray(scores, dtype=np.float64), rew[:, 0].copy()


def cpu_df_dict():
    """df of the bench corpus as the oracle's dict (vectorised: the Python
    compute_doc_freq of the oracle takes ~8 s on this corpus and is not what is timed)."""
    from oracle import ciderd_oracle
    synth = load_synth()
    rng = np.random.default_rng(synth.SEED)
    refs, nw = synth.make_refs(rng, DF_VIDEOS, N_REFS)
    df, n = ciderd_oracle.build_df_from_strings(synth.df_corpus_strings(refs, nw))
    return df, n


def run_cpu_sample(batch, n_videos, procs, df_pair, keep=None):
    """Oracle port of utils.get_cst_reward on the first n_videos of the batch -> seconds.
    keep (a dict) receives the per-caption scores and rewards of a single-process run."""
    sampled = batch['sampled'][:n_videos * SEQ_PER_IMG]
    refs = batch['refs'][:n_videos]
    if procs <= 1:
        _cpu_init(*df_pair)
        t0 = time.perf_counter()
        scores, rewards = _cpu_shard((sampled, refs))
        secs = time.perf_counter() - t0
        if keep is not None:
            keep['scores'], keep['rewards'] = scores, rewards
        return secs
    import multiprocessing as mp
    bounds = np.linspace(0, n_videos, procs + 1).astype(int)
    shards = [(sampled[a * SEQ_PER_IMG:b * SEQ_PER_IMG], refs[a:b]) for a, b in zip(bounds[:-1], bounds[1:]) if b > a]
    with mp.get_context('fork').Pool(procs, initializer=_cpu_init, initargs=df_pair) as pool:
        pool.map(_cpu_shard, [(s[:SEQ_PER_IMG], r[:1]) for s, r in shards])      # warm the workers
        t0 = time.perf_counter()
        pool.map(_cpu_shard, shards)
        return time.perf_counter() - t0


def reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    synth = load_synth()
    cores = os.cpu_count() or 1
    df_pair = cpu_df_dict()
    # size one step so that steps + warmup finish in about two minutes
    probe_videos = max(cores, 8)
    batch = synth.make_batch(max(4096, probe_videos), SEQ_PER_IMG, N_REFS, seed=1000)
    t = run_cpu_sample(batch, probe_videos, cores, df_pair)
    rate = probe_videos * SEQ_PER_IMG / t
    budget = 120.0 / max(1, args.steps + args.warmup)
    n_videos = int(min(4096, max(cores, rate * budget / SEQ_PER_IMG)))
    for _ in range(args.warmup):
        run_cpu_sample(batch, n_videos, cores, df_pair)
    times = [run_cpu_sample(batch, n_videos, cores, df_pair) for _ in range(args.steps)]
    total = float(np.sum(times))
    value = n_videos * SEQ_PER_IMG * args.steps / total
    sample = 'first %d videos (%d captions) of the step batch per step, %d worker processes' % (
        n_videos, n_videos * SEQ_PER_IMG, cores)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * total / args.steps,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': workload_config(args.videos, args.gpus),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
        'note': 'oracle port of utils.get_cst_reward + restated CiderD (the scorer is an absent, unpinned '
                'dependency of the reference; nothing to compile into oracle/_ref)',
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------
# the CUDA arm
# ---------------------------------------------------------------------------
def event_ms(fn, steps, barrier):
    """steps calls of fn bracketed by CUDA events on the current stream -> ms per call."""
    import torch
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    barrier()
    return a.elapsed_time(b) / steps


def product_arm(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as graft
    graft.build()
    import cst_captioning_b200 as cst
    from cst_captioning_b200 import dist as cdist
    from cst_captioning_b200 import synth

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the reward path has no CPU fallback')
    torch.cuda.set_device(local)
    bind_to_gpu_numa_node(local)       # before any pinned allocation: first touch decides the NUMA node
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    # df table: built on the device from the 6513 x 20 corpus (words + <end>, use_eos=1)
    rng = np.random.default_rng(synth.SEED)
    corpus, _ = synth.make_refs(rng, DF_VIDEOS, N_REFS)
    table = cst.DfTable.from_rows(torch.from_numpy(corpus).to(dev), use_eos=1)
    scorer = cst.CiderD(df=table)

    n_videos = args.videos
    n_cap = n_videos * SEQ_PER_IMG
    batch = synth.make_batch(n_videos, SEQ_PER_IMG, N_REFS, seed=1000 + rank)
    h_cand = torch.from_numpy(batch['sampled']).pin_memory()
    h_refs = torch.from_numpy(batch['refs']).pin_memory()
    d_cand = h_cand.to(dev)
    d_refs = cst.RefSet(h_refs.to(dev).reshape(n_videos * N_REFS, REF_W), per_video=N_REFS)
    kw = dict(expand_feat=1, seq_per_img=SEQ_PER_IMG, scb_captions=20, scb_baseline=2, use_eos=1, return_stats=True)

    def step():
        rew, m, b, stats = cst.get_cst_reward(d_cand, d_refs, scorer, **kw)
        if world > 1:
            dist.all_reduce(stats)           # [sum_m, sum_b, n_m, n_b, bad]: the only exchange
        return rew, stats

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    # ---- device-resident timing (the headline `value`) -----------------------------
    launches0 = cst.lib.cst_kernel_launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        t_all0.record()
        for a, b in ev:
            a.record()
            rew, stats = step()
            b.record()
        t_all1.record()
        barrier()
    launches = cst.lib.cst_kernel_launches() - launches0
    total_ms = max_over_ranks(t_all0.elapsed_time(t_all1))
    per_step = [a.elapsed_time(b) for a, b in ev]
    stats = stats.cpu().numpy()
    value = world * n_cap * args.steps / (total_ms * 1e-3)
    clock_summary = clocks.summary()

    # the dominant kernel of a step is the scoring kernel: one launch per step (plus the 5 us fold of the
    # statistics), bracketed by events on the stream it runs on
    kern_ms = float(np.mean(per_step))
    peak, peak_src = measured_peak()
    achieved = n_cap * BYTES_PER_CAPTION / (kern_ms * 1e-3) / 1e9
    traffic = ncu_traffic() or {}
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    sm_hz = 1e6 * float(clock_summary.get('sm_mhz') or clock_summary.get('sm_max_mhz') or 1965.0)
    warp_inst = traffic.get('warp_inst_per_launch')
    roofline = {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                'traffic': traffic.get('bytes_per_launch'), 'traffic_capture': traffic.get('capture'),
                'kernel': traffic.get('kernel', 'k_score2'), 'kernel_ms': kern_ms,
                'bytes_per_caption': BYTES_PER_CAPTION, 'peak_source': peak_src,
                'frac_of_nominal_8TBs': achieved / 8000.0,
                # what actually binds: warp-instruction issue (the HBM roofline is the one BASELINE.json names)
                'limiter': 'instruction issue and dependent-load latency, not HBM',
                'warp_inst_per_caption': (warp_inst / traffic['captions_per_launch']) if warp_inst else None,
                'issue_frac': (warp_inst / (sms * 4 * sm_hz * kern_ms * 1e-3)) if warp_inst else None,
                'issue_note': 'executed warp instructions of the ncu capture / (SMs x 4 schedulers x SM clock x '
                              'live kernel time)'}

    # ---- the reference's own row dtype: int64 (model.py:255-266, create_sequencelabel.py:93) ----
    d_cand64 = d_cand.long()
    d_refs64 = cst.RefSet(d_refs.rows.long(), per_video=N_REFS)
    for _ in range(3):
        rew64, _, _, _ = cst.get_cst_reward(d_cand64, d_refs64, scorer, **kw)
    ms64 = event_ms(lambda: cst.get_cst_reward(d_cand64, d_refs64, scorer, **kw), args.steps, barrier)
    rew64, _, _, _ = cst.get_cst_reward(d_cand64, d_refs64, scorer, **kw)
    int64_rows = {'value': n_cap / (ms64 * 1e-3), 'unit': UNIT + ' per GPU', 'ms_per_step': ms64,
                  'bytes_per_caption': 2 * (BYTES_PER_CAPTION - 4) + 4,
                  'same_bits_as_int32_rows': bool(torch.equal(rew64, rew))}
    del d_cand64, d_refs64, rew64

    # ---- references staged once (CookedRefs) -----------------------------------------
    # The ground-truth captions of a training set never change, so their half of the kernel can be
    # run once and kept in HBM; a step then passes candidates + video indices.  NOT the headline
    # (that one, like the reference, works the refs up from raw rows every step).
    staged = None
    e2e_staged = None
    if not args.no_staged:
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cooked = cst.CookedRefs.cook(d_refs, scorer, use_eos=1)      # first call: module load
        del cooked
        c0.record()
        cooked = cst.CookedRefs.cook(d_refs, scorer, use_eos=1)
        c1.record()
        cooked_b = cooked.batch(torch.arange(n_videos, device=dev))
        for _ in range(3):
            rew_c, _, _, stats_c = cst.get_cst_reward(d_cand, cooked_b, scorer, **kw)
        ms_c = event_ms(lambda: cst.get_cst_reward(d_cand, cooked_b, scorer, **kw), args.steps, barrier)
        rew_c, _, _, stats_c = cst.get_cst_reward(d_cand, cooked_b, scorer, **kw)
        staged = {'value': n_cap / (ms_c * 1e-3), 'unit': UNIT + ' per GPU', 'ms_per_step': ms_c,
                  'cook_ms_once': c0.elapsed_time(c1), 'image_bytes_per_video': cooked.stride,
                  'images_bytes_total': cooked.nbytes(),
                  'same_bits_as_raw_rows': bool(torch.equal(rew_c, rew)),
                  'note': 'cst_refs_cook once outside the timed region, then cst_reward_cooked per step'}
        # ... and end to end: pinned host candidates in, pinned rewards out, refs already staged
        pipe_s = cst.HostRewardPipeline(scorer, seq_per_img=SEQ_PER_IMG, scb_captions=20, scb_baseline=2, use_eos=1,
                                        slices=args.e2e_slices)
        out_s = torch.empty(n_cap, dtype=torch.float32).pin_memory()

        def e2e_staged_step():
            pipe_s(h_cand, cooked_b, out=out_s)
            torch.cuda.synchronize()

        e2e_staged_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_staged_step()
        barrier()
        e2e_staged_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
        e2e_staged = {'value': world * n_cap / e2e_staged_s, 'unit': UNIT, 'ms_per_step': 1e3 * e2e_staged_s,
                      'h2d_bytes_per_step': int(h_cand.numel() * 4), 'd2h_bytes_per_step': int(out_s.numel() * 4),
                      'steps': args.steps,
                      'same_bits_as_raw_rows': bool(torch.equal(out_s.to(dev), rew[:, 0])),
                      'note': 'HostRewardPipeline with the references staged on the device once (a training set\'s '
                              'references are static, dataloader.py:138-139): only candidates go up'}
        del cooked, cooked_b, pipe_s

    # ---- end to end through the public API with HOST buffers -----------------------
    out_pin = torch.empty(n_cap, dtype=torch.float32).pin_memory()
    st_pin = torch.empty(7, dtype=torch.float64).pin_memory()       # CST_N_STATS

    pipe = cst.HostRewardPipeline(scorer, seq_per_img=SEQ_PER_IMG, scb_captions=20, scb_baseline=2, use_eos=1,
                                  slices=args.e2e_slices)

    def e2e_step():
        # public API for host-resident batches: pinned rows in, pinned rewards + statistics out
        _, st = pipe(h_cand, h_refs, out=out_pin)
        if world > 1:
            dist.all_reduce(st)
        st_pin.copy_(st, non_blocking=True)
        torch.cuda.synchronize()

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = {'value': world * n_cap * args.steps / e2e_s, 'unit': UNIT,
           'h2d_bytes_per_step': int(h_cand.numel() * 4 + h_refs.numel() * 4),
           'd2h_bytes_per_step': int(out_pin.numel() * 4 + st_pin.numel() * 8), 'steps': args.steps,
           'api': 'HostRewardPipeline (pinned host rows, %d slices, upload overlapped with scoring)' % args.e2e_slices,
           'ms_per_step': 1e3 * e2e_s / args.steps,
           'same_bits_as_device_resident': bool(torch.equal(out_pin.to(dev), rew[:, 0]))}

    # the platform's ceiling for that leg: the same pinned bytes uploaded by plain copies, every rank at
    # once (N ranks share the host's PCIe fabric), no kernels of ours -- max over ranks like every timing
    d_c, d_r = torch.empty_like(d_cand), torch.empty_like(h_refs, device=dev)
    for _ in range(2):
        d_c.copy_(h_cand, non_blocking=True)
        d_r.copy_(h_refs, non_blocking=True)
    plain_ms = max_over_ranks(event_ms(lambda: (d_c.copy_(h_cand, non_blocking=True), d_r.copy_(h_refs, non_blocking=True)),
                                       10, barrier))
    h2d_bytes = e2e['h2d_bytes_per_step']
    e2e['plain_copy'] = {'ms_per_step': plain_ms, 'GBs_per_gpu': h2d_bytes / plain_ms / 1e6,
                         'GBs_all_gpus': world * h2d_bytes / plain_ms / 1e6,
                         'e2e_over_plain_copy': plain_ms / e2e['ms_per_step'],
                         'note': 'upload of the step\'s pinned rows alone, all ranks concurrently: what the host '
                                 'platform allows; the end-to-end step cannot be shorter'}
    del d_c, d_r

    # informational: the same pipeline fed int16 host rows (ids <= 32767 fit; half the upload)
    h_cand16, h_refs16 = h_cand.to(torch.int16).pin_memory(), h_refs.to(torch.int16).pin_memory()

    def e2e16_step():
        _, st = pipe(h_cand16, h_refs16, out=out_pin)
        st_pin.copy_(st, non_blocking=True)
        torch.cuda.synchronize()

    e2e16_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e16_step()
    barrier()
    e2e16_s = time.perf_counter() - t0
    e2e['int16_host_rows'] = {'value': n_cap * args.steps / e2e16_s, 'unit': UNIT + ' per GPU',
                              'ms_per_step': 1e3 * e2e16_s / args.steps,
                              'h2d_bytes_per_step': int(h_cand16.numel() * 2 + h_refs16.numel() * 2)}
    del h_cand16, h_refs16

    # ---- the fused consumer of the reward (model.py:7-23): streaming kernels, HBM-bound ----
    ok_other = True
    criterion = None
    if world == 1 and not args.no_criterion:
        # (timed through the C ABI: at 0.08 ms per launch the autograd wrapper's host time would be the number)
        from cst_captioning_b200 import _lib
        from cst_captioning_b200.criterion import _scratch_for as crit_scratch
        t_w = CAND_W
        seq = d_cand.long()
        logp = -torch.rand(n_cap, t_w, device=dev)
        col = rew[:, 0].contiguous()
        out2 = torch.empty(2, dtype=torch.float32, device=dev)
        grad = torch.empty(n_cap, t_w, dtype=torch.float32, device=dev)
        go = torch.ones(1, dtype=torch.float32, device=dev)
        scr = crit_scratch(dev)
        stream = torch.cuda.current_stream(dev).cuda_stream

        def crit_fwd():
            _lib.check(cst.lib.cst_reward_criterion_fwd(
                seq.data_ptr(), _lib.DTYPE_I64, seq.stride(0), logp.data_ptr(), logp.stride(0), col.data_ptr(), 1, 0,
                n_cap, t_w, out2.data_ptr(), out2.data_ptr() + 4, scr.data_ptr(), scr.numel(), dev.index, stream))

        def crit_bwd():
            _lib.check(cst.lib.cst_reward_criterion_bwd(
                seq.data_ptr(), _lib.DTYPE_I64, seq.stride(0), col.data_ptr(), 1, 0, out2.data_ptr() + 4,
                go.data_ptr(), n_cap, t_w, grad.data_ptr(), grad.stride(0), dev.index, stream))

        for _ in range(3):
            crit_fwd()
            crit_bwd()
        f_ms = event_ms(crit_fwd, args.steps, barrier)
        b_ms = event_ms(crit_bwd, args.steps, barrier)
        # ... and the module gives the same loss (the path a training loop takes)
        loss = cst.RewardCriterion()(seq, logp, rew)
        nbytes = n_cap * t_w * (8 + 4) + n_cap * 4       # int64 seq + fp32 logp (bwd: fp32 grad) + reward column
        criterion = {'rows': n_cap, 'width': t_w, 'fwd_ms': f_ms, 'bwd_ms': b_ms,
                     'fwd_GBs': nbytes / f_ms / 1e6, 'bwd_GBs': nbytes / b_ms / 1e6,
                     'fwd_frac': nbytes / f_ms / 1e6 / peak, 'bwd_frac': nbytes / b_ms / 1e6 / peak,
                     'bound': 'hbm', 'bytes_per_token': 12,
                     'module_loss_equals_abi_loss': bool(float(loss) == float(out2[0]))}
        del seq, logp, loss, grad

    # ---- the other reward scorers of train.py:119-124 on the same batch ---------
    # BLEU-4 (k_bleu) and ROUGE-L (k_rouge), each followed by the SCB* baseline (k_scb), timed on the device;
    # with the CPU leg on, the first videos' scores are held to the restated scorers as well.
    other = None
    if world == 1 and not args.no_other:
        other = {}
        n_o = min(64, n_videos)
        for name, sc in (('bleu4', cst.Bleu(device=dev)), ('rouge_l', cst.Rouge(device=dev))):
            for _ in range(3):
                cst.get_cst_reward(d_cand, d_refs, sc, **kw)
            ms_o = event_ms(lambda: cst.get_cst_reward(d_cand, d_refs, sc, **kw), args.steps, barrier)
            other[name] = {'value': n_cap / (ms_o * 1e-3), 'unit': UNIT, 'ms_per_step': ms_o,
                           'GBs': n_cap * BYTES_PER_CAPTION / ms_o / 1e6,
                           'frac': n_cap * BYTES_PER_CAPTION / ms_o / 1e6 / peak,
                           'kernels_per_step': 2}
            if not args.no_cpu:
                from oracle import bleu_rouge_oracle, reward_oracle
                ref_sc = bleu_rouge_oracle.Bleu() if name == 'bleu4' else bleu_rouge_oracle.Rouge()
                gts_o = [batch['refs'][v].astype(np.int64) for v in range(n_o)]
                want = reward_oracle.score_rows(list(batch['sampled'][:n_o * SEQ_PER_IMG].astype(np.int64)), gts_o, ref_sc,
                                                1, SEQ_PER_IMG, 1)
                sub_o = cst.RefSet(d_refs.rows[:n_o * N_REFS], per_video=N_REFS)
                got = sc.score_rows(d_cand[:n_o * SEQ_PER_IMG], sub_o, SEQ_PER_IMG, use_eos=1).cpu().numpy()
                err = float(np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)))
                other[name]['check'] = {'n_compared': int(want.size), 'max_rel_err': err,
                                        'bit_exact': bool(np.array_equal(got, want)), 'ok': bool(err <= 1e-9)}
                ok_other = ok_other and err <= 1e-9

    # ---- multi-GPU legs BASELINE.json lists beside the weak sweep (N > 1 only) --------------
    strong = None
    config4 = None
    if world > 1:
        # strong scaling + hardware parity: ONE global batch (seed 1000, args.videos videos) cut by
        # dist.shard_range; rank 0 also scores all of it alone, and every rank holds its slice of the
        # rewards and the all-reduced means against that single-GPU result
        g = synth.make_batch(args.videos, SEQ_PER_IMG, N_REFS, seed=1000)
        lo, hi = cdist.shard_range(args.videos, rank, world)
        s_cand = torch.from_numpy(g['sampled'][lo * SEQ_PER_IMG:hi * SEQ_PER_IMG]).to(dev)
        s_refs = cst.RefSet(torch.from_numpy(g['refs'][lo:hi]).to(dev).reshape((hi - lo) * N_REFS, REF_W),
                            per_video=N_REFS)

        def strong_step():
            r_, _, _, st_ = cst.get_cst_reward(s_cand, s_refs, scorer, **kw)
            dist.all_reduce(st_)
            return r_, st_

        for _ in range(3):
            strong_step()
        ms_s = max_over_ranks(event_ms(strong_step, args.steps, barrier))
        r_s, st_s = strong_step()
        full = torch.empty(args.videos * SEQ_PER_IMG, dtype=torch.float32, device=dev)
        ref_stats = torch.empty(7, dtype=torch.float64, device=dev)
        if rank == 0:
            f_cand = torch.from_numpy(g['sampled']).to(dev)
            f_refs = cst.RefSet(torch.from_numpy(g['refs']).to(dev).reshape(args.videos * N_REFS, REF_W),
                                per_video=N_REFS)
            r_f, _, _, st_f = cst.get_cst_reward(f_cand, f_refs, scorer, **kw)
            full.copy_(r_f[:, 0])
            ref_stats.copy_(st_f)
            del f_cand, f_refs
        dist.broadcast(full, 0)
        dist.broadcast(ref_stats, 0)
        bad = torch.tensor([0 if torch.equal(r_s[:, 0], full[lo * SEQ_PER_IMG:hi * SEQ_PER_IMG]) else 1],
                           device=dev, dtype=torch.float64)
        dist.all_reduce(bad)
        m_n, b_n = float(st_s[0] / st_s[2]), float(st_s[1] / st_s[3])
        m_1, b_1 = float(ref_stats[0] / ref_stats[2]), float(ref_stats[1] / ref_stats[3])
        strong = {'value': args.videos * SEQ_PER_IMG / (ms_s * 1e-3), 'unit': UNIT, 'ms_per_step': ms_s,
                  'global_videos': args.videos, 'videos_per_gpu': hi - lo, 'scaling': 'strong',
                  'm_score': m_n, 'b_score': b_n, 'm_score_1gpu': m_1, 'b_score_1gpu': b_1,
                  'm_rel_diff': abs(m_n - m_1) / abs(m_1), 'b_rel_diff': abs(b_n - b_1) / abs(b_1),
                  'ranks_whose_rewards_differ_from_1gpu': int(bad.item()),
                  'ok': bool(bad.item() == 0 and abs(m_n - m_1) <= 1e-12 * abs(m_1) and abs(b_n - b_1) <= 1e-12 * abs(b_1))}
        del s_cand, s_refs, full

        # BASELINE configs[3]: SCB with the baseline from precomputed GT consensus scores, batch 512
        # sharded over the ranks (64 videos per GPU at 8): per-step latency, where launch + all-reduce
        # are what is left
        g4 = synth.make_batch(512, SEQ_PER_IMG, N_REFS, seed=13)
        lo4, hi4 = cdist.shard_range(512, rank, world)
        c4 = torch.from_numpy(g4['sampled'][lo4 * SEQ_PER_IMG:hi4 * SEQ_PER_IMG]).to(dev)
        r4 = torch.from_numpy(g4['refs'][lo4:hi4]).to(dev)
        bc4 = torch.from_numpy(np.random.default_rng(4).random((512, SEQ_PER_IMG))[lo4:hi4]).to(dev)
        kw4 = dict(bcmrscores=bc4, expand_feat=1, seq_per_img=SEQ_PER_IMG, scb_captions=20, scb_baseline=1, use_eos=1,
                   use_mixer=1, return_stats=True)

        def c4_step(reduce=True):
            r_, _, _, st_ = cst.get_cst_reward(c4, r4, scorer, **kw4)
            if reduce:
                dist.all_reduce(st_)
            return st_

        for _ in range(10):
            c4_step()
        ms_red = max_over_ranks(event_ms(c4_step, 200, barrier))
        ms_lazy = max_over_ranks(event_ms(lambda: c4_step(False), 200, barrier))
        side = torch.cuda.Stream(dev)
        with torch.cuda.stream(side):
            c4_step(False)
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            keep4 = c4_step(False)
        ms_graph = max_over_ranks(event_ms(graph.replay, 200, barrier))
        config4 = {'workload': 'BASELINE.json configs[3]: SCB (GT-consensus baseline), batch 512 sharded over the ranks',
                   'videos_per_gpu': hi4 - lo4, 'captions_per_step': 512 * SEQ_PER_IMG,
                   'ms_per_step_with_allreduce': ms_red, 'ms_per_step_stats_reduced_lazily': ms_lazy,
                   'ms_per_step_cuda_graph_replay_no_collective': ms_graph,
                   'captions_per_s_with_allreduce': 512 * SEQ_PER_IMG / (ms_red * 1e-3),
                   'note': 'the statistics are only read every print_log_interval steps (train.py:229-233), so the '
                           'all-reduce can wait until then: the lazy figures are a step without it'}
        del graph, keep4

    # ---- CPU baseline: oracle port, one core, bounded sample (rank 0, N = 1 only) --
    # ... and a parity check of the GPU step at the bench's own size against that sample
    cpu = None
    check = {'m_score': float(stats[0] / stats[2]), 'b_score': float(stats[1] / stats[3]),
             'bad_rows': float(stats[4]), 'df_ngrams': len(table)}
    ok = True
    if rank == 0 and world == 1 and not args.no_cpu:
        n_s = min(args.cpu_videos, n_videos)
        keep = {}
        secs = run_cpu_sample(batch, n_s, 1, cpu_df_dict(), keep)
        cpu = {'value': n_s * SEQ_PER_IMG / secs, 'unit': UNIT, 'cores': 1, 'kind': 'port',
               'sample': 'first %d videos (%d captions) of the step batch, oracle port of utils.get_cst_reward '
                         '+ restated CiderD, %.1f s' % (n_s, n_s * SEQ_PER_IMG, secs),
               'host_cores_available': os.cpu_count()}
        # float64 scores of the same videos from the CUDA path (cst_cider_scores), and the float32 rewards of
        # the timed step itself
        sub = cst.RefSet(d_refs.rows[:n_s * N_REFS], per_video=N_REFS)
        g_scores = scorer.score_rows(d_cand[:n_s * SEQ_PER_IMG], sub, SEQ_PER_IMG, use_eos=1).cpu().numpy()
        g_rew = rew[:n_s * SEQ_PER_IMG, 0].double().cpu().numpy()
        c_scores, c_rew = keep['scores'], keep['rewards']
        nz = np.abs(c_scores) > 0
        rel = np.zeros_like(c_scores)
        rel[nz] = np.abs(g_scores[nz] - c_scores[nz]) / np.abs(c_scores[nz])
        zeros_agree = bool(np.all(g_scores[~nz] == 0.0))
        rew_err = float(np.max(np.abs(g_rew - c_rew.astype(np.float32).astype(np.float64))))
        ok = bool(rel.max() <= 1e-5 and zeros_agree and rew_err <= 1e-5)
        check.update({'n_compared': int(c_scores.size), 'max_rel_err': float(rel.max()), 'tolerance': 1e-5,
                      'zero_scores_agree': zeros_agree, 'max_abs_err_reward_f32': rew_err, 'ok': ok,
                      'against': 'float64 scores of the oracle port on the first %d videos of this very batch' % n_s})

    if rank == 0:
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': total_ms / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': workload_config(n_videos, world), 'roofline': roofline, 'cpu_baseline': cpu, 'e2e': e2e,
            'gpu_launches': int(launches), 'clocks': clock_summary, 'int64_rows': int64_rows,
            'staged_refs': staged, 'e2e_staged': e2e_staged, 'criterion': criterion, 'strong': strong,
            'config4': config4, 'other_scorers': other, 'check': check,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    if not ok or not ok_other:
        sys.stderr.write('bench.py: the GPU step does not match the CPU port on the sampled videos\n')
        return 1
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--videos', type=int, default=65536, help='videos per GPU per step')
    ap.add_argument('--cpu-videos', type=int, default=640, help='CPU-baseline sample size (videos)')
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-staged', action='store_true', help='skip the informational staged-references leg')
    ap.add_argument('--no-criterion', action='store_true', help='skip the RewardCriterion leg')
    ap.add_argument('--no-other', action='store_true', help='skip the BLEU-4 / ROUGE-L scorer leg')
    ap.add_argument('--e2e-slices', type=int, default=16, help='video slices of the host pipeline')
    args = ap.parse_args()
    if args.impl == 'reference':
        return reference_arm(args)
    return product_arm(args)


if __name__ == '__main__':
    sys.exit(main())
