#!/usr/bin/env python
"""Benchmark of the fused planar-flow free-energy forward+backward (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (the configuration BASELINE's metric is quoted on): K=32 planar
layers, D=2, target U1 (`p1`), 2^20 Monte-Carlo samples per GPU per step, float32, script
semantics (get_data.py:97-115).  One "step" = one fused forward+backward over one batch of
2^20 fresh samples: ONE launch of the objective kernel, whose last block adds the per-SM rows
in a fixed order and, when N>1, exchanges the 3+2KD+2K result sums with the peers over NVLink.
Samples are resident in HBM for `value`; `e2e` goes through the NumPy-facing objective with
pinned host samples (H2D of the samples, D2H of F and the gradient inside the timed region);
`e2e_pageable` is the same call on a pageable float64 NumPy array, `e2e_rng` the opt-in call that
draws the samples inside the kernel.

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU stand-in for the reference's
autograd path (torch float64 op-for-op restatement from oracle/, all host threads) on the
same config and metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, 'vi-normflows_b200')
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

K_LAYERS, DIM, N_PER_GPU, TARGET = 32, 2, 1 << 20, 'p1'
N_ROTATE = 32          # distinct sample buffers: 32 x 8 MiB = 256 MiB > 126 MB L2
METRIC = 'MC samples/sec for K=32 planar-flow ELBO fwd+bwd'
UNIT = 'samples/s'

# algorithmic work per sample, SURVEY.md §8d convention (FMA = 2 flop; tanh = 2 MUFU, log/rcp/exp/sqrt = 1)
FLOPS_PER_SAMPLE = K_LAYERS * (14 * DIM + 25) + 2 * DIM + 64        # 1764
MUFU_PER_SAMPLE = 4 * K_LAYERS + 10                                 # 138
BYTES_PER_SAMPLE = 4 * DIM                                          # 8
NCU_TRAFFIC_BYTES = 8.620800e6
NCU_TRAFFIC_SOURCE = 'ncu capture profiles/r02_ncu_elbo_pair_kernel.txt (not measured in this run)'


def params_np(seed=7):
    rs = np.random.RandomState(seed)
    W = 0.1 + 0.01 * rs.randn(K_LAYERS, DIM)
    U = 0.1 + 0.01 * rs.randn(K_LAYERS, DIM)
    B = 0.1 * np.ones(K_LAYERS)
    return W, U, B


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace('.', '').isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace('.', '').isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[5:9]):
                    if v.lower().startswith('active'):
                        reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def cpu_step_fn(n_cpu):
    """One fwd+bwd of the CPU stand-in on n_cpu samples (torch float64 autograd over the
    op-for-op restatement of get_data.py:97-113 -- the role autograd 1.3 plays in the reference)."""
    from oracle import planar_torch as pt
    from nfcore import semantics as sem
    W, U, B = params_np()
    params = sem.pack(W, U, B)
    z0 = torch.randn(n_cpu, DIM, dtype=torch.float64, generator=torch.Generator().manual_seed(1234))

    def step():
        p = torch.tensor(params, dtype=torch.float64, requires_grad=True)
        F = pt.script_objective(p, z0, pt.p1, 1e-7, K_LAYERS)
        g, = torch.autograd.grad(F, p)
        return float(F.detach()), g
    return step


def time_cpu(n_cpu, steps, warmup):
    step = cpu_step_fn(n_cpu)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return n_cpu / dt, dt


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1; the reference arm is entitled to every host core
    try:
        torch.set_num_threads(max(1, os.cpu_count() or 1))
    except RuntimeError:
        pass
    # bounded sample: largest power of two <= 2^18 samples/step that keeps the whole run near 90 s
    n_cpu = 1 << 18
    _, t1 = time_cpu(n_cpu, 1, 1)
    while n_cpu > (1 << 14) and t1 * (args.steps + args.warmup) > 90.0:
        n_cpu >>= 1
        t1 /= 2
    value, dt = time_cpu(n_cpu, args.steps, args.warmup)
    cores = torch.get_num_threads()
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': dict(config_dict(args.gpus), reference_samples_per_step=n_cpu,
                       note='the CPU arm times a bounded sample of %d samples per step of this workload; the rate is what is compared' % n_cpu),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': '%d samples per step of the same K=32/D=2/U1 workload, torch float64 autograd '
                                   'over oracle/planar_torch.py (stand-in for autograd 1.3, not installable); '
                                   'os.cpu_count()=%d' % (n_cpu, os.cpu_count())},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line))


def config_dict(n_gpus):
    return {'workload': 'C2: shared-parameter planar-flow free energy fwd+bwd, K=32, D=2, target U1 (p1), '
                        '2^20 MC samples per GPU per step, script semantics (get_data.py:97-115)',
            'K': K_LAYERS, 'D': DIM, 'target': 'U1', 'samples_per_gpu_per_step': N_PER_GPU,
            'global_samples_per_step': N_PER_GPU * n_gpus, 'parallelism': 'sample-sharded dp%d' % n_gpus,
            'l2': 'rotating %d distinct sample buffers (%d MiB > 126 MB L2)'
                  % (N_ROTATE, N_ROTATE * N_PER_GPU * BYTES_PER_SAMPLE >> 20)}


def run_ours(args):
    import torch.distributed as dist
    from nfcore import ops, semantics as sem
    from nfcore.objective import SharedFlowObjective

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    ops.require_cuda()

    W, U, B = params_np()
    S, c = sem.forward(W, U)
    w_d, s_d, b_d, c_d = (torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (W, S, B, c))
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    eps = [torch.randn(N_PER_GPU, DIM, dtype=torch.float32, device=dev, generator=gen) for _ in range(N_ROTATE)]
    nres = 3 + 2 * K_LAYERS * DIM + 2 * K_LAYERS
    res = torch.empty(nres, dtype=torch.float32, device=dev)
    KD = K_LAYERS * DIM
    out = (res[:3], res[3:3 + KD].view(K_LAYERS, DIM), res[3 + KD:3 + 2 * KD].view(K_LAYERS, DIM),
           res[3 + 2 * KD:3 + 2 * KD + K_LAYERS], res[3 + 2 * KD + K_LAYERS:])

    small_ar = None
    if world > 1 and os.environ.get('NF_BENCH_NCCL', '0') != '1':
        ok = 1
        try:
            from nfcore.peer import SmallAllReduce
            small_ar = SmallAllReduce(dist.group.WORLD, dev)      # one-shot exchange over NVLink peer memory
        except Exception as e:                                    # no P2P mapping: NCCL carries the vector
            if rank == 0:
                print('bench: peer all-reduce unavailable (%s); using NCCL' % e, file=sys.stderr)
            small_ar, ok = None, 0
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)  # all ranks take the same path
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            small_ar = None
    resx = torch.zeros(nres + 3, dtype=torch.float32, device=dev)   # exchange form: [results | n_hi, n_lo, status]

    def local_step(i):
        ops.planar_elbo(eps[i % N_ROTATE], w_d, s_d, b_d, c_d, TARGET, None, 1e-7, True, out)

    def step(i):
        if world == 1:
            local_step(i)
        elif small_ar is not None:
            # objective + exchange: the last block of the objective kernel pushes, waits and adds (one launch)
            ops.planar_elbo_exchange(eps[i % N_ROTATE], w_d, s_d, b_d, c_d, TARGET, None, 1e-7, resx, small_ar)
        else:
            local_step(i)
            dist.all_reduce(res)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n_warm = max(args.warmup, 3)
    for i in range(n_warm):
        step(i)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    # keep every GPU under this load while the clock sampler comes up (an idle GPU drops its clocks within a fraction
    # of a second, and the timed region below is only milliseconds long); these steps are untimed warm-up too
    t_busy = time.time() + 0.4
    extra_warm = 0
    while True:
        for i in range(50):
            step(i)
        extra_warm += 50
        stop = torch.tensor([1 if time.time() >= t_busy else 0], dtype=torch.int32, device=dev)
        if world > 1:
            dist.all_reduce(stop, op=dist.ReduceOp.MIN)      # all ranks leave the loop in the same round (lock-step exchange)
        if int(stop.item()) == 1:
            break
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # keep the GPU busy a little longer so the sampler sees clocks under this load
    t_end = time.time() + 0.6
    while rank == 0 and time.time() < t_end:      # rank-local kernels only: no collective outside the lock-step loop
        for i in range(20):
            local_step(i)
        torch.cuda.synchronize()
    clocks = sampler.stop() if sampler else None
    barrier()
    if small_ar is not None and float(resx[nres + 2].item()) != 0.0:
        raise RuntimeError('bench: the peer exchange reported a missing rank')
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms = float(tms.item())
    ms_per_step = ms / args.steps
    value = N_PER_GPU * world / (ms_per_step * 1e-3)

    # ---- end to end through the NumPy-facing objective: pinned host samples in, F + gradient out ----
    obj = SharedFlowObjective(TARGET, K_LAYERS, DIM, dtype='float32',
                              process_group=dist.group.WORLD if world > 1 else None)
    params = sem.pack(W, U, B)
    host = [torch.randn(N_PER_GPU, DIM, dtype=torch.float32).pin_memory() for _ in range(4)]
    # warm-up: the first ~100 uploads after idle run at a fraction of the PCIe rate (the link trains up under load:
    # 540 -> 230 us per 8 MiB call over the first 100 calls, tools/e2e_stages.py with NF_SESSION_TIMING=1), so the host-API
    # legs are warmed until 50 calls take the same time as the 50 before them (within 2 %)
    prev = None
    for chunk in range(40):                        # at most 2000 calls (~0.6 s); every rank runs the same count
        t0 = time.perf_counter()
        for i in range(50):
            obj.value_and_grad(params, host[i % 4])
        cur = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(cur, op=dist.ReduceOp.MAX)
        cur = float(cur.item())
        if chunk >= 2 and prev is not None and abs(cur - prev) <= 0.02 * prev:
            break
        prev = cur
    # five timed regions of exactly K steps each (barrier + synchronize on both sides, max over ranks), the median is
    # reported: a K = 20 region lasts 5 ms, and one host hiccup of a millisecond would otherwise decide the number
    e2e_regions = []
    for rep in range(5):
        barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(args.steps):
            F_e2e, g_e2e = obj.value_and_grad(params, host[i % 4])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tdt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tdt, op=dist.ReduceOp.MAX)
        e2e_regions.append(float(tdt.item()))
    e2e_value = N_PER_GPU * world * args.steps / float(np.median(e2e_regions))
    # the same call with what a drop-in caller holds: a pageable NumPy array (what np.random.multivariate_normal returns,
    # get_data.py:100), converted to float32 inside the call; the draw itself is outside the timed region
    host_np = [h.numpy().astype(np.float64) for h in host[:2]]
    for i in range(2):
        obj.value_and_grad(params, host_np[i % 2])
    barrier()
    t0 = time.perf_counter()
    n_pg = max(4, min(args.steps, 20))
    for i in range(n_pg):
        obj.value_and_grad(params, host_np[i % 2])
    torch.cuda.synchronize()
    tpg = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tpg, op=dist.ReduceOp.MAX)
    e2e_pageable = N_PER_GPU * world * n_pg / float(tpg.item())
    # opt-in: the samples drawn inside the kernel (nf_planar_elbo_fwd_bwd_rng_f32; the library's counter-based stream,
    # not NumPy's): only the K x D constants go up, only F and the gradient come back
    e2e_rng = None
    if world == 1:
        for i in range(3):
            obj.value_and_grad_rng(params, N_PER_GPU, 1234, i)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(args.steps):
            F_rng, _ = obj.value_and_grad_rng(params, N_PER_GPU, 1234, 3 + i)
        torch.cuda.synchronize()
        e2e_rng = {'value': N_PER_GPU * args.steps / (time.perf_counter() - t0), 'unit': UNIT, 'F': float(F_rng),
                   'h2d_bytes_per_step': 4 * (2 * KD + 2 * K_LAYERS), 'd2h_bytes_per_step': 4 * (nres + 3),
                   'input': 'none: samples drawn in the kernel (Philox-4x32-10 keyed by seed, counter = (sample pair, step))'}
    h2d = N_PER_GPU * BYTES_PER_SAMPLE + 4 * (2 * KD + 2 * K_LAYERS)
    d2h = 4 * nres

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except OSError:
            pass
        prop = torch.cuda.get_device_properties(dev)
        sms = prop.multi_processor_count
        f_max = (peaks.get('sm_max_mhz') or (clocks or {}).get('sm_max_mhz') or 1965.0) * 1e6
        per_gpu = value / world
        mufu_peak = sms * 16 * f_max
        fp32_peak = sms * 128 * 2 * f_max
        hbm_peak = (peaks.get('hbm_gbs') or 6650.0) * 1e9
        fracs = {'mufu': per_gpu * MUFU_PER_SAMPLE / mufu_peak, 'fp32': per_gpu * FLOPS_PER_SAMPLE / fp32_peak,
                 'hbm': per_gpu * BYTES_PER_SAMPLE / hbm_peak}
        f_load = ((clocks or {}).get('sm_mhz') or 0) * 1e6
        roofline = {
            'bound': 'mufu', 'achieved': per_gpu * MUFU_PER_SAMPLE / 1e9, 'peak': mufu_peak / 1e9, 'unit': 'G MUFU-op/s',
            'frac': fracs['mufu'],
            # dram__bytes_read.sum + dram__bytes_write.sum of one objective-kernel launch at this workload: a constant
            # copied from the committed `ncu --set full` capture named in traffic_source, NOT measured by this run
            # (algorithmic bytes: 8.39e6)
            'traffic': NCU_TRAFFIC_BYTES, 'traffic_source': NCU_TRAFFIC_SOURCE,
            'per_unit': {'flops': FLOPS_PER_SAMPLE, 'mufu': MUFU_PER_SAMPLE, 'bytes': BYTES_PER_SAMPLE},
            'peak_source': 'formula SMs*16*sm_max_mhz (SURVEY.md §8d; MEASURED_PEAKS.json has no MUFU figure); '
                           'sm_max_mhz and hbm_gbs %s' % ('of measured' if peaks else 'of fallback'),
            'frac_fp32': fracs['fp32'], 'frac_hbm': fracs['hbm'],
            'achieved_hbm_gbs': per_gpu * BYTES_PER_SAMPLE / 1e9, 'peak_hbm_gbs': hbm_peak / 1e9,
            'frac_mufu_at_clock_under_load': (per_gpu * MUFU_PER_SAMPLE / (sms * 16 * f_load)) if f_load else None,
            # what the kernel issues, from the committed ncu capture (not measured by this run): the instruction-issue bound
            # sits below the MUFU bound and is the one the kernel runs into
            'issue_model': {'lane_instructions_per_sample': 1026, 'issue_bound_samples_per_s': sms * 128 * f_max / 1026.0,
                            'smsp_issue_active_pct': 51.3, 'source': 'profiles/r02_ncu_elbo_pair_kernel.txt'},
            'kernel_ms': ms_per_step,
            'kernel_note': 'events bracket one launch per step: the objective kernel, whose last block adds the per-SM rows'
                           + ((' and does the NVLink peer exchange' if small_ar is not None else '; + NCCL all-reduce') if world > 1 else ''),
        }
        n_cpu = 1 << 18
        cpu_value, cpu_dt = time_cpu(n_cpu, 6, 1) if world == 1 else (None, None)
        # informational: the same kernel on BASELINE config 3's batch size (2^24 samples per step, U1 target),
        # where per-launch fixed costs (tile quantisation, prologue, row kernel) are amortised
        big = None
        if world == 1:
            nb = 1 << 24
            eb = [torch.randn(nb, DIM, dtype=torch.float32, device=dev, generator=gen) for _ in range(2)]
            for i in range(2):
                ops.planar_elbo(eb[i % 2], w_d, s_d, b_d, c_d, TARGET, None, 1e-7, True, out)
            torch.cuda.synchronize()
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b0.record()
            for i in range(10):
                ops.planar_elbo(eb[i % 2], w_d, s_d, b_d, c_d, TARGET, None, 1e-7, True, out)
            b1.record()
            torch.cuda.synchronize()
            vb = nb * 10 / (b0.elapsed_time(b1) * 1e-3)
            big = {'samples_per_step': nb, 'value': vb, 'unit': UNIT, 'frac_mufu': vb * MUFU_PER_SAMPLE / mufu_peak}
            del eb
        # informational: one more BASELINE configuration, device-timed (C5 at D = 64: K = 64, 2^19 samples, diagonal
        # Gaussian target; the matrix-product path).  Not part of `value`.
        other = None
        if world == 1:
            try:
                Kc, Dc, Nc = 64, 64, 1 << 19
                rsx = np.random.RandomState(5)
                Wc = (0.1 + 0.01 * rsx.randn(Kc, Dc)) / np.sqrt(Dc / 2); Uc = (0.1 + 0.01 * rsx.randn(Kc, Dc)) / np.sqrt(Dc / 2)
                Sc, cc = sem.forward(Wc, Uc)
                ac = [torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (Wc, Sc, 0.1 * np.ones(Kc), cc)]
                tpc = torch.as_tensor(np.concatenate([rsx.uniform(-1, 1, Dc), rsx.uniform(-1, 1, Dc)]), dtype=torch.float32).to(dev)
                ec = [torch.randn(Nc, Dc, dtype=torch.float32, device=dev, generator=gen) for _ in range(2)]
                oc = ops.planar_elbo(ec[0], *ac, 'gaussdiag', tpc)
                torch.cuda.synchronize()
                c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                c0.record()
                for i in range(10):
                    ops.planar_elbo(ec[i % 2], *ac, 'gaussdiag', tpc, out=oc)
                c1.record()
                torch.cuda.synchronize()
                msc = c0.elapsed_time(c1) / 10
                other = {'c5_d64': {'K': Kc, 'D': Dc, 'samples_per_step': Nc, 'ms_per_step': msc, 'value': Nc / msc * 1e3, 'unit': UNIT,
                                    'path': 'matrix-product formulation (csrc/nf_elbo_gram.cu)'}}
                del ec
            except Exception as e:          # informational only
                other = {'c5_d64': {'error': str(e)}}
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': n_warm + extra_warm,
            'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': 'f32', 'data': 'synthetic', 'config': config_dict(world),
            'roofline': roofline,
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                    'F': float(F_e2e), 'input': 'pinned host float32 samples (best case of the host API)',
                    'regions_ms': [round(x * 1e3, 3) for x in e2e_regions], 'stat': 'median of 5 timed regions of `steps` calls each'},
            'e2e_pageable': {'value': e2e_pageable, 'unit': UNIT, 'steps': n_pg,
                             'input': 'pageable float64 NumPy array, as np.random.multivariate_normal returns it (get_data.py:100); '
                                      'the RNG draw itself is not timed'},
            'gpu_launches': (1 if (world == 1 or small_ar is not None) else 2) * args.steps,
            'exchange': None if world == 1 else ('in the objective kernel\'s last block: NVLink peer stores + flags, %d floats' % (nres + 2)
                                                 if small_ar is not None else 'ncclAllReduce (%d floats)' % nres),
            'clocks': clocks,
        }
        if e2e_rng is not None:
            line['e2e_rng'] = e2e_rng
        if big is not None:
            line['steady_state_2p24'] = big
        if other is not None:
            line['other_configs'] = other
        if cpu_value is not None:
            line['cpu_baseline'] = {
                'value': cpu_value, 'unit': UNIT, 'cores': torch.get_num_threads(), 'kind': 'port',
                'sample': '6 steps of 2^18 samples of the same workload, torch float64 autograd over '
                          'oracle/planar_torch.py (stand-in for autograd 1.3); os.cpu_count()=%d' % os.cpu_count()}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
