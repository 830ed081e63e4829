#!/usr/bin/env python
"""bench.py -- learner transitions/s of the RelayRL hot path on B200 (BASELINE.json metric).

A step is ONE full learner epoch on one batch of synthetic trajectories:
  ingest -> reward-to-go/GAE scan + statistics -> policy step -> train_vf_iters value steps
(REINFORCE.receive_trajectory + train_model, reference REINFORCE.py:70-125).
Workload (config.workload): REINFORCE with baseline, d=4, A=2, H=128, 80 value iterations, C3-style
variable episode lengths, N = --transitions per GPU (default 64 Mi = BASELINE's target batch).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--transitions N_PER_GPU] [--impl reference]
Under torchrun (N > 1) every rank owns its own shard (weak scaling); gradients/statistics are
all-reduced over NCCL inside the library.  One JSON line is printed by rank 0.
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
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

D, A, H, ITERS = 4, 2, 128, 80
GAMMA, LAM = 0.98, 0.97


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p["hbm_gbs"], p["bf16_tflops"], p["bf16_tflops_sustained"], "measured"
    except Exception:
        return 6650.0, 1590.0, 1400.0, "fallback"


def episode_lengths(rng, n):
    """C3 mixture: 90 % U{1..1000}, 9 % U{1000..20000}, 1 % U{20000..2000000}; the last
    episode is truncated to fit."""
    parts, tot = [], 0
    while tot < n:
        k = 4096
        u = rng.random(k)
        L = np.where(u < 0.9, rng.integers(1, 1001, k), np.where(u < 0.99, rng.integers(1000, 20001, k),
                                                                 rng.integers(20000, 2000001, k)))
        parts.append(L)
        tot += int(L.sum())
    L = np.concatenate(parts).astype(np.int64)
    c = np.cumsum(L)
    m = int(np.searchsorted(c, n, side="left"))      # first episode whose end reaches n
    L = L[:m + 1].copy()
    L[m] = n - (int(c[m - 1]) if m > 0 else 0)
    return L


def make_stream(n, seed, pinned):
    """Synthetic action stream of n transitions + one terminal marker per episode, SoA, host."""
    import torch
    rng = np.random.default_rng(seed)
    lens = episode_lengths(rng, n)
    E = len(lens)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    assert offsets[-1] == n
    rows = n + E
    g = torch.Generator().manual_seed(seed)

    def buf(shape, dtype=torch.float32):
        return torch.empty(shape, dtype=dtype, pin_memory=pinned)

    obs = buf((rows, D)).normal_(generator=g)
    act = buf((rows,)).copy_(torch.randint(0, A, (rows,), generator=g).float())
    mask = buf((rows, A)).fill_(1.0)
    # 10 % of rows have the action that was NOT taken masked out
    sel = torch.rand(rows, generator=g) < 0.1
    idx = torch.nonzero(sel).squeeze(1)
    mask[idx, (1 - act[idx]).long()] = 0.0
    rew = buf((rows,)).uniform_(-1, 1, generator=g)
    val = buf((rows,)).normal_(generator=g)
    logp = buf((rows,)).normal_(generator=g)
    done = torch.zeros(rows, dtype=torch.uint8, pin_memory=pinned)
    marker_rows = offsets[1:] + np.arange(E)          # row index of each terminal marker
    done[torch.from_numpy(marker_rows)] = 1
    return dict(obs=obs, act=act, mask=mask, rew=rew, val=val, logp=logp, done=done, rows=rows, n=n, E=E,
                offsets=offsets, marker_rows=marker_rows)


def device_columns(st, dev):
    """Transitions only (markers removed), resident in HBM, + host offsets / last_vals."""
    import torch
    keep = torch.from_numpy(np.flatnonzero(st["done"].numpy() == 0))
    cols = {k: st[k][keep].to(dev) for k in ("obs", "act", "mask", "rew", "val", "logp")}
    last_vals = st["rew"].numpy()[st["marker_rows"]].copy()
    return cols, st["offsets"], last_vals


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            self._halt.wait(0.2)

    def finish(self):
        self._halt.set()
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def cpu_baseline(sample_n, iters, threads, seed=7, store_prefix=0):
    """The oracle port of the reference CPU learner (oracle/learner_oracle.py) on a bounded sample of
    the same workload: ingest (segment + store) -> finish_path per episode -> get() -> train_model.
    Returns (transitions/s, seconds, stage split).  The port's ingest is VECTORISED (one numpy gather per column) where
    the reference loops per transition in Python (REINFORCE.py:72-87): that is generous to the CPU; with store_prefix
    > 0 the reference-style loop (oracle.store_loop) is timed on that many actions beside it."""
    import torch
    from oracle import learner_oracle as lo
    torch.set_num_threads(threads)
    st = make_stream(sample_n, seed, pinned=False)
    O = lo.OracleLearner(D, A, sample_n + 1, hidden=H, with_baseline=True, gamma=GAMMA, lam=LAM,
                         train_vf_iters=iters, seed=1)
    cols = [st[k].numpy() for k in ("obs", "act", "mask", "rew", "val", "logp", "done")]
    t0 = time.perf_counter()
    O.ingest(*cols)
    t1 = time.perf_counter()
    data, _ = O.get()
    t2 = time.perf_counter()
    out = lo.train_epoch(O.pi, O.v, data, O.pi_lr, O.vf_lr, O.iters)
    t3 = time.perf_counter()
    split = {"ingest_vectorised_s": t1 - t0, "get_s": t2 - t1, "train_model_s": t3 - t2,
             "transitions_per_s_without_ingest": sample_n / (t3 - t1)}
    if store_prefix:
        k = min(store_prefix, st["rows"])
        t4 = time.perf_counter()
        lo.store_loop(*[c[:k] for c in cols], lo_f32(GAMMA), lo_f32(LAM), True)
        split["reference_style_store_loop"] = {"actions": int(k), "actions_per_s": k / (time.perf_counter() - t4)}
    dt = t3 - t0
    return sample_n / dt, dt, split


def lo_f32(x):
    return float(np.float32(x))


def run_reference(args):
    """--impl reference: the reference's CPU learner (oracle port; the reference is Python + Rust and cannot travel to
    the GPU box) on the host cores.  Each timed step is one full epoch on a BOUNDED SAMPLE of the workload
    (--ref-sample transitions, not the GPU arm's 64 Mi): `config` says so and `extrapolated_to` names the size the
    GPU arm runs.  After the timed steps the epoch is run once at each --ref-linearity size: the per-transition cost is
    flat to slightly rising with N, so the sample does not flatter the GPU."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = args.ref_sample
    times, split = [], None
    for i in range(args.warmup + args.steps):
        tps, dt, sp = cpu_baseline(sample, ITERS, threads, seed=7 + i,
                                   store_prefix=(1 << 20) if i == args.warmup + args.steps - 1 else 0)
        if i >= args.warmup:
            times.append(dt)
            split = sp
    ms = 1000.0 * sum(times) / len(times)
    value = sample / (ms / 1000.0)
    linearity = [{"transitions": sample, "transitions_per_s": value, "us_per_transition": 1e6 / value}]
    for n in args.ref_linearity:
        if n > 0:
            tps, dt, _ = cpu_baseline(n, ITERS, threads, seed=99)
            linearity.append({"transitions": n, "transitions_per_s": tps, "us_per_transition": 1e6 / tps})
    sample_desc = "%d transitions (C3 length mixture), full epoch with %d value iterations, per step" % (sample, ITERS)
    cfg = workload_config(args, sample)
    cfg["sampled"] = True
    cfg["extrapolated_to"] = args.n
    cfg["note"] = "CPU arm timed on a bounded sample of the GPU arm's workload; per-transition cost vs size in `linearity`"
    print(json.dumps({
        "impl": "reference", "metric": "learner transitions/sec", "value": value, "unit": "transitions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": "transitions/s", "cores": threads, "kind": "port",
                         "sample": sample_desc, "stages": split, "linearity": linearity,
                         "ingest_note": "the port's ingest is vectorised; the reference's per-transition store loop is "
                                        "timed separately (stages.reference_style_store_loop)"},
        "e2e": {"value": value, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(args, n):
    return {"workload": "REINFORCE_with_baseline epoch: ingest+scan+normalise+1 policy step+%d value steps; "
                        "d=%d A=%d H=%d; %d transitions per GPU; C3 episode-length mixture" % (ITERS, D, A, H, n),
            "transitions_per_gpu": n, "obs_dim": D, "act_dim": A, "hidden": H, "train_vf_iters": ITERS,
            "l2": "inputs (%.1f GB per epoch pass) exceed the 126 MB L2" % (n * 4 * (D + A + 6) / 1e9),
            "parallelism": "dp%d" % args.gpus}


def measure_scan(cols, offsets, last_vals, n, dev, with_baseline=True, reps=20):
    """K2 alone: median CUDA-event time of rrl_scan_device on the stream it is launched on
    (torch's current stream).  Inputs (0.5 GB at 64 Mi) exceed the 126 MB L2."""
    import torch
    from relayrl_b200 import scan_device, scan_workspace_bytes
    ws = torch.empty(scan_workspace_bytes(n), dtype=torch.uint8, device=dev)
    adv, ret = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    sstats = torch.zeros(8, dtype=torch.float64, device=dev)
    d_off, d_lv = torch.from_numpy(offsets).to(dev), torch.from_numpy(last_vals).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    e0 = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
    e1 = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
    for i in range(3 + reps):
        if i >= 3:
            e0[i - 3].record()
        scan_device(cols["rew"], cols["val"], d_off, d_lv, n, len(last_vals), GAMMA, LAM, with_baseline, adv, ret,
                    sstats, ws, stream=stream)
        if i >= 3:
            e1[i - 3].record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in zip(e0, e1))[reps // 2]


# (21.150976 MB + 0.020480 MB) / 1 048 576 transitions, profiles/r2_tc_kernel_metrics.csv
NCU_VALUE_KERNEL_BYTES_PER_TRANSITION = (21150976 + 20480) / 1048576.0
# (551.635456 MB + 487.865856 MB) / 67 108 864 transitions, profiles/r2_scan_kernel_metrics.csv
NCU_SCAN_KERNEL_BYTES_PER_TRANSITION = (551635456 + 487865856) / 67108864.0


def linear_init(d, hidden, out, seed):
    """torch.nn.Linear's default init (U(-1/sqrt(fan_in), 1/sqrt(fan_in))) for a d->hidden->hidden->out MLP, as a list
    of (W [out,in], b) arrays in torch order and as the flat vector the C ABI takes.  bench.py's own copy: the
    product side of the bench must not reach into oracle/."""
    rng = np.random.default_rng(seed)
    ps = []
    for fan_in, fan_out in ((d, hidden), (hidden, hidden), (hidden, out)):
        b = 1.0 / np.sqrt(fan_in)
        ps.append(rng.uniform(-b, b, (fan_out, fan_in)).astype(np.float32))
        ps.append(rng.uniform(-b, b, fan_out).astype(np.float32))
    return ps, np.concatenate([p.reshape(-1) for p in ps])


def measure_normalise(n, dev, reps=20):
    """K3 alone (the get()-style tap): 8 B per transition, median CUDA-event time on the launching stream."""
    import torch
    from relayrl_b200 import normalise_device
    adv, out = torch.randn(n, device=dev), torch.empty(n, device=dev)
    stats = torch.tensor([0.0, 0.0, float(n), 0.25, 1.5, 0.0, 0.0, 0.0], dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i in range(3 + reps):
        if i >= 3:
            ev[i - 3][0].record()
        normalise_device(adv, n, stats, out, stream=stream)
        if i >= 3:
            ev[i - 3][1].record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[reps // 2]


def measure_actor(dev, batch=65536, d=128, A_=18, H_=256, reps=20, cpu_batch=65536):
    """C4: one batched PolicyWithBaseline.step per call, device-resident inputs, CUDA events on
    the actor's own stream; beside it the oracle's step (torch CPU, all host threads) on the same shape."""
    import torch
    from relayrl_b200 import Actor
    (pi_np, pi_flat), (vf_np, vf_flat) = linear_init(d, H_, A_, 11), linear_init(d, H_, 1, 12)
    act = Actor(d, A_, hidden=H_, with_baseline=True, device=dev.index or 0, max_batch=batch)
    act.set_params(0, pi_flat)
    act.set_params(1, vf_flat)
    g = torch.Generator(device="cpu").manual_seed(4)
    obs_h = torch.randn(batch, d, generator=g)
    mask_h = (torch.rand(batch, A_, generator=g) > 0.2).float()
    mask_h[:, 0] = 1.0
    obs, mask = obs_h.to(dev), mask_h.to(dev)
    a_out = torch.empty(batch, dtype=torch.int64, device=dev)
    lp, v = torch.empty(batch, device=dev), torch.empty(batch, device=dev)
    st = torch.cuda.ExternalStream(act.stream, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i in range(3 + reps):
        if i >= 3:
            ev[i - 3][0].record(st)
        act.step_device(obs, mask, a_out, lp, v, seed=4, offset=i)
        if i >= 3:
            ev[i - 3][1].record(st)
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)[reps // 2]
    # end to end: the host-buffer call an agent makes (rrl_actor_step): obs and mask from pinned host memory, act / logp_a /
    # v back on the host, wall clock around the (synchronous) call
    obs_np, mask_np = obs_h.pin_memory().numpy(), mask_h.pin_memory().numpy()
    e2e = []
    for i in range(3 + reps):
        t0 = time.perf_counter()
        act.step(obs_np, mask_np, seed=4, offset=100 + i)
        e2e.append(time.perf_counter() - t0)
    e2e_ms = sorted(e2e[3:])[reps // 2] * 1e3
    act.close()
    # CPU baseline leg: the oracle's restatement of PolicyWithBaseline.step on the same weights
    from oracle import learner_oracle as lo
    pi, vf = [torch.from_numpy(p) for p in pi_np], [torch.from_numpy(p) for p in vf_np]
    torch.set_num_threads(os.cpu_count() or 1)
    lo.actor_step(pi, vf, obs_h[:cpu_batch], mask_h[:cpu_batch])
    t0 = time.perf_counter()
    for _ in range(3):
        lo.actor_step(pi, vf, obs_h[:cpu_batch], mask_h[:cpu_batch])
    cpu_s = (time.perf_counter() - t0) / 3
    flops = 2.0 * (d * H_ + H_ * H_ + H_ * A_) + 2.0 * (d * H_ + H_ * H_ + H_)
    return {"workload": "C4: batch %d, d=%d, A=%d, H=%d, with baseline (policy + value + sampling per env)" % (batch, d, A_, H_),
            "env_steps_per_s": batch / (ms * 1e-3), "ms_per_call": ms, "tflops": flops * batch / (ms * 1e-3) / 1e12,
            "executed_bf16_tflops": 6.0 * 4.0 * (d * H_ + H_ * H_) * batch / (ms * 1e-3) / 1e12,
            "kernel": "actor_tc_kernel<256> (tcgen05, bf16x3)",
            "e2e": {"env_steps_per_s": batch / (e2e_ms * 1e-3), "ms_per_call": e2e_ms,
                    "h2d_bytes_per_call": batch * (d + A_) * 4, "d2h_bytes_per_call": batch * 16},
            "cpu_env_steps_per_s": cpu_batch / cpu_s, "cpu_cores": os.cpu_count() or 1, "cpu_kind": "port"}


def measure_without_baseline(dev, local, n=1 << 20, reps=5):
    """C2: REINFORCE without baseline on a 1 Mi-transition batch: ingest from pinned host memory,
    scan, statistics, one policy step; CUDA events on the learner's stream.  Beside it the oracle port on the host."""
    import torch
    from relayrl_b200 import Learner, _lib
    from relayrl_b200._lib import BatchView
    st = make_stream(n, 1234, pinned=True)
    L = Learner(D, A, n + 1, hidden=H, with_baseline=False, gamma=GAMMA, lam=LAM, train_vf_iters=ITERS, device=local)
    L.set_params(0, linear_init(D, H, A, 1)[1])
    view = BatchView(st["obs"].data_ptr(), st["act"].data_ptr(), st["mask"].data_ptr(), st["rew"].data_ptr(),
                     st["val"].data_ptr(), st["logp"].data_ptr(), st["done"].data_ptr(), st["rows"], _lib.RRL_F32,
                     _lib.RRL_F32)
    ls = torch.cuda.ExternalStream(L.stream, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i in range(2 + reps):
        if i >= 2:
            ev[i - 2][0].record(ls)
        L.ingest_raw(view)
        stats = L.train_epoch()
        if i >= 2:
            ev[i - 2][1].record(ls)
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)[reps // 2]
    L.close()
    # CPU leg: the oracle port, same stream, all host threads
    from oracle import learner_oracle as lo
    torch.set_num_threads(os.cpu_count() or 1)
    O = lo.OracleLearner(D, A, n + 1, hidden=H, with_baseline=False, gamma=GAMMA, lam=LAM, train_vf_iters=ITERS, seed=1)
    t0 = time.perf_counter()
    O.ingest(st["obs"].numpy(), st["act"].numpy(), st["mask"].numpy(), st["rew"].numpy(), st["val"].numpy(),
             st["logp"].numpy(), st["done"].numpy())
    O.train()
    cpu_s = time.perf_counter() - t0
    return {"workload": "C2: REINFORCE without baseline, %d transitions, host ingest + scan + statistics + 1 policy step" % n,
            "transitions_per_s": n / (ms * 1e-3), "ms_per_epoch": ms, "stage_ms": {"scan": stats["ms_scan"], "policy": stats["ms_policy"]},
            "cpu_transitions_per_s": n / cpu_s, "cpu_cores": os.cpu_count() or 1, "cpu_kind": "port"}


def correctness_check(L, cols, offsets, last_vals, n, dev, dist, rank, world, final=True):
    """Outside every timed region: the result of the path the bench times, checked on the spot -- once at the initial
    weights (final=False) and once after the timed steps.
    (1) value and policy loss and the value gradient at the CURRENT weights, from the library (rrl_eval_grad: the same
        kernels as a training step, gradient all-reduced over the ranks) against a float64 evaluation of the same device
        buffers by plain torch (tests/f64_eval.py; sums all-reduced over the ranks): rel 1e-5 on the losses, always; 3e-5 in
        the L2 norm on the gradient (tests/test_gpu_scale.py explains the bound) at the initial weights.  After the timed
        steps the synthetic targets (noise) are fitted as far as they can be: the per-row terms of the gradient cancel to
        ~1e-4 of their size, so its relative error is float32 row noise times that amplification (measured 4e-4 .. 2e-3
        on one and on two GPUs alike, csrc/mlp_tc.cu "Accuracy note") -- reported, not asserted;
    (2) at N > 1: every rank holds bit-identical weights after the timed steps (same all-reduced float64 gradient, same
        Adam) -- the data-parallel replicas never need a broadcast.
    A failure raises: a SCALE run that diverged does not print a line."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import f64_eval
    pi, vf = L.get_params(0), L.get_params(1)
    out = {}
    if world > 1 and final:
        mine = torch.from_numpy(np.concatenate([pi, vf])).to(dev)
        ref = mine.clone()
        dist.broadcast(ref, 0)
        same = torch.tensor([int(torch.equal(mine, ref))], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        out["replicas_bit_identical"] = bool(same.item())
        assert out["replicas_bit_identical"], "data-parallel replicas hold different weights"
    L.ingest_device(cols["obs"], cols["act"], cols["mask"], cols["rew"], cols["val"], cols["logp"], n, offsets, last_vals)
    L.finish()
    adv = torch.from_numpy(L.advantages(n, True).copy()).to(dev)
    ret = torch.from_numpy(L.returns(n).copy()).to(dev)
    gp, lp = L.eval_grad(0)
    gv, lv = L.eval_grad(1)
    L.reset_buffer()
    got = torch.tensor([lp[0], lv[0]], dtype=torch.float64, device=dev)   # this rank's sums over the global N
    r = f64_eval.evaluate(pi, vf, D, H, A, cols, adv, ret, n_global=n * world, dtype=torch.float64)
    want = torch.tensor([r["loss_pi"], r["loss_v"]], dtype=torch.float64, device=dev)
    g64 = torch.from_numpy(r["v_grad"]).to(dev)
    if world > 1:
        dist.all_reduce(got), dist.all_reduce(want), dist.all_reduce(g64)
    got, want = got.tolist(), want.tolist()
    g64 = g64.cpu().numpy()
    out.update(loss_v=got[1], loss_v_float64=want[1], loss_pi=got[0], loss_pi_float64=want[0],
               value_grad_rel_l2=float(np.linalg.norm(gv - g64) / np.linalg.norm(g64)))
    blocks, o = {}, 0
    for name, k in (("W1", H * D), ("b1", H), ("W2", H * H), ("b2", H), ("W3", H), ("b3", 1)):
        blocks[name] = float(np.linalg.norm(gv[o:o + k] - g64[o:o + k]) / max(np.linalg.norm(g64[o:o + k]), 1e-300))
        o += k
    out["value_grad_rel_l2_by_block"] = blocks
    assert abs(got[1] - want[1]) <= 1e-5 * abs(want[1]), out
    assert abs(got[0] - want[0]) <= 1e-5 * abs(want[0]) + 1e-7, out
    if not final:
        assert out["value_grad_rel_l2"] <= 3e-5, out
    out["passed"] = True
    return out


def measure_c1(dev, local, hidden, epochs=20):
    """C1: CartPole-shaped epochs -- 8 trajectories of U{8..500} transitions, f64 obs/mask on the
    wire, with baseline, 80 value iterations -- through rrl_ingest_packed + rrl_train_epoch (the calls the plug-in
    makes): latency per epoch, host wall clock around the calls (the regime is launch/latency-bound, not bandwidth-
    bound).  Beside it the oracle port on the host cores, same epochs."""
    import torch
    from relayrl_b200 import Learner
    rng = np.random.default_rng(1)
    L = Learner(D, A, 8 * 500 + 8, hidden=hidden, with_baseline=True, gamma=GAMMA, lam=LAM, train_vf_iters=ITERS, device=local)
    L.set_params(0, linear_init(D, hidden, A, 1)[1])
    L.set_params(1, linear_init(D, hidden, 1, 2)[1])
    eps = []
    for _ in range(epochs + 3):
        lens = rng.integers(8, 501, 8)
        done = np.concatenate([np.r_[np.zeros(k, np.uint8), np.uint8(1)] for k in lens])
        n = len(done)
        eps.append((rng.standard_normal((n, D)), rng.integers(0, A, n).astype(np.float32), np.ones((n, A)),
                    np.ones(n, np.float32), rng.standard_normal(n).astype(np.float32),
                    rng.standard_normal(n).astype(np.float32), done))
    times, launches = [], 0
    for i, e in enumerate(eps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        L.ingest(*e)
        st = L.train_epoch()          # returns after the stream has drained (stats come back to the host)
        if i >= 3:
            times.append(time.perf_counter() - t0)
            launches = st["kernel_launches"]
    L.close()
    from oracle import learner_oracle as lo
    torch.set_num_threads(os.cpu_count() or 1)
    O = lo.OracleLearner(D, A, 8 * 500 + 8, hidden=hidden, with_baseline=True, gamma=GAMMA, lam=LAM, train_vf_iters=ITERS, seed=1)
    ctimes = []
    for i, e in enumerate(eps[:6]):
        t0 = time.perf_counter()
        O.ingest(*e)
        O.train()
        if i >= 1:
            ctimes.append(time.perf_counter() - t0)
    ms, cms = 1e3 * sorted(times)[len(times) // 2], 1e3 * sorted(ctimes)[len(ctimes) // 2]
    mean_n = float(np.mean([int((e[6] == 0).sum()) for e in eps]))
    return {"workload": "C1: 8 trajectories per epoch (%.0f transitions on average), d=4 A=2 H=%d, with baseline, 80 value iterations" % (mean_n, hidden),
            "ms_per_epoch": ms, "kernel_launches_per_epoch": launches, "transitions_per_s": mean_n / (ms * 1e-3),
            "cpu_ms_per_epoch": cms, "cpu_cores": os.cpu_count() or 1, "cpu_kind": "port"}


def run_scan_only(args):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n = args.n
    rng = np.random.default_rng(64)
    lens = episode_lengths(rng, n)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    cols = {"rew": torch.empty(n, device=dev).uniform_(-1, 1), "val": torch.randn(n, device=dev)}
    last_vals = rng.uniform(-1, 1, len(lens)).astype(np.float32)
    hbm, _, _, which = peaks()
    out = {}
    for wb, bpt in ((True, 16), (False, 12)):
        ms = measure_scan(cols, offsets, last_vals, n, dev, with_baseline=wb)
        gbs = bpt * n / (ms * 1e-3) / 1e9
        out["with_baseline" if wb else "without_baseline"] = {"launch_ms": ms, "achieved": gbs, "unit": "GB/s",
                                                              "peak": hbm, "frac": gbs / hbm, "bytes_per_transition": bpt}
    print(json.dumps({"scan_only": out, "n": n, "episodes": int(len(lens)), "peak_source": which}))


def run_ours(args):
    import torch
    from relayrl_b200 import Learner, _lib, nccl_unique_id, scan_device, scan_workspace_bytes
    from relayrl_b200._lib import BatchView
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist = None
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.tensor(list(nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        nccl_id = bytes(idt.cpu().tolist())

    n = args.n
    st = make_stream(n, 64 + rank, pinned=True)
    L = Learner(D, A, n + 1, hidden=H, with_baseline=True, gamma=GAMMA, lam=LAM, train_vf_iters=ITERS,
                device=local, rank=rank, nranks=world, nccl_id=nccl_id)
    L.set_params(0, linear_init(D, H, A, 1)[1])
    L.set_params(1, linear_init(D, H, 1, 2)[1])
    cols, offsets, last_vals = device_columns(st, dev)
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def step_device():
        L.ingest_device(cols["obs"], cols["act"], cols["mask"], cols["rew"], cols["val"], cols["logp"], n, offsets,
                        last_vals)
        return L.train_epoch()

    view = BatchView(st["obs"].data_ptr(), st["act"].data_ptr(), st["mask"].data_ptr(), st["rew"].data_ptr(),
                     st["val"].data_ptr(), st["logp"].data_ptr(), st["done"].data_ptr(), st["rows"], _lib.RRL_F32,
                     _lib.RRL_F32)

    def step_e2e():
        L.ingest_raw(view)          # pinned host -> HBM inside the timed region
        return L.train_epoch()      # stats come back to the host (D2H) inside the call

    # every kernel and copy of a step runs on the learner's own stream: the timed region is bracketed by CUDA
    # events recorded on THAT stream (torch's current stream would see nothing), with a barrier +
    # synchronize on both sides; the job's time is the max over ranks
    lib_stream = torch.cuda.ExternalStream(L.stream, device=dev)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(lib_stream)
        stats = [fn() for _ in range(steps)]
        e1.record(lib_stream)
        barrier()
        dt = e0.elapsed_time(e1) / 1000.0
        if dist is not None:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = t.item()
        return dt, stats

    check0 = None if args.no_check else correctness_check(L, cols, offsets, last_vals, n, dev, dist, rank, world, final=False)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dt, stats = timed(step_device, args.steps, args.warmup)
    clocks = sampler.finish() if rank == 0 else None
    dt_e2e, stats_e2e = timed(step_e2e, args.steps, 1)   # the same number of timed steps as `value`
    e2e_steps = args.steps
    check = None if args.no_check else {"initial_weights": check0,
                                        "after_timed_steps": correctness_check(L, cols, offsets, last_vals, n, dev, dist, rank, world)}

    scan_ms = measure_scan(cols, offsets, last_vals, n, dev)
    extras = {}
    if rank == 0 and world == 1 and not args.no_extras:
        del cols                      # the learner keeps its own copy
        torch.cuda.empty_cache()
        norm_ms = measure_normalise(n, dev)
        extras["roofline_normalise"] = {"kernel": "normalise_kernel (K3, get()-style tap only)", "bound": "hbm",
                                        "achieved": 8.0 * n / (norm_ms * 1e-3) / 1e9, "unit": "GB/s", "launch_ms": norm_ms,
                                        "bytes_per_transition": 8}
        extras["actor"] = measure_actor(dev)
        extras["without_baseline"] = measure_without_baseline(dev, local)
        extras["c1"] = [measure_c1(dev, local, 64), measure_c1(dev, local, 128)]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    hbm, bf16_burst, bf16_sust, which = peaks()
    ms = 1000.0 * dt / args.steps
    total_n = n * world
    value = total_n / (dt / args.steps)
    ms_e2e = 1000.0 * dt_e2e / e2e_steps
    s = stats[-1]
    # dominant kernel: the fused value-net forward+backward pass (80 launches per step)
    flops_per_launch = 6.0 * (D * H + H * H + H) * n
    value_launch_ms = sum(x["ms_value"] for x in stats) / len(stats) / ITERS
    achieved_tf = flops_per_launch / (value_launch_ms * 1e-3) / 1e12
    scan_gbs = 16.0 * n / (scan_ms * 1e-3) / 1e9
    cpu_threads = os.cpu_count() or 1
    # the reported baseline is taken on rank 0 at N = 1 only (with N ranks on one host the oracles would fight for the
    # cores); the driver's separate --impl reference run is the comparable figure at every N
    cpu_tps, cpu_dt, cpu_split = cpu_baseline(args.cpu_sample, ITERS, cpu_threads) if world == 1 else (None, 0.0, None)
    line = {
        "metric": "learner transitions/sec", "value": value, "unit": "transitions/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, n),
        "e2e": {"value": total_n / (ms_e2e / 1000.0), "unit": "transitions/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": int(n * 4 * (D + A + 4)), "d2h_bytes_per_step": int(C.sizeof(_lib.EpochStats) + 112)},
        "gpu_launches": int(sum(x["kernel_launches"] for x in stats)),
        "clocks": clocks,
        "roofline": {"kernel": "mlp_tc_kernel<VALUE> (fused value fwd+bwd: 3 GEMMs per 64-row tile on tcgen05, bf16x3)",
                     "bound": "tensor", "achieved": achieved_tf, "peak": bf16_sust, "unit": "TFLOP/s",
                     "frac": achieved_tf / bf16_sust, "traffic": NCU_VALUE_KERNEL_BYTES_PER_TRANSITION * n,
                     "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture at 1 Mi "
                                       "transitions (profiles/r2_tc_kernel_metrics.csv), scaled per transition; "
                                       "algorithmic bytes are 20 per transition",
                     "peak_source": which + " bf16 sustained",
                     "flops_per_launch": flops_per_launch, "launch_ms": value_launch_ms,
                     "executed_bf16_tflops": 6.0 * 3 * 2 * H * H * n / (value_launch_ms * 1e-3) / 1e12,
                     "note": "achieved = algorithmic fp32 flops (6*(D*H+H*H+H) per transition); each fp32 GEMM "
                             "is six bf16 MMAs (bf16x3 split), executed_bf16_tflops counts those"},
        "roofline_scan": {"kernel": "scan_chunk_kernel<with_baseline>", "bound": "hbm", "achieved": scan_gbs, "peak": hbm,
                          "unit": "GB/s", "frac": scan_gbs / hbm, "traffic": NCU_SCAN_KERNEL_BYTES_PER_TRANSITION * n,
                          "traffic_source": "ncu --set full at 64 Mi transitions (profiles/r2_scan_kernel_metrics.csv), "
                                            "scaled per transition", "peak_source": which,
                          "bytes_per_transition": 16, "launch_ms": scan_ms},
        "stage_ms": {"scan": s["ms_scan"], "policy": s["ms_policy"], "value": s["ms_value"]},
        "cpu_baseline": {"value": cpu_tps, "unit": "transitions/s", "cores": cpu_threads, "kind": "port",
                         "sample": "%d transitions, full epoch with %d value iterations, %.1f s" % (args.cpu_sample, ITERS, cpu_dt),
                         "stages": cpu_split} if world == 1 else None,
        "losses": {"loss_pi": s["loss_pi"], "loss_v": s["loss_v"], "entropy": s["entropy"]},
        "check": check,
    }
    if "roofline_normalise" in extras:
        extras["roofline_normalise"].update(peak=hbm, frac=extras["roofline_normalise"]["achieved"] / hbm, peak_source=which)
    line.update(extras)
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--transitions", dest="n", type=int, default=1 << 26, help="transitions per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=1 << 17)
    ap.add_argument("--ref-sample", type=int, default=1 << 17)
    ap.add_argument("--ref-linearity", type=int, nargs="*", default=[1 << 20, 1 << 22],
                    help="sizes at which the CPU arm runs one more epoch each (per-transition cost vs N)")
    ap.add_argument("--no-check", action="store_true", help="skip the float64 correctness check after the timed steps")
    ap.add_argument("--no-extras", action="store_true", help="skip the K3 and actor (C4) side measurements")
    ap.add_argument("--scan-only", action="store_true", help="measure K2 alone (roofline_scan) and exit")
    args = ap.parse_args()
    if args.scan_only:
        run_scan_only(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
