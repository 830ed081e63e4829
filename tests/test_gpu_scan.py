"""GPU parity of K1/K2/K3 (ingest, segmented scan, statistics, normalisation) through the C ABI,
against the oracle and the golden vectors generated from the reference.

Tolerances: offsets bit-exact; ret/adv: the kernel computes in float64 like the reference
 but associates differently, so the f32-rounded outputs may differ by one ulp on rare
elements -> rtol 1e-6 (the north-star bound is 1e-5) with atol 1e-6 for values that cancel to ~0."""
import numpy as np
import pytest

from conftest import load_golden
from oracle import learner_oracle as lo
from oracle import scan_ref as sr

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-6, 1e-6


def _learner(cap, wb, d=3, A=2, **kw):
    from relayrl_b200 import Learner
    return Learner(d, A, cap, hidden=64, with_baseline=wb, **kw)


def _feed(L, rew, val, offsets, last_vals, n_total, d=3, A=2, chunk=None, f64=False):
    """Re-creates the action stream (transitions + terminal markers) and ingests it."""
    n_ep = len(offsets) - 1
    done, r, v = [], [], []
    for e in range(n_ep):
        s, t = int(offsets[e]), int(offsets[e + 1])
        done += [0] * (t - s) + [1]
        r += list(rew[s:t]) + [last_vals[e]]
        v += list(val[s:t]) + [0.0]
    s = int(offsets[-1])
    done += [0] * (n_total - s)
    r += list(rew[s:n_total])
    v += list(val[s:n_total])
    done = np.array(done, np.uint8)
    n = len(done)
    r, v = np.array(r, np.float32), np.array(v, np.float32)
    dt = np.float64 if f64 else np.float32
    obs, mask = np.zeros((n, d), dt), np.ones((n, A), dt)
    act, logp = np.zeros(n, np.float32), np.zeros(n, np.float32)
    step = chunk or n
    for i in range(0, max(n, 1), step):
        sl = slice(i, i + step)
        L.ingest(obs[sl], act[sl], mask[sl], r[sl], v[sl], logp[sl], done[sl])


@pytest.mark.parametrize("name", ["scan_with_baseline", "scan_without_baseline"])
@pytest.mark.parametrize("chunk", [None, 997])
def test_scan_golden(name, chunk):
    g = load_golden(name)
    wb = bool(g["with_baseline"])
    n = len(g["rew"])
    L = _learner(n + 5, wb, gamma=float(g["gamma"]), lam=float(g["lam"]))
    _feed(L, g["rew"], g["val"], g["offsets"], g["last_vals"], n, chunk=chunk, f64=chunk is not None)
    nt, ne = L.finish()
    assert nt == n and ne == len(g["offsets"]) - 1
    assert np.array_equal(L.offsets(), g["offsets"])          # bit-exact segmentation
    np.testing.assert_allclose(L.returns(n), g["ret_raw"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(L.advantages(n, normalised=False), g["adv_raw"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(L.advantages(n, normalised=True), g["adv_norm"], rtol=1e-5, atol=1e-5)
    mean, std = L.adv_stats()
    m0, s0 = lo.adv_statistics(g["adv_raw"])
    assert abs(mean - m0) <= 2e-6 * max(1.0, abs(s0)) and abs(std - s0) <= 2e-6 * s0
    L.close()


def _random_case(rng, n_target, kind):
    if kind == "short":
        lens = rng.integers(1, 40, size=n_target // 20)
    elif kind == "ones":
        lens = np.ones(n_target, np.int64)
    elif kind == "long":  # forces carries across many 2048-wide tiles
        lens = np.array([n_target // 2, 1, n_target // 3, 7, 4097, 2048, 2047, 2049])
    else:  # C3-like mixture
        lens = np.where(rng.random(n_target // 300) < 0.9, rng.integers(1, 1000, n_target // 300),
                        rng.integers(1000, 20000, n_target // 300))
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n = int(offs[-1])
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, len(lens)).astype(np.float32)
    return rew, val, offs, lvs, n


@pytest.mark.parametrize("kind", ["short", "ones", "long", "mix"])
@pytest.mark.parametrize("wb", [True, False])
def test_scan_vs_oracle_random(kind, wb):
    rng = np.random.default_rng(["short", "ones", "long", "mix"].index(kind) * 2 + int(wb))
    rew, val, offs, lvs, n = _random_case(rng, 60000, kind)
    tail = 13
    rew_t = np.concatenate([rew, rng.uniform(-1, 1, tail).astype(np.float32)])
    val_t = np.concatenate([val, np.zeros(tail, np.float32)])
    g, l = float(np.float32(0.99)), float(np.float32(0.95))
    L = _learner(n + tail + 1, wb, gamma=0.99, lam=0.95)
    _feed(L, rew_t, val_t, offs, lvs, n + tail)
    L.finish()
    a_ref, r_ref = sr.finish_paths(rew, val, offs, lvs, g, l, wb)
    assert np.array_equal(L.offsets(), offs)
    got_r, got_a = L.returns(n + tail), L.advantages(n + tail, normalised=False)
    np.testing.assert_allclose(got_r[:n], r_ref, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(got_a[:n], a_ref, rtol=RTOL, atol=ATOL)
    # the unfinished tail keeps what the buffers held: zeros in a fresh learner
    assert not got_a[n:].any() and not got_r[n:].any()
    ulp_frac = np.mean(got_a[:n] != a_ref)
    assert ulp_frac < 1e-3, "more than 0.1%% of advantages differ from the f64 oracle: %g" % ulp_frac
    # statistics include the tail
    full = np.concatenate([a_ref, np.zeros(tail, np.float32)])
    np.testing.assert_allclose(L.advantages(n + tail), lo.normalise(full), rtol=1e-5, atol=1e-5)
    L.close()


def test_empty_paths_and_empty_buffer():
    L = _learner(100, True)
    z = np.zeros(0, np.float32)
    # two terminal markers and nothing else: two empty paths
    L.ingest(np.zeros((2, 3)), np.zeros(2), np.ones((2, 2)), np.array([0.5, -0.5], np.float32), np.zeros(2), np.zeros(2),
             np.array([1, 1], np.uint8))
    assert L.finish() == (0, 2)
    assert np.array_equal(L.offsets(), [0, 0, 0])
    L.reset_buffer()
    assert L.finish() == (0, 0)
    del z
    L.close()


def test_capacity_error_matches_reference_assert():
    from relayrl_b200 import RelayRLB200Error
    L = _learner(10, False)
    n = 11
    with pytest.raises(RelayRLB200Error) as e:
        L.ingest(np.zeros((n, 3)), np.zeros(n), np.ones((n, 2)), np.zeros(n, np.float32), None, np.zeros(n),
                 np.zeros(n, np.uint8))
    assert e.value.status == -3
    L.close()


def test_stale_tail_across_epochs():
    """Epoch 2's unfinished tail keeps epoch 1's adv/ret at those positions."""
    rng = np.random.default_rng(3)
    L = _learner(400, False, gamma=0.98, lam=0.97)
    n1 = 300
    rew1 = rng.uniform(-1, 1, n1).astype(np.float32)
    offs1 = np.array([0, 100, 300], np.int64)
    _feed(L, rew1, np.zeros(n1, np.float32), offs1, np.zeros(2, np.float32), n1)
    L.finish()
    a1 = L.advantages(n1, normalised=False).copy()
    L.reset_buffer()
    n2 = 250
    rew2 = rng.uniform(-1, 1, n2).astype(np.float32)
    offs2 = np.array([0, 200], np.int64)
    _feed(L, rew2, np.zeros(n2, np.float32), offs2, np.zeros(1, np.float32), n2)
    L.finish()
    a2 = L.advantages(n2, normalised=False)
    ref, _ = sr.finish_paths(rew2[:200], None, offs2, np.zeros(1, np.float32), float(np.float32(.98)), float(np.float32(.97)), False)
    np.testing.assert_allclose(a2[:200], ref, rtol=RTOL, atol=ATOL)
    assert np.array_equal(a2[200:], a1[200:250])
    L.close()


def test_stateless_scan_device_large_properties():
    """BASELINE-scale (C3-like lengths) through rrl_scan_device: sampled episodes against the C
    oracle, linearity of the without-baseline scan, and mean/std against a torch f64 reduction."""
    import torch
    from relayrl_b200 import normalise_device, scan_device, scan_workspace_bytes
    rng = np.random.default_rng(64)
    n = 1 << 24
    lens = []
    tot = 0
    while tot < n:
        u = rng.random()
        L = int(rng.integers(1, 1001)) if u < 0.9 else int(rng.integers(1000, 20001)) if u < 0.99 else int(rng.integers(20000, 2000001))
        L = min(L, n - tot)
        lens.append(L)
        tot += L
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    E = len(lens)
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, E).astype(np.float32)
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a).to(dev)
    d_rew, d_val, d_offs, d_lvs = t(rew), t(val), t(offs), t(lvs)
    ws = torch.empty(scan_workspace_bytes(n), dtype=torch.uint8, device=dev)
    g, l = float(np.float32(0.98)), float(np.float32(0.97))
    for wb in (True, False):
        adv, ret = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        stats = torch.zeros(8, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        scan_device(d_rew, d_val, d_offs, d_lvs, n, E, g, l, wb, adv, ret, stats, ws)
        torch.cuda.synchronize()
        a_ref, r_ref = sr.finish_paths(rew, val, offs, lvs, g, l, wb)
        np.testing.assert_allclose(ret.cpu().numpy(), r_ref, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(adv.cpu().numpy(), a_ref, rtol=RTOL, atol=ATOL)
        st = stats.cpu().numpy()
        a64 = adv.double()
        assert st[2] == n
        np.testing.assert_allclose(st[3], a64.mean().item(), rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(st[4], a64.std(unbiased=False).item(), rtol=1e-6)
        out = torch.empty(n, device=dev)
        normalise_device(adv, n, stats, out)
        torch.cuda.synchronize()
        # numpy on the host divides in IEEE f32, as the reference does (replay_buffer.py:105)
        ref = (adv.cpu().numpy() - np.float32(st[3])) / np.float32(st[4])
        assert np.array_equal(out.cpu().numpy(), ref)
    # linearity (no baseline): scan(a*x) == a*scan(x) for a power of two, exactly
    adv2, ret2 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    scan_device(d_rew * 4, d_val, d_offs, d_lvs, n, E, g, l, False, adv2, ret2, None, ws)
    torch.cuda.synchronize()
    assert torch.equal(adv2, adv * 4) and torch.equal(ret2, ret * 4)


@pytest.mark.parametrize("wb", [True, False])
def test_scan_special_values_take_the_slow_path(wb):
    """Zeros, f32 denormals, results in the denormal range and results that overflow to Inf: the scan converts f32 <-> f64 with
    branch-free integer arithmetic and sends a thread whose run holds such a value to the exact conversion;
    both copies must agree with the C oracle (sequential f64 recurrence, IEEE narrowing)."""
    import torch
    from relayrl_b200 import scan_device, scan_workspace_bytes
    rng = np.random.default_rng(11)
    n = 40000
    lens = []
    while sum(lens) < n:
        lens.append(int(rng.integers(1, 700)))
    lens[-1] -= sum(lens) - n
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    E = len(lens)
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, E).astype(np.float32)
    rew[1000:9000] = 0.0                                  # exact zeros in, exact zeros out (without baseline)
    val[1000:9000] = 0.0
    lvs[(offs[1:] > 1000) & (offs[1:] <= 9000)] = 0.0
    rew[10000:10400] = np.float32(1e-41)                  # f32 denormal inputs
    val[10200:10300] = np.float32(-3e-42)
    rew[12000:16000] *= np.float32(1e-37)                 # results in the f32 denormal range
    val[12000:16000] *= np.float32(1e-38)
    rew[20000:20300] = np.float32(3e38)                   # discounted sums overflow to +Inf
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a).to(dev)
    ws = torch.empty(scan_workspace_bytes(n), dtype=torch.uint8, device=dev)
    g, l = float(np.float32(0.98)), float(np.float32(0.97))

    def run(rew_, val_):
        adv, ret = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        scan_device(t(rew_), t(val_), t(offs), t(lvs), n, E, g, l, wb, adv, ret, None, ws)
        torch.cuda.synchronize()
        return adv.cpu().numpy(), ret.cpu().numpy()

    got_a, got_r = run(rew, val)
    with np.errstate(all="ignore"):
        a_ref, r_ref = sr.finish_paths(rew, val, offs, lvs, g, l, wb)
    for got, ref, name in ((got_a, a_ref, "adv"), (got_r, r_ref, "ret")):
        assert not np.isnan(got).any() and not np.isnan(ref).any(), name
        assert np.array_equal(np.isinf(got), np.isinf(ref)), name
        fin = np.isfinite(ref)
        assert np.array_equal(np.sign(got[~fin]), np.sign(ref[~fin])), name
        # denormal results: absolute agreement to one f32 denormal step; elsewhere the usual 1e-6
        np.testing.assert_allclose(got[fin], ref[fin], rtol=RTOL, atol=2e-45, err_msg=name)
        assert np.mean(got[fin] != ref[fin]) < 1e-3, name
    assert np.isinf(got_r).any()
    if not wb:
        assert (got_r[1000:9000] == 0).any() and (np.abs(got_r[12000:16000]) < 1.2e-38).any()
    # Non-finite INPUTS are outside the parity contract (in the reference one NaN reward turns the advantage
    # statistics, and with them every weight, into NaN; here 0 * NaN additionally crosses the episode boundary to
    # the left because a path end is a zero slope of the affine map).  What must hold: the call completes (a data
    # NaN is never mistaken for the look-back's "unset" sentinel) and the poisoned element itself is NaN.
    rew2, val2 = rew.copy(), val.copy()
    rew2[25000] = np.float32("nan")
    val2[30000] = np.float32("inf")
    got_a, got_r = run(rew2, val2)
    assert np.isnan(got_r[25000]) and np.isnan(got_a[25000])
    assert np.isfinite(got_r[39000:]).all()


@pytest.mark.parametrize("gamma,lam", [(0.98, 0.97), (0.9995, 1.0), (1.0, 1.0), (0.5, 0.0)])
@pytest.mark.parametrize("wb", [True, False])
def test_scan_both_kernels_long_paths_vs_c_oracle(gamma, lam, wb):
    """3 Mi transitions with paths of up to 1.2 Mi transitions (hundreds of tiles without a path end) through the
    stateless operator.  gamma = 0.98 and 0.5 run the chunk kernel (every CTA recomputes the few tiles to its right
    instead of waiting for them), gamma = 0.9995 and 1 -- whose influence never decays below the cut-off within a run --
    the look-back kernel.  Both against the sequential float64 recurrence of oracle/scan_ref.c."""
    import torch
    from relayrl_b200 import scan_device, scan_workspace_bytes
    rng = np.random.default_rng(int(gamma * 1e4) + int(wb))
    n = 3 * 1024 * 1024 + 77
    lens, tot = [], 0
    while tot < n:
        k = int(rng.integers(1, 500)) if rng.random() < 0.8 else int(rng.integers(100_000, 1_200_000))
        k = min(k, n - tot)
        lens.append(k)
        tot += k
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    rew = (rng.uniform(-1, 1, n) * (1e-3 if gamma == 1.0 else 1.0)).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, len(lens)).astype(np.float32)
    g, l = float(np.float32(gamma)), float(np.float32(lam))
    dev = torch.device("cuda:0")
    d = {k: torch.from_numpy(v).to(dev) for k, v in dict(rew=rew, val=val, offs=offs, lvs=lvs).items()}
    adv, ret = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    stats = torch.zeros(8, dtype=torch.float64, device=dev)
    ws = torch.empty(scan_workspace_bytes(n), dtype=torch.uint8, device=dev)
    scan_device(d["rew"], d["val"], d["offs"], d["lvs"], n, len(lens), g, l, wb, adv, ret, stats, ws,
                stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    a_ref, r_ref = sr.finish_paths(rew, val, offs, lvs, g, l, wb)
    got_a, got_r = adv.cpu().numpy(), ret.cpu().numpy()
    np.testing.assert_allclose(got_r, r_ref, rtol=RTOL, atol=ATOL * max(1.0, np.abs(r_ref).max() * 1e-3))
    np.testing.assert_allclose(got_a, a_ref, rtol=RTOL, atol=ATOL * max(1.0, np.abs(a_ref).max() * 1e-3))
    assert np.mean(got_a != a_ref) < 1e-3 and np.mean(got_r != r_ref) < 1e-3
    s = stats.cpu().numpy()
    np.testing.assert_allclose(s[0], got_a.astype(np.float64).sum(), rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(s[1], (got_a.astype(np.float64) ** 2).sum(), rtol=1e-9)
