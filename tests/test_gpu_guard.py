"""Out-of-bounds guard bands (compute-sanitizer is closed on this GPU pool, profiles/r2_compute_sanitizer_closed.txt, so
the memcheck SURVEY.md section 5 asks for is done with bounds checks of our own): every output and scratch buffer of the
device-pointer entry points is carved out of a larger allocation whose margins hold a sentinel pattern, at sizes that
leave partial tiles (2048-transition scan tiles, 128- and 32-environment actor tiles, 64-row learner tiles); after the
call the margins must be untouched and the interior equal to a run on plain buffers."""
import numpy as np
import pytest
import torch

from oracle import learner_oracle as lo

pytestmark = pytest.mark.gpu

GUARD = 4096          # elements on either side (keeps 16-byte alignment of the interior)
SENT = -7.25e33


def guarded(n, dtype=torch.float32, dev="cuda:0"):
    sent = SENT if dtype.is_floating_point else -(1 << 40) + 12345
    full = torch.full((n + 2 * GUARD,), sent, dtype=dtype, device=dev)
    return full, full[GUARD:GUARD + n]


def intact(full, n):
    sent = full[0].item()
    return bool((full[:GUARD] == sent).all().item()) and bool((full[GUARD + n:] == sent).all().item())


@pytest.mark.parametrize("n,with_baseline", [(1, True), (2047, True), (2049, False), (1000003, True), (3 * 2048 * 444 + 5, True)])
def test_scan_stays_inside_its_buffers(n, with_baseline):
    from relayrl_b200 import scan_device, scan_workspace_bytes
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(n)
    lens = np.minimum(rng.geometric(1 / 300.0, size=n // 200 + 2), 5000)
    offs = np.concatenate([[0], np.cumsum(lens)])
    offs = offs[offs < n]
    offs = np.concatenate([offs, [n]]).astype(np.int64) if rng.random() < 0.5 else offs.astype(np.int64)   # (un)finished tail
    n_ep = len(offs) - 1
    rew = torch.empty(n, device=dev).uniform_(-1, 1)
    val = torch.randn(n, device=dev)
    lvs = torch.empty(max(n_ep, 1), device=dev).uniform_(-1, 1)
    d_offs = torch.from_numpy(offs).to(dev)
    wsb = scan_workspace_bytes(n)
    assert wsb % 8 == 0
    out = {}
    for name in ("plain", "guarded"):
        if name == "plain":
            adv_f, adv = None, torch.zeros(n, device=dev)
            ret_f, ret = None, torch.zeros(n, device=dev)
            ws_f, ws = None, torch.zeros(wsb // 8, dtype=torch.float64, device=dev)
            st_f, st = None, torch.zeros(6, dtype=torch.float64, device=dev)
        else:
            (adv_f, adv), (ret_f, ret) = guarded(n), guarded(n)
            ws_f, ws = guarded(wsb // 8, torch.float64)
            st_f, st = guarded(6, torch.float64)
            adv.zero_(), ret.zero_(), st.zero_()     # (slot 5 of the statistics is written by the multi-rank merge only)
        scan_device(rew, val, d_offs, lvs, n, n_ep, 0.98, 0.95, with_baseline, adv, ret, st, ws)
        torch.cuda.synchronize()
        out[name] = (adv.clone(), ret.clone(), st.clone())
        if name == "guarded":
            assert intact(adv_f, n), "adv written outside [0, n)"
            assert intact(ret_f, n), "ret written outside [0, n)"
            assert intact(ws_f, wsb // 8), "scan workspace written beyond rrl_scan_workspace_bytes"
            assert intact(st_f, 6), "statistics written beyond 6 doubles"
    for a, b in zip(out["plain"], out["guarded"]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("d,A,H,B", [(128, 18, 256, 129), (128, 18, 256, 148 * 128 + 1), (20, 5, 256, 1000), (4, 2, 128, 127),
                                     (4, 2, 64, 33), (3, 1, 128, 4097)])
def test_actor_stays_inside_its_buffers(d, A, H, B):
    from relayrl_b200 import Actor
    dev = torch.device("cuda:0")
    pi, v = lo.init_params(d, H, A, 5), lo.init_params(d, H, 1, 6)
    a = Actor(d, A, hidden=H, with_baseline=True, max_batch=B)
    a.set_params(0, lo.flatten(pi).numpy()), a.set_params(1, lo.flatten(v).numpy())
    g = torch.Generator(device="cpu").manual_seed(B)
    obs = torch.randn(B, d, generator=g).to(dev)
    mask = torch.ones(B, A, device=dev)
    res = {}
    for name in ("plain", "guarded"):
        if name == "plain":
            bufs = [(None, torch.zeros(B, dtype=torch.int64, device=dev)), (None, torch.zeros(B, device=dev)),
                    (None, torch.zeros(B, device=dev)), (None, torch.zeros(B * A, device=dev)), (None, torch.zeros(B * A, device=dev))]
        else:
            bufs = [guarded(B, torch.int64), guarded(B), guarded(B), guarded(B * A), guarded(B * A)]
        (_, act), (_, lp), (_, vv), (_, pr), (_, lg) = bufs
        a.step_device(obs, mask, act, lp, vv, probs=pr.view(B, A), logits=lg.view(B, A), seed=3, offset=1)
        torch.cuda.synchronize()
        res[name] = [t.clone() for _, t in bufs]
        if name == "guarded":
            for (full, t), what in zip(bufs, ("act", "logp_a", "v", "probs", "logits")):
                assert intact(full, t.numel()), "%s written outside its %d elements" % (what, t.numel())
    for x, y in zip(res["plain"], res["guarded"]):
        assert torch.equal(x, y)
    a.close()


@pytest.mark.parametrize("n,flags", [(65, 0), (4097, 0), (40000, 0), (70001, 1)])
def test_learner_taps_stay_inside_their_buffers(n, flags):
    """The learner owns its device buffers; what a caller hands it are host arrays (taps) and device inputs.  A capacity of
    exactly n + 1 rows puts the last tile at the very end of every column: an over-read would leave the allocation (caught
    as a CUDA fault by the synchronising calls), an over-write of adv/ret would land in the next column and change it."""
    from relayrl_b200 import Learner
    d, A = 4, 2
    rng = np.random.default_rng(n)
    lens = rng.integers(1, 400, n // 150 + 2)
    offs = np.concatenate([[0], np.cumsum(lens)])
    offs = np.concatenate([offs[offs < n], [n]]).astype(np.int64)
    E = len(offs) - 1
    dev = torch.device("cuda:0")
    cols = dict(obs=torch.randn(n, d, device=dev), act=torch.randint(0, A, (n,), device=dev).float(), mask=torch.ones(n, A, device=dev),
                rew=torch.empty(n, device=dev).uniform_(-1, 1), val=torch.randn(n, device=dev), logp=torch.randn(n, device=dev))
    lvs = rng.uniform(-1, 1, E).astype(np.float32)
    stats = []
    for cap in (n + 1, 2 * n + 4099):
        L = Learner(d, A, cap, hidden=128, with_baseline=True, train_vf_iters=2, flags=flags)
        L.set_params(0, lo.flatten(lo.init_params(d, 128, A, 1)).numpy())
        L.set_params(1, lo.flatten(lo.init_params(d, 128, 1, 2)).numpy())
        L.ingest_device(cols["obs"], cols["act"], cols["mask"], cols["rew"], cols["val"], cols["logp"], n, offs, lvs)
        torch.cuda.synchronize()
        L.finish()
        ret, adv = L.returns(n).copy(), L.advantages(n, False).copy()
        st = L.train_epoch()
        stats.append((ret, adv, st["loss_pi"], st["loss_v"], L.get_params(0), L.get_params(1)))
        L.close()
    for x, y in zip(*stats):
        assert np.array_equal(np.asarray(x), np.asarray(y))
