"""Learner parity AT THE BENCHMARKED SIZE (4 Mi and 64 Mi transitions, the bench's C3 episode mixture): losses and
gradients of both networks from the CUDA path (rrl_eval_grad, through the C ABI) against a float64 evaluation of the
same device buffers by plain torch (tests/f64_eval.py) -- SURVEY.md section 7's tie-breaker at sizes where the
CPU oracle would take hours.  What this pins that the small cases cannot: ~7 000 tiles per CTA, the 2048-row f32
accumulation window spilling into the per-CTA float64 partials thousands of times, and 1/N at N = 2^22 .. 2^26.

Criterion (the one of test_tc_gradients_many_tiles_vs_float64_autograd at 200x the size): losses rel 1e-5; each
gradient at least as close to the float64 result as twice the distance of a float32 evaluation of the same formulas
(torch, TF32 off), or 3e-6 of its largest entry, whichever is larger."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import f64_eval  # noqa: E402

pytestmark = pytest.mark.gpu

D, A, H = 4, 2, 128


def _c3_lengths(rng, n):
    lens, tot = [], 0
    while tot < n:
        u = rng.random()
        k = int(rng.integers(1, 1001)) if u < 0.9 else int(rng.integers(1000, 20001)) if u < 0.99 \
            else int(rng.integers(20000, 2000001))
        k = min(k, n - tot)
        lens.append(k)
        tot += k
    return np.asarray(lens, np.int64)


def _linear_init(rng, d, hidden, out):
    ps = []
    for fi, fo in ((d, hidden), (hidden, hidden), (hidden, out)):
        b = 1.0 / np.sqrt(fi)
        ps += [rng.uniform(-b, b, (fo, fi)).astype(np.float32).reshape(-1), rng.uniform(-b, b, fo).astype(np.float32)]
    return np.concatenate(ps)


@pytest.mark.parametrize("n", [1 << 22, 1 << 26])
def test_losses_and_gradients_vs_float64_at_bench_size(n):
    from relayrl_b200 import Learner
    dev = torch.device("cuda", 0)
    free, _ = torch.cuda.mem_get_info()
    if free < 40 * n * 4:
        pytest.skip("not enough free device memory for %d transitions" % n)
    rng = np.random.default_rng(n % 1000)
    lens = _c3_lengths(rng, n)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    last_vals = rng.uniform(-1, 1, len(lens)).astype(np.float32)
    g = torch.Generator(device=dev).manual_seed(n % 1000)
    cols = {"obs": torch.randn(n, D, device=dev, generator=g),
            "act": torch.randint(0, A, (n,), device=dev, generator=g).float(),
            "mask": torch.ones(n, A, device=dev),
            "rew": torch.empty(n, device=dev).uniform_(-1, 1, generator=g),
            "val": torch.randn(n, device=dev, generator=g),
            "logp": torch.randn(n, device=dev, generator=g)}
    sel = torch.nonzero(torch.rand(n, device=dev, generator=g) < 0.1).squeeze(1)     # 10 %: the other action masked
    cols["mask"][sel, (1 - cols["act"][sel]).long()] = 0.0
    pi0, v0 = _linear_init(rng, D, H, A), _linear_init(rng, D, H, 1)
    L = Learner(D, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=1)
    L.set_params(0, pi0), L.set_params(1, v0)
    torch.cuda.synchronize()
    L.ingest_device(cols["obs"], cols["act"], cols["mask"], cols["rew"], cols["val"], cols["logp"], n, offsets, last_vals)
    assert L.finish() == (n, len(lens))
    assert np.array_equal(L.offsets(), offsets)
    adv = torch.from_numpy(L.advantages(n, True).copy()).to(dev)      # ReplayBuffer.get()'s normalised advantages
    ret = torch.from_numpy(L.returns(n).copy()).to(dev)
    gp, (lpi, kl, ent) = L.eval_grad(0)
    gv, (lv, _, _) = L.eval_grad(1)
    L.close()
    r64 = f64_eval.evaluate(pi0, v0, D, H, A, cols, adv, ret, dtype=torch.float64)
    torch.backends.cuda.matmul.allow_tf32 = False
    r32 = f64_eval.evaluate(pi0, v0, D, H, A, cols, adv, ret, dtype=torch.float32)
    np.testing.assert_allclose(lpi, r64["loss_pi"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(kl, r64["kl"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(ent, r64["entropy"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(lv, r64["loss_v"], rtol=1e-5)
    for name, got, key in (("policy", gp, "pi_grad"), ("value", gv, "v_grad")):
        want, w32 = r64[key], r32[key]
        e_gpu, e_ref, scale = np.abs(got - want).max(), np.abs(w32 - want).max(), np.abs(want).max()
        print("n=%d %s gradient: max |err| vs float64 %.3g (torch float32: %.3g), largest entry %.3g"
              % (n, name, e_gpu, e_ref, scale))
        assert e_gpu <= max(2.0 * e_ref, 3e-6 * scale), (name, e_gpu, e_ref, scale)
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=max(2.0 * e_ref, 3e-6 * scale), err_msg=name)
