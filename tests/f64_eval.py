"""Test infrastructure: losses and flat gradients of the two networks evaluated by plain torch on the GPU, in
float64 (the tie-breaker SURVEY.md section 7 asks for at sizes the CPU oracle cannot reach) or float32, in chunks so
that 64 Mi transitions fit.  Restates REINFORCE.compute_loss_pi / compute_loss_vf (reference
src/native/python/algorithms/REINFORCE/REINFORCE.py:141-160) and the networks of kernel.py:12-46,78-84.

Used by tests/test_gpu_scale.py and by `bench.py --check` (outside the timed region).  Never imported by the product.
"""
import torch


def unflatten(flat, d, H, out, dtype, dev):
    """torch.nn.Linear order W1,b1,W2,b2,W3,b3 (kernel.py:15-21)."""
    flat = torch.as_tensor(flat).to(dev).to(dtype)
    ps, o = [], 0
    for s in ((H, d), (H,), (H, H), (H,), (out, H), (out,)):
        n = 1
        for k in s:
            n *= k
        ps.append(flat[o:o + n].reshape(s).clone().requires_grad_(True))
        o += n
    return ps


def _mlp3(ps, x):
    W1, b1, W2, b2, W3, b3 = ps
    h = torch.relu(torch.nn.functional.linear(x, W1, b1))
    h = torch.relu(torch.nn.functional.linear(h, W2, b2))
    return torch.nn.functional.linear(h, W3, b3)


def evaluate(pi_flat, v_flat, d, H, A, cols, adv, ret, n_global=None, dtype=torch.float64, chunk=1 << 20):
    """cols: device f32 tensors obs [n,d], act [n], mask [n,A], logp [n]; adv (normalised), ret: device f32 [n].
    Returns dict: loss_pi, kl, entropy, pi_grad (flat), loss_v, v_grad (flat) -- means over n_global (default n);
    sums are formed per chunk and accumulated in `dtype`."""
    dev = cols["obs"].device
    n = cols["obs"].shape[0]
    N = float(n_global if n_global is not None else n)
    pi = unflatten(pi_flat, d, H, A, dtype, dev)
    vf = unflatten(v_flat, d, H, 1, dtype, dev) if v_flat is not None else None
    acc = torch.zeros(4, dtype=torch.float64, device=dev)
    mx = torch.zeros(3, dtype=torch.float64, device=dev)      # max |h1|, max |adv|, max |2 (v - ret)|
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        obs = cols["obs"][lo:hi].to(dtype)
        mask = cols["mask"][lo:hi].to(dtype)
        act = cols["act"][lo:hi].long()
        logits = _mlp3(pi, obs) + (mask - 1) * 1e8
        with torch.no_grad():
            mx[0] = torch.maximum(mx[0], torch.relu(torch.nn.functional.linear(obs, pi[0], pi[1])).max().double())
            mx[1] = torch.maximum(mx[1], adv[lo:hi].abs().max().double())
        logp = logits.gather(-1, act.unsqueeze(-1)).squeeze(-1)          # the RAW masked logit
        l_pi = -(logp * adv[lo:hi].to(dtype)).sum()
        l_pi.backward()
        with torch.no_grad():
            probs = torch.softmax(logits, -1)
            lg = torch.clamp(logits, min=torch.finfo(logits.dtype).min)
            acc[0] += l_pi.double()
            acc[1] += (cols["logp"][lo:hi].to(dtype) - logp).sum().double()
            acc[2] += -(lg * probs).sum().double()
        if vf is not None:
            v = _mlp3(vf, obs).squeeze(-1)
            l_v = ((v - ret[lo:hi].to(dtype)) ** 2).sum()
            l_v.backward()
            acc[3] += l_v.detach().double()
            with torch.no_grad():
                mx[0] = torch.maximum(mx[0], torch.relu(torch.nn.functional.linear(obs, vf[0], vf[1])).max().double())
                mx[2] = torch.maximum(mx[2], (2 * (v - ret[lo:hi].to(dtype))).abs().max().double())
        del obs, mask, logits, logp
    a = (acc / N).tolist()
    m = mx.tolist()
    # Largest change ONE (transition, hidden unit) pair can make to a W2-gradient entry: |dhead| |w3| |h1| / N.  A float32
    # evaluation may take the other branch of a ReLU for a pre-activation within rounding of zero; each such flip moves
    # gradient entries by up to this much, which at N = 2^20 .. 2^22 is above float32 rounding noise.
    flip_pi = m[1] * float(pi[4].abs().max()) * m[0] / N
    out = {"flip_bound_pi": flip_pi, "loss_pi": a[0], "kl": a[1], "entropy": a[2],
           "pi_grad": (torch.cat([p.grad.reshape(-1) for p in pi]).double() / N).cpu().numpy()}
    if vf is not None:
        out["loss_v"] = a[3]
        out["flip_bound_v"] = m[2] * float(vf[4].abs().max()) * m[0] / N
        out["v_grad"] = (torch.cat([p.grad.reshape(-1) for p in vf]).double() / N).cpu().numpy()
    return out
