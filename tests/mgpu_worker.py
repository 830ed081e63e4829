"""Run under torchrun (one rank per GPU): data-parallel learner vs the single-GPU learner on the
same data.  Rank 0 prints 'MGPU_OK' on success.  Used by tests/test_gpu_multi.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
for p in (ROOT, os.path.join(ROOT, "relayrl-prototype_b200")):
    sys.path.insert(0, p)
from oracle import learner_oracle as lo  # noqa: E402
from relayrl_b200 import Learner, nccl_unique_id  # noqa: E402
from relayrl_b200.shard import merge_offsets, shard_episodes, shard_rows  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.tensor(list(nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    nccl_id = bytes(idt.cpu().tolist())

    # the reference's defaults: 80 value iterations (src/sys_utils/config_loader.rs:68-78)
    d, A, H, iters = 4, 2, 128, int(os.environ.get("MGPU_ITERS", "80"))
    n_eps = int(os.environ.get("MGPU_EPISODES", "400"))     # 400 episodes ~ 120 k transitions; 28000 ~ 8 Mi
    with_oracle = os.environ.get("MGPU_ORACLE", "1") == "1"
    rng = np.random.default_rng(21)
    lens = rng.integers(1, 600, n_eps)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n, E = int(offs[-1]), len(lens)
    obs = rng.standard_normal((n, d)).astype(np.float32)
    act = rng.integers(0, A, n).astype(np.float32)
    mask = np.ones((n, A), np.float32)
    # rewards the value net can learn (a smooth function of the observation plus a little noise): over the 80 Adam steps
    # the gradient then stays well above float32 summation noise.  With pure-noise rewards the net fits the mean within a
    # few steps, every gradient entry becomes noise, Adam normalises that noise to steps of +-lr, and two correct float32
    # runs random-walk up to 80 lr apart (measured: 0.059 between the 1-GPU and the 2-GPU run, with all losses within
    # 1e-5) -- that would test Adam's conditioning, not the data-parallel path.
    w_true = rng.standard_normal(d).astype(np.float32)
    rew = (0.5 * np.tanh(obs @ w_true) + 0.05 * rng.uniform(-1, 1, n)).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    logp = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, E).astype(np.float32)
    pi0 = lo.flatten(lo.init_params(d, H, A, 1)).numpy()
    v0 = lo.flatten(lo.init_params(d, H, 1, 2)).numpy()

    def to(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    def run(L, lo_, hi, loc_offs, loc_lvs, epochs=1, grads=None):
        out = []
        for _ in range(epochs):
            def ingest():
                L.ingest_device(to(obs[lo_:hi]), to(act[lo_:hi]), to(mask[lo_:hi]), to(rew[lo_:hi]), to(val[lo_:hi]),
                                to(logp[lo_:hi]), hi - lo_, loc_offs, loc_lvs)
                torch.cuda.synchronize()
                L.finish()
            if grads is not None:      # the gradients at the initial weights (all-reduced when data parallel), then start over
                ingest()
                grads.append((L.eval_grad(0)[0], L.eval_grad(1)[0]))
                L.reset_buffer()
            ingest()
            local_offsets = L.offsets()
            out.append((L.train_epoch(), local_offsets))
        return out

    eb = shard_episodes(offs, world)
    lo_, hi, loc = shard_rows(offs, eb, rank)
    # the same kernel on every shard size: at 4 and 8 GPUs a shard (< 32 768 rows) would otherwise take the CUDA-core kernel
    # while the 1-GPU run takes the tensor-core one (measured: the two paths' policy gradients differ by 8e-6 in rel L2 at
    # this size -- ReLU decisions -- which is a statement about the kernels, not about the data-parallel exchange)
    flags = int(os.environ.get("MGPU_FLAGS", "1"))     # RRL_FLAG_TC_ANY_SIZE
    L = Learner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, device=local, rank=rank,
                nranks=world, nccl_id=nccl_id, flags=flags)
    L.set_params(0, pi0), L.set_params(1, v0)
    g_dp = []
    res = run(L, lo_, hi, loc, lvs[eb[rank]:eb[rank + 1]], grads=g_dp)
    pi_dp, v_dp = L.get_params(0), L.get_params(1)
    # replicas must be bit-identical (same all-reduced gradient, same Adam)
    t = to(np.concatenate([pi_dp, v_dp]))
    ref_t = t.clone()
    dist.broadcast(ref_t, 0)
    assert torch.equal(t, ref_t), "replicas diverged on rank %d" % rank
    gathered = [None] * world
    dist.all_gather_object(gathered, res[0][1])
    L.close()
    if rank == 0:
        assert np.array_equal(merge_offsets(gathered), offs)       # offsets identical after concatenation
        S = Learner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, device=local, flags=flags)
        S.set_params(0, pi0), S.set_params(1, v0)
        g_one = []
        ref = run(S, 0, n, offs, lvs, grads=g_one)
        for (a, _), (b, _) in zip(res, ref):
            assert a["n_transitions"] == b["n_transitions"] == n and a["n_episodes"] == b["n_episodes"] == E
            for k in ("loss_pi", "kl", "entropy", "loss_v", "loss_v_old", "adv_mean", "adv_std"):
                np.testing.assert_allclose(a[k], b[k], rtol=1e-5, atol=1e-7, err_msg=k)
        # (1) ONE evaluation: every rank sums its CTA partials in float64, NCCL adds the ranks in float64, and the result is
        # rounded to f32 once -- the same construction as the single-GPU gradient, so the two differ by float64
        # re-association and by where the 2048-row f32 accumulation windows fall (rank-local tiling).
        for name, gd, go in (("policy", g_dp[0][0], g_one[0][0]), ("value", g_dp[0][1], g_one[0][1])):
            rel = float(np.linalg.norm(gd.astype(np.float64) - go) / np.linalg.norm(go))
            print("MGPU %s gradient at the initial weights, data parallel vs 1 GPU: rel L2 %.3g, max abs %.3g of max |g| %.3g"
                  % (name, rel, np.abs(gd - go).max(), np.abs(go).max()), flush=True)
            assert rel <= 2e-6, (name, rel)

        def strict(a, b):
            e = np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))
            return float((e <= 5e-7 + 1e-5 * np.abs(b)).mean()), float(e.max())

        # (2) after the epoch.  A few Adam steps keep the 1e-7 differences of (1) small: strict agreement.  The reference's
        # 80 value steps do not: Adam divides by sqrt(v_hat), most of the 17 k value weights see gradients that are sums of
        # 120 k cancelling terms, and float32 noise in those sums is normalised into steps of a fraction of lr, every step.
        # Measured on B200 (this worker, 2 GPUs): after 80 steps the float32 CPU reference (torch) keeps 9.5 % of the value
        # weights within rel 1e-5 + abs 5e-7 of its own float64 run, max 16.5 lr; the 1-GPU CUDA run 14.2 %, max 20 lr; the
        # 2-GPU run 14.1 %, max 8 lr; all three losses agree to 4e-7.  So at 80 steps the criterion is the tie-breaker of
        # the data-parallel run must be as close to float64 as the float32 reference and the 1-GPU run.
        few = iters <= 5
        verdict = []
        for which, name, got, want, lr in ((0, "policy", pi_dp, S.get_params(0), 3e-4), (1, "value", v_dp, S.get_params(1), 1e-3)):
            frac, emax = strict(got, want)
            print("MGPU %s vs 1 GPU after %d value steps: strict %.5f max abs err %.3g (%.3g lr)"
                  % (name, iters, frac, emax, emax / lr), flush=True)
            if few or which == 0:
                assert frac >= 0.999, "%s: strict agreement with the 1-GPU run %.5f" % (name, frac)
                np.testing.assert_allclose(got, want, rtol=1e-5, atol=0.1 * lr)
        if with_oracle:
            torch.set_num_threads(8)
            done = np.zeros(n + E, np.uint8)
            rows = np.ones(n + E, bool)
            mark = offs[1:] + np.arange(E)
            done[mark], rows[mark] = 1, False

            def widen(a, fill=0.0):
                out = np.full((n + E,) + a.shape[1:], fill, a.dtype)
                out[rows] = a
                return out

            rew_m = widen(rew)
            rew_m[mark] = lvs
            O = lo.OracleLearner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, seed=1)
            O.ingest(widen(obs), widen(act), widen(mask, 1.0), rew_m, widen(val), widen(logp), done)
            assert np.array_equal(np.asarray(O.offsets), offs)
            data, _ = O.get()
            pi_l, v_l = lo.unflatten(pi0, d, H, A), lo.unflatten(v0, d, H, 1)
            ref = lo.train_epoch(pi_l, v_l, data, O.pi_lr, O.vf_lr, iters)
            ref64 = lo.train_epoch([p.double() for p in pi_l], [p.double() for p in v_l],
                                   {k: t.double() for k, t in data.items()}, O.pi_lr, O.vf_lr, iters)
            a0 = res[0][0]
            print("MGPU losses (data parallel | oracle f32 | oracle f64): pi %.9g %.9g %.9g  v_old %.9g %.9g %.9g  v %.9g %.9g %.9g"
                  % (a0["loss_pi"], ref["LossPi"], ref64["LossPi"], a0["loss_v_old"], ref["LossVOld"], ref64["LossVOld"],
                     a0["loss_v"], ref["LossV"], ref64["LossV"]), flush=True)
            for k, r in (("loss_pi", "LossPi"), ("loss_v", "LossV"), ("loss_v_old", "LossVOld")):
                np.testing.assert_allclose(a0[k], ref[r], rtol=1e-5, atol=1e-7, err_msg=k)
            for which, name, got, key, lr in ((0, "policy", pi_dp, "pi_params", 3e-4), (1, "value", v_dp, "v_params", 1e-3)):
                w32, w64 = lo.flatten(ref[key]).numpy(), lo.flatten(ref64[key]).numpy()
                f_ref, m_ref = strict(w32, w64)
                f_gpu, m_gpu = strict(got, w64)
                f_one, m_one = strict(S.get_params(which), w64)
                print("MGPU %s vs float64 oracle: strict %.5f (float32 reference: %.5f, 1-GPU run: %.5f), max %.3g (%.3g, %.3g)"
                      % (name, f_gpu, f_ref, f_one, m_gpu, m_ref, m_one), flush=True)
                assert f_gpu >= min(f_ref, f_one, 0.999) - 0.02, (name, f_gpu, f_ref, f_one)
                assert m_gpu <= 2.0 * max(m_ref, m_one) + 0.01 * lr, (name, m_gpu, m_ref, m_one)
                if f_ref >= 0.999:   # (where the float32 reference itself is well-conditioned; after 3 value steps it keeps
                    # 97.5 % within the strict bound of its float64 run, max 0.56 lr: gradient entries below Adam's eps)
                    np.testing.assert_allclose(got, w32, rtol=1e-5, atol=0.1 * lr)
        S.close()
        print("MGPU_OK world=%d n=%d" % (world, n))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
