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

    d, A, H, iters = 4, 2, 128, 2   # few Adam steps: later ones amplify float32-level differences (see below)
    rng = np.random.default_rng(21)
    lens = rng.integers(1, 600, 400)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    n, E = int(offs[-1]), len(lens)
    obs = rng.standard_normal((n, d)).astype(np.float32)
    act = rng.integers(0, A, n).astype(np.float32)
    mask = np.ones((n, A), np.float32)
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    logp = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, E).astype(np.float32)
    pi0 = lo.flatten(lo.init_params(d, H, A, 1)).numpy()
    v0 = lo.flatten(lo.init_params(d, H, 1, 2)).numpy()

    def to(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    def run(L, lo_, hi, loc_offs, loc_lvs, epochs=1):
        out = []
        for _ in range(epochs):
            L.ingest_device(to(obs[lo_:hi]), to(act[lo_:hi]), to(mask[lo_:hi]), to(rew[lo_:hi]), to(val[lo_:hi]),
                            to(logp[lo_:hi]), hi - lo_, loc_offs, loc_lvs)
            torch.cuda.synchronize()
            L.finish()
            local_offsets = L.offsets()
            out.append((L.train_epoch(), local_offsets))
        return out

    eb = shard_episodes(offs, world)
    lo_, hi, loc = shard_rows(offs, eb, rank)
    L = Learner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, device=local, rank=rank,
                nranks=world, nccl_id=nccl_id)
    L.set_params(0, pi0), L.set_params(1, v0)
    res = run(L, lo_, hi, loc, lvs[eb[rank]:eb[rank + 1]])
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
        S = Learner(d, A, n + 1, hidden=H, with_baseline=True, train_vf_iters=iters, device=local)
        S.set_params(0, pi0), S.set_params(1, v0)
        ref = run(S, 0, n, offs, lvs)
        for (a, _), (b, _) in zip(res, ref):
            assert a["n_transitions"] == b["n_transitions"] == n and a["n_episodes"] == b["n_episodes"] == E
            for k in ("loss_pi", "kl", "entropy", "loss_v", "loss_v_old", "adv_mean", "adv_std"):
                np.testing.assert_allclose(a[k], b[k], rtol=1e-5, atol=1e-7, err_msg=k)
        # The all-reduce adds two ranks' f32 gradients where one rank sums everything in f64 first,
        # i.e. the gradients differ at the 1e-7 level; Adam turns that into lr-sized differences on
        # the parameters whose gradient is ~eps.  The reference shows the same sensitivity between
        # MKL thread counts (tests/conftest.py), hence >= 99 % strict and everything within 0.1 lr.
        for got, want, lr in ((pi_dp, S.get_params(0), 3e-4), (v_dp, S.get_params(1), 1e-3)):
            err = np.abs(got - want)
            frac = (err <= 5e-7 + 1e-5 * np.abs(want)).mean()
            # two float32-level evaluations of the same epoch: the reference's own float32 run sits at
            # 98.7 % strict agreement with its float64 run (tests/test_gpu_learner.py), so two of them
            # may differ by about twice that
            assert frac >= 0.97, "strict agreement %.4f" % frac
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=0.1 * lr)
        S.close()
        print("MGPU_OK world=%d n=%d" % (world, n))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
