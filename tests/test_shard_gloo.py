"""Host-side logic of the data-parallel path on CPU: episode-aligned sharding, and a world_size-2
gloo run in which every rank derives its shard, the ranks exchange what the GPU path exchanges
(advantage statistics as (sum, sumsq, n), gradient sums scaled by 1/N_global) and arrive at the
single-process result."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG, ROOT
from oracle import learner_oracle as lo
from oracle import scan_ref as sr
from relayrl_b200.shard import merge_offsets, shard_episodes, shard_rows


def _offsets(rng, E):
    lens = rng.integers(0, 300, E)
    return np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)


@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 8])
def test_shard_whole_episodes_balanced(nranks):
    rng = np.random.default_rng(nranks)
    offs = _offsets(rng, 500)
    eb = shard_episodes(offs, nranks)
    assert eb[0] == 0 and eb[-1] == len(offs) - 1 and (np.diff(eb) >= 0).all()
    locals_ = [shard_rows(offs, eb, r)[2] for r in range(nranks)]
    assert np.array_equal(merge_offsets(locals_), offs)          # bit-exact after concatenation
    sizes = np.array([l[-1] for l in locals_])
    assert sizes.sum() == offs[-1]
    assert sizes.max() - sizes.min() <= 2 * 300                   # within one episode of the ideal split


def test_shard_degenerate():
    assert np.array_equal(shard_episodes(np.array([0]), 4), [0, 0, 0, 0, 0])
    offs = np.array([0, 10])
    eb = shard_episodes(offs, 4)
    assert eb[0] == 0 and eb[-1] == 1 and (np.diff(eb) >= 0).all()


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    rng = np.random.default_rng(5)
    offs = _offsets(rng, 60)
    n = int(offs[-1])
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, 60).astype(np.float32)
    obs = rng.standard_normal((n, 4)).astype(np.float32)
    g, l = float(np.float32(.98)), float(np.float32(.97))
    eb = shard_episodes(offs, world)
    lo_, hi, loc = shard_rows(offs, eb, rank)
    adv, ret = sr.finish_paths(rew[lo_:hi], val[lo_:hi], loc, lvs[eb[rank]:eb[rank + 1]], g, l, True)
    # (1) statistics: all-gather (sum, sumsq, n), merge in rank order -> identical on every rank
    mine = torch.tensor([adv.astype(np.float64).sum(), (adv.astype(np.float64) ** 2).sum(), float(hi - lo_)],
                        dtype=torch.float64)
    gathered = [torch.zeros(3, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, mine)
    S, Q, N = (sum(t[i].item() for t in gathered) for i in range(3))
    mean = np.float32(S / N)
    std = np.float32(np.sqrt(max(Q - 2 * float(mean) * S + N * float(mean) ** 2, 0.0) / N))
    # (2) gradient: local sum scaled by 1/N_global, all-reduce(sum)
    params = [p.requires_grad_(True) for p in lo.init_params(4, 64, 1, 3)]
    loss_local = ((lo.value_forward(params, torch.from_numpy(obs[lo_:hi])) - torch.from_numpy(ret)) ** 2).sum() / N
    loss_local.backward()
    grad = torch.cat([p.grad.reshape(-1) for p in params])
    dist.all_reduce(grad)
    if rank == 0:
        np.savez(out, mean=mean, std=std, grad=grad.numpy(), N=N)
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_matches_single_process(tmp_path):
    out = str(tmp_path / "r0.npz")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    rng = np.random.default_rng(5)
    offs = _offsets(rng, 60)
    n = int(offs[-1])
    rew = rng.uniform(-1, 1, n).astype(np.float32)
    val = rng.standard_normal(n).astype(np.float32)
    lvs = rng.uniform(-1, 1, 60).astype(np.float32)
    obs = rng.standard_normal((n, 4)).astype(np.float32)
    g, l = float(np.float32(.98)), float(np.float32(.97))
    adv, ret = sr.finish_paths(rew, val, offs, lvs, g, l, True)
    m0, s0 = lo.adv_statistics(adv)
    assert got["N"] == n
    np.testing.assert_allclose(got["mean"], m0, rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(got["std"], s0, rtol=1e-6)
    params = [p.requires_grad_(True) for p in lo.init_params(4, 64, 1, 3)]
    lo.loss_v(params, torch.from_numpy(obs), torch.from_numpy(ret)).backward()
    ref = torch.cat([p.grad.reshape(-1) for p in params]).numpy()
    np.testing.assert_allclose(got["grad"], ref, rtol=1e-4, atol=1e-5 * np.abs(ref).max())
