"""world_size-2 gloo test (CPU) of the host-side multi-GPU logic: shard ranges, the partial
exchange over torch.distributed, and the one-shot softmin combine.  The
per-shard partials come from the oracle (no GPU here); the combine is restated in numpy and
must reproduce the unsharded oracle result."""
import os
import sys

import numpy as np
import pytest

import oracle_lib as orc
from motionplanning_b200 import _capi as capi
from motionplanning_b200 import dist as mdist
from motionplanning_b200 import scenarios as sc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
K, N = 1024, 20


def test_shard_ranges_cover_and_are_contiguous():
    for Kt in (1, 7, 1024, 65536, 1048576 + 3):
        for G in (1, 2, 3, 4, 8):
            spans = [mdist.shard_range(Kt, g, G) for g in range(G)]
            assert spans[0][0] == 0 and sum(n for _, n in spans) == Kt
            for (o0, n0), (o1, _) in zip(spans, spans[1:]):
                assert o0 + n0 == o1
            assert max(n for _, n in spans) - min(n for _, n in spans) <= 1


def _scene():
    args, kw, circles, blocks = sc.config(0, K=K, N=N)
    kw = dict(kw, noise_mode=capi.NOISE_SUPPLIED, precision=capi.FP64)
    eps = orc.generate_noise(sc.SIGMA, sc.SEED, 0, K, N)
    rng = np.random.RandomState(2)
    U = rng.uniform(-0.1, 0.1, (N, 2))
    return args, kw, circles, blocks, eps, U


def local_partial(rank, world):
    """What mppi_step_local produces on this rank's shard, computed with the oracle."""
    args, kw, circles, blocks, eps, U = _scene()
    off, kl = mdist.shard_range(K, rank, world)
    a = args[:5] + (kl,) + args[6:]
    r = orc.plan(orc.make_config(a, dict(kw, FeasibilityCount=kl)), circles, blocks, args[3], U,
                 np.ascontiguousarray(eps[:, off:off + kl, :]), capi.FP64)
    w = np.exp(-(r["costs"] - r["rho"]) / sc.LAMBDA)
    u = np.clip(U[:, None, :] + eps[:, off:off + kl, :], sc.CL, sc.CU)
    p = capi.Partial()
    p.rho, p.eta = r["rho"], w.sum()
    p.argmin_global, p.rollouts, p.feasible = off + r["argmin"] - 1, kl, r["feasible_count"]
    S = np.einsum("k,jkc->jc", w, u)
    for c in range(2):
        for j in range(N):
            p.S[c * N + j] = S[j, c]
    return p


def combine(partials):
    """numpy statement of finalize_kernel<PH_COMBINE> (csrc/mppi_update.cuh)."""
    rho = min(p.rho for p in partials)
    best = [p for p in partials if p.rho == rho][0]
    scale = [np.exp(-(p.rho - rho) / sc.LAMBDA) for p in partials]
    eta = sum(p.eta * s for p, s in zip(partials, scale))
    S = sum(np.array(p.S[:2 * N]) * s for p, s in zip(partials, scale))
    return (S / eta).reshape(2, N).T, rho, eta, best.argmin_global + 1, sum(p.rollouts for p in partials) + 1


def _worker(rank, world, port, out):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = local_partial(rank, world)
        parts = mdist.exchange_partials(mine, N)
        assert len(parts) == world and parts[rank].rho == mine.rho and parts[rank].argmin_global == mine.argmin_global
        U, rho, eta, argmin, rc = combine(parts)
        np.save(os.path.join(out, f"U{rank}.npy"), U)
        np.save(os.path.join(out, f"s{rank}.npy"), np.array([rho, eta, argmin, rc]))
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_and_combine_equals_unsharded(built, tmp_path):
    import torch.multiprocessing as mp

    world, port = 2, 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    args, kw, circles, blocks, eps, U = _scene()
    ref = orc.plan(orc.make_config(args, kw), circles, blocks, args[3], U, eps, capi.FP64)
    U0, U1 = np.load(tmp_path / "U0.npy"), np.load(tmp_path / "U1.npy")
    s0, s1 = np.load(tmp_path / "s0.npy"), np.load(tmp_path / "s1.npy")
    assert np.array_equal(U0, U1) and np.array_equal(s0, s1)      # identical on every rank
    np.testing.assert_allclose(U0, ref["Control"], rtol=0, atol=1e-13)
    assert s0[0] == ref["rho"] and s0[1] == pytest.approx(ref["eta"], rel=1e-13)
    assert int(s0[2]) == ref["argmin"] and int(s0[3]) == ref["rollout_count"]


def test_combine_is_invariant_to_shard_count(built):
    args, kw, circles, blocks, eps, U = _scene()
    ref = orc.plan(orc.make_config(args, kw), circles, blocks, args[3], U, eps, capi.FP64)
    for G in (1, 2, 4, 8):
        Uc, rho, eta, argmin, rc = combine([local_partial(g, G) for g in range(G)])
        np.testing.assert_allclose(Uc, ref["Control"], rtol=0, atol=1e-13)
        assert (rho, argmin, rc) == (ref["rho"], ref["argmin"], ref["rollout_count"])
