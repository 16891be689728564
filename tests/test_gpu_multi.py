"""Sample sharding over 2 real GPUs, one process per GPU, exchange in-kernel over NVLink peer
memory (mppi_peer_export / mppi_peer_init) or by the in-library NCCL all-gather (mppi_comm_init): the sharded result must equal the single-GPU result
on the same global Philox stream.  Skipped on boxes with fewer than 2 GPUs (run it with
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`)."""
import multiprocessing as mp

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

K, N, STEPS = 16384, 30, 3


def _gpu_count():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


def _plan_single(device, precision):
    import motionplanning_b200 as m
    from motionplanning_b200 import scenarios as sc

    args, kw, circles, _ = sc.config(0, K=K, N=N)
    kw.update(precision=precision, noise_mode=m.capi.NOISE_PHILOX, device=device)
    mppi = m.defineMPPI(*args, **kw)
    m.defineMPPIobs(mppi, circles)
    out = []
    U = np.zeros((N, 2))
    for _ in range(STEPS):
        m.defineMPPINominalControl(mppi, U)
        m.MPPIPlan(mppi)
        U = np.array(mppi.r.Control, copy=True)
        out.append((U, mppi.r.argmin, mppi.r.rho, mppi.r.cost, mppi.r.RolloutCount))
    return out


def _worker(rank, world, precision, conn, exchange="peer"):
    try:
        import motionplanning_b200 as m
        from motionplanning_b200 import scenarios as sc
        from motionplanning_b200.dist import shard_range

        args, kw, circles, _ = sc.config(0, K=K, N=N)
        off, kl = shard_range(K, rank, world)
        kw.update(precision=precision, noise_mode=m.capi.NOISE_PHILOX, device=rank, k_offset=off, K_local=kl)
        mppi = m.defineMPPI(*args, **kw)
        m.defineMPPIobs(mppi, circles)
        if exchange == "peer":
            conn.send(mppi.peer_export())
            handles = conn.recv()
            mppi.peer_init(handles, rank, world)
        else:   # in-library NCCL all-gather captured in the step graph (mppi_comm_init); rank 0 makes the unique id
            from motionplanning_b200.mppi import comm_unique_id

            conn.send(comm_unique_id() if rank == 0 else b"")
            uid = conn.recv()[0]
            mppi.comm_init(uid, rank, world)
        out = []
        U = np.zeros((N, 2))
        for _ in range(STEPS):
            m.defineMPPINominalControl(mppi, U)
            m.MPPIPlan(mppi)
            U = np.array(mppi.r.Control, copy=True)
            out.append((U, mppi.r.argmin, mppi.r.rho, mppi.r.cost, mppi.r.RolloutCount))
        conn.send(("ok", out))
    except Exception as e:  # surface the failure in the parent
        conn.send(("error", repr(e)))


@pytest.mark.skipif(_gpu_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("exchange", ["peer", "nccl"])
@pytest.mark.parametrize("precision", [0, 1])  # FP32, FP64
def test_sharded_exchange_matches_single_gpu(precision, exchange):
    ctx = mp.get_context("spawn")
    world = min(_gpu_count(), 4) if exchange == "peer" else 2
    pipes, procs = [], []
    for r in range(world):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_worker, args=(r, world, precision, b, exchange))
        p.start()
        pipes.append(a)
        procs.append(p)
    handles = [c.recv() for c in pipes]
    for c in pipes:
        c.send(handles)
    results = [c.recv() for c in pipes]
    for p in procs:
        p.join(120)
    for status, payload in results:
        assert status == "ok", payload
    ref = _plan_single(0, precision)
    tol = 1e-5 if precision == 0 else 1e-11
    for step in range(STEPS):
        U_ref, arg_ref, rho_ref, cost_ref, rc_ref = ref[step]
        for status, out in results:
            U, arg, rho, cost, rc = out[step]
            assert arg == arg_ref and rc == rc_ref
            assert abs(rho - rho_ref) <= tol * abs(rho_ref)
            assert np.abs(U - U_ref).max() <= tol * max(1.0, np.abs(U_ref).max())
            assert abs(cost - cost_ref) <= 10 * tol * abs(cost_ref)
        # all ranks hold bit-identical results (same fold order on every rank)
        for other in results[1:]:
            assert np.array_equal(results[0][1][step][0], other[1][step][0])
