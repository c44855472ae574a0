"""Multi-GPU (NCCL; 2, 4 and 8 ranks) run of the row-sharded paths against the single-GPU path of the same model classes.

Needs as many CUDA devices as ranks (``gpurun --gpus N``); skipped otherwise.  One process per GPU; every rank seeds ``np.random``
identically, so the replicated host state must stay identical and equal to the unsharded run: block labels and
integer counts bit-exact, fp parameters to round-off (different summation order across shards)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return port


def _run_models(comm, world=2):
    """The three model families on a (possibly sharded) network; returns the sampled state."""
    from graphistician_b200.stochastic_block_model import GaussianStochasticBlockModel
    from graphistician_b200.eigenmodel import LogisticEigenmodel
    from graphistician_b200.adjacency import LatentDistanceAdjacencyDistribution
    from graphistician_b200.old_abstractions import FixedGaussianNetwork, network_for
    out = {}
    # --- Gaussian SBM, shared c (config 2 model)
    N, C, B = max(640, 128 * world + 128), 6, 2
    np.random.seed(3)
    truth = GaussianStochasticBlockModel(N, B, C=C, dtype="f64")
    net = truth.rvs()
    network = FixedGaussianNetwork(net.A.astype(np.uint8), net.W)
    if comm is not None:
        network.shard(comm)
    model = GaussianStochasticBlockModel(N, B, C=C, dtype="f64")
    for it in range(3):
        np.random.seed(100 + it)
        model.resample(network)
    out["sbm_c"], out["sbm_p"], out["sbm_mu"] = model.c.copy(), model.p.copy(), model.mus.copy()
    out["sbm_ll"] = model.log_likelihood(network)
    # --- logistic eigenmodel, injected omega / normals, then the device PG path
    N, D = max(512, 128 * world), 4
    np.random.seed(5)
    rng = np.random.RandomState(9)
    eg = LogisticEigenmodel(N, D, p=0.1, sigma_F=1.0, lmbda=np.ones(D), dtype="f64")
    A = eg.rvs().astype(np.uint8)
    network = FixedGaussianNetwork(A, None)
    if comm is not None:
        network.shard(comm)
    eg2 = LogisticEigenmodel(N, D, p=0.1, sigma_F=1.0, lmbda=np.ones(D), dtype="f64")
    O = rng.gamma(2.0, 0.1, (N, N)) + 0.02
    for it in range(2):
        np.random.seed(200 + it)
        eg2.resample(network, omega=O, normals=rng.randn(N, D))
    out["eig_F"], out["eig_mu0"] = eg2.F.copy(), eg2.mu_0
    np.random.seed(300)
    eg2.resample(network)                                   # device Polya-gamma draw (counter-based: shard invariant)
    out["eig_F_pg"], out["eig_ll"] = eg2.F.copy(), eg2.log_likelihood(network)
    # --- latent distance, injected HMC draws
    N, dim = max(384, 128 * world), 2
    np.random.seed(7)
    lat = LatentDistanceAdjacencyDistribution(N, dim=dim, dtype="f64")
    A = lat.rvs().astype(np.uint8)
    network = network_for(A)
    if comm is not None:
        network.shard(comm)
    lat2 = LatentDistanceAdjacencyDistribution(N, dim=dim, dtype="f64")
    rng = np.random.RandomState(11)
    for it in range(3):
        draws = dict(L=(rng.randn(N, dim), np.log(rng.rand())), mu_0=(np.array(rng.randn()), np.log(rng.rand())),
                     mu_self=(np.array(rng.randn()), np.log(rng.rand())))
        lat2.resample(A, draws=draws)
    out["lat_L"], out["lat_mu0"], out["lat_ll"] = lat2.L.copy(), lat2.mu_0, lat2.log_likelihood(A)
    network.invalidate()
    network._comm = None
    return out


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    from graphistician_b200 import _lib
    _lib.set_default_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world)
    try:
        from graphistician_b200.sharding import Comm
        comm = Comm()
        assert comm.device == "cuda" and comm.world == world
        sharded = _run_models(comm, world)
        dist.barrier()
        if rank == 0:
            single = _run_models(None, world)
            res = {}
            for k in single:
                a, b = np.asarray(sharded[k]), np.asarray(single[k])
                res[k] = bool(np.array_equal(a, b)) if a.dtype.kind in "iu" else float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
            q.put(res)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_models_match_single_gpu(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs (gpurun --gpus %d)" % (world, world))
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res["sbm_c"] is True, res
    for k, v in res.items():
        if k != "sbm_c":
            assert v <= 1e-8, (k, res)
