"""Angle-sharded multi-rank logic on CPU: world_size 2 over gloo.  Checks the
shard partition, the all-reducing dot product, that sum_k A_k (one all_reduce
of the per-rank dose volumes) equals the unsharded projection (via the oracle),
and that LinearLBFGS with sharded variables follows the unsharded iterates."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from drtvam_b200.parallel import angle_ranges, shard_bounds


def test_angle_ranges_balance():
    assert angle_ranges([10] * 8, 2) == [(0, 4), (4, 8)]
    assert angle_ranges([10] * 8, 3)[0][0] == 0 and angle_ranges([10] * 8, 3)[-1][1] == 8
    r = angle_ranges([100, 1, 1, 1, 1, 100], 2)
    assert r[0][1] == r[1][0] and 1 <= r[0][1] <= 5
    # uneven counts (filter_radon-like): work, not angle count, is balanced
    counts = [1, 1, 1, 1, 50, 50]
    (a0, a1), (b0, b1) = angle_ranges(counts, 2)
    assert abs(sum(counts[a0:a1]) - sum(counts[b0:b1])) <= 50


def test_slab_grid_keeps_a_slab_one_block_of_layers():
    """parallel.slab_grid: (angle groups, z-slabs) with A * S = world size; a slab is one mostly-full block of
    96..128 layers, the rest of the ranks become angle groups, and where no such cut exists the scene is sharded
    by angle alone (the grids measured in DESIGN.md section 7)."""
    from drtvam_b200.parallel import slab_grid
    assert slab_grid(1, 400) == (1, 1)
    assert slab_grid(2, 400) == (2, 1)       # two slabs of 200 layers would be three blocks each
    assert slab_grid(4, 400) == (1, 4)
    assert slab_grid(8, 400) == (2, 4)
    assert slab_grid(8, 128) == (8, 1)       # one block already
    assert slab_grid(2, 256) == (1, 2)
    for world in (1, 2, 3, 4, 6, 8):
        for layers in (50, 128, 200, 256, 400, 512, 1000):
            a, s_ = slab_grid(world, layers)
            assert a * s_ == world and (s_ == 1 or 96 <= layers // s_ <= 128)


def test_shard_bounds_cover_everything():
    per = 12
    pix = torch.cat([torch.arange(k * per, k * per + (k % 3 + 1) * 3) for k in range(7)]).to(torch.int32)
    spans = [shard_bounds(pix, 7, per, r, 3) for r in range(3)]
    assert spans[0][0] == 0 and spans[-1][1] == pix.numel()
    assert all(spans[i][1] == spans[i + 1][0] for i in range(2))
    with pytest.raises(ValueError):
        shard_bounds(torch.flip(pix, [0]), 7, per, 0, 2)


def _worker(rank, world, port, workdir, result):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "oracle"), os.path.join(root, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import oracle as orc
    from drtvam_b200.lbfgs import LinearLBFGS
    from drtvam_b200.parallel import make_consensus, make_dot
    from helpers import build, pixels_np

    scene, integ, sensor, d, _ = build("C3", workdir, 16)
    pix_all = pixels_np(scene)
    p_all = np.random.default_rng(0).random(pix_all.size).astype(np.float32)
    lo, hi, partition = scene.shard(rank, world)
    assert partition == "angle" and scene.shard_index.tolist() == list(range(lo, hi))
    o = orc.Oracle(d.as_dict(), "f64", threads=1)
    dose = torch.from_numpy(o.forward(p_all[lo:hi], pixels_np(scene)))
    scene.reduce_dose(dose)                           # the one all_reduce of the forward pass
    full = o.forward(p_all, pix_all)
    err_fwd = float(np.linalg.norm(dose.numpy() - full) / np.linalg.norm(full))

    # adjoint needs no collective: the local slice of A^T g is the slice of the global one
    g = np.random.default_rng(1).standard_normal(d.film_shape())
    adj_local = o.adjoint(g, pixels_np(scene))
    adj_full = o.adjoint(g, pix_all)[lo:hi]
    err_adj = float(np.abs(adj_local - adj_full).max())

    dot = make_dot(None, world)
    a = torch.from_numpy(p_all[lo:hi].astype(np.float64))
    err_dot = abs(float(dot(a, a)) - float(np.dot(p_all.astype(np.float64), p_all.astype(np.float64))))

    # sharded L-BFGS on a linear least-squares problem == unsharded
    torch.manual_seed(0)
    A = torch.randn(30, 10, dtype=torch.float64)
    b = A @ torch.rand(10, dtype=torch.float64)
    cols = slice(0, 5) if rank == 0 else slice(5, 10)
    key = 'projector.active_data'

    def run(Amat, dotf, reduce):
        def render_fn(v):
            y = Amat @ v[key]
            if reduce:
                dist.all_reduce(y)
            return y
        opt = LinearLBFGS(render_fn=render_fn, loss_fn=lambda y, p: torch.sum((y - b) ** 2), dot=dotf)
        opt[key] = torch.zeros(Amat.shape[1], dtype=torch.float64)
        for _ in range(12):
            x = opt[key]
            y = Amat @ x
            if reduce:
                y = y.detach()
                yy = Amat @ x
                with torch.no_grad():
                    tot = yy.clone()
                    dist.all_reduce(tot)
                loss = torch.sum((tot - b) ** 2)
                x.grad = 2 * Amat.T @ (tot - b)     # replicated dose gradient, local adjoint
                opt.step(tot, loss)
            else:
                loss = torch.sum((y - b) ** 2)
                loss.backward()
                opt.step(y.detach(), loss.detach())
        return opt[key].detach()

    x_shard = run(A[:, cols], dot, True)
    x_full = run(A, None, False)
    err_lbfgs = float((x_shard - x_full[cols]).abs().max())
    # the pattern shards come back as ONE vector / image stack in the order of the unsharded pixel list
    # (what optimize() saves as patterns.npz at world > 1)
    scene.projector.active_data = torch.from_numpy(p_all[lo:hi].copy())
    err_gather = float((scene.gather_patterns() - torch.from_numpy(p_all)).abs().max())
    imgs = scene.full_patterns().reshape(-1).numpy()
    dense = np.zeros(imgs.size, np.float32)
    dense[pix_all.astype(np.int64)] = p_all
    err_gather = max(err_gather, float(np.abs(imgs - dense).max()))
    # consensus: redundantly computed scalars that differ in the last bits become identical on every rank
    c = make_consensus(None, world)(torch.tensor([1.0 + rank * 1e-15, 2.0], dtype=torch.float64))
    both = [torch.zeros_like(c) for _ in range(world)]
    dist.all_gather(both, c)
    same = all(torch.equal(both[0], b_) for b_ in both)
    if rank == 0:
        result.put((err_fwd, err_adj, err_dot, err_lbfgs, err_gather, same))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_angle_sharding_gloo(workdir):
    ctx = mp.get_context("spawn")
    result = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, workdir, result)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    err_fwd, err_adj, err_dot, err_lbfgs, err_gather, same = result.get(timeout=10)
    assert err_fwd < 1e-12 and err_adj < 1e-12 and err_dot < 1e-9 and err_lbfgs < 1e-8
    assert err_gather == 0.0 and same


def _optimize_worker(rank, world, port, cfg, result):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)   # both ranks on cuda:0; gloo stages CUDA tensors
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from drtvam_b200.optimize import optimize
    vol = optimize(cfg, progress=False)
    if rank == 0:
        result.put(vol.cpu().numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_rank_optimize_saves_the_same_files_as_one_rank(workdir):
    """optimize() at world size 2 (angle shards, both ranks on the one GPU of the test box, gloo): the run ends
    with the files of a one-rank run (optimize.py:331-358) - patterns.npz holds the gathered pattern stack,
    not a shard - and prints the same part."""
    import copy
    from test_optimization import analytic_reference, fixture_config
    from drtvam_b200.optimize import optimize
    cfg1 = fixture_config("box_hole_index_matched", workdir)
    cfg1["regular_sampling"] = True          # same rays on one and on two ranks
    cfg1["n_steps"] = 12
    cfg2 = copy.deepcopy(cfg1)
    cfg1["output"] += "_w1"
    cfg2["output"] += "_w2"
    vol1 = optimize(cfg1, progress=False).cpu().numpy()
    ctx = mp.get_context("spawn")
    result = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_optimize_worker, args=(r, 2, port, cfg2, result)) for r in range(2)]
    for p in procs:
        p.start()
    vol2 = result.get(timeout=600)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    pat1 = np.load(os.path.join(cfg1["output"], "patterns.npz"))["patterns"]
    pat2 = np.load(os.path.join(cfg2["output"], "patterns.npz"))["patterns"]
    assert pat1.shape == pat2.shape == (200, 20, 200)
    assert os.path.exists(os.path.join(cfg2["output"], "patterns_normalized_uint8.npz"))
    # same iterates up to the summation order of the all-reduced dose and dots
    rel = np.linalg.norm(pat2 - pat1) / np.linalg.norm(pat1)
    relv = np.linalg.norm(vol2 - vol1) / np.linalg.norm(vol1)
    print("two-rank optimize: patterns", rel, "final dose", relv)
    assert rel < 1e-2 and relv < 1e-3
    l1, l2 = np.load(os.path.join(cfg1["output"], "loss.npy")), np.load(os.path.join(cfg2["output"], "loss.npy"))
    assert np.allclose(l1, l2, rtol=1e-3, atol=1e-6 * l1[0])
    thr = (cfg2["loss"]["tl"] + cfg2["loss"]["tu"]) / 2
    # 12 iterations do not finish the print; what counts is that both runs print the same part
    agree1 = np.mean(np.isclose(analytic_reference(), vol1[:, :, :, 0] > thr)) * 100
    agree2 = np.mean(np.isclose(analytic_reference(), vol2[:, :, :, 0] > thr)) * 100
    assert abs(agree1 - agree2) < 0.01
