"""Spatial slab partition of one scene (BASELINE.json configs[4]): bit-exact against the
unpartitioned codec.  The ranks here are virtual (threads sharing cuda:0, slab.ThreadComm); the
same code runs over NCCL with slab.DistComm (bench.py --scene, checked on 2 GPUs)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _model(pcc, config, seed=0):
    model = pcc.ColorModel(config)
    pcc.init_synthetic(model, seed)
    model.eval().to(DEV)
    model.update()
    return model


def _scene(pcc, radius=60.0, grid=192, seed=0, q=(0.5, 0.5)):
    x = pcc.synthetic.shell_frame(radius, center=[grid // 2] * 3, grid=grid, seed=seed)
    x = x[torch.randperm(x.shape[0], generator=torch.Generator().manual_seed(seed))]
    qc, qf = pcc.synthetic.uniform_q_map(x, *q)
    return x, (qc, qf)


def test_slab_map_halo_and_kernel_map(pcc):
    """Per level: the local rows are owned + the neighbours' one-voxel ring, and a convolution over
    the slab equals the same rows of the convolution over the whole map."""
    from conftest import random_cloud
    slab = pcc.slab
    c = random_cloud(20000, 96, 3)
    c[:, 1:] *= 2                                            # a stride-2 level
    whole, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), 2)
    rng = np.random.default_rng(0)
    F = torch.from_numpy(rng.standard_normal((whole.n, 8)).astype(np.float32)).to(DEV)
    W = torch.from_numpy((rng.standard_normal((27, 8, 16)) / 5).astype(np.float32)).to(DEV)
    ref = pcc.ops.conv(F, whole.kernel_map(whole, 3), W, None)
    ref_s2 = whole.stride2().keys
    ref_ex = whole.expand(3).keys
    bounds = [0, 64, 96, 192]

    def rank(comm):
        sl = slab.Slab(comm, bounds)
        a, b = sl.row_range(whole.keys, sl.x0, sl.x1)
        m = slab.SlabMap(whole.keys[a:b].clone(), 2, sl)
        # halo = whole-map keys with x in [x0 - 2, x0) and [x1, x1 + 2)
        la, lb = sl.row_range(whole.keys, sl.x0 - 2, sl.x0)
        ra, rb = sl.row_range(whole.keys, sl.x1, sl.x1 + 2)
        assert torch.equal(m.full.keys[m.n:m.n + m.n_halo_left], whole.keys[la:lb])
        assert torch.equal(m.full.keys[m.n + m.n_halo_left:], whole.keys[ra:rb])
        Fl = m.lift(F[a:b])
        assert torch.equal(Fl[m.n:m.n + m.n_halo_left], F[la:lb]) and torch.equal(Fl[m.n + m.n_halo_left:], F[ra:rb])
        out = pcc.ops.conv(Fl, m.kernel_map(m, 3), W, None)
        return (a, b, out, m.stride2().keys, m.expand(3).keys, m.count_total())

    res = slab.ThreadComm.run(3, rank)
    assert [r[0] for r in res] == [0] + [r[1] for r in res[:-1]] and res[-1][1] == whole.n
    assert torch.equal(torch.cat([r[2] for r in res]).view(torch.int32), ref.view(torch.int32))
    assert torch.equal(torch.cat([r[3] for r in res]), ref_s2)
    assert torch.equal(torch.cat([r[4] for r in res]), ref_ex)
    assert all(r[5] == whole.n for r in res)


def test_slab_topk_equals_global(pcc):
    slab = pcc.slab
    from conftest import random_cloud
    c = random_cloud(30000, 128, 5)
    whole, _ = pcc.ops.canonical_map(torch.from_numpy(c).to(DEV), 1)
    rng = np.random.default_rng(1)
    logits = torch.from_numpy(np.round(rng.standard_normal((whole.n, 4)) * 4).astype(np.float32) / 4).to(DEV)  # many ties
    k = whole.n // 3
    ref = pcc.ops.topk_rows(logits, k)
    bounds = [0, 32, 96, 128]

    def rank(comm):
        sl = slab.Slab(comm, bounds)
        a, b = sl.row_range(whole.keys, sl.x0, sl.x1)
        m = slab.SlabMap(whole.keys[a:b].clone(), 1, sl)
        return m.topk(m.lift(logits[a:b]), [k][0]) + a

    got = torch.cat(slab.ThreadComm.run(3, rank))
    assert torch.equal(got.int(), ref.int())


@pytest.mark.parametrize("config_name,world,q", [("OURS_CONFIG", 2, (0.5, 0.5)), ("OURS_CONFIG", 4, (0.0, 1.0)),
                                                 ("ABLATION_CONFIG", 3, (1.0, 0.0))])
def test_scene_partition_bit_exact(pcc, config_name, world, q):
    model = _model(pcc, getattr(pcc, config_name))
    x, Q = _scene(pcc, q=q)
    strings, shape, k, coords = model.compress(x.to(DEV), (Q[0].to(DEV), Q[1].to(DEV)))
    ref = model.decompress(coordinates=coords, strings=strings, shape=shape, k=k)
    bounds = pcc.scene.balanced_bounds(x, world)
    parts = pcc.scene.split_scene(x, Q, bounds)
    assert sum(p[0].shape[0] for p in parts) == x.shape[0]

    def rank(comm):
        sl = pcc.slab.Slab(comm, bounds)
        xs, (qc, qf) = parts[comm.rank]
        s, sh, kk, c = pcc.scene.compress_scene(model, xs.to(DEV), (qc.to(DEV), qf.to(DEV)), sl)
        out = pcc.scene.decompress_scene(model, c, s, sh, kk, sl)
        return s, sh, kk, c, out

    for s, sh, kk, c, out in pcc.slab.ThreadComm.run(world, rank):
        assert sh == shape and kk == k
        assert s[0][0] == strings[0][0] and s[1][0] == strings[1][0], "bitstreams differ from the one-GPU codec"
        assert torch.equal(c, coords)
        assert torch.equal(out.view(torch.int32), ref.view(torch.int32)), "decoded scene differs"
