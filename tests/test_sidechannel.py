"""Side channel (octree occupancy code) and file container — SURVEY.md §8f rows 1-2.

CPU half: the numpy oracle (round trips, hand-worked vectors, malformed streams) and the product's
pure-host container against it.  GPU half: pcc_octree_encode / pcc_octree_decode byte-exact against
the oracle, and ColorModel.compress(path=...) / decompress(path=...) against the in-memory path.
"""
import os
import struct

import numpy as np
import pytest
import torch

from conftest import random_cloud

DEV = "cuda:0"


@pytest.fixture(scope="module")
def sc():
    from oracle import sidechannel
    return sidechannel


# ------------------------------------------------------------------------------ oracle (CPU)
def test_oracle_hand_worked_stream(sc):
    # depth 2, cells (0,0,0), (0,0,1), (3,3,3): root has children 0 and 7; child 0 holds cells 0 and 1
    # (z bit), child 7 holds cell 7.
    q = np.array([[0, 0, 0], [0, 0, 1], [3, 3, 3]])
    assert sc.octree_encode(q, 2) == bytes([0b10000001, 0b00000011, 0b10000000])
    # x is the most significant bit of the child index
    assert sc.octree_encode(np.array([[1, 0, 0]]), 1) == bytes([1 << 4])
    assert sc.octree_encode(np.array([[0, 1, 0]]), 1) == bytes([1 << 2])
    assert sc.octree_encode(np.array([[0, 0, 1]]), 1) == bytes([1 << 1])
    assert sc.octree_encode(np.zeros((0, 3), dtype=np.int64), 3) == b""


@pytest.mark.parametrize("n,extent,seed", [(1, 1, 0), (8, 2, 1), (500, 16, 2), (4000, 200, 3)])
def test_oracle_round_trip(sc, n, extent, seed):
    c = random_cloud(n, extent, seed)[:, 1:].astype(np.int64)
    depth = max(1, int(c.max()).bit_length())
    stream = sc.octree_encode(c, depth)
    back = sc.octree_decode(stream, depth)
    assert np.array_equal(back[np.lexsort((back[:, 2], back[:, 1], back[:, 0]))],
                          c[np.lexsort((c[:, 2], c[:, 1], c[:, 0]))])
    # Morton order on the way out
    m = sc.morton(back, depth)
    assert np.all(m[1:] > m[:-1])


def test_oracle_full_block_and_malformed(sc):
    g = np.stack(np.meshgrid(range(4), range(4), range(4), indexing="ij"), -1).reshape(-1, 3)
    stream = sc.octree_encode(g, 2)
    assert stream == bytes([255] * 9)
    with pytest.raises(ValueError):
        sc.octree_decode(stream[:-1], 2)
    with pytest.raises(ValueError):
        sc.octree_decode(stream + b"\x01", 2)


def test_oracle_side_channel_origin_and_stride(sc):
    xyz = (random_cloud(300, 40, 5, lo=-20)[:, 1:].astype(np.int64)) * 8
    data = sc.gpcc_encode(xyz, 8)
    depth, stride, ox, oy, oz = struct.unpack_from(">BB3i", data)
    assert stride == 8 and (ox, oy, oz) == tuple(xyz.min(0)) and depth == 6
    back = sc.gpcc_decode(data)
    assert np.array_equal(back, xyz[np.lexsort((xyz[:, 2], xyz[:, 1], xyz[:, 0]))])


def test_container_header_layout_and_product_host_side(sc, pcc):
    # field order of model/model.py:243-251 / :293-301, big-endian int32
    pts, y, z = b"\x01\x02\x03", b"yy", b"zzzzz"
    blob = sc.pack_container(pts, [[y], [z]], [17], [[100], [200], [300]])
    assert blob[:28] == struct.pack(">7i", 17, 3, 2, 5, 100, 200, 300)
    assert blob[28:] == pts + y + z
    assert sc.unpack_container(blob) == (pts, [[y], [z]], [17], [[100], [200], [300]])
    # the product's host side writes the same bytes and reads them back
    from pcc_b200 import bitstream
    assert bitstream.pack_container(pts, [[y], [z]], [17], [[100], [200], [300]]) == blob
    assert bitstream.unpack_container(blob) == (pts, [[y], [z]], [17], [[100], [200], [300]])
    with pytest.raises(pcc.PccError):
        bitstream.unpack_container(blob[:-1])
    with pytest.raises(pcc.PccError):
        bitstream.unpack_container(blob[:10])
    with pytest.raises(pcc.PccError):
        bitstream.pack_container(pts, [[y]], [17], [[1], [2], [3]])


# ------------------------------------------------------------------------------ product (GPU)
def _keys(pcc, c):
    keys, status = pcc.ops.pack_coords(torch.from_numpy(c).to(DEV))
    assert int(status.item()) == 0
    return keys


@pytest.mark.gpu
@pytest.mark.parametrize("n,extent,stride,seed", [(1, 1, 8, 0), (8, 2, 8, 1), (700, 16, 8, 2), (20000, 128, 8, 3),
                                                  (3000, 512, 1, 4), (1500, 3, 2, 5), (120000, 512, 8, 6)])
def test_octree_bytes_match_oracle(sc, pcc, n, extent, stride, seed):
    c = random_cloud(n, extent, seed)
    c[:, 1:] *= stride
    depth = max(1, int(c[:, 1:].max() // stride).bit_length())
    stream = pcc.ops.octree_encode(_keys(pcc, c), stride, depth)
    assert stream == sc.octree_encode(c[:, 1:] // stride, depth)
    keys = pcc.ops.octree_decode(stream, stride, depth, DEV)
    got = pcc.ops.unpack_keys(keys).cpu().numpy()
    want = sc.octree_decode(stream, depth) * stride
    assert np.array_equal(got[:, 1:], want) and not got[:, 0].any()


@pytest.mark.gpu
def test_octree_deeper_than_needed_and_full_block(sc, pcc):
    g = np.stack(np.meshgrid(range(4), range(4), range(4), indexing="ij"), -1).reshape(-1, 3).astype(np.int32)
    c = np.concatenate([np.zeros((64, 1), np.int32), g], 1)
    assert pcc.ops.octree_encode(_keys(pcc, c), 1, 2) == bytes([255] * 9)
    for depth in (3, 12):
        s = pcc.ops.octree_encode(_keys(pcc, c), 1, depth)
        assert s == sc.octree_encode(g, depth)
        back = pcc.ops.unpack_keys(pcc.ops.octree_decode(s, 1, depth, DEV)).cpu().numpy()[:, 1:]
        assert np.array_equal(back, sc.octree_decode(s, depth))
    assert pcc.ops.octree_encode(torch.empty(0, dtype=torch.int64, device=DEV), 8, 4) == b""
    assert pcc.ops.octree_decode(b"", 8, 4, DEV).numel() == 0


@pytest.mark.gpu
def test_octree_rejects_bad_input(pcc):
    c = np.array([[0, 8, 8, 8], [0, 16, 8, 3]], dtype=np.int32)      # z = 3 is off the stride-8 grid
    with pytest.raises(pcc.PccError):
        pcc.ops.octree_encode(_keys(pcc, c), 8, 4)
    c = np.array([[0, 8, 8, 8], [0, 256, 8, 8]], dtype=np.int32)     # 256 / 8 = 32 does not fit depth 5
    with pytest.raises(pcc.PccError):
        pcc.ops.octree_encode(_keys(pcc, c), 8, 5)
    c = np.array([[1, 8, 8, 8]], dtype=np.int32)                     # batch 1
    with pytest.raises(pcc.PccError):
        pcc.ops.octree_encode(_keys(pcc, c), 8, 5)
    good = pcc.ops.octree_encode(_keys(pcc, np.array([[0, 8, 8, 8], [0, 16, 8, 8]], dtype=np.int32)), 8, 3)
    with pytest.raises(pcc.PccError):
        pcc.ops.octree_decode(good[:-1], 8, 3, DEV)                  # truncated
    with pytest.raises(pcc.PccError):
        pcc.ops.octree_decode(good + b"\x00", 8, 3, DEV)             # trailing byte
    with pytest.raises(pcc.PccError):
        pcc.ops.octree_decode(bytes([255] * 9), 1, 2, DEV, cap=63)   # more leaves than capacity


@pytest.mark.gpu
def test_side_channel_matches_oracle_with_negative_origin(sc, pcc):
    c = random_cloud(5000, 100, 7, lo=-37)
    c[:, 1:] *= 8
    from pcc_b200 import bitstream
    data = bitstream.gpcc_encode(torch.from_numpy(c).to(DEV))
    assert data == sc.gpcc_encode(c[:, 1:], 8)
    back = bitstream.gpcc_decode(data, DEV)
    assert back.dtype == torch.int32
    assert np.array_equal(back.cpu().numpy(), sc.gpcc_decode(data))
    # float [N,3] input, as gpcc_decode's output is fed back by the reference (model/model.py:184-186)
    data3 = bitstream.gpcc_encode(torch.from_numpy(c[:, 1:]).to(DEV).double())
    assert data3 == data


@pytest.mark.gpu
def test_file_round_trip_equals_in_memory_path(pcc, tmp_path):
    config = pcc.OURS_CONFIG
    model = pcc.ColorModel(config)
    pcc.codec.init_synthetic(model, 3)
    model.eval().to(DEV)
    model.update()
    x = pcc.synthetic.shell_frame(24.0, center=[64, 64, 64], grid=128, seed=3)
    qc, qf = pcc.synthetic.uniform_q_map(x, 0.5, 0.5)
    Q = (qc.to(DEV), qf.to(DEV))
    strings, shape, k, coords = model.compress(x.to(DEV), Q)
    ref = model.decompress(coordinates=coords, strings=strings, shape=shape, k=k)
    path = os.path.join(tmp_path, "frame.bin")
    assert model.compress(x.to(DEV), Q, path=path) is None
    size = os.path.getsize(path)
    side = len(model.gpcc_encode(coords))
    assert size == 28 + side + len(strings[0][0]) + len(strings[1][0])
    c2, s2, sh2, k2 = model.load_bitstream(path)
    assert s2 == strings and sh2 == list(shape) and k2 == [[int(v[0])] for v in k]
    assert torch.equal(c2, coords[:, 1:].to(torch.int32))
    out = model.decompress(path=path)
    assert torch.equal(out, ref)
