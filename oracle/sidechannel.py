"""ORACLE — TEST INFRASTRUCTURE ONLY.  Side channel + file container, numpy restatement.

**Parity unpinned for these two rows**: the reference's side channel is the G-PCC `tmc3` binary
(model/model.py:318-395; un-vendored submodule dependencies/mpeg-pcc-tmc13, .SUBMODULES.json) and its
container goes through the `bitstream` package (model/model.py:9,236-260,285-311; absent, unpinned
in requirements.txt).  Neither can run here, so there is no reference output to pin against.  What
is restated is (a) the container's field order from the reference's own call sites
(model/model.py:243-257 write, :293-311 read) with the byte order `bitstream` documents for numpy
integers (big-endian), and (b) this repo's own octree occupancy code, written independently of the
CUDA kernels (np.unique per level here, sorted Morton codes + block scans there) so that equal
bytes mean something.
"""
import struct

import numpy as np

HEADER = struct.Struct(">7i")
SIDE = struct.Struct(">BB3i")


def morton(q, depth):
    """q [N,3] non-negative ints < 2^depth -> codes with bit 3 l + (2 - axis) = bit l of q[:, axis]."""
    q = np.asarray(q, dtype=np.uint64)
    code = np.zeros(q.shape[0], dtype=np.uint64)
    for l in range(depth):
        for axis in range(3):
            code |= ((q[:, axis] >> np.uint64(l)) & np.uint64(1)) << np.uint64(3 * l + (2 - axis))
    return code


def demorton(code, depth):
    code = np.asarray(code, dtype=np.uint64)
    q = np.zeros((code.shape[0], 3), dtype=np.int64)
    for l in range(depth):
        for axis in range(3):
            q[:, axis] |= (((code >> np.uint64(3 * l + (2 - axis))) & np.uint64(1)) << np.uint64(l)).astype(np.int64)
    return q


def octree_encode(q, depth):
    """Breadth-first occupancy bytes of the octree over the cells q [N,3] (unique)."""
    q = np.asarray(q)
    if q.shape[0] == 0:
        return b""
    codes = morton(q, depth)
    out = []
    for l in range(depth):
        s = np.uint64(3 * (depth - l))
        child = np.unique(codes >> (s - np.uint64(3)))       # occupied nodes of level l + 1, sorted
        parent = child >> np.uint64(3)
        bit = (np.uint64(1) << (child & np.uint64(7))).astype(np.uint8)
        nodes, first = np.unique(parent, return_index=True)  # occupied nodes of level l, sorted
        out.append(np.bitwise_or.reduceat(bit, first).astype(np.uint8))
    return np.concatenate(out).tobytes()


def octree_decode(stream, depth):
    """Occupancy bytes -> cells [N,3] in Morton order.  Raises ValueError on a malformed stream."""
    data = np.frombuffer(bytes(stream), dtype=np.uint8)
    if data.size == 0:
        return np.zeros((0, 3), dtype=np.int64)
    nodes = np.zeros(1, dtype=np.uint64)
    pos = 0
    for _ in range(depth):
        if pos + nodes.size > data.size:
            raise ValueError("truncated")
        occ = data[pos:pos + nodes.size]
        pos += nodes.size
        bits = np.unpackbits(occ[:, None], axis=1, bitorder="little").astype(bool)  # [n, 8], column c = child c
        idx_node, idx_child = np.nonzero(bits)
        nodes = (nodes[idx_node] << np.uint64(3)) | idx_child.astype(np.uint64)
    if pos != data.size:
        raise ValueError("trailing bytes")
    return demorton(nodes, depth)


def gpcc_encode(xyz, stride=8):
    """xyz [N,3] ints on the stride grid -> side-channel bytes (header + octree)."""
    xyz = np.asarray(xyz, dtype=np.int64)
    if xyz.shape[0] == 0:
        return SIDE.pack(1, stride, 0, 0, 0)
    lo = xyz.min(axis=0)
    q = (xyz - lo) // stride
    depth = max(1, int(q.max()).bit_length())
    return SIDE.pack(depth, stride, *[int(v) for v in lo]) + octree_encode(q, depth)


def gpcc_decode(data):
    """-> xyz [N,3] int64 sorted lexicographically."""
    depth, stride, ox, oy, oz = SIDE.unpack_from(data)
    q = octree_decode(data[SIDE.size:], depth)
    xyz = q * stride + np.array([ox, oy, oz], dtype=np.int64)
    order = np.lexsort((xyz[:, 2], xyz[:, 1], xyz[:, 0]))
    return xyz[order]


def pack_container(points_bitstream, strings, shape, k):
    """model/model.py:243-257."""
    y, z = strings[0][0], strings[1][0]
    head = HEADER.pack(int(shape[0]), len(points_bitstream), len(y), len(z), int(k[0][0]), int(k[1][0]), int(k[2][0]))
    return head + bytes(points_bitstream) + bytes(y) + bytes(z)


def unpack_container(data):
    """model/model.py:293-311."""
    n_z, n_pts, n_y, n_zs, k1, k2, k3 = HEADER.unpack_from(data)
    p = HEADER.size
    pts = data[p:p + n_pts]
    y = data[p + n_pts:p + n_pts + n_y]
    z = data[p + n_pts + n_y:p + n_pts + n_y + n_zs]
    return pts, [[y], [z]], [n_z], [[k1], [k2], [k3]]
