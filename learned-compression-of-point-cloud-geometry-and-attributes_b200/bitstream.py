"""File container and latent-coordinate side channel.

Container (model/model.py:214-315).  The reference builds the file through the `bitstream`
package (absent here, unpinned): a header of seven 32-bit integers

    shape[0], len(points_bitstream), len(strings[0][0]), len(strings[1][0]), k[0][0], k[1][0], k[2][0]

(:243-251, read back :293-301) followed by the three payloads (:254-257).  `bitstream` writes numpy
integers big-endian; that byte order is restated here with `struct` ("parity unpinned": no
reference-written file exists to pin it).

Side channel (model/model.py:318-395).  The reference pipes the stride-8 coordinates through the
G-PCC `tmc3` binary (un-vendored submodule).  Here the payload is this library's own lossless code:

    depth u8 | stride u8 | origin x, y, z int32 big-endian | octree occupancy bytes

produced on the GPU by pcc_octree_encode / pcc_octree_decode (csrc/octree.cu).  It is not G-PCC
compatible; files written by the reference cannot be read and vice versa.
"""
import struct

import torch

from . import ops
from ._lib import PccError

_HEADER = struct.Struct(">7i")
_SIDE = struct.Struct(">BB3i")


# ---------------------------------------------------------------------------- side channel
def gpcc_encode(points, stride=8):
    """model/model.py:318-358.  points: [N,4] (b,x,y,z) or [N,3] integer-valued CUDA tensor of unique
    coordinates on the `stride` grid -> bytes."""
    if not points.is_cuda:
        raise PccError("points must be a CUDA tensor: the product path has no CPU fallback")
    xyz = points[:, -3:]
    if xyz.is_floating_point():
        xyz = torch.floor(xyz)
    xyz = xyz.to(torch.int32)
    n = xyz.shape[0]
    if n == 0:
        return _SIDE.pack(1, stride, 0, 0, 0)
    lo = xyz.min(dim=0).values
    hi = xyz.max(dim=0).values
    lo_h, hi_h = lo.tolist(), hi.tolist()
    span = max((h - l) // stride for l, h in zip(lo_h, hi_h))
    depth = max(1, int(span).bit_length())
    if depth > 16:
        raise PccError("coordinate span exceeds 2^16 stride cells")
    shifted = torch.cat([torch.zeros((n, 1), dtype=torch.int32, device=xyz.device), xyz - lo], dim=1)
    keys, status = ops.pack_coords(shifted)
    if int(status.item()):
        raise PccError("coordinate span exceeds the 16-bit key range")
    return _SIDE.pack(depth, stride, *lo_h) + ops.octree_encode(keys, stride, depth)


def gpcc_decode(data, device):
    """model/model.py:360-395 -> [N,3] int32 coordinates in canonical (x,y,z lexicographic) order."""
    data = bytes(data)
    if len(data) < _SIDE.size:
        raise PccError("side channel truncated")
    depth, stride, ox, oy, oz = _SIDE.unpack_from(data)
    if not (1 <= depth <= 16) or stride < 1:
        raise PccError("side channel header corrupt")
    keys = ops.octree_decode(data[_SIDE.size:], stride, depth, device)
    keys, _ = ops.sort_keys(keys, want_perm=False)
    xyz = ops.unpack_keys(keys)[:, 1:]
    return xyz + torch.tensor([ox, oy, oz], dtype=torch.int32, device=xyz.device)


# ---------------------------------------------------------------------------- container
def pack_container(points_bitstream, strings, shape, k):
    """Header + payloads as bytes (model/model.py:243-257)."""
    if len(strings) != 2 or len(shape) != 1 or len(k) != 3:
        raise PccError("the container holds one hyperprior pair: strings=[[y],[z]], shape=[N_z], k=[[k1],[k2],[k3]]")
    y, z = strings[0][0], strings[1][0]
    ks = [int(kk[0]) if isinstance(kk, (list, tuple)) else int(kk) for kk in k]
    head = _HEADER.pack(int(shape[0]), len(points_bitstream), len(y), len(z), *ks)
    return head + bytes(points_bitstream) + bytes(y) + bytes(z)


def unpack_container(data):
    """bytes -> (points_bitstream, strings, shape, k) (model/model.py:293-311)."""
    if len(data) < _HEADER.size:
        raise PccError("bitstream truncated (header)")
    n_z, n_pts, n_y, n_zs, k1, k2, k3 = _HEADER.unpack_from(data)
    if min(n_z, n_pts, n_y, n_zs, k1, k2, k3) < 0 or _HEADER.size + n_pts + n_y + n_zs != len(data):
        raise PccError("bitstream header does not match the file length")
    p = _HEADER.size
    pts = data[p:p + n_pts]
    p += n_pts
    y = data[p:p + n_y]
    p += n_y
    z = data[p:p + n_zs]
    return pts, [[y], [z]], [n_z], [[k1], [k2], [k3]]


def save_bitstream(path, points, strings, shape, k):
    """model/model.py:214-263."""
    data = pack_container(gpcc_encode(points), strings, shape, k)
    with open(path, "wb") as f:
        f.write(data)
    return len(data)


def load_bitstream(path, device):
    """model/model.py:267-315 -> (coordinates [N,3] int32 on `device`, strings, shape, k)."""
    with open(path, "rb") as f:
        data = f.read()
    pts, strings, shape, k = unpack_container(data)
    return gpcc_decode(pts, device), strings, shape, k
