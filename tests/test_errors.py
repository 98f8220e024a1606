"""Error behaviour of the host layer that needs no device: malformed inputs are refused before any launch."""
import pytest
import torch


def test_empty_and_malformed_clouds_are_refused(pcc):
    model = pcc.ColorModel(pcc.OURS_CONFIG).eval()
    q = (torch.zeros((0, 4), dtype=torch.int32), torch.zeros((0, 2)))
    with pytest.raises(pcc._lib.PccError, match="non-empty"):
        model.compress(torch.zeros((0, 6)), q)
    with pytest.raises(pcc._lib.PccError, match="non-empty"):
        model.compress(torch.zeros((5, 3)), q)
    with pytest.raises(pcc._lib.PccError, match="non-empty"):
        model.decompress(coordinates=torch.zeros((0, 4), dtype=torch.int32), strings=[[b""], [b""]], shape=[0],
                         k=[[], [], []])


def test_product_path_has_no_cpu_fallback(pcc):
    """Device operators refuse host tensors instead of computing on the CPU."""
    with pytest.raises(pcc._lib.PccError):
        pcc.ops.pack_coords(torch.zeros((4, 4), dtype=torch.int32))
