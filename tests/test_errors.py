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


@pytest.mark.gpu
def test_corrupt_headers_are_refused_before_any_out_of_bounds_read(pcc):
    """shape[0] (N_z) and k come from the container header; values that disagree with the
    coordinates must raise instead of sizing tensors the kernels then index past."""
    dev = "cuda:0"
    model = pcc.ColorModel(pcc.OURS_CONFIG)
    pcc.init_synthetic(model, 2)
    model.eval().to(dev)
    model.update()
    x = pcc.synthetic.shell_frame(16.0, center=[64] * 3, grid=128, seed=2)
    qc, qf = pcc.synthetic.uniform_q_map(x, 0.5, 0.5)
    strings, shape, k, coords = model.compress(x.to(dev), (qc.to(dev), qf.to(dev)))
    good = model.decompress(coordinates=coords, strings=strings, shape=shape, k=k)
    assert good.shape[0] == x.shape[0]
    for bad_shape in ([shape[0] - 1], [shape[0] + 5], [0]):
        with pytest.raises(pcc._lib.PccError, match="corrupt stream"):
            model.decompress(coordinates=coords, strings=strings, shape=bad_shape, k=k)
    for stage in range(3):
        bad_k = [list(kk) for kk in k]
        bad_k[stage][0] = 10 ** 9
        with pytest.raises(pcc._lib.PccError, match="corrupt stream"):
            model.decompress(coordinates=coords, strings=strings, shape=shape, k=bad_k)
    with pytest.raises(pcc._lib.PccError, match="corrupt stream"):     # truncated y string
        model.decompress(coordinates=coords, strings=[[strings[0][0][:len(strings[0][0]) // 8 * 4]], strings[1]],
                         shape=shape, k=k)
    # the model still works afterwards
    assert torch.equal(model.decompress(coordinates=coords, strings=strings, shape=shape, k=k), good)
