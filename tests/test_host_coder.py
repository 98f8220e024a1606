"""CPU tests of the product's host-side entropy coder / table builders (C-ABI host entry points;
no GPU needed) against the oracle's independent restatement: byte-identical bitstreams."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def gc(oracle):
    return oracle.entropy.GaussianConditionalOracle()


def test_gaussian_tables_match_oracle(pcc, gc):
    cdf, length, offset = pcc.ops.gaussian_cdf_tables(gc.scale_table, 1e-9)
    assert np.array_equal(length, gc.cdf_length)
    assert np.array_equal(offset, gc.offset)
    assert cdf.shape == gc.quantized_cdf.shape
    assert np.array_equal(cdf, gc.quantized_cdf)
    # every row: starts at 0, strictly increasing up to 2^16 (all frequencies positive)
    for i in range(cdf.shape[0]):
        row = cdf[i, :length[i]]
        assert row[0] == 0 and row[-1] == 65536 and (np.diff(row) > 0).all()


def test_factorized_tables_match_oracle(pcc, oracle):
    import torch
    model = pcc.ColorModel(pcc.OURS_CONFIG)
    pcc.codec.init_synthetic(model, 5)
    eb = model.entropy_model.entropy_bottleneck
    with torch.no_grad():  # non-trivial quantiles / factors so the tables differ per channel
        eb.quantiles[:, 0, 0] -= torch.rand(128) * 6
        eb.quantiles[:, 0, 2] += torch.rand(128) * 9
        eb.quantiles[:, 0, 1] += torch.rand(128) - 0.5
        for i in range(4):
            getattr(eb, f"_factor{i}").uniform_(-0.5, 0.5)
    assert eb.update(force=True)
    params = {k: v.detach().numpy() for k, v in eb.named_parameters()}
    ebo = oracle.entropy.EntropyBottleneckOracle(params)
    assert np.array_equal(eb._cdf_length.numpy(), ebo.cdf_length)
    assert np.array_equal(eb._offset.numpy(), ebo.offset)
    assert np.array_equal(eb._quantized_cdf.numpy(), ebo.quantized_cdf)


@pytest.mark.parametrize("n", [0, 1, 7, 1000, 300000])
@pytest.mark.parametrize("spread", [1.0, 40.0, 3000.0])
def test_rans_bytes_match_oracle_and_roundtrip(pcc, oracle, gc, n, spread):
    rng = np.random.default_rng(n + int(spread))
    idx = rng.integers(0, 64, size=n).astype(np.int32)
    sym = np.rint(rng.standard_normal(n) * spread).astype(np.int32)  # large spread -> many escapes
    if n > 5:
        sym[:3] = [0, -2 ** 20, 2 ** 20]  # long bypass sequences (> 15 nibbles is impossible for int32)
    ours = pcc.ops.rans_encode(sym, idx, gc.quantized_cdf, gc.cdf_length, gc.offset)
    ref = oracle.entropy.rans_encode(sym, idx, gc.quantized_cdf, gc.cdf_length, gc.offset)
    assert ours == ref
    assert len(ours) % 4 == 0 and len(ours) >= 8
    assert np.array_equal(pcc.ops.rans_decode(ours, idx, gc.quantized_cdf, gc.cdf_length, gc.offset), sym)
    assert np.array_equal(oracle.entropy.rans_decode(ours, idx, gc.quantized_cdf, gc.cdf_length, gc.offset), sym)


def test_rans_rejects_bad_index(pcc, gc):
    with pytest.raises(pcc.PccError):
        pcc.ops.rans_encode(np.zeros(3, np.int32), np.array([0, 64, 0], np.int32), gc.quantized_cdf,
                            gc.cdf_length, gc.offset)


def test_pmf_to_quantized_cdf_matches_oracle(pcc, oracle):
    rng = np.random.default_rng(0)
    for trial in range(200):
        n = int(rng.integers(1, 60))
        pmf = rng.random(n).astype(np.float32) ** 8  # many near-zero entries -> exercises the steal loop
        pmf /= max(pmf.sum(), 1e-30)
        a = pcc.ops.pmf_to_quantized_cdf(pmf)
        b = oracle.entropy.pmf_to_quantized_cdf(pmf)
        assert np.array_equal(a, b)
        assert a[0] == 0 and a[-1] == 65536 and (np.diff(a) > 0).all()
    with pytest.raises(pcc.PccError):
        pcc.ops.pmf_to_quantized_cdf(np.array([0.5, np.nan], np.float32))


def test_rans_decoder_refuses_corrupt_streams(pcc, gc):
    """a truncated / tampered stream is an error, not garbage symbols returned as PCC_OK."""
    rng = np.random.default_rng(11)
    n = 20000
    idx = rng.integers(0, 64, size=n).astype(np.int32)
    sym = np.rint(rng.standard_normal(n) * 30.0).astype(np.int32)
    good = pcc.ops.rans_encode(sym, idx, gc.quantized_cdf, gc.cdf_length, gc.offset)
    assert np.array_equal(pcc.ops.rans_decode(good, idx, gc.quantized_cdf, gc.cdf_length, gc.offset), sym)
    with pytest.raises(pcc.PccError, match="corrupt stream"):          # cut in half: the decoder runs past the end
        pcc.ops.rans_decode(good[:len(good) // 2 // 4 * 4], idx, gc.quantized_cdf, gc.cdf_length, gc.offset)
    # garbage of the right length: either an over-long bypass run or an exhausted stream is hit, or (rarely) the
    # noise decodes to in-range symbols — but never a crash, and never silently more words than the stream has
    noise = rng.integers(0, 256, size=len(good), dtype=np.uint8).tobytes()
    try:
        out = pcc.ops.rans_decode(noise, idx, gc.quantized_cdf, gc.cdf_length, gc.offset)
        assert out.shape == (n,)
    except pcc.PccError as e:
        assert "corrupt stream" in str(e)
