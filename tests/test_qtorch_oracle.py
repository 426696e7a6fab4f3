"""CPU: what a qtorch-equipped reference would compute in FloatQuantizer.quant (quant.py:1063-1070) against the rounding the
product uses for the same call (fp32 -> fp8 round-to-nearest-even, saturating at the format's finite maximum: the rounding
of `Tensor.to(float8_*)`, of the reference's own real-quant path quant.py:1183/1211 and of its Triton kernels).

`oracle/qtorch_oracle.py` restates QPyTorch's published bit-level algorithm; qtorch itself cannot be installed here, so
this is a QUANTIFICATION of the unpinned corner, not a parity pin. The assertions are the facts DESIGN.md
quotes."""
import numpy as np
import torch

from oracle import qtorch_oracle as qt


def _torch_cast(x, dtype):
    fin = torch.finfo(dtype)
    return torch.from_numpy(x).clamp(fin.min, fin.max).to(dtype).float().numpy()


def _grid(dtype):
    codes = torch.arange(256, dtype=torch.int32).to(torch.uint8).view(dtype).float()
    return np.unique(codes[torch.isfinite(codes)].numpy())


def test_e4m3_grid_and_saturation():
    g = _grid(torch.float8_e4m3fn)
    assert qt.max_value(4, 3) == 240.0 and g.max() == 448.0
    inside = g[np.abs(g) <= 240.0]
    # every e4m3fn code up to 240 (subnormals included) is a fixed point of qtorch's (4, 3) rounding ...
    assert np.array_equal(qt.float_quantize(inside, 4, 3), inside)
    # ... and the 14 codes above it (256 .. 448, both signs) do not exist there: they saturate to +-240
    above = g[np.abs(g) > 240.0]
    assert above.size == 14
    assert np.array_equal(qt.float_quantize(above, 4, 3), np.sign(above) * 240.0)


def test_e5m2_grid_matches():
    g = _grid(torch.float8_e5m2)
    assert qt.max_value(5, 2) == 57344.0 == g.max()
    assert np.array_equal(qt.float_quantize(g, 5, 2), g)


def test_difference_to_round_to_nearest_even_is_ties_and_e4m3_saturation():
    rng = np.random.default_rng(0)
    for (e, m, dtype) in ((4, 3, torch.float8_e4m3fn), (5, 2, torch.float8_e5m2)):
        top = qt.max_value(e, m)
        x = (rng.standard_normal(400000) * np.exp(rng.uniform(-8, np.log(top), 400000))).astype(np.float32)
        x = x[np.abs(x) <= top]
        a, b = qt.float_quantize(x, e, m), _torch_cast(x, dtype)
        assert float((a != b).mean()) == 0.0, 'continuous inputs inside the common range: the two roundings agree'
        # exact midpoints between neighbouring codes: qtorch rounds away from zero, the cast to even
        g = _grid(dtype)
        g = g[(g > 0) & (g < top)]
        mid = ((g[:-1] + g[1:]) / 2).astype(np.float32)
        a, b = qt.float_quantize(mid, e, m), _torch_cast(mid, dtype)
        assert np.array_equal(a, g[1:]), 'ties away from zero'
        assert 0.45 < float((a != b).mean()) < 0.55, 'half of the midpoints round the other way under ties-to-even'


def test_share_of_a_quantized_weight_that_would_differ():
    """A bf16 weight scaled to the e4m3fn range the way the reference does (s = absmax / 448 per channel): how many
    elements land above 240 (saturated by qtorch, representable for the cast) and how many on exact ties."""
    g = torch.Generator().manual_seed(0)
    w = (torch.randn(256, 4096, generator=g) * 0.02).to(torch.bfloat16).float()
    s = w.abs().amax(dim=1, keepdim=True) / 448.0
    t = (w / s).numpy().astype(np.float32)
    a, b = qt.float_quantize(t, 4, 3), _torch_cast(t, torch.float8_e4m3fn)
    differ = a != b
    over = np.abs(t) > 248.0          # rounds to 256 or more under the cast
    frac, frac_over, frac_tie = float(differ.mean()), float(over.mean()), float((differ & ~over).mean())
    print(f'e4m3 per_channel weight: {frac:.4%} of the elements differ: {frac_over:.4%} lie above qtorch\'s maximum, '
          f'{frac_tie:.4%} are exact ties (bf16 / bf16 quotients hit midpoints)')
    assert np.all(differ[over]), 'everything above 248 saturates to 240 under qtorch and not under the cast'
    assert 0.02 < frac_over < 0.06 and frac_tie < 0.01       # a Gaussian row: ~4 % beyond 0.55 * absmax
    assert np.array_equal(a[~differ], b[~differ])
