"""CPU: host-side mirror of the reference interface (ctor kwargs, qmin/qmax, unit grid, error types)."""
import pytest
import torch

from lightcompress_b200 import quant as Q


def test_int_ranges_match_reference_conventions():
    q = Q.IntegerQuantizer(4, True, 'per_group', group_size=128)
    assert int(q.qmin) == -8 and int(q.qmax) == 7 and q.qmin.dim() == 0
    q = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    assert q.qmin.dtype == torch.float32 and float(q.qmin) == 0.0 and int(q.qmax) == 15  # quant.py:673-677
    q = Q.IntegerQuantizer(8, True, 'per_channel', int_range=[-127, 127])
    assert int(q.qmin) == -127
    assert q._code_dtype()[0] == torch.int8
    assert Q.IntegerQuantizer(8, False, 'per_channel')._code_dtype()[0] == torch.uint8
    assert Q.IntegerQuantizer(4, False, 'per_channel')._code_dtype()[0] == torch.int32


def test_float_ranges():
    q = Q.FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
    assert float(q.qmax) == 448.0 and q.sym
    q = Q.FloatQuantizer('e5m2', True, 'per_channel', use_qtorch=True)
    assert float(q.qmax) == 57344.0
    with pytest.raises(NotImplementedError):
        Q.FloatQuantizer('e2m1', True, 'per_channel', use_qtorch=True)


def test_unit_grid():
    q = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    u = q._units(torch.empty(8, 512))
    assert (u.rows, u.cols, u.ur, u.uc, u.n_units, u.qshape) == (8, 512, 1, 128, 32, (32, 1))
    u = q._units(torch.empty(8, 64))  # last dim < group: untouched -> one unit per row (quant.py:631-632)
    assert (u.ur, u.uc, u.qshape) == (1, 64, (8, 1))
    with pytest.raises(ValueError):
        q._units(torch.empty(5, 300))
    q = Q.IntegerQuantizer(8, True, 'per_token')
    u = q._units(torch.empty(2, 12, 1024))
    assert (u.rows, u.cols, u.uc, u.qshape) == (24, 1024, 1024, (2, 12, 1))
    q = Q.IntegerQuantizer(8, True, 'per_tensor')
    assert q._units(torch.empty(64, 320)).qshape == ()
    q = Q.FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
    u = q._units(torch.empty(576, 2048))
    assert u.qshape == (5, 1, 16, 1) and u.n_units == 80


def test_no_cpu_fallback():
    q = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    with pytest.raises(RuntimeError, match='no CPU path'):
        q.fake_quant_weight_dynamic(torch.randn(4, 256))


def test_unsupported_modes_fail_loudly():
    with pytest.raises(NotImplementedError):
        Q.IntegerQuantizer(4, True, 'per_channel', ste=True)
    with pytest.raises(AssertionError):
        Q.IntegerQuantizer(8, False, 'per_block', block_size=128)  # quant.py:60
