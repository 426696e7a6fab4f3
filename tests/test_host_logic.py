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


def test_fp8_infer_mirror_has_no_cpu_path_and_keeps_the_reference_assertions():
    """lightcompress_b200.fp8_infer (mirror of kernel.py / module_utils.py:41-46, 223-261) on CPU tensors: the reference's
    own assertions fire first, then the no-CPU-path error; the module keeps the reference's parameter layout."""
    from lightcompress_b200 import fp8_infer as fi
    with pytest.raises(AssertionError, match='divisible by block_size'):
        fi.act_quant(torch.zeros(4, 100))
    with pytest.raises(AssertionError, match='contiguous'):
        fi.act_quant(torch.zeros(256, 4).t())
    with pytest.raises(RuntimeError, match='no CPU path'):
        fi.act_quant(torch.zeros(4, 128))
    a = torch.zeros(4, 128).to(torch.float8_e4m3fn)
    b = torch.zeros(128, 128).to(torch.float8_e4m3fn)
    with pytest.raises(RuntimeError, match='no CPU path'):
        fi.fp8_gemm(a, torch.ones(4, 1), b, torch.ones(1, 1))
    with pytest.raises(AssertionError, match='contiguous'):
        fi.fp8_gemm(a, torch.ones(4, 1), b.t(), torch.ones(1, 1))
    m = fi.LlmcFp8Linear(300, 200, True, 128)
    assert m.weight.dtype == torch.float8_e4m3fn and tuple(m.weight.shape) == (200, 300)
    assert m.weight_scale_inv.dtype == torch.float32 and tuple(m.weight_scale_inv.shape) == (2, 3)
    assert tuple(m.bias.shape) == (200,)


def test_weight_lists_outside_the_multi_launch_envelope_loop_over_the_reference_api():
    """fake / real_quant_weights_dynamic fall back to the per-tensor calls (which have no CPU path) instead of guessing."""
    q = Q.FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
    assert not q._block_fp8_multi_ok([torch.zeros(128, 128, dtype=torch.bfloat16)])          # CPU tensor
    assert not q._block_fp8_multi_ok([])
    with pytest.raises(RuntimeError, match='no CPU path'):
        q.real_quant_weights_dynamic([torch.zeros(128, 128, dtype=torch.bfloat16)])
    assert q.real_quant_weights_dynamic([]) == [] and q.fake_quant_weights_dynamic([]) == []


def test_patch_fp8_inference_rebinds_only_what_a_module_defines():
    import types

    from lightcompress_b200 import fp8_infer as fi
    from lightcompress_b200.algo import patch_fp8_inference
    kernel = types.ModuleType('ref_kernel')
    kernel.act_quant = kernel.fp8_gemm = kernel.weight_cast_to_fp8 = kernel.weight_cast_to_bf16 = object()
    kernel.fp8_gemm_configs = [1, 2]
    mu = types.ModuleType('ref_module_utils')
    mu.act_quant = mu.fp8_gemm = mu.block_wise_fp8_forward_func = object()
    done = patch_fp8_inference(kernel, mu)
    assert kernel.act_quant is fi.act_quant and kernel.fp8_gemm is fi.fp8_gemm and mu.fp8_gemm is fi.fp8_gemm
    assert mu.block_wise_fp8_forward_func is fi.block_wise_fp8_forward_func
    assert kernel.weight_cast_to_fp8 is fi.weight_cast_to_fp8 and kernel.fp8_gemm_configs == [1, 2]
    assert not hasattr(kernel, 'block_wise_fp8_forward_func') and not hasattr(mu, 'weight_cast_to_fp8')
    assert sorted(done['ref_module_utils']) == ['act_quant', 'block_wise_fp8_forward_func', 'fp8_gemm']


def test_float_quantizer_without_qtorch_host_side():
    """use_qtorch unset: any eXmY format constructs (quant.py:963-1003 sets no qmin / qmax on that path), the 1-D
    granularities are accepted, per_tensor raises like the reference (maxval.shape[0] of a 0-dim tensor), real quant and
    CPU tensors are refused loudly."""
    for bit in ('e4m3', 'e5m2', 'e2m1', 'e3m2', 'e4m7'):
        q = Q.FloatQuantizer(bit, True, 'per_channel')
        assert q.qmin is None and q.qmax is None and q.e_bits == int(bit[1]) and q.m_bits == int(bit[-1])
    with pytest.raises(NotImplementedError):
        Q.FloatQuantizer('e2m1', True, 'per_channel', use_qtorch=True)        # qtorch grid formats are not on the hot path
    q = Q.FloatQuantizer('e4m3', True, 'per_group', group_size=128)
    assert q._elem_units(torch.zeros(4, 256)) == (8, 128)
    assert Q.FloatQuantizer('e4m3', True, 'per_token')._elem_units(torch.zeros(2, 3, 64)) == (6, 64)
    with pytest.raises(IndexError):
        Q.FloatQuantizer('e4m3', True, 'per_tensor')._elem_units(torch.zeros(4, 8))
    with pytest.raises(NotImplementedError):
        Q.FloatQuantizer('e4m3', True, 'per_block', block_size=128)._elem_units(torch.zeros(128, 128))
    with pytest.raises(NotImplementedError):
        Q.FloatQuantizer('e4m3', True, 'per_channel', calib_algo='mse')._elem_units(torch.zeros(4, 8))
    with pytest.raises(RuntimeError, match='no CPU path'):
        Q.FloatQuantizer('e4m3', True, 'per_channel').fake_quant_weight_dynamic(torch.zeros(4, 8, dtype=torch.bfloat16))
    with pytest.raises(NotImplementedError):
        Q.FloatQuantizer('e4m3', True, 'per_channel').real_quant_weight_dynamic(torch.zeros(4, 8, dtype=torch.bfloat16))


def test_fp8_gemm_split_k_plan_is_a_host_function():
    """lcb_fp8_gemm_workspace_bytes is the split-K decision (a fitted cost model, no device call): decode shapes with a long
    K get a workspace (fp32 sums + one counter per tile, 16-byte multiples), prefill shapes and short K do not."""
    from lightcompress_b200 import _cabi
    lib = _cabi.load()
    f = lib.lcb_fp8_gemm_workspace_bytes
    assert f(16, 2048, 7168) == (16 * 2048 + 8) * 4          # 8 output tiles
    assert f(1, 1536, 7168) == (1536 + 8) * 4                # 6 tiles: counters padded to 8
    assert f(128, 7168, 2048) == 0 and f(64, 24576, 1536) == 0 and f(4096, 4096, 4096) == 0 and f(16, 2048, 256) == 0
    assert f(0, 128, 128) == 0
    assert f(128, 2048, 7168) > 0 and f(128, 4096, 4096) > 0
