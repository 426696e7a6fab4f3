"""Real-quant export modules: the mirrors of module_utils.py:855-1158 that `deploy('<backend>_quant')` puts in place of
the block linears (base_bq.py:933-962 -> model.replace_module_all -> <Cls>.new(module, w_q, quant_config)).

Same buffer names, shapes and dtypes as the reference classes (they are what `save_model` writes and what the serving
back ends load), but quantization and packing run on the GPU: `w_q` = IntegerQuantizer / FloatQuantizer
.real_quant_weight_{dynamic,static} (lcb_fakequant with a code output mode), then lcb_pack_rows (the reference packs with
numpy on the host, module_utils.py:931-955) or lcb_awq_gemm_pack (a Python loop per input channel in the reference,
:1099-1158).
"""
import torch
import torch.nn as nn

from .. import pack as lcpack


def _cfg(quant_config, key, default=None):
    w = quant_config['weight']
    return w.get(key, default) if hasattr(w, 'get') else getattr(w, key, default)


class VllmRealQuantLinear(nn.Module):
    """module_utils.py:855-955 (also the class behind lightllm_quant / sgl_quant / lightx2v_quant)."""

    def __init__(self, weight, bias, scales, input_scale, need_pack, scales_name):
        super().__init__()
        self.register_buffer('weight_packed' if need_pack else 'weight', weight)
        if bias is not None:
            self.register_buffer('bias', bias)
        else:
            self.bias = None
        self.register_buffer(scales_name, scales)
        self.register_buffer('input_scale', input_scale)

    @torch.no_grad()
    def forward(self, x):
        raise NotImplementedError

    @classmethod
    @torch.no_grad()
    def new(cls, module, w_q, quant_config):
        weight, scales = cls.quant_pack(module, w_q, quant_config)
        input_scale = getattr(module, 'buf_act_scales_0', None)
        act = quant_config.get('act') if hasattr(quant_config, 'get') else None
        if act and act.get('static', False) and quant_config.get('quant_type', 'int-quant') == 'int-quant':
            input_scale = input_scale.unsqueeze(0)
        bias = module.bias.data if module.bias is not None else None
        need_pack = _cfg(quant_config, 'need_pack', False)
        scales_name = 'weight_scale_inv' if _cfg(quant_config, 'granularity') == 'per_block' else 'weight_scale'
        m = cls(weight, bias, scales, input_scale, need_pack, scales_name)
        m.in_features, m.out_features = module.in_features, module.out_features
        m.weight_shape, m.weight_dtype = weight.shape, weight.dtype
        m.scales_shape, m.scales_dtype = scales.shape, scales.dtype
        m.zeros_shape = m.zeros_dtype = None
        return m

    @classmethod
    @torch.no_grad()
    def quant_pack(cls, module, w_q, quant_config):
        if module.weight.data.dtype == torch.float8_e4m3fn:      # an FP8 block-wise checkpoint: back to bf16 first (:920-925)
            from ..quant import weight_cast_to_bf16
            module.weight.data = weight_cast_to_bf16(module.weight.data, module.weight_scale_inv.data, module.block_size)
        weight, scales, zeros = w_q(module)
        if _cfg(quant_config, 'need_pack', False):
            weight, scales = lcpack.vllm_pack(weight, scales, _cfg(quant_config, 'bit'))
        return weight, scales

    def __repr__(self):
        return (f'{type(self).__name__}(in_features={self.in_features}, out_features={self.out_features}, '
                f'bias={self.bias is not None}, weight_shape={self.weight_shape}, weight_dtype={self.weight_dtype}, '
                f'scales_shape={self.scales_shape}, scales_dtype={self.scales_dtype}, zeros_shape={self.zeros_shape}, '
                f'zeros_dtype={self.zeros_dtype})')


class LightllmRealQuantLinear(VllmRealQuantLinear):
    pass


class SglRealQuantLinear(VllmRealQuantLinear):
    pass


class Lightx2vRealQuantLinear(VllmRealQuantLinear):
    pass


class AutoawqRealQuantLinear(nn.Module):
    """module_utils.py:1029-1158 (also mlcllm_quant)."""

    def __init__(self, weight, bias, scales, zeros):
        super().__init__()
        self.register_buffer('qweight', weight)
        if bias is not None:
            self.register_buffer('bias', bias)
        else:
            self.bias = None
        self.register_buffer('scales', scales)
        if zeros is not None:
            self.register_buffer('qzeros', zeros)
        else:
            self.qzeros = None

    @torch.no_grad()
    def forward(self, x):
        raise NotImplementedError

    @classmethod
    @torch.no_grad()
    def new(cls, module, w_q, quant_config):
        weight, scales, zeros = cls.quant_pack(module, w_q, quant_config)
        bias = module.bias.data if module.bias is not None else None
        m = cls(weight, bias, scales, zeros)
        m.in_features, m.out_features = module.in_features, module.out_features
        m.weight_shape, m.weight_dtype = weight.shape, weight.dtype
        m.scales_shape, m.scales_dtype = scales.shape, scales.dtype
        m.zeros_shape = zeros.shape if zeros is not None else None
        m.zeros_dtype = zeros.dtype if zeros is not None else None
        return m

    @classmethod
    @torch.no_grad()
    def quant_pack(cls, module, w_q, quant_config):
        if module.weight.data.dtype == torch.float8_e4m3fn:
            from ..quant import weight_cast_to_bf16
            module.weight.data = weight_cast_to_bf16(module.weight.data, module.weight_scale_inv.data, module.block_size)
        _, scales, zeros = w_q(module)
        pack_version = _cfg(quant_config, 'pack_version')
        if pack_version != 'gemm_pack':
            raise NotImplementedError(f'Not support {pack_version}.')
        return lcpack.awq_gemm_pack(module.weight.data, scales, zeros, _cfg(quant_config, 'bit'), _cfg(quant_config, 'group_size'))

    __repr__ = VllmRealQuantLinear.__repr__


class MlcllmRealQuantLinear(AutoawqRealQuantLinear):
    pass


REALQUANT_LINEAR_MAP = {     # module_utils.py:1222-1229
    'vllm_quant': VllmRealQuantLinear,
    'lightllm_quant': LightllmRealQuantLinear,
    'sgl_quant': SglRealQuantLinear,
    'autoawq_quant': AutoawqRealQuantLinear,
    'mlcllm_quant': MlcllmRealQuantLinear,
    'lightx2v_quant': Lightx2vRealQuantLinear,
}
