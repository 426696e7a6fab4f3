"""Algorithm layer of the B200 hot path: kernel-level pieces (GPTQLayer, awq_search_scale, smoothquant_scale,
rtn_quantize), the block-loop plug-ins (ALGO_REGISTRY['RTN' | 'GPTQ' | 'Awq' | 'SmoothQuant']) and
`patch_reference`, which rebinds the reference classes' hot methods onto the same kernels."""
from . import adapters  # noqa: F401
from .algos import GPTQ, Awq  # noqa: F401
from .awq import awq_search_scale, awq_search_scale_module, get_act_scale, get_scales, get_weight_scale  # noqa: F401
from .blockwise import ALGO_REGISTRY, RTN, BaseBlockwiseQuantization, SmoothQuant, make_quantizer  # noqa: F401
from .gptq import GPTQLayer  # noqa: F401
from .patch import patch_autoclipper, patch_fp8_inference, patch_reference  # noqa: F401
from .realquant import REALQUANT_LINEAR_MAP, AutoawqRealQuantLinear, VllmRealQuantLinear  # noqa: F401
from .rtn import rtn_quantize, rtn_real_quantize, rtn_real_quantize_sharded  # noqa: F401
from .smoothquant import ActMax, smoothquant_scale  # noqa: F401
