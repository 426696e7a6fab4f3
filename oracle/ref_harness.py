"""Drive the UNMODIFIED reference pipeline (oracle/_ref, assembled by oracle/build_ref.py) on an in-memory HF model.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and tests/golden/make_e2e_golden.py); never by lightcompress_b200/ or a
timed region of bench.py. Follows SURVEY.md §8c recipe 3 = the flow of llmc/__main__.py:28-78:

    model = MODEL_REGISTRY[type](config)                      -> `wrap_model` (an in-memory HF model, no from_pretrained)
    model.collect_first_block_input(calib_data, padding_mask)
    opt = ALGO_REGISTRY[method](model, quant_cfg, model.get_first_block_input(), model.get_padding_mask(), config)
    opt.run_block_loop(); opt.deploy('fake_quant')

On a CUDA machine the copy under oracle/_ref/llmc is the reference verbatim (CUDA eager); with device='cpu' the tree under
oracle/_ref/cpu (the reference's own ci_check CPU rewrite) is used instead.
"""
import importlib
import os
import sys

import torch
import torch.distributed as dist

from . import build_ref

_ACTIVE = {}


def available():
    return build_ref.available()


def activate(device='cuda'):
    """Put the reference on sys.path (once per process; one variant per process) and return its modules."""
    variant = 'cpu' if str(device) == 'cpu' else ''
    if _ACTIVE:
        if _ACTIVE['variant'] != variant:
            raise RuntimeError('one reference variant per process (cpu rewrite vs verbatim)')
        return _ACTIVE['mods']
    if not available():
        raise RuntimeError('oracle/_ref is missing: run `python oracle/build_ref.py` where /root/reference exists')
    import transformers  # noqa: F401  (must be imported before the accelerate stand-in is visible)
    import transformers.models.llama.modeling_llama  # noqa: F401
    import transformers.models.opt.modeling_opt  # noqa: F401
    root = os.path.join(build_ref.OUT, variant) if variant else build_ref.OUT
    shims = os.path.join(build_ref.OUT, 'shims')
    for name in ('accelerate', 'easydict'):
        try:
            importlib.import_module(name)
        except Exception:
            sys.path.append(shims)   # appended: a real install wins
            importlib.import_module(name)
    sys.path.insert(0, root)
    os.environ.setdefault('RANK', '0')
    os.environ.setdefault('LOCAL_RANK', '0')
    os.environ.setdefault('WORLD_SIZE', '1')
    if not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', str(29500 + os.getpid() % 2000))
        dist.init_process_group('gloo', rank=0, world_size=1)
    from loguru import logger
    logger.remove()
    logger.add(sys.stderr, level=os.environ.get('LCB_REF_LOGLEVEL', 'ERROR'))
    mods = {
        'registry': importlib.import_module('llmc.utils.registry_factory'),
        'quantization': importlib.import_module('llmc.compression.quantization'),
        'base_model': importlib.import_module('llmc.models.base_model'),
        'llama': importlib.import_module('llmc.models.llama'),
        'opt': importlib.import_module('llmc.models.opt'),
        'easydict': importlib.import_module('easydict'),
    }
    for k in ('quant', 'gptq', 'awq', 'auto_clip', 'smoothquant', 'module_utils', 'base_blockwise_quantization'):
        mods[k] = importlib.import_module('llmc.compression.quantization.' + k)
    _ACTIVE.update(variant=variant, mods=mods)
    return mods


def wrap_model(hf_model, model_type, device='cuda'):
    """A reference `llmc.models.{Llama,Opt}` object around an in-memory HF model: everything BaseModel.__init__
    (base_model.py:23-54) sets, minus build_tokenizer / build_model (no checkpoints on the box)."""
    mods = activate(device)
    cls = {'Llama': mods['llama'].Llama, 'Opt': mods['opt'].Opt}[model_type]
    ED = mods['easydict'].EasyDict
    m = object.__new__(cls)
    dtype = next(hf_model.parameters()).dtype
    m.config = ED({'model': {'type': model_type, 'path': None, 'torch_dtype': str(dtype)}})
    m.model_type, m.model_path, m.tokenizer_mode = model_type, None, 'fast'
    m.use_cpu_to_save_cuda_mem_for_catcher = False
    m.torch_dtype = dtype
    m.block_wise_quant = False
    m.device_map, m.use_cache = None, False
    m.mm_model = m.vision_model = m.vision_projector = m.audio_model = m.audio_projector = None
    m.modality = 'language'
    m.kvcache_buffer = []
    m.tokenizer = None
    m.model = hf_model
    m.model_config = hf_model.config
    if hasattr(m.model_config, 'use_cache'):
        m.model_config.use_cache = False
    hf_model.eval()
    m.update_key_info()
    return m


def make_config(model_type, quant_cfg, seq_len):
    mods = activate(_ACTIVE.get('variant') == 'cpu' and 'cpu' or 'cuda')
    ED = mods['easydict'].EasyDict
    cfg = ED({'model': {'type': model_type, 'path': None, 'torch_dtype': 'auto'},
              'calib': {'name': 'synthetic', 'seq_len': seq_len, 'bs': -1, 'n_samples': 0},
              'quant': dict(quant_cfg)})
    q = cfg.quant
    q['modality'] = 'language'   # __main__.get_modality (llmc/utils/utils.py) sets this on the language config
    return cfg, q


def run_pipeline(hf_model, model_type, quant_cfg, calib_batches, device='cuda', deploy='fake_quant', patch=None):
    """calib_batches: list of {'input_ids': LongTensor[b, T]} (+ 'attention_mask'). Returns (blockwise_opt, model_adapter).
    `patch(mods)` runs after the reference is imported and before the algorithm object is created."""
    mods = activate(device)
    if patch is not None:
        patch(mods)
    model = wrap_model(hf_model, model_type, device)
    seq_len = int(calib_batches[0]['input_ids'].shape[1])
    config, qcfg = make_config(model_type, quant_cfg, seq_len)
    model.collect_first_block_input([dict(b) for b in calib_batches], None)
    opt = mods['registry'].ALGO_REGISTRY[qcfg['method']](model, qcfg, model.get_first_block_input(),
                                                         model.get_padding_mask(), config)
    opt.run_block_loop()
    if deploy:
        opt.deploy(deploy)
    return opt, model


def load_quant_only():
    """The reference's quantizer module (llmc/compression/quantization/quant.py) WITHOUT importing the package __init__
    (which pulls in every algorithm, vllm, transformers: ~30 s): a stub package whose __path__ is the reference
    directory, so that quant.py's `from .utils import ceil_div` resolves. Used by bench.py's CPU arm."""
    import types
    name = '_lcb_refq'
    if name + '.quant' in sys.modules:
        return sys.modules[name + '.quant']
    if not available():
        raise RuntimeError('oracle/_ref is missing')
    pkg = types.ModuleType(name)
    pkg.__path__ = [os.path.join(build_ref.OUT, 'llmc', 'compression', 'quantization')]
    sys.modules[name] = pkg
    from loguru import logger
    logger.remove()
    logger.add(sys.stderr, level='ERROR')
    return importlib.import_module(name + '.quant')
