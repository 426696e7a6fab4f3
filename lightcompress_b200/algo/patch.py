"""`patch_reference(ALGO_REGISTRY)`: rebind the hot methods of the reference's own algorithm classes onto the sm_100a
kernels, leaving their block loop, hooks, bookkeeping and deploy untouched.

Exercised on the REAL reference classes in tests/test_ref_pipeline_gpu.py (the unmodified reference ships to the GPU box
as test infrastructure under oracle/_ref) and on stand-in classes in tests/test_algo_gpu.py.
"""
import os

import torch
import torch.nn as nn

from .. import ops
from .. import quant as lcquant
from . import awq as lcawq
from . import dist as lcdist


def _lookup(registry, name):
    """The reference's Register is a dict SUBCLASS that keeps its entries in `self._dict` and overrides __getitem__ /
    __contains__ only (registry_factory.py:1-37): `registry.get(name)` answers from the empty base dict."""
    try:
        return registry[name] if name in registry else None
    except KeyError:
        return None


def as_lcb_quantizer(q):
    """The product quantizer that computes what `q` computes: `q` itself after the class swap of INTEGRATION.md §3(a), or a
    twin built from the reference quantizer's constructor arguments (quant.py:47-101 keeps them: bit, sym, granularity,
    kwargs) when only the algorithms were patched."""
    if q is None or isinstance(q, lcquant.BaseQuantizer):
        return q
    twin = q.__dict__.get('_lcb_twin')
    if twin is None:
        cls = lcquant.FloatQuantizer if type(q).__name__ == 'FloatQuantizer' else lcquant.IntegerQuantizer
        twin = cls(q.bit, q.sym, q.granularity, **dict(getattr(q, 'kwargs', {})))
        q.__dict__['_lcb_twin'] = twin
    return twin


# ---- GPTQ (gptq.py) ------------------------------------------------------------------------------------------------------
def _gptq_add_batch(orig):
    def add_batch(self, layer, name, inp, out):
        if not isinstance(layer, nn.Linear) or type(layer).__name__ == 'RotateLinear':
            return orig(self, layer, name, inp, out)      # Conv2d / Conv1D / online rotation: reference path
        if inp.dim() == 2:
            inp = inp.unsqueeze(0)
        cache = self.layers_cache[name]
        cache['nsamples'] += inp.shape[0]
        cache['lcb_unnormalised'] = True                   # H holds sum x x^T until process_hessian_and_weights
        ops.hessian_accum(cache['H'], inp)
    return add_batch


class _FusedCholeskyChain(object):
    """While active, the reference's `H = torch.linalg.cholesky(H); H = torch.cholesky_inverse(H);
    H = torch.linalg.cholesky(H, upper=True)` (gptq.py:172-174) runs as ONE ops.cholinv_upper call: the first call
    computes the final factor and hands back a token, the other two pass it through. Anything that is not that exact
    chain on an fp32 CUDA matrix falls through to torch."""

    def __enter__(self):
        self._chol, self._cinv = torch.linalg.cholesky, torch.cholesky_inverse
        st = {}

        def cholesky(A, *a, upper=False, **k):
            if upper and st.get('tok') is A:
                st.clear()
                return A
            if (not upper and not a and not k and A.is_cuda and A.dtype == torch.float32 and A.dim() == 2
                    and A.shape[0] == A.shape[1]):
                st['tok'] = ops.cholinv_upper(A.contiguous().clone(), 0.0)
                st['stage'] = 1
                return st['tok']
            return self._chol(A, *a, upper=upper, **k)

        def cholesky_inverse(L, *a, **k):
            if st.get('tok') is L and st.get('stage') == 1:
                st['stage'] = 2
                return L
            return self._cinv(L, *a, **k)

        torch.linalg.cholesky, torch.cholesky_inverse = cholesky, cholesky_inverse
        return self

    def __exit__(self, *exc):
        torch.linalg.cholesky, torch.cholesky_inverse = self._chol, self._cinv
        return False


def _gptq_process(orig):
    def process_hessian_and_weights(self, layer, name):
        cache = self.layers_cache[name]
        if cache.pop('lcb_unnormalised', False):
            n = lcdist.allreduce_hessian(cache['H'], cache['nsamples'])   # one collective per linear (gptq.py:292-295)
            ops.hessian_finalize(cache['H'], n)
        with _FusedCholeskyChain():
            return orig(self, layer, name)
    return process_hessian_and_weights


def _gptq_weight_transform(orig):
    def weight_transform(self, W, Hinv, Losses, tmp):
        wq = as_lcb_quantizer(self.wquantizer)
        if type(wq).__name__ != 'IntegerQuantizer' or self.blocksize > 128 or self.blocksize % 8 or W.shape[1] % 8:
            return orig(self, W, Hinv, Losses, tmp)
        R = W.shape[0]
        per_group = wq.granularity == 'per_group'
        gs = wq.group_size if per_group else 0
        static_scales = static_zeros = col_group = None
        if per_group and self.static_groups:
            static_scales = torch.stack([g['scale'].reshape(-1).float() for g in self.groups])
            if not wq.sym:
                static_zeros = torch.stack([g['zero'].reshape(-1).float() for g in self.groups])
            if self.actorder:
                col_group = (self.perm // gs).to(torch.int32)
        elif not per_group:
            static_scales = self.qparams['scale'].reshape(1, R).float()
            if not wq.sym:
                static_zeros = self.qparams['zero'].reshape(1, R).float()
        Wc = W if W.is_contiguous() else W.contiguous()
        t, l, sc, zp = ops.gptq_weight_transform(Wc, Hinv.contiguous(), self.blocksize, gs, float(wq.qmin), float(wq.qmax),
                                                 wq.sym, wq.round_zp, n_nonout=self.n_nonout, static_scales=static_scales,
                                                 static_zeros=static_zeros, col_group=col_group)
        if Wc is not W:
            W.copy_(Wc)
        tmp.copy_(t)
        Losses.copy_(l)
        if per_group and not self.static_groups:            # search_column_qparams bookkeeping (gptq.py:359-366)
            dev = W.device
            for g in range(sc.shape[0]):
                q = {'scale': sc[g].reshape(R, 1), 'zero': zp[g].reshape(R, 1) if zp is not None else torch.tensor(0.0),
                     'qmax': wq.qmax.to(dev), 'qmin': wq.qmin.to(dev)}
                if g < len(self.groups):
                    self.groups[g] = q
                else:
                    self.groups.append(q)
            self.qparams = dict(self.groups[sc.shape[0] - 1])
    return weight_transform


# ---- AWQ (awq.py) ----------------------------------------------------------------------------------------------------------
def _awq_search(orig):
    def search_scale_subset(self, prev_op, layers_dict, input, inspect_module, is_gqa, subset_kwargs):
        layers = list(layers_dict.values())
        fused = (len(layers) == 1 and inspect_module is layers[0] and isinstance(inspect_module, nn.Linear)
                 and inspect_module.bias is None and len(input) == 1 and not is_gqa and self.w_only
                 and not self.padding_mask and getattr(self, 'awq_bs', None) is None
                 and input[0].dtype in (torch.bfloat16, torch.float16)
                 and inspect_module.weight.dtype == input[0].dtype and int(os.environ.get('WORLD_SIZE', '1')) == 1)
        if not fused:
            return orig(self, prev_op, layers_dict, input, inspect_module, is_gqa, subset_kwargs)
        x = input[0].to(inspect_module.weight.device)
        scales, _, _ = lcawq.awq_search_scale(x, inspect_module.weight.data, as_lcb_quantizer(self.wquantizer),
                                              self.trans_version)
        return scales
    return search_scale_subset


# ---- SmoothQuant (smoothquant.py) ----------------------------------------------------------------------------------------
def _sq_act_scale(orig):
    def get_act_scale(self, tensors):
        acc = None
        for x in tensors:
            x = x.cuda()
            if acc is None:
                acc = torch.zeros(x.shape[-1], dtype=torch.float32, device=x.device)
            ops.col_absstat(x, acc, mean=False)
        return acc
    return get_act_scale


# ---- AutoClip (auto_clip.py) ----------------------------------------------------------------------------------------------
def _autoclip_layer(orig):
    def auto_clip_layer(self, block_idx, layer_name, w, inputs, n_grid=20, max_shrink=0.5, n_sample_token=512, eps=0.0):
        wq = as_lcb_quantizer(self.wquantizer)
        gs = wq.group_size if wq.granularity == 'per_group' else w.shape[1]
        v1 = self.clip_version == 'v1'
        ok = ((v1 or (self.clip_version == 'v2' and wq.granularity == 'per_channel' and (self.w_only or eps == 0.0)))
              and len(inputs) == 1 and not self.padding_mask and w.dim() == 2
              and type(wq).__name__ == 'IntegerQuantizer' and wq.granularity in ('per_group', 'per_channel')
              and w.dtype in (torch.bfloat16, torch.float16) and inputs[0].dtype == w.dtype
              and w.shape[1] % gs == 0 and gs % 8 == 0 and (gs <= 128 or gs % 128 == 0) and w.shape[0] % 64 == 0)
        aq = None
        if ok and not self.w_only:
            aq = as_lcb_quantizer(self.aquantizer)
            ok = type(aq).__name__ == 'IntegerQuantizer' and aq.granularity == 'per_token'
        if not ok:
            return orig(self, block_idx, layer_name, w, inputs, n_grid, max_shrink, n_sample_token, eps)
        from . import autoclip as lcac
        x = inputs[0].to(w.device)
        wd = w.data if hasattr(w, 'data') else w
        if v1 and aq is None:
            return lcac.auto_clip_layer(wd, x, wq, self.clip_sym, n_grid, max_shrink, n_sample_token)
        return lcac.auto_clip_layer_general(wd, x, wq, self.clip_sym, self.clip_version, aq, n_grid, max_shrink, n_sample_token)
    return auto_clip_layer


def _rebind(cls, name, wrap):
    orig = getattr(cls, name)
    saved = cls.__dict__.get('_lcb_originals') or {}
    saved[name] = (name in cls.__dict__, orig)     # (defined on this class itself?, function)
    cls._lcb_originals = saved
    setattr(cls, name, wrap(orig))


def _restore(cls):
    """Undo `_rebind` on a class. Returns True when something was restored."""
    saved = cls.__dict__.get('_lcb_originals')
    if not saved:
        return False
    for name, (own, orig) in saved.items():
        if own:
            setattr(cls, name, orig)
        else:
            delattr(cls, name)                      # inherited: fall back to the base class again
    cls._lcb_originals = None
    cls._lcb_patched = False
    return True


def patch_autoclipper(cls):
    """cls: llmc.compression.quantization.auto_clip.AutoClipper. Routes auto_clip_layer to lcb_autoclip_search (v1, w_only)
    or to the split lcb_autoclip_init / _step search (v2 per_channel, w+a); anything else keeps the reference path.
    Returns True when newly patched."""
    if cls.__dict__.get('_lcb_patched', False):
        return False
    _rebind(cls, 'auto_clip_layer', _autoclip_layer)
    cls._lcb_patched = True
    return True


def unpatch_autoclipper(cls):
    return _restore(cls)


def unpatch_reference(registry):
    """Restore the reference methods `patch_reference` rebound. Returns the restored names."""
    return [n for n in ('GPTQ', 'Awq', 'SmoothQuant') if _lookup(registry, n) is not None and _restore(_lookup(registry, n))]


def patch_reference(registry):
    """registry: llmc.utils.registry_factory.ALGO_REGISTRY (or any dict name -> class). Returns the patched names."""
    done = []
    g = _lookup(registry, 'GPTQ')
    if g is not None and not g.__dict__.get('_lcb_patched', False):
        _rebind(g, 'add_batch', _gptq_add_batch)
        _rebind(g, 'process_hessian_and_weights', _gptq_process)
        _rebind(g, 'weight_transform', _gptq_weight_transform)
        g._lcb_patched = True
        done.append('GPTQ')
    a = _lookup(registry, 'Awq')
    if a is not None and not a.__dict__.get('_lcb_patched', False):
        _rebind(a, 'search_scale_subset', _awq_search)
        a._lcb_patched = True
        done.append('Awq')
    s = _lookup(registry, 'SmoothQuant')
    if s is not None and not s.__dict__.get('_lcb_patched', False):
        _rebind(s, 'get_act_scale', _sq_act_scale)
        s._lcb_patched = True
        done.append('SmoothQuant')
    return done


def patch_fp8_inference(*modules):
    """Rebind the reference's block-wise FP8 inference functions — the Triton module `kernel.py` (`act_quant` :31-54,
    `fp8_gemm` :218-242, `weight_cast_to_fp8 / weight_cast_to_bf16` :73-138) and `module_utils.py`
    (`block_wise_fp8_forward_func` :41-46, which `LlmcFp8Linear.forward` and the FP8 fake / real linears call) — onto
    `lightcompress_b200.fp8_infer`. Pass every module object that holds one of those names
    (`llmc.compression.quantization.kernel`, `.module_utils`, `.quant`); names a module does not define are left alone.
    Returns {module name: [rebound names]}."""
    from .. import fp8_infer
    names = ('act_quant', 'fp8_gemm', 'block_wise_fp8_forward_func', 'weight_cast_to_fp8', 'weight_cast_to_bf16')
    done = {}
    for mod in modules:
        hit = []
        for n in names:
            if hasattr(mod, n):
                setattr(mod, n, getattr(fp8_infer, n))
                hit.append(n)
        done[getattr(mod, '__name__', repr(mod))] = hit
    return done
