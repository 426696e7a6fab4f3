"""GPTQ and Awq as block-loop plug-ins (gptq.py:22-478, awq.py:29-372) on top of the kernel-level pieces."""
import torch
import torch.nn as nn

from .. import ops
from . import awq as lcawq
from .blockwise import ALGO_REGISTRY, BaseBlockwiseQuantization
from .gptq import GPTQLayer, HessianState


@ALGO_REGISTRY
class GPTQ(BaseBlockwiseQuantization):
    def __init__(self, model, quant_config, input, padding_mask, config, modality='language'):
        super().__init__(model, quant_config, input, padding_mask, config)
        sp = self.special
        self.true_sequential = sp.get('true_sequential', True)
        self.static_groups = sp.get('static_groups', False)
        self.actorder = sp.get('actorder', False)
        self.percdamp = sp.get('percdamp', 0.01)
        self.blocksize = sp.get('blocksize', 128)
        if sp.get('owq', False):
            raise NotImplementedError('OWQ outlier columns: use ops.gptq_weight_transform(n_nonout=...) directly')
        self.model_dtype = next(self.blocks[0].parameters()).dtype
        self.layers_cache = {}
        self.qparams = {}     # module -> buf_scales / buf_zeros (+ act-order perm) for deploy (gptq.py:398-452)

    def _new_layer(self, m, hessian=None):
        return GPTQLayer(m.weight, self.wquantizer, self.blocksize, self.percdamp, self.actorder, self.static_groups,
                         hessian=hessian)

    def block_init(self, block):                       # gptq.py:318-322
        self.named_layers = self.model.get_block_linears(block)
        # collect_model_qparams (gptq.py:325-331, base_bq.py:338-365): buf_scales / buf_zeros of the ORIGINAL weight in the
        # model dtype. The reference collects them for every block at construction; a block's weights do not change
        # before its own turn, so collecting at block_init is the same arithmetic.
        for m in self.named_layers.values():
            if isinstance(m, nn.Linear) and m not in self.qparams:
                _, sc, zp, _, _ = self.wquantizer.get_tensor_qparams(m.weight.data)
                self.qparams[m] = {'scales': sc, 'zeros': None if self.wquantizer.sym else zp, 'perm': None, 'invperm': None}
        self.layers_cache, self.hessian_owner = {}, {}
        subsets = self.model.get_subsets_in_block(block)
        for i, subset in enumerate(subsets):
            self.subset_init(subset)
            if self.true_sequential and i > 0:
                # the reference accumulates these in the first forward too and throws them away when the subset is
                # re-hooked after its predecessors were quantized (gptq.py:311-315, base_bq.py:505-526): skip the work
                for n in subset['layers']:
                    self.hessian_owner[n] = False
        for n, m in self.named_layers.items():
            if n not in self.layers_cache:
                self.layers_cache[n] = self._new_layer(m)
                self.hessian_owner[n] = True

    def subset_init(self, subset):                     # gptq.py:311-315: the next subset's Hessians start over
        # The linears of one subset read the same input (subset['input'], models/llama.py:30-75), so their Hessians and
        # inverse factors are identical: one HessianState per (subset, in_features) is accumulated by its first linear
        # and factorised once. `share_hessian: False` in `special` restores one state per linear.
        shared = {}
        for n, m in subset['layers'].items():
            key = m.weight.shape[1] if self.special.get('share_hessian', True) else n
            first = key not in shared
            if first:
                shared[key] = HessianState(m.weight.shape[1], m.weight.device)
            self.layers_cache[n] = self._new_layer(m, shared[key])
            self.hessian_owner[n] = first

    graph_safe_hooks = True   # the hook launches one kernel into a persistent buffer and keeps no tensor

    def needs_input(self, name):   # only the owner of a shared Hessian accumulates (cache_input_hook below)
        return name in self.layers_cache and self.hessian_owner.get(name, True)

    def cache_input_hook(self, m, x, y, name, feat_dict):   # gptq.py:247-251: accumulate, do not keep the input
        if name in self.layers_cache and self.hessian_owner.get(name, True):
            if self._capturing:   # block_forward is recording a CUDA graph: the batch is counted once per replay
                st = self.layers_cache[name].hessian
                b = st.add_batch(x[0].detach(), count=False)
                self._on_replay.append(lambda st=st, b=b: setattr(st, 'nsamples', st.nsamples + b))
            else:
                self.layers_cache[name].add_batch(x[0].detach())

    @torch.no_grad()
    def subset_transform(self, subset, input_feat, subset_kwargs):   # gptq.py:97-126
        wq = self.wquantizer
        per_group = wq.granularity == 'per_group'
        for name, layer in subset['layers'].items():
            if not isinstance(layer, nn.Linear):
                continue
            g = self.layers_cache.pop(name)
            buf = self.qparams[layer]
            R = layer.weight.shape[0]
            ss = sz = None
            if per_group and self.static_groups:      # search_group_qparams (gptq.py:380-395): split the model-dtype buffers
                ss = buf['scales'].float().reshape(R, -1).t().contiguous()
                sz = None if wq.sym else buf['zeros'].float().reshape(R, -1).t().contiguous()
            elif not per_group:                       # search_layer_qparams (gptq.py:368-378)
                ss = buf['scales'].float().reshape(1, R)
                sz = None if wq.sym else buf['zeros'].float().reshape(1, R)
            layer.weight.data = g.quantize(static_scales=ss, static_zeros=sz)
            if per_group and not self.static_groups:  # update_model_qparams (gptq.py:398-409): the searched fp32 groups
                buf['scales'], buf['zeros'] = g.buf_qparams()
                buf['perm'], buf['invperm'] = g.perm, g.invperm
            self.last_losses = g.losses

    def w_qdq(self, module, wquantizer):               # gptq.py:425-452: always the stored (static) qparams
        buf = self.qparams.get(module)
        if buf is None:
            raise RuntimeError('GPTQ.w_qdq: the module was never calibrated (no buf_scales)')
        w = module.weight.data
        need_perm = buf['perm'] is not None and wquantizer.granularity == 'per_group' and not self.static_groups
        if need_perm:
            w = w[:, buf['perm']]
        args = {'scales': buf['scales'], 'zeros': buf['zeros'], 'qmax': wquantizer.qmax, 'qmin': wquantizer.qmin}
        q = wquantizer.fake_quant_weight_static(w, args).to(self.model_dtype)
        return q[:, buf['invperm']] if need_perm else q

    def w_q(self, module, wquantizer):                 # gptq.py:412-422
        buf = self.qparams[module]
        assert buf['perm'] is None or self.static_groups or wquantizer.granularity != 'per_group', \
            'act-order with dynamic groups cannot be exported (gptq.py:455-457)'
        args = {'scales': buf['scales'].to(self.model_dtype), 'zeros': buf['zeros'], 'qmax': wquantizer.qmax,
                'qmin': wquantizer.qmin}
        return wquantizer.real_quant_weight_static(module.weight.data, args)


@ALGO_REGISTRY
class Awq(BaseBlockwiseQuantization):
    def __init__(self, model, quant_config, input, padding_mask, config):
        super().__init__(model, quant_config, input, padding_mask, config)
        sp = self.special
        self.trans = sp.get('trans', True)
        self.trans_version = sp.get('trans_version', 'v2')
        self.weight_clip = sp.get('weight_clip', False)
        self.clip_sym = sp.get('clip_sym', self.wquantizer.sym)
        self.clip_version = sp.get('clip_version', 'v1')
        if self.weight_clip and self.clip_version == 'v2':     # base_blockwise_quantization.py:228-230
            assert self.wquantizer.calib_algo == 'learnable', "clip_version v2 needs weight calib_algo: learnable"
        self.save_clip = sp.get('save_clip', False)
        self.weight_clips = {}  # block -> {'<layer>.weight_quantizer.upbound_factor' | 'lowbound_factor': cpu tensor | None}
        self.true_sequential = False
        self.awq_bs = sp.get('awq_bs', None)
        self.best_index = {}   # (block, first layer name) -> selected candidate, for parity reports
        self.act_scales = {}   # '<prefix>.<block>.<layer>' -> applied scale (awq.py:365-368, kept whether or not save_scale)
        self.clips = {}        # (block, layer name) -> (best_max_val, best_min_val) of AutoClip

    @torch.no_grad()
    def block_transform(self, block, input_feat, block_kwargs):   # awq.py:280-298
        if self.trans:
            super().block_transform(block, input_feat, block_kwargs)
        if self.weight_clip:
            from .autoclip import auto_clip_block
            rec, factors = {}, {}
            auto_clip_block(self.model.get_block_linears(block), input_feat, self.wquantizer, self.clip_sym,
                            n_sample_token=(self.config or {}).get('calib', {}).get('seq_len', None)
                            if isinstance(self.config, dict) else None, record=rec, clip_version=self.clip_version,
                            aquantizer=None if self.w_only else self.aquantizer, padding_mask=self.padding_mask,
                            factors=factors)
            for n, v in rec.items():
                self.clips[(self.block_idx, n)] = v
            if self.save_clip:                                    # auto_clip.py:219-229
                for n, (up, low) in factors.items():
                    d = self.weight_clips.setdefault(self.block_idx, {})
                    d[f'{n}.weight_quantizer.upbound_factor'] = up.cpu()
                    d[f'{n}.weight_quantizer.lowbound_factor'] = None if low is None else low.cpu()

    @torch.no_grad()
    def subset_transform(self, subset, input_feat, subset_kwargs):   # awq.py:300-372
        layers_dict, prev_op = subset['layers'], subset['prev_op']
        if not subset.get('do_trans', True):
            return
        if len(prev_op) == 0 or prev_op[0] is None:
            return
        p = prev_op[0]
        layers = list(layers_dict.values())
        if isinstance(p, nn.Linear) and p.out_features not in (layers[0].in_features, layers[0].in_features * 2,
                                                               layers[0].in_features * 3):
            return   # GQA v -> o: "Cannot apply scale" unless do_gqa_trans (not covered)
        xs = input_feat[subset['input'][0]]
        inspect = subset['inspect']
        own = self.special.get('own_gemm', True)
        if len(xs) != 1 or self.padding_mask or self.awq_bs is not None:
            # several batches / awq_bs chunks / padding masks: the reference loop as written (awq.py:179-278)
            scale, best = lcawq.awq_search_scale_general(xs, layers, inspect, self.wquantizer, subset_kwargs, self.trans_version,
                                                         aquantizer=None if self.w_only else self.aquantizer,
                                                         padding_mask=self.padding_mask, awq_bs=self.awq_bs,
                                                         n_samples=self.n_samples, own_gemm=own)
        else:
            x = xs[0]
            kwargs = subset_kwargs[0] if isinstance(subset_kwargs, list) and subset_kwargs else (subset_kwargs or {})
            if isinstance(inspect, nn.Linear) and len(layers) == 1 and inspect is layers[0] and inspect.bias is None \
                    and self.w_only and x.dtype in (torch.bfloat16, torch.float16):
                scale, best, _ = lcawq.awq_search_scale(x, inspect.weight.data, self.wquantizer, self.trans_version)
            else:
                scale, best, _ = lcawq.awq_search_scale_module(x, layers, inspect, self.wquantizer, kwargs, self.trans_version,
                                                               own_gemm=own, aquantizer=None if self.w_only else self.aquantizer)
        self.best_index[(self.block_idx, next(iter(layers_dict)))] = best
        prefix = getattr(self.model, 'block_name_prefix', 'model.layers')
        for n in layers_dict:
            self.act_scales[f'{prefix}.{self.block_idx}.{n}'] = scale
        self.apply_scale(scale, prev_op, layers)
        self.update_input_feat(scale, input_feat, layers_dict)
