"""Block loop and plug-in registry: the standalone counterpart of blockwise_optimization.py:8-114 and
base_blockwise_quantization.py:41-1029, reduced to what RTN / GPTQ / Awq / SmoothQuant need.

Same plug-in contract as the reference: `ALGO_REGISTRY[name](model, quant_config, input, padding_mask, config)`, then
`.run_block_loop()` and `.deploy('fake_quant')`. `model` is any object with the reference adapter protocol
(`get_blocks()`, `get_block_linears(block)`, `get_subsets_in_block(block)`): the reference's own `llmc.models.*` classes
qualify, and `lightcompress_b200.algo.adapters` has minimal Llama / OPT ones. All tensors stay on the GPU (the
reference shuttles cached inputs and state dicts through host memory: blockwise_optimization.py:53-61, awq.py:199).
"""
import functools
from collections import defaultdict

import torch
import torch.nn as nn

from .. import ops
from ..quant import FloatQuantizer, IntegerQuantizer


class Register(dict):
    """registry_factory.py:1-49: decorator registry keyed by class __name__."""

    def __call__(self, target):
        self[target.__name__] = target
        return target


ALGO_REGISTRY = Register()


def _vllm_cls():
    from .realquant import VllmRealQuantLinear
    return VllmRealQuantLinear


class _ChildCache(object):
    """Outputs of a block's modules (norms, attention, MLP, and the modules inside them), kept per forward chunk across
    the passes that `true_sequential` + `quant_out` make over one block (base_bq.py:438-449, 505-526: up to five forwards
    of the same calibration data through the same block). A module whose parameters did not change since its outputs
    were recorded, and before which — in execution order, at every enclosing level — every module was itself reused, sees
    bit-identical inputs and would recompute bit-identical outputs: its forward is replaced by a lookup for that pass.
    The first module of a level that cannot be reused is entered and the same rule applies to ITS children (its input is
    still identical); everything after it is recomputed and re-recorded. No model knowledge: the modules, their order
    and their outputs are whatever the block executes. Outputs are references to tensors the forward produces anyway;
    keeping them alive costs memory, bounded by `budget` bytes for the modules below the block's direct children."""

    def __init__(self, block, data, budget):
        self.block_id, self.data_id, self.n = id(block), id(data), len(data)
        self.order = {}          # parent name ('' = the block) -> child names (dotted) in execution order, as discovered
        self.out = {}            # name -> {chunk index: output}
        self.sig = {}            # name -> parameter signature the outputs were recorded under
        self.calls = {}          # name -> forwards seen in the current chunk (a module invoked twice is never reused)
        self.bytes = {}          # name -> bytes its recorded outputs keep alive
        self.own = {'': _ChildCache.own_signature(block)}   # name -> own_signature at recording time
        self.budget = budget
        self.deep = True

    @staticmethod
    def signature(mod):
        return tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in list(mod.parameters()) + list(mod.buffers()))

    @staticmethod
    def own_signature(mod):
        """the module's OWN tensors (not its descendants'): what the code BETWEEN its children may read"""
        return tuple((t.data_ptr(), t._version, tuple(t.shape))
                     for t in list(mod.parameters(recurse=False)) + list(mod.buffers(recurse=False)))

    @staticmethod
    def nbytes(o):
        if torch.is_tensor(o):
            return o.numel() * o.element_size()
        if isinstance(o, (tuple, list)):
            return sum(_ChildCache.nbytes(v) for v in o)
        return 0

    def drop(self, name):
        self.out.pop(name, None)
        self.bytes.pop(name, None)

    def over_budget(self):
        return sum(v for k, v in self.bytes.items() if '.' in k) > self.budget


class _ChunkedList(list):
    """The per-sample tensors of a calibration pass (what `self.input['data']` holds) together with the chunk tensors they
    are views of: a later pass over the same data feeds the chunks to the block directly instead of concatenating
    `forward_batch` samples again (32 `torch.cat` of 268 MB per Llama-3-8B block: 14 ms of device time)."""
    chunks = None     # list of [B, T, C] tensors; sample i is chunks[i // B][i % B : i % B + 1]
    B = 0


class _BlockPrefetcher(object):
    """Copies the next block's parameters and buffers to the device on a side stream from a worker thread while the
    current block is calibrated (base_bq.py:397 does the same copy synchronously at the block's turn)."""

    def __init__(self, dev):
        import threading
        self.dev = dev
        self.stream = torch.cuda.Stream(device=dev)
        self.thread = None
        self.block = None
        self.error = None
        self._threading = threading

    @staticmethod
    def on_device(block, dev):
        return all(t.device.type == dev.type for t in list(block.parameters()) + list(block.buffers()))

    def start(self, block):
        if self.on_device(block, self.dev):
            return
        self.block, self.error = block, None

        def work():
            try:
                with torch.cuda.device(self.stream.device), torch.cuda.stream(self.stream):
                    block.to(self.dev)
            except BaseException as e:     # noqa: BLE001  (re-raised on the calling thread by wait())
                self.error = e
        self.thread = self._threading.Thread(target=work, name='lcb-block-prefetch', daemon=True)
        self.thread.start()

    def wait(self, block):
        if self.thread is None or self.block is not block:
            return
        self.thread.join()
        self.thread = self.block = None
        if self.error is not None:
            raise self.error
        main = torch.cuda.current_stream(self.stream.device)
        main.wait_stream(self.stream)
        for t in list(block.parameters()) + list(block.buffers()):
            if t.is_cuda:
                t.data.record_stream(main)    # allocated on the side stream, used (and one day freed) under the main one

    def close(self):
        if self.thread is not None:
            self.thread.join()
            self.thread = self.block = None


class _StopForward(Exception):
    """Raised by the re-hook pre-hooks once every layer of the next subset has seen its input: the rest of the block
    forward (whose output nobody reads, base_bq.py:505-526) is not computed."""


def make_quantizer(cfg):
    """base_bq.py:152-179: quant_config['weight' | 'act'] -> quantizer instance."""
    cfg = dict(cfg)
    bit, sym, gran = cfg.pop('bit'), cfg.pop('symmetric'), cfg.pop('granularity')
    qtype = cfg.pop('quant_type', 'int-quant')
    cls = FloatQuantizer if qtype == 'float-quant' else IntegerQuantizer
    return cls(bit, sym, gran, **cfg)


class BaseBlockwiseQuantization(object):
    def __init__(self, model, quant_config, input, padding_mask, config):
        self.model = model
        self.blocks = model.get_blocks()
        self.quant_config = quant_config
        self.input = input
        self.padding_mask = padding_mask
        self.config = config
        self.data_free = not input
        self.block_idx = None
        self.dev = torch.device('cuda')
        self.wquantizer = make_quantizer(quant_config['weight'])
        self.w_only = not quant_config.get('act')
        self.aquantizer = None if self.w_only else make_quantizer(quant_config['act'])
        self.act_static = (not self.w_only) and quant_config['act'].get('static', False)
        if self.act_static:
            raise NotImplementedError('static activation quantization is not on the CUDA hot path')
        self.special = quant_config.get('special', {}) or {}
        self.true_sequential = self.special.get('true_sequential', False)
        self.quant_out = quant_config.get('quant_out', False)
        # per-sample block forwards are launch-bound on a B200 (≈1 ms of host work against ≈0.3-0.6 ms of kernels for a
        # 2048-token sample): with capture-safe hooks the forward is recorded once per pass and replayed per sample
        self.use_graph = self.special.get('cuda_graph', True)
        # ... and several calibration samples ride one block forward when the hooks only accumulate sums over tokens
        # (GPTQ: H is a sum, nsamples counts the samples of a batch like gptq.py:258, 281-285) or when there is no hook
        self.forward_batch = int(self.special.get('forward_batch', 16))
        self._capturing = False
        self._on_replay = []
        self._uniform = None
        self._child_cache = False        # _ChildCache of the block being calibrated (False: none yet)
        self._hooked_modules = set()     # modules whose input hooks must fire in the running pass
        self._done_q = {}                # module -> its fake-quantized weight, computed once per block
        if self.input:
            for kw in self.input['kwargs']:   # blockwise_optimization.py:20-25
                kw.pop('use_cache', None)
                for k in ('past_key_value', 'past_key_values'):
                    if k in kw:
                        kw[k] = None
            self.n_samples = sum(d.shape[0] for d in self.input['data'])

    # ---- closures bound into modules at deploy time (base_bq.py:46-82) ------------------------------------------------
    def w_qdq(self, module, wquantizer):
        args = {}
        lo, up = getattr(module, 'buf_lowbound_factor', None), getattr(module, 'buf_upbound_factor', None)
        if lo is not None or up is not None:   # AWQ clip v2 / learnable ranges (base_bq.py:46-52)
            args = {'lowbound_factor': lo, 'upbound_factor': up}
        return wquantizer.fake_quant_weight_dynamic(module.weight.data, args)

    def w_q(self, module, wquantizer):
        return wquantizer.real_quant_weight_dynamic(module.weight.data)

    def a_qdq(self, act, module, aquantizer):
        return aquantizer.fake_quant_act_dynamic(act)

    # ---- block loop (blockwise_optimization.py:31-51, base_bq.py:367-452) ----------------------------------------------
    def run_block_loop(self):
        """blockwise_optimization.py:31-51. The reference moves a block to the GPU when its turn comes (base_bq.py:397) and
        back afterwards (:418); here the weights stay on the device once they are there (180 GB of HBM3e), and while block i
        is calibrated block i + 1 is copied in by a worker thread on a side stream (`special.prefetch_blocks`, default on; a
        host-to-device copy from pageable memory blocks its calling thread, hence the thread). A block that already lives on
        the device costs nothing."""
        blocks = list(self.blocks)
        pre = _BlockPrefetcher(self.dev) if self.special.get('prefetch_blocks', True) and self.dev.type == 'cuda' else None
        try:
            for i, block in enumerate(blocks):
                self.block_idx = i
                if pre is not None:
                    pre.wait(block)                       # no-op unless a copy of this block is in flight
                    if i + 1 < len(blocks):
                        pre.start(blocks[i + 1])
                self.block_opt(block)
        finally:
            if pre is not None:
                pre.close()

    @torch.no_grad()
    def block_forward(self, block, input_data=None, may_stop=False):
        data = self.input['data'] if input_data is None else input_data
        dev = next(block.parameters()).device
        for kw in self.input['kwargs']:
            for k, v in list(kw.items()):
                if torch.is_tensor(v):
                    kw[k] = v.to(dev)
                elif isinstance(v, tuple):
                    kw[k] = tuple(t.to(dev) if torch.is_tensor(t) else t for t in v)
        if self.graph_safe_hooks and dev.type == 'cuda' and len(data) >= 2 and self._uniform_samples(data, dev):
            if self.forward_batch > 1:
                if input_data is None and not isinstance(data, _ChunkedList):
                    # the first block's inputs arrive as separate samples: concatenate the chunks ONCE for all its passes
                    B = self.forward_batch
                    cl = _ChunkedList()
                    cl.chunks, cl.B = [], B
                    for i in range(0, len(data), B):
                        part = data[i:i + B]
                        x = part[0] if len(part) == 1 else torch.cat(part, 0)
                        cl.chunks.append(x)
                        cl.extend(x.split([p.shape[0] for p in part], 0))
                    data = self.input['data'] = cl
                return self._block_forward_batched(block, data, may_stop)
            if self.use_graph and len(data) >= 4:
                return self._block_forward_graph(block, data, may_stop)
        out = []
        for i, x in enumerate(data):
            kw = self.input['kwargs'][i]
            try:
                y = block(x.to(dev), **kw)
            except _StopForward:
                if not may_stop:
                    raise
                continue
            out.append(y[0] if isinstance(y, tuple) else y)
        return out

    graph_safe_hooks = False   # the default hook keeps references to the inputs: not replayable

    def _uniform_samples(self, data, dev):
        """True when every sample has one shape and the same keyword tensors (one recorded forward serves them all)."""
        if self._uniform is None:
            kws = self.input['kwargs']

            def same(a, b):
                if torch.is_tensor(a) and torch.is_tensor(b):
                    return a is b or (a.shape == b.shape and a.dtype == b.dtype and a.device == b.device
                                      and bool(torch.equal(a, b)))
                if isinstance(a, (tuple, list)) and isinstance(b, (tuple, list)) and len(a) == len(b):
                    return all(same(u, v) for u, v in zip(a, b))
                return (a is b) or (not torch.is_tensor(a) and not torch.is_tensor(b) and a == b)
            self._uniform = all(kw.keys() == kws[0].keys() and all(same(kw[k], kws[0][k]) for k in kw) for kw in kws[1:])
        return self._uniform and all(x.shape == data[0].shape and x.dtype == data[0].dtype and x.device == dev
                                     for x in data)

    def _block_forward_batched(self, block, data, may_stop):
        """`forward_batch` samples per block forward: the same per-token arithmetic in fewer, larger GEMMs (the keyword
        tensors — position ids / embeddings, masks — are identical for every sample, so they broadcast over the batch).
        Outputs are handed back per sample (views), in order. Children of the block whose outputs an earlier pass over
        the same data already produced under the same parameters are looked up instead of recomputed (_ChildCache)."""
        kw = self.input['kwargs'][0]
        out = []
        B = self.forward_batch
        n_chunks = (len(data) + B - 1) // B
        cache = self._child_cache if self.special.get('reuse_child_outputs', True) else None
        if cache is not None and (cache is False or cache.block_id != id(block) or cache.data_id != id(data)
                                  or cache.n != len(data)):
            cache = self._child_cache = _ChildCache(block, data, int(self.special.get('reuse_budget_gb', 64)) << 30)
            cache.deep = self.special.get('reuse_child_outputs', True) != 'direct'

        handles, stubbed = [], []
        state = {'chunk': 0}
        if cache is not None:
            modules = dict(block.named_modules())
            modules.pop('', None)
            hooked = self._hooked_modules
            reuse = set()

            def plan(parent):
                """stub the reusable prefix of `parent`'s children; enter the first one that is not reusable"""
                pmod = block if parent == '' else modules.get(parent)
                if pmod is None or cache.own.get(parent) != _ChildCache.own_signature(pmod):
                    return     # the parent's own tensors changed: what flows between its children may have changed too
                for name in cache.order.get(parent, ()):
                    mod = modules.get(name)
                    ok = (mod is not None and len(cache.out.get(name, {})) == n_chunks
                          and cache.sig.get(name) == _ChildCache.signature(mod)
                          and not any(m in hooked for m in mod.modules()))
                    if ok:
                        reuse.add(name)
                        continue
                    if mod is not None and cache.deep:
                        plan(name)
                    break
            plan('')

            def inside_reused(name):
                return any(name.startswith(r + '.') for r in reuse)
            for name in list(cache.out):
                if name not in reuse and not inside_reused(name):
                    cache.drop(name)           # recomputed (or not reached) in this pass: stale from here on
            for name in reuse:
                def lookup(*a, _o=cache.out[name], **k):
                    return _o[state['chunk']]
                modules[name].forward = lookup
                stubbed.append(modules[name])

            def pre(mod, args, name):
                parent = name.rpartition('.')[0]
                lst = cache.order.setdefault(parent, [])
                if name not in lst:
                    lst.append(name)
                cache.calls[name] = cache.calls.get(name, 0) + 1
                cache.own[name] = _ChildCache.own_signature(mod)    # at ENTRY: a pass may stop inside this module

            def post(mod, args, output, name):
                if cache.calls.get(name, 0) == 1 and (cache.deep or '.' not in name):
                    cache.out.setdefault(name, {})[state['chunk']] = output
                    cache.sig[name] = _ChildCache.signature(mod)
                    cache.bytes[name] = cache.bytes.get(name, 0) + _ChildCache.nbytes(output)
                else:
                    cache.drop(name)
            for name, mod in modules.items():
                if name in reuse or inside_reused(name) or not (cache.deep or '.' not in name):
                    continue
                handles.append(mod.register_forward_pre_hook(functools.partial(pre, name=name)))
                handles.append(mod.register_forward_hook(functools.partial(post, name=name)))
        chunks = data.chunks if isinstance(data, _ChunkedList) and data.B == B and data.chunks is not None else None
        y_chunks = []
        try:
            for ci, i in enumerate(range(0, len(data), B)):
                part = data[i:i + B]
                if chunks is not None:
                    x = chunks[ci]
                else:
                    x = part[0] if len(part) == 1 else torch.cat(part, 0)
                state['chunk'] = ci
                if cache is not None:
                    cache.calls = {}
                try:
                    y = block(x, **kw)
                except _StopForward:
                    if not may_stop:
                        raise
                    continue
                y = y[0] if isinstance(y, tuple) else y
                out.extend(y.split([p.shape[0] for p in part], 0))
                y_chunks.append(y)
        finally:
            for h in handles:
                h.remove()
            for child in stubbed:
                child.__dict__.pop('forward', None)
            if cache is not None and cache.deep and cache.over_budget():
                cache.deep = False             # too much memory kept alive: keep only the block's direct children
                for name in [k for k in cache.out if '.' in k]:
                    cache.drop(name)
        if len(y_chunks) == n_chunks and len(out) == len(data):   # a complete pass: its outputs remember their chunks
            res = _ChunkedList(out)
            res.chunks, res.B = y_chunks, B
            return res
        return out

    def _block_forward_graph(self, block, data, may_stop):
        """Sample 0 runs eagerly (real work, and the warm-up a capture needs); the forward of sample 1 is recorded —
        hooks included — and replayed for samples 1.. from one static input buffer. Same kernels, same order, same
        results as the eager loop."""
        kw = self.input['kwargs'][0]

        def call(x):
            y = block(x, **kw)
            return y[0] if isinstance(y, tuple) else y

        out = []
        try:
            out.append(call(data[0]))
        except _StopForward:
            if not may_stop:
                raise
        static_x = data[1].clone()
        static_y = None
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        self._on_replay = []
        try:
            with torch.cuda.stream(side):
                g.capture_begin()
                self._capturing = True
                try:
                    static_y = call(static_x)
                except _StopForward:
                    if not may_stop:
                        raise
                finally:
                    self._capturing = False
                    g.capture_end()
        finally:
            self._capturing = False
        torch.cuda.current_stream().wait_stream(side)
        for x in data[1:]:
            static_x.copy_(x)
            g.replay()
            for note in self._on_replay:
                note()
            if static_y is not None:
                out.append(static_y.clone())
        self._on_replay = []
        return out

    def cache_input_hook(self, m, x, y, name, feat_dict):
        inp = x[0].detach()
        feat_dict[name].append(inp.unsqueeze(0) if inp.dim() == 2 else inp)

    def register_hooks(self, modules, feat_dict):
        return [m.register_forward_hook(functools.partial(self.cache_input_hook, name=n, feat_dict=feat_dict))
                for n, m in modules.items()]

    def block_init(self, block):
        pass

    def subset_init(self, subset):
        pass

    def block_opt(self, block):
        block = block.to(self.dev)
        named_linears = self.model.get_block_linears(block)
        input_feat = defaultdict(list)
        self._child_cache, self._done_q = False, {}
        if not self.data_free:
            self.block_init(block)
            if self.quant_out and self.true_sequential:
                # the float output is discarded (base_bq.py:423-426) and every subset after the first re-collects its
                # inputs once its predecessors are quantized (:505-526): the first pass only has to reach subset 0
                subsets = self.model.get_subsets_in_block(block)
                input_feat.update(self._collect_inputs(block, subsets[0]['layers']))
            else:
                handles = self.register_hooks(named_linears, input_feat)
                out = self.block_forward(block)
                if not self.quant_out:
                    self.input['data'] = out     # next block sees the float output (base_bq.py:423-428)
                del out
                for h in handles:
                    h.remove()
            self.block_transform(block, input_feat, self.input['kwargs'])
            if self.quant_out:               # base_bq.py:438-449: the next block sees the fake-quantized block's output
                self.input['data'] = self._forward_fake_quant(block, named_linears)
            self._child_cache, self._done_q = False, {}
        else:
            self.block_transform(block, input_feat, None)

    def _forward_fake_quant(self, block, linears):
        """One block forward with every linear fake-quantized (FakeQuantLinear: w_qdq weight, a_qdq input when w+a),
        the float weights restored afterwards."""
        mdt = getattr(self, 'model_dtype', None)
        saved = {n: m.weight.data for n, m in linears.items()}
        handles = []
        for n, m in linears.items():
            if m not in self._done_q:
                self._done_q[m] = self.w_qdq(m, self.wquantizer).to(mdt or m.weight.dtype)
            m.weight.data = self._done_q[m]
            if not self.w_only:
                handles.append(m.register_forward_pre_hook(self._act_hook))
        hooks, self.graph_safe_hooks = self.graph_safe_hooks, self.w_only
        try:
            out = self.block_forward(block)
        finally:
            self.graph_safe_hooks = hooks
            for h in handles:
                h.remove()
            for n, m in linears.items():
                m.weight.data = saved[n]
        return out

    def block_transform(self, block, input_feat, block_kwargs):
        subsets = self.model.get_subsets_in_block(block)
        self._done_layers = {}
        for index, subset in enumerate(subsets):
            if subset['has_kwargs']:
                if 'sub_keys' in subset:
                    subset_kwargs = [{k: kw[v] for k, v in subset['sub_keys'].items()} for kw in block_kwargs]
                else:
                    subset_kwargs = block_kwargs
            else:
                subset_kwargs = {}
            self.subset_transform(subset, input_feat, subset_kwargs)
            if self.true_sequential and index != len(subsets) - 1:
                input_feat.update(self.rehook_next_subset(block, subset, subsets[index + 1]))

    def rehook_next_subset(self, block, subset, next_subset):
        """base_bq.py:505-526: the transformed subset runs fake-quantised while the next subset's inputs are re-collected."""
        self.subset_init(next_subset)
        self._done_layers.update(subset['layers'])   # every subset transformed so far stays fake-quantised (FakeQuantLinear)
        saved = {n: m.weight.data for n, m in self._done_layers.items()}
        mdt = getattr(self, 'model_dtype', None)
        handles = []
        for n, m in self._done_layers.items():
            if m not in self._done_q:    # FakeQuantLinear caches tmp_weight on its first forward too (module_utils.py:722-725)
                self._done_q[m] = self.w_qdq(m, self.wquantizer).to(mdt or m.weight.dtype)
            m.weight.data = self._done_q[m]
            if not self.w_only:                      # FakeQuantLinear.forward applies a_qdq too (module_utils.py:712-737)
                handles.append(m.register_forward_pre_hook(self._act_hook))
        hooks, self.graph_safe_hooks = self.graph_safe_hooks, self.graph_safe_hooks and self.w_only
        try:
            feat = self._collect_inputs(block, next_subset['layers'])
        finally:
            self.graph_safe_hooks = hooks
            for h in handles:
                h.remove()
            for n, m in self._done_layers.items():
                m.weight.data = saved[n]
        return feat

    def _collect_inputs(self, block, layers):
        """Run the block until every module of `layers` has seen its input (cache_input_hook on each), and no further:
        pre-hooks — the input is all the hook needs — so neither these modules' own products nor anything after them is
        computed; the forward is cut as soon as the last of them has been reached."""
        feat = defaultdict(list)
        pending = set()
        # only the layers whose hook does something (GPTQ: the ONE linear that owns the Hessian its subset shares): the
        # forward stops in front of the first of them that is the last — the products of the linears before it (q / k
        # in front of v, gate in front of up: 40 of 330 ms per Llama-3-8B block) are never formed
        layers = {n: m for n, m in layers.items() if self.needs_input(n)}
        if not layers:
            return feat

        def pre_hook(m, x, name):
            self.cache_input_hook(m, x, None, name, feat)
            pending.discard(name)
            if not pending:
                pending.update(layers)
                raise _StopForward()

        pending.update(layers)
        handles = [m.register_forward_pre_hook(functools.partial(pre_hook, name=n)) for n, m in layers.items()]
        self._hooked_modules = set(layers.values())
        try:
            self.block_forward(block, may_stop=True)
        finally:
            self._hooked_modules = set()
            for h in handles:
                h.remove()
        return feat

    def subset_transform(self, subset, input_feat, subset_kwargs):
        pass

    def needs_input(self, name):
        """Does `cache_input_hook` of this layer consume its input? (False: the hook is a no-op, nothing waits for it.)"""
        return True

    # ---- scale application (base_bq.py:597-778, 877-897) --------------------------------------------------------------
    @torch.no_grad()
    def apply_scale(self, scales, prev_op, layers):
        assert len(prev_op) == 1, 'Only support single prev_op.'
        p = prev_op[0]
        if isinstance(p, nn.Linear):
            assert len(layers) == 1
            self.scale_fc_fc(p, layers[0], scales)
        elif hasattr(p, 'weight') and p.weight.dim() == 1:   # any LayerNorm / RMSNorm flavour
            self.scale_ln_fcs(p, layers, scales)
        else:
            raise NotImplementedError(f'prev_op {type(p)} not supported yet!')

    @torch.no_grad()
    def scale_ln_fcs(self, ln, fcs, scales):
        s = scales.to(device=ln.weight.device, dtype=ln.weight.dtype)
        ln.weight.div_(s)
        if getattr(ln, 'bias', None) is not None:
            ln.bias.div_(s)
        for fc in fcs:
            ops.scale_cols(fc.weight.data, s.to(fc.weight.dtype), out=fc.weight.data)   # fc.weight *= s[None, :]

    @torch.no_grad()
    def scale_fc_fc(self, fc1, fc2, scales):
        s = scales.to(fc1.weight.device)
        if fc1.out_features == fc2.in_features * 2:
            half = fc1.weight.data.shape[0] // 2
            fc1.weight.data[half:].div_(s.view(-1, 1))
            if fc1.bias is not None:
                fc1.bias.data[half:].div_(s.view(-1))
        elif fc1.out_features == fc2.in_features:
            if fc1.bias is not None:
                fc1.bias.div_(s.view(-1))
            fc1.weight.div_(s.view(-1, 1))
        else:
            raise NotImplementedError('Can not scale this fc-fc (fused qkv / GQA repeat are not covered).')
        ops.scale_cols(fc2.weight.data, s.to(fc2.weight.dtype), out=fc2.weight.data)

    @torch.no_grad()
    def update_input_feat(self, scale, input_feat, layers_dict):
        for name in layers_dict:
            input_feat[name] = [ops.scale_cols(x, scale.to(x.dtype), div=True) for x in input_feat[name]]

    # ---- deploy (base_bq.py:933-962) ------------------------------------------------------------------------------------
    @torch.no_grad()
    def deploy(self, quant_format='fake_quant'):
        """'fake_quant': every block linear keeps its class and gets the fake-quantized weight (+ the activation
        fake-quant in front when w+a) — what EffcientFakeQuantLinear computes. '<backend>_quant' (vllm / lightllm / sgl /
        lightx2v / autoawq / mlcllm): every block linear is REPLACED by the backend's real-quant module holding codes,
        scales (and zero-points), quantized by `w_q` and packed on the GPU (algo/realquant.py)."""
        from .realquant import REALQUANT_LINEAR_MAP
        if quant_format == 'origin_float':
            return
        if quant_format in REALQUANT_LINEAR_MAP:
            return self._deploy_real(REALQUANT_LINEAR_MAP[quant_format])
        if quant_format not in ('fake_quant', 'fake_quant_wo_kv'):
            raise NotImplementedError(f"Quant format '{quant_format}' is not implemented.")
        for block in self.blocks:
            block.to(self.dev)
            linears = list(self.model.get_block_linears(block).values())
            if type(self).w_qdq is BaseBlockwiseQuantization.w_qdq and not any(
                    hasattr(m, 'buf_upbound_factor') or hasattr(m, 'buf_lowbound_factor') for m in linears):
                # plain dynamic w_qdq: all linears of the block in one launch (lcb_fakequant_multi)
                for m, q in zip(linears, self.wquantizer.fake_quant_weights_dynamic([m.weight.data for m in linears])):
                    m.weight.data = q
            else:
                for m in linears:
                    m.weight.data = self.w_qdq(m, self.wquantizer).to(getattr(self, 'model_dtype', None) or m.weight.dtype)
            if not self.w_only:
                for m in linears:
                    m.register_forward_pre_hook(self._act_hook)

    def _deploy_real(self, cls):
        """model.replace_module_all(cls, {'w_q': ..., 'quant_config': ...}) (base_model.py:371-411)."""
        w_q = functools.partial(self.w_q, wquantizer=self.wquantizer)
        for block in self.blocks:
            block.to(self.dev)
            named = self.model.get_block_linears(block)
            plain = type(self).w_q is BaseBlockwiseQuantization.w_q and hasattr(self.wquantizer, 'real_quant_weights_dynamic')
            pre = {}
            if plain and issubclass(cls, _vllm_cls()):
                # dynamic w_q of every linear of the block at once: FP8 per_block expert matrices go 64 per launch
                mods = [m for m in named.values() if not getattr(m, 'no_quant', False)
                        and m.weight.dtype != torch.float8_e4m3fn]
                for m, res in zip(mods, self.wquantizer.real_quant_weights_dynamic([m.weight.data for m in mods])):
                    pre[m] = res
            for name, m in named.items():
                if getattr(m, 'no_quant', False):
                    continue
                new = cls.new(m, (lambda mod, r=pre.get(m): r) if m in pre else w_q, self.quant_config)
                parent_name, _, child = name.rpartition('.')
                parent = block.get_submodule(parent_name) if parent_name else block
                setattr(parent, child, new)

    def _act_hook(self, module, args):
        return (self.a_qdq(args[0], module, self.aquantizer),) + tuple(args[1:])


@ALGO_REGISTRY
class RTN(BaseBlockwiseQuantization):
    """rtn.py:10-28: nothing to calibrate; quantization happens at deploy."""

    def block_opt(self, block):
        pass


@ALGO_REGISTRY
class SmoothQuant(BaseBlockwiseQuantization):
    """smoothquant.py:14-79."""

    def __init__(self, model, quant_config, input, padding_mask, config):
        super().__init__(model, quant_config, input, padding_mask, config)
        self.alpha = self.special.get('alpha', 0.5)

    @torch.no_grad()
    def subset_transform(self, subset, input_feat, subset_kwargs):
        from . import smoothquant as sq
        prev_op = subset['prev_op']
        if not (len(prev_op) == 1 and hasattr(prev_op[0], 'weight') and prev_op[0].weight.dim() == 1):
            return   # filter_subset: LN -> FC subsets only (smoothquant.py:21-24)
        layers = list(subset['layers'].values())
        tensors = input_feat[subset['input'][0]]
        am = sq.ActMax(layers[0].in_features, layers[0].weight.device)
        for x in tensors:
            am.update(x)
        scale = sq.smoothquant_scale(am.result(), [m.weight.data for m in layers], self.alpha)
        self.apply_scale(scale, prev_op, layers)
