"""Minimal model adapters with the reference's adapter protocol (llmc/models/base_model.py:22, llama.py, opt.py):
blocks, block linears, subsets, first-block input capture. Enough to drive the block loop on a HF model object."""
import torch
import torch.nn as nn


class _Stop(Exception):
    pass


class BaseAdapter(object):
    block_name_prefix = 'model.layers'

    def __init__(self, hf_model):
        self.model = hf_model
        self.blocks = self.find_blocks()
        self.first_block_input = None

    def get_blocks(self):
        return self.blocks

    def get_block_linears(self, block):   # base_model.py:361-366
        return {n: m for n, m in block.named_modules() if isinstance(m, nn.Linear)}

    @torch.no_grad()
    def collect_first_block_input(self, calib_data):
        """base_model.py:174-192 / 280-336 (Catcher): run the model until the first block, keep its args / kwargs."""
        data, kwargs = [], []
        first = self.blocks[0]

        def pre_hook(module, args, kw):
            data.append(args[0] if args else kw.pop('hidden_states'))
            kwargs.append(dict(kw))
            raise _Stop()

        h = first.register_forward_pre_hook(pre_hook, with_kwargs=True)
        for batch in calib_data:
            try:
                self.model(**batch)
            except _Stop:
                pass
        h.remove()
        self.first_block_input = {'data': data, 'kwargs': kwargs}
        return self.first_block_input

    def get_first_block_input(self):
        return self.first_block_input


class LlamaAdapter(BaseAdapter):
    """llmc/models/llama.py."""

    def find_blocks(self):
        return self.model.model.layers

    def get_subsets_in_block(self, block):
        a, m = block.self_attn, block.mlp
        return [
            {'layers': {'self_attn.q_proj': a.q_proj, 'self_attn.k_proj': a.k_proj, 'self_attn.v_proj': a.v_proj},
             'prev_op': [block.input_layernorm], 'input': ['self_attn.q_proj'], 'inspect': a, 'has_kwargs': True},
            {'layers': {'self_attn.o_proj': a.o_proj}, 'prev_op': [a.v_proj], 'input': ['self_attn.o_proj'],
             'inspect': a.o_proj, 'has_kwargs': False},
            {'layers': {'mlp.gate_proj': m.gate_proj, 'mlp.up_proj': m.up_proj},
             'prev_op': [block.post_attention_layernorm], 'input': ['mlp.gate_proj'], 'inspect': m, 'has_kwargs': False,
             'is_mlp': True},
            {'layers': {'mlp.down_proj': m.down_proj}, 'prev_op': [m.up_proj], 'input': ['mlp.down_proj'],
             'inspect': m.down_proj, 'has_kwargs': False, 'is_mlp': True},
        ]


class OptAdapter(BaseAdapter):
    """llmc/models/opt.py."""
    block_name_prefix = 'model.decoder.layers'

    def find_blocks(self):
        return self.model.model.decoder.layers

    def get_subsets_in_block(self, block):
        a = block.self_attn
        return [
            {'layers': {'self_attn.q_proj': a.q_proj, 'self_attn.k_proj': a.k_proj, 'self_attn.v_proj': a.v_proj},
             'prev_op': [block.self_attn_layer_norm], 'input': ['self_attn.q_proj'], 'inspect': a, 'has_kwargs': True},
            {'layers': {'self_attn.out_proj': a.out_proj}, 'prev_op': [a.v_proj], 'input': ['self_attn.out_proj'],
             'inspect': a.out_proj, 'has_kwargs': False},
            {'layers': {'fc1': block.fc1}, 'prev_op': [block.final_layer_norm], 'input': ['fc1'], 'inspect': block.fc1,
             'has_kwargs': False, 'is_mlp': True},
            {'layers': {'fc2': block.fc2}, 'prev_op': [block.fc1], 'input': ['fc2'], 'inspect': block.fc2,
             'has_kwargs': False, 'is_mlp': True, 'do_trans': False},
        ]
