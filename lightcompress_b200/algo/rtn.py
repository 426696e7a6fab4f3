"""RTN (rtn.py:10-28): no calibration; quantization is the deploy-time w_qdq / a_qdq (base_bq.py:46-82)."""
import torch


@torch.no_grad()
def rtn_quantize(weight, wquantizer):
    """w_qdq: fake-quantised weight of one linear."""
    return wquantizer.fake_quant_weight_dynamic(weight)


@torch.no_grad()
def rtn_real_quantize(weight, wquantizer):
    """w_q: (codes, scales, zeros|None) for export."""
    return wquantizer.real_quant_weight_dynamic(weight)


@torch.no_grad()
def rtn_real_quantize_sharded(named_weights, wquantizer, quantize=None, gather=True):
    """Real-quant export of many independent weights (MoE experts, the linears of a block) sharded over the ranks
    (DeepSeek-V2-Lite, 64 routed + 2 shared experts x 3 linears per MoE block -> weight i on rank
    i mod world). Every rank quantizes its own weights with the fused kernel; with gather=True the owner broadcasts
    (codes, scales, zeros) so that every rank ends with the full dict (checkpoint writers usually want gather=False on
    all ranks but 0 — then only the local shard is returned). No collective touches the data path of the kernel.

    named_weights: ordered dict name -> tensor (same order on every rank). Returns dict name -> (codes, scales, zeros|None).
    `quantize` defaults to wquantizer.real_quant_weight_dynamic (injectable for the CPU tests of the host logic)."""
    import torch.distributed as dist

    from . import dist as lcdist
    rank, ws = lcdist.world()
    names = list(named_weights)
    out = {}
    mine = [n for i, n in enumerate(names) if i % ws == rank]
    if quantize is None and hasattr(wquantizer, 'real_quant_weights_dynamic'):
        # FP8 per_block experts: 64 weights per launch (lcb_fakequant_block_fp8_multi); otherwise the same per-tensor calls
        out.update(zip(mine, wquantizer.real_quant_weights_dynamic([named_weights[n] for n in mine])))
    else:
        quantize = quantize or wquantizer.real_quant_weight_dynamic
        for n in mine:
            out[n] = quantize(named_weights[n])
    if ws == 1 or not gather:
        return out
    for i, n in enumerate(names):
        owner = i % ws
        meta = [None]
        if owner == rank:
            c, s, z = out[n]
            meta[0] = [(tuple(t.shape), t.dtype) if t is not None else None for t in (c, s, z)]
        dist.broadcast_object_list(meta, src=owner)
        parts = []
        for j, m in enumerate(meta[0]):
            if m is None:
                parts.append(None)
                continue
            shape, dtype = m
            dev = named_weights[n].device
            t = out[n][j] if owner == rank else torch.empty(shape, dtype=dtype, device=dev)
            # fp8 tensors travel as bytes (NCCL / gloo have no fp8 type)
            wire = t.view(torch.uint8) if t.dtype in (torch.float8_e4m3fn, torch.float8_e5m2) else t
            wire = wire.contiguous()
            dist.broadcast(wire, src=owner)
            parts.append(wire.view(dtype) if wire.dtype != dtype else wire)
        out[n] = tuple(parts)
    return out
