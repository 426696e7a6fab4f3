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
