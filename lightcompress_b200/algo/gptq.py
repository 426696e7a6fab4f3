"""GPTQ per-linear calibration on the GPU kernels: the compute of gptq.py:128-295 with explicit tensors.

    g = GPTQLayer(layer.weight, wquantizer, blocksize=128, percdamp=0.01)
    for x in calibration_inputs: g.add_batch(x)          # GPTQ.add_batch           (gptq.py:254-295)
    W_new = g.quantize()                                  # process_hessian_and_weights + weight_transform (:128-244)
    g.scales, g.zeros                                     # per-group qparams for deploy (update_model_qparams :398-409)
"""
import torch

from .. import ops
from . import dist as lcdist


class HessianState(object):
    """Running Hessian of one layer INPUT and its damped inverse factor. Linears that see the same input (q/k/v, gate/up:
    one `subset` of the reference, models/llama.py) may share one state: the reference accumulates and factorises an
    identical H once per linear (gptq.py:254-295, 128-176); sharing is the same arithmetic done once."""

    def __init__(self, columns, device):
        self.columns = columns
        # layer_init (gptq.py:297-309): H zeros, nsamples 0. Here S is the UN-NORMALISED sum until finalize().
        self.S = torch.zeros((columns, columns), dtype=torch.float32, device=device)
        self.nsamples = 0
        self.H = None
        self._factor = None     # (key, dead, perm, Hinv)

    @torch.no_grad()
    def add_batch(self, inp, count=True):
        """gptq.py:254-295 for Linear layers: `nsamples` counts batches (inp.shape[0]), not tokens.
        count=False: launch only (CUDA-graph capture of the block forward); the caller adds the batch count per replay."""
        if inp.dim() == 2:
            inp = inp.unsqueeze(0)
        if count:
            self.nsamples += inp.shape[0]
        ops.hessian_accum(self.S, inp)
        return inp.shape[0]

    @torch.no_grad()
    def finalize(self):
        """One all-reduce of (S, nsamples) over the ranks, then H = (2 / n) * sym(S)."""
        if self.H is None:
            self.H = lcdist.finalize_sharded_hessian(self.S, self.nsamples)   # nsamples stays this rank's own count
            self.S = None
        return self.H

    @torch.no_grad()
    def factor(self, percdamp, actorder, splits):
        """gptq.py:143-174 without the weight: dead columns, act-order permutation, damping, Cholesky-inverse upper
        factor (ops.cholinv_upper: SURVEY.md §8f N1; raises torch.linalg.LinAlgError like torch.linalg.cholesky).
        Returns (dead mask, perm | None, Hinv); computed once per (percdamp, actorder)."""
        key = (float(percdamp), bool(actorder), int(splits))
        if self._factor is not None and self._factor[0] == key:
            return self._factor[1:]
        H = self.finalize()
        dead = torch.diag(H) == 0
        H[dead, dead] = 1
        perm = None
        if actorder:
            perm = torch.argsort(torch.diag(H), descending=True)
            H = H[perm][:, perm]
        Hinv = ops.cholinv_upper(H.contiguous(), percdamp, splits=splits)
        self.H = None
        self._factor = (key, dead, perm, Hinv)
        return dead, perm, Hinv


class GPTQLayer(object):
    def __init__(self, weight, wquantizer, blocksize=128, percdamp=0.01, actorder=False, static_groups=False, hessian=None):
        if not weight.is_cuda:
            raise RuntimeError('GPTQLayer: weight must live on a CUDA device (no CPU path)')
        self.weight = weight
        self.wquantizer = wquantizer
        self.blocksize = blocksize
        self.percdamp = percdamp
        self.actorder = actorder
        self.static_groups = static_groups
        self.rows, self.columns = weight.shape[0], weight.shape[1]
        self.hessian = hessian if hessian is not None else HessianState(self.columns, weight.device)
        assert self.hessian.columns == self.columns
        self.scales = self.zeros = None
        self.losses = None
        self.perm = self.invperm = None
        self.chol_splits = 3   # bf16 pieces per fp32 operand in the factorisation GEMMs (3 = fp32-class accuracy)

    # the Hessian-side state, kept reachable under the reference's names
    S = property(lambda self: self.hessian.S)
    H = property(lambda self: self.hessian.H)
    nsamples = property(lambda self: self.hessian.nsamples)

    def add_batch(self, inp, count=True):
        return self.hessian.add_batch(inp, count)

    def finalize(self):
        return self.hessian.finalize()

    def _qrange(self):
        return self.wquantizer._qrange()

    @torch.no_grad()
    def process_hessian_and_weights(self):
        """gptq.py:128-176: dead columns, optional act-order permutation, damping, Cholesky inverse upper factor."""
        dead, self.perm, Hinv = self.hessian.factor(self.percdamp, self.actorder, self.chol_splits)
        W = self.weight.data.clone().float()
        W[:, dead] = 0
        if self.actorder:
            W = W[:, self.perm].contiguous()
            self.invperm = torch.argsort(self.perm)
        return W, Hinv

    @torch.no_grad()
    def quantize(self, static_scales=None, static_zeros=None):
        """Returns the error-compensated float weight `tmp` in fp32 like the reference (gptq.py:178-197: the layer keeps
        fp32 weights until deploy's w_qdq casts to the model dtype); per-group scales /
        zeros are left in self.scales / self.zeros as [C/g, R] fp32 (`groups` of the reference, :380-409)."""
        W, Hinv = self.process_hessian_and_weights()
        wq = self.wquantizer
        qmin, qmax = self._qrange()
        gs = wq.group_size if wq.granularity == 'per_group' else 0
        col_group = None
        if gs > 0 and not self.static_groups and static_scales is None and wq.calib_algo != 'minmax':
            raise NotImplementedError('GPTQ with on-the-fly group qparams fuses the min/max search into the column kernel; '
                                      f"calib_algo={wq.calib_algo!r} needs static_groups=True or per_channel")
        if gs > 0 and self.static_groups:
            if static_scales is None:   # search_group_qparams (gptq.py:380-395): qparams of the original weight IN THE
                # MODEL DTYPE (collect_block_qparams, base_bq.py:338-365), upcast exactly when they meet the fp32 column
                _, s, z, _, _ = wq.get_tensor_qparams(self.weight.data)
                static_scales = s.float().reshape(self.rows, -1).t().contiguous()
                static_zeros = None if wq.sym else z.float().reshape(self.rows, -1).t().contiguous()
            if self.actorder:
                col_group = (self.perm // gs).to(torch.int32)
        elif gs <= 0 and static_scales is None:   # per_channel: search_layer_qparams (gptq.py:368-378), model dtype
            _, s, z, _, _ = wq.get_tensor_qparams(self.weight.data)
            static_scales = s.float().reshape(1, self.rows)
            static_zeros = None if wq.sym else z.float().reshape(1, self.rows)
        tmp, losses, scales, zeros = ops.gptq_weight_transform(
            W, Hinv, self.blocksize, gs, qmin, qmax, wq.sym, wq.round_zp, static_scales=static_scales,
            static_zeros=static_zeros, col_group=col_group)
        rank, world = lcdist.world()
        if world > 1:
            # every rank factorises the same all-reduced H, but split-K partial sums land in L2 in a different order on
            # every GPU (fp32 reduce-adds, ~1e-7 apart), and GPTQ's error feedback turns that into other grid points:
            # rank 0's solution is the one every rank keeps (the reference's ranks agree because cuSOLVER is deterministic)
            import torch.distributed as dist
            for t in (tmp, scales, zeros):          # `losses` is per-rank bookkeeping
                if t is not None:
                    dist.broadcast(t, src=0)
        self.losses = losses
        self.scales, self.zeros = scales, zeros
        if self.actorder:
            tmp = tmp[:, self.invperm]
        return tmp.reshape(self.weight.shape)

    def buf_qparams(self):
        """Per-group qparams in the layout `FakeQuantLinear` buffers use: [R * C/g, 1] (merge_qparams, gptq.py:345-356)."""
        s = self.scales.t().reshape(-1, 1)
        z = self.zeros.t().reshape(-1, 1) if self.zeros is not None else None
        return s, z
