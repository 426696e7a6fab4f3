"""Drop-in mirror of `llmc.compression.quantization.quant` for the B200 hot path.

`IntegerQuantizer` / `FloatQuantizer` keep the reference's constructor kwargs, method names, argument
meaning, return conventions and error behaviour (quant.py:46-1221) but every tensor pass runs in the
hand-written sm_100a kernels behind the C ABI (include/lcb200.h). PyTorch is only used for device
memory, views and the stream handle. There is no CPU / eager fallback: tensors must live on a CUDA
device and the extension must be built.
"""
import torch

from . import _cabi

_DT = {torch.float32: _cabi.LCB_F32, torch.float16: _cabi.LCB_F16, torch.bfloat16: _cabi.LCB_BF16}
_RANK = {torch.bfloat16: 0, torch.float16: 0, torch.float32: 1}


def _promote(*dtypes):
    """torch type promotion restricted to {bf16, f16, f32} for tensors with dim >= 1."""
    kinds = set(dtypes)
    if len(kinds) == 1:
        return dtypes[0]
    return torch.float32  # any f32 operand, or a bf16/f16 mix, promotes to f32


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)


_cur_dev = torch.cuda.current_device


def _stream(index=None):
    """cudaStream_t of torch's current stream on device `index` (raw handle: no Stream object is built)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device() if index is None else index)
    return torch.cuda.current_stream(index).cuda_stream


class _NoCtx(object):
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NOCTX = _NoCtx()


def _on_device(dev):
    """Context that makes `dev` the current CUDA device; free when it already is."""
    if dev.index is None or dev.index == torch.cuda.current_device():
        return _NOCTX
    return torch.cuda.device(dev)


def _require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f'{what}: tensor is on {t.device}; lightcompress_b200 has no CPU path '
                           '(the sm_100a extension is the only implementation)')
    if t.dtype not in _DT:
        raise TypeError(f'{what}: unsupported dtype {t.dtype}')


class _Units:
    """How a tensor maps onto the [rows, cols] / unit_rows x unit_cols grid of lcb_fakequant."""
    __slots__ = ('rows', 'cols', 'ur', 'uc', 'n_units', 'qshape')


class BaseQuantizer(object):
    """Mirror of quant.py:46-658 (ctor kwargs = the YAML `quant.weight|act` schema)."""

    def __init__(self, bit, symmetric, granularity, **kwargs):
        self.bit = bit
        self.sym = symmetric
        self.granularity = granularity
        self.kwargs = kwargs
        self.calib_algo = kwargs.get('calib_algo', 'minmax')
        if granularity == 'per_group':
            self.group_size = kwargs['group_size']
        elif granularity == 'per_head':
            self.head_num = kwargs['head_num']
        elif granularity == 'per_block':
            assert self.calib_algo == 'minmax' and self.sym
            self.block_size = kwargs['block_size']
        if kwargs.get('ste', False) or kwargs.get('ste_all', False):
            raise NotImplementedError('straight-through estimators need autograd; not on the CUDA hot path')
        self.round_func = torch.round
        self.ste_all = False
        self.round_zp = kwargs.get('round_zp', True)
        self.mse_b_num = kwargs.get('mse_b_num', 1)
        self.maxshrink = kwargs.get('maxshrink', 0.8)
        self.mse_grid = kwargs.get('mse_grid', 100)
        self.bins = kwargs.get('bins', 2048)

    def __setattr__(self, name, value):
        # any public attribute (bit, sym, qmin, qmax, granularity, group_size, ...) may change what a call plan encodes
        object.__setattr__(self, name, value)
        if name[0] != '_':
            plans = self.__dict__.get('_plans')
            if plans:
                self._drop_plans()

    def _drop_plans(self):
        plans = self.__dict__.get('_plans')
        if plans:
            lib = _cabi._lib
            for p in plans.values():
                if lib is not None and p[0]:
                    lib.lcb_fakequant_plan_destroy(p[0])
            plans.clear()

    def __del__(self):
        try:
            self._drop_plans()
        except Exception:   # noqa: BLE001  (interpreter shutdown)
            pass

    def _make_plan(self, key, dev_index, u, spec):
        """Remember the argument bundle of a dynamic fake-quant call (lcb_fakequant_plan_*): the next call with a tensor of
        the same shape / dtype / device goes from Python to the launch in one 7-pointer ctypes call."""
        lib = _cabi._lib or _cabi.load()
        plans = self.__dict__.setdefault('_plans', {})
        if len(plans) >= 64:
            self._drop_plans()
        h = lib.lcb_fakequant_plan_create(u.rows, u.cols, u.ur, u.uc, spec, 1)
        if h:
            plans[key] = (h, dev_index, lib.lcb_fakequant_plan_run)

    # ---- layout ---------------------------------------------------------------------------------
    def _units(self, tensor):
        shape = tuple(tensor.shape)
        d = self.__dict__
        key = (shape, self.granularity, d.get('group_size'), d.get('block_size'), d.get('head_num'))
        cache = d.setdefault('_units_cache', {})
        u = cache.get(key)
        if u is None:
            if len(cache) > 256:
                cache.clear()
            u = cache[key] = self._units_uncached(tensor, shape)
        return u

    def _units_uncached(self, tensor, shape):
        u = _Units()
        g = self.granularity
        numel = tensor.numel()
        last = shape[-1] if len(shape) else 1
        if g == 'per_group' and last >= self.group_size:
            if last % self.group_size != 0:
                raise ValueError(f'Dimension {last} not divisible by group size {self.group_size}')
            u.rows, u.cols, u.ur, u.uc = numel // last, last, 1, self.group_size
            u.qshape = (numel // self.group_size, 1)
        elif g == 'per_head':
            u.rows, u.cols = self.head_num, numel // self.head_num
            u.ur, u.uc = 1, u.cols
            u.qshape = (self.head_num, 1)
        elif g == 'per_block':
            assert len(shape) == 2, 'per_block expects a 2-D weight'
            bs = self.block_size
            u.rows, u.cols, u.ur, u.uc = shape[0], shape[1], bs, bs
            u.qshape = (-(-shape[0] // bs), 1, -(-shape[1] // bs), 1)
        elif g == 'per_tensor':
            u.rows, u.cols = (numel // last, last) if numel else (0, 0)
            u.ur, u.uc = max(u.rows, 1), max(u.cols, 1)
            u.qshape = ()
        else:  # per_channel / per_token / short per_group rows: one unit per last-dim row
            u.rows, u.cols, u.ur, u.uc = (numel // last if last else 0), last, 1, max(last, 1)
            u.qshape = shape[:-1] + (1,)
        u.n_units = 1
        for d in u.qshape:
            u.n_units *= d
        return u

    def reshape_tensor(self, tensor, allow_padding=False):
        """quant.py:612-642 (views only; per_block pads like the reference)."""
        g = self.granularity
        if g == 'per_group':
            if tensor.shape[-1] >= self.group_size:
                if tensor.shape[-1] % self.group_size != 0:
                    raise ValueError(f'Dimension {tensor.shape[-1]} not divisible by group size {self.group_size}')
                return tensor.reshape(-1, self.group_size)
            return tensor
        if g == 'per_head':
            return tensor.reshape(self.head_num, -1)
        if g == 'per_block':
            bs = self.block_size
            m, n = tensor.shape
            mp, np_ = -(-m // bs) * bs, -(-n // bs) * bs
            if (mp, np_) != (m, n):
                t = tensor.new_zeros((mp, np_))
                t[:m, :n] = tensor
            else:
                t = tensor
            return t.view(-1, bs, np_ // bs, bs)
        return tensor

    def restore_tensor(self, tensor, shape):
        """quant.py:644-658."""
        if tensor.shape == shape:
            return tensor
        if self.granularity == 'per_block':
            return tensor.reshape(-1, tensor.shape[2] * tensor.shape[3])[:shape[0], :shape[1]]
        return tensor.reshape(shape)

    # ---- the one kernel call ------------------------------------------------------------------------
    def _spec(self, dtype, math_dtype, qparam_dtype, out_mode, qmin, qmax, osf=1.0, clip=1.0, qp_plus1=0):
        key = (dtype, math_dtype, qparam_dtype, out_mode, qmin, qmax, osf, clip, qp_plus1, self.sym, self.round_zp)
        cache = self.__dict__.setdefault('_spec_cache', {})
        s = cache.get(key)
        if s is None:
            if len(cache) > 256:
                cache.clear()
            s = cache[key] = self._spec_uncached(dtype, math_dtype, qparam_dtype, out_mode, qmin, qmax, osf, clip)
            s.qp_math_plus1 = qp_plus1
        return s

    def _spec_uncached(self, dtype, math_dtype, qparam_dtype, out_mode, qmin, qmax, osf=1.0, clip=1.0):
        s = _cabi.QuantSpec()
        s.dtype, s.math_dtype, s.qparam_dtype = _DT[dtype], _DT[math_dtype], _DT[qparam_dtype]
        s.kind = self._kind()
        s.sym = 1 if self.sym else 0
        s.round_zp = 1 if self.round_zp else 0
        s.out_mode = out_mode
        s.qmin, s.qmax = float(qmin), float(qmax)
        s.output_scale_factor = float(osf)
        s.clip_ratio = float(clip)
        return s

    def _run(self, x, rows, cols, ur, uc, spec, dynamic, out=None, scales=None, zeros=None):
        lib = _cabi._lib or _cabi.load()
        ws = None
        if dynamic and ur >= rows and uc >= cols:
            ws = torch.empty(int(lib.lcb_fakequant_workspace_bytes()), dtype=torch.uint8, device=x.device)
        st = lib.lcb_fakequant(
            x.data_ptr(), out.data_ptr() if out is not None else None,
            scales.data_ptr() if scales is not None else None,
            zeros.data_ptr() if zeros is not None else None,
            rows, cols, ur, uc, spec, 1 if dynamic else 0, ws.data_ptr() if ws is not None else None,
            _stream(x.device.index))
        if st:
            _cabi.check(st, 'lcb_fakequant')

    def _math_dtype_dynamic(self, dtype):
        return torch.float32 if self.granularity == 'per_block' else dtype

    def _qrange(self):
        # float(0-dim tensor) costs ~1 us each: cache until qmin / qmax are re-assigned
        c = self.__dict__.get('_qrange_cache')
        if c is None or c[0] is not self.qmin or c[1] is not self.qmax:
            c = self.__dict__['_qrange_cache'] = (self.qmin, self.qmax, float(self.qmin), float(self.qmax))
        return c[2], c[3]

    def _dynamic(self, tensor, out_mode, out_dtype, want_qparams, osf=1.0, clip=1.0, out=None, plan_key=None):
        """get_tensor_qparams [+ quant / quant_dequant] in one pass. Returns (out, scales, zeros)."""
        _require_cuda(tensor, 'dynamic quant')
        if self.calib_algo == 'hqq':   # every other value ranges by min/max here, as get_tensor_range does (quant.py:119-130)
            raise NotImplementedError(f'calib_algo={self.calib_algo!r} is not on the CUDA hot path')
        if self.calib_algo == 'mse':
            assert clip == 1.0
            return self._dynamic_mse(tensor, out_mode, out_dtype, want_qparams, osf)
        x = tensor.contiguous()
        u = self._units(x)
        md = self._math_dtype_dynamic(x.dtype)
        qmin, qmax = self._qrange()
        # per_tensor: the 0-dim min/max divided by a 0-dim *float* tensor (qmax - qmin with qmin = tensor(0.0), or a
        # float qmax) promotes to fp32 under torch's 0-dim rules -> fp32 scales / zeros (quant.py:552-556)
        qd = md
        if self.granularity == 'per_tensor' and (self._kind() != _cabi.KIND_INT or not self.sym):
            qd = torch.float32
        spec = self._spec(x.dtype, md, qd, out_mode, qmin, qmax, osf, clip, _DT[qd] + 1)
        if out_mode == _cabi.OUT_NONE:
            out = None
        elif out is not None:
            assert out.is_contiguous() and out.shape == x.shape and out.dtype == out_dtype and out.device == x.device
        elif out_dtype == x.dtype:
            out = torch.empty_like(x)
        else:
            out = torch.empty(x.shape, dtype=out_dtype, device=x.device)
        scales = zeros = None
        if want_qparams:
            scales = torch.empty(u.qshape, dtype=qd, device=x.device)
            if not self.sym and self._kind() == _cabi.KIND_INT:
                zeros = torch.empty(u.qshape, dtype=qd, device=x.device)
        with _on_device(x.device):
            self._run(x, u.rows, u.cols, u.ur, u.uc, spec, True, out, scales, zeros)
        if plan_key is not None and out_mode == _cabi.OUT_FAKE and not want_qparams and osf == 1.0 and clip == 1.0 \
                and not (u.ur >= u.rows and u.uc >= u.cols) and x.device.index is not None:
            self._make_plan(plan_key, x.device.index, u, spec)
        return out, scales, zeros

    def _dynamic_mse(self, tensor, out_mode, out_dtype, want_qparams, osf=1.0):
        """calib_algo='mse': searched fp32 range (lcb_mse_range) -> fp32 qparams -> the static kernel, whose math dtype
        promotes to fp32 exactly as the reference's quant_dequant does with fp32 scales (quant.py:690-717)."""
        x = tensor.contiguous()
        u = self._units(x)
        rng = self.get_mse_range(x)
        scales, zeros, qmax, qmin = self.get_qparams(rng, x.device)
        asym = not self.sym and self._kind() == _cabi.KIND_INT
        out = None
        if out_mode != _cabi.OUT_NONE:
            out = self._static(x, scales, zeros if asym else None, qmax, qmin, out_mode, out_dtype, osf, units=u)
        if not want_qparams:
            return out, None, None
        return out, scales.reshape(u.qshape), (zeros.reshape(u.qshape) if asym else None)

    def _static(self, tensor, scales, zeros, qmax, qmin, out_mode, out_dtype, osf=1.0, units=None):
        _require_cuda(tensor, 'static quant')
        x = tensor.contiguous()
        u = units or self._units(x)
        dts = [x.dtype]
        if torch.is_tensor(scales) and scales.dim() > 0:
            dts.append(scales.dtype)
        zt = torch.is_tensor(zeros) and zeros.numel() > 1
        if zt:
            dts.append(zeros.dtype)
        md = _promote(*dts)
        scales = torch.as_tensor(scales, device=x.device)
        if scales.numel() == 1 and u.n_units != 1:
            u = _Units()
            last = x.shape[-1]
            u.rows, u.cols = x.numel() // last, last
            u.ur, u.uc, u.n_units = u.rows, u.cols, 1
        if scales.numel() != u.n_units:
            raise ValueError(f'static qparams hold {scales.numel()} values, tensor has {u.n_units} units')
        sc = scales.to(device=x.device, dtype=torch.float32).contiguous()  # exact upcast; math dtype stays `md`
        zp = None
        if not self.sym and self._kind() == _cabi.KIND_INT and zeros is not None:
            zp = torch.as_tensor(zeros, device=x.device).to(dtype=torch.float32)
            if zp.numel() == 1 and u.n_units != 1:
                zp = zp.expand(u.n_units)
            zp = zp.contiguous()
        qmin_f = float(qmin) if qmin is not None else float(self.qmin)
        qmax_f = float(qmax) if qmax is not None else float(self.qmax)
        spec = self._spec(x.dtype, md, torch.float32, out_mode, qmin_f, qmax_f, osf)
        out = torch.empty(x.shape, dtype=out_dtype, device=x.device)
        with _on_device(x.device):
            self._run(x, u.rows, u.cols, u.ur, u.uc, spec, False, out, sc, zp)
        return out

    # ---- reference API ------------------------------------------------------------------------------
    def get_tensor_qparams(self, tensor, args={}):
        """quant.py:690-697 / :1043-1059 → (reshaped_tensor, scales, zeros, qmax, qmin)."""
        dev = tensor.device
        if self.calib_algo == 'learnable':   # range -> sigmoid bounds -> qparams (quant.py:119-130, 205-219)
            t = self.reshape_tensor(tensor)
            rng = self.get_learnable_range(tensor, **{k: v for k, v in args.items()
                                                      if k in ('lowbound_factor', 'upbound_factor')})
            scales, zeros, qmax, qmin = self.get_qparams(rng, dev)
            return t, scales, zeros, qmax, qmin
        _, scales, zeros, = self._dynamic(tensor, _cabi.OUT_NONE, None, True)
        if zeros is None:
            zeros = torch.tensor(0.0)
        return self.reshape_tensor(tensor), scales, zeros, self.qmax.to(dev), self.qmin.to(dev)

    def get_minmax_range(self, tensor):
        """quant.py:132-143: (min_val, max_val) of every unit — per_block: of |t|.float() (fp32, zero padding counted);
        otherwise amin / amax in the tensor dtype, shaped [units, 1] (per_tensor: 0-dim). Accepts the tensor in its
        original or in its reshape_tensor() layout (lcb_minmax)."""
        _require_cuda(tensor, 'get_minmax_range')
        x = tensor.contiguous()
        if self.granularity == 'per_block' and x.dim() == 4:
            raise NotImplementedError('4-D per_block views: pass the 2-D weight')
        u = self._units(x)
        per_block = self.granularity == 'per_block'
        rd = torch.float32 if per_block else x.dtype
        mins = torch.empty(u.qshape, dtype=rd, device=x.device)
        maxs = torch.empty(u.qshape, dtype=rd, device=x.device)
        lib = _cabi.load()
        ws = None
        if u.n_units == 1:
            ws = torch.empty(int(lib.lcb_fakequant_workspace_bytes()), dtype=torch.uint8, device=x.device)
        with _on_device(x.device):
            st = lib.lcb_minmax(x.data_ptr(), mins.data_ptr(), maxs.data_ptr(), u.rows, u.cols, u.ur, u.uc, _DT[x.dtype],
                                _DT[rd], 1 if per_block else 0, ws.data_ptr() if ws is not None else None,
                                _stream(x.device.index))
        _cabi.check(st, 'lcb_minmax')
        return mins, maxs

    def get_learnable_range(self, tensor, lowbound_factor=None, upbound_factor=None):
        """quant.py:205-219: the min/max search on the GPU kernel, the sigmoid bounds on the [units, 1] results."""
        min_val, max_val = self.get_minmax_range(tensor)
        if self.sym:
            if upbound_factor is not None:
                abs_max = torch.max(max_val.abs(), min_val.abs()).clamp(min=1e-5)
                abs_max = torch.sigmoid(upbound_factor) * abs_max
                min_val, max_val = -abs_max, abs_max
        elif upbound_factor is not None and lowbound_factor is not None:
            min_val = torch.sigmoid(lowbound_factor) * min_val
            max_val = torch.sigmoid(upbound_factor) * max_val
        return min_val, max_val

    def get_tensor_range(self, tensor, args={}):
        """quant.py:119-130."""
        if self.calib_algo == 'learnable':
            return self.get_learnable_range(tensor, **args)
        if self.calib_algo == 'mse':
            return self.get_mse_range(tensor)
        return self.get_minmax_range(tensor)

    def get_mse_range(self, tensor, norm=2.4, bs=256):
        """quant.py:145-203 on the GPU (lcb_mse_range): fp32 (min_val, max_val) [units, 1] of the shrink-grid search.
        Accepts the tensor in its original or in its reshape_tensor() layout. Units are rows (per_group / per_channel /
        per_token / per_head) of a 2-D tensor, the only layouts the reference's search handles."""
        _require_cuda(tensor, 'get_mse_range')
        if self.granularity in ('per_tensor', 'per_block'):
            raise NotImplementedError(f'calib_algo=mse with granularity={self.granularity}: the reference search indexes '
                                      '[units, 1] ranges (quant.py:157-161)')
        if self._kind() != _cabi.KIND_INT:
            raise NotImplementedError('calib_algo=mse with FloatQuantizer: a shrunk range overflows the fp8 grid and the '
                                      'outcome is defined by qtorch.float_quantize (not pinned); IntegerQuantizer only')
        if tensor.dim() != 2:
            raise RuntimeError('get_mse_range handles 2-D tensors (the error sum runs over dim 1, quant.py:187)')
        x = tensor.contiguous()
        u = self._units(x)
        assert self.mse_b_num >= 1 and u.n_units % self.mse_b_num == 0, \
            'Batch number must be divisible by tensor.shape[0],'
        mins = torch.empty((u.n_units, 1), dtype=torch.float32, device=x.device)
        maxs = torch.empty((u.n_units, 1), dtype=torch.float32, device=x.device)
        qmin, qmax = self._qrange()
        spec = self._spec(x.dtype, torch.float32, torch.float32, _cabi.OUT_FAKE, qmin, qmax)
        lib = _cabi.load()
        with _on_device(x.device):
            st = lib.lcb_mse_range(x.data_ptr(), mins.data_ptr(), maxs.data_ptr(), u.n_units, u.uc, spec,
                                   int(self.maxshrink * self.mse_grid), float(self.mse_grid), float(norm),
                                   _stream(x.device.index))
        _cabi.check(st, 'lcb_mse_range')
        return mins, maxs

    # ---- static activation calibration (quant.py:103-117, 221-264, 524-586): host glue over the GPU range kernel -------
    def reshape_batch_tensors(self, act_tensors):
        assert len(act_tensors) > 0, (
            'Calibration data is insufficient. Please provide more data to ensure '
            'all experts in the MOE receive an adequate number of tokens.')
        if isinstance(act_tensors[0], tuple):
            act_tensors = [torch.stack(tl) for tl in zip(*act_tensors)]
        elif len(act_tensors) == 1:
            act_tensors[0] = [act_tensors[0][i] for i in range(act_tensors[0].size(0))]
        else:
            act_tensors = [act_tensors]
        return act_tensors

    def get_minmax_stats(self, act_tensors):
        stats = {}
        for idx, tensors in enumerate(act_tensors):
            mins, maxs = [], []
            for tensor in tensors:
                mn, mx = self.get_minmax_range(tensor)
                mins.append(mn.reshape(-1)[:1].float())
                maxs.append(mx.reshape(-1)[:1].float())
            # the reference builds these with torch.tensor([min_val]) one .item() at a time; one transfer here
            stats[idx] = {'min': torch.cat(mins).cpu(), 'max': torch.cat(maxs).cpu()}
        return stats

    def get_static_minmax_range(self, act_tensors):
        stats = self.get_minmax_stats(self.reshape_batch_tensors(act_tensors))
        return [r['min'].mean() for r in stats.values()], [r['max'].mean() for r in stats.values()]

    def get_static_moving_minmax_range(self, act_tensors, alpha):
        mins, maxs = [], []
        for tensors in self.reshape_batch_tensors(act_tensors):
            mmin = mmax = None
            for tensor in tensors:
                mn, mx = self.get_minmax_range(tensor)
                if mmin is None:
                    mmin, mmax = mn, mx
                else:
                    mmin = mmin + alpha * (mn - mmin)
                    mmax = mmax + alpha * (mx - mmax)
            mins.append(mmin)
            maxs.append(mmax)
        return mins, maxs

    def get_batch_tensors_qparams(self, act_tensors, alpha=0.01, args={}):
        if self.calib_algo == 'static_minmax':
            min_vals, max_vals = self.get_static_minmax_range(act_tensors)
        elif self.calib_algo == 'static_moving_minmax':
            min_vals, max_vals = self.get_static_moving_minmax_range(act_tensors, alpha)
        elif self.calib_algo == 'static_hist':
            raise NotImplementedError('calib_algo=static_hist (histogram search) is not on the CUDA hot path')
        else:
            raise ValueError(f'Unsupported calibration algorithm: {self.calib_algo}')
        out = ([], [], [], [])
        for mn, mx in zip(min_vals, max_vals):
            sc, zp, qmax, qmin = self.get_qparams((mn, mx), mn.device)
            for lst, v in zip(out, (sc, zp, qmin, qmax)):
                lst.append(v)
        return out

    def get_qparams(self, tensor_range, device):
        """quant.py:545-559 on given ranges (lcb_qparams_from_range) → (scales, zeros, qmax, qmin)."""
        min_val, max_val = tensor_range[0], tensor_range[1]
        if not min_val.is_cuda:
            # static calibration (quant.py:253-264, 561-586) averages per-batch ranges on the host: scalars travel to the
            # GPU for the arithmetic and come back to where the reference leaves them
            dev = torch.device('cuda', torch.cuda.current_device())
            sc, zp, qmax, qmin = self.get_qparams((min_val.to(dev), max_val.to(dev)), dev)
            return sc.to(min_val.device), zp.to(min_val.device), qmax.to(device), qmin.to(device)
        md = _promote(min_val.dtype, max_val.dtype)
        mn, mx = min_val.to(md).contiguous(), max_val.to(md).contiguous()
        qd = md
        if mn.dim() == 0 and (self._kind() != _cabi.KIND_INT or not self.sym):
            qd = torch.float32      # 0-dim range / 0-dim float divisor promotes (quant.py:552-556)
        qmin, qmax = self._qrange()
        spec = self._spec(md, md, qd, _cabi.OUT_NONE, qmin, qmax, 1.0, 1.0, _DT[qd] + 1)
        scales = torch.empty(mn.shape, dtype=qd, device=mn.device)
        asym = not self.sym and self._kind() == _cabi.KIND_INT
        zeros = torch.empty(mn.shape, dtype=qd, device=mn.device) if asym else None
        lib = _cabi.load()
        with _on_device(mn.device):
            st = lib.lcb_qparams_from_range(mn.data_ptr(), mx.data_ptr(), scales.data_ptr(),
                                            zeros.data_ptr() if asym else None, mn.numel(), spec, _stream(mn.device.index))
        _cabi.check(st, 'lcb_qparams_from_range')
        return scales, (zeros if asym else torch.tensor(0.0)), self.qmax.to(device), self.qmin.to(device)

    def _bcast_units(self, tensor, scales):
        """units for direct quant()/dequant()/quant_dequant() calls on already reshaped tensors."""
        u = _Units()
        n = scales.numel() if torch.is_tensor(scales) else 1
        if tensor.dim() == 4 and torch.is_tensor(scales) and scales.dim() == 4:
            rb, bs, cb, bs2 = tensor.shape
            raise NotImplementedError('4-D per_block views: call the fake_quant_* / real_quant_* entry points')
        last = tensor.shape[-1]
        u.rows, u.cols = tensor.numel() // last, last
        if n == 1:
            u.ur, u.uc, u.n_units = u.rows, u.cols, 1
        elif n == u.rows:
            u.ur, u.uc, u.n_units = 1, u.cols, u.rows
        else:
            raise ValueError(f'cannot broadcast {n} qparams over tensor of shape {tuple(tensor.shape)}')
        return u

    def quant_dequant(self, tensor, scales, zeros, qmax, qmin, output_scale_factor=1):
        """quant.py:714-717 on a reshaped [units, unit_len] tensor with [units, 1] (or scalar) qparams."""
        md = _promote(tensor.dtype, *( [scales.dtype] if torch.is_tensor(scales) and scales.dim() > 0 else []))
        x = tensor if md == tensor.dtype else tensor.to(md)
        return self._static(x, scales, zeros, qmax, qmin, _cabi.OUT_FAKE, x.dtype, output_scale_factor,
                            units=self._bcast_units(x, scales))

    def quant(self, tensor, scales, zeros, qmax, qmin):
        """quant.py:699-708: integer codes kept in the (promoted) floating dtype."""
        md = _promote(tensor.dtype, *( [scales.dtype] if torch.is_tensor(scales) and scales.dim() > 0 else []))
        x = tensor if md == tensor.dtype else tensor.to(md)
        return self._static(x, scales, zeros, qmax, qmin, _cabi.OUT_CODE_T, x.dtype,
                            units=self._bcast_units(x, scales))

    # fake quant --------------------------------------------------------------------------------------
    def _split_mixed(self, t, args, dim):
        if 'int_indices' in args:
            if self.granularity == 'per_group':
                assert len(args['int_indices']) % self.group_size == 0
            return t.index_select(dim, torch.as_tensor(args['int_indices'], device=t.device))
        return None

    def _fake_dynamic(self, tensor, args, is_weight):
        if not args:
            # hot path (deploy-time w_qdq, one a_qdq per linear per forward): a remembered call plan, no Python bookkeeping
            key = (tensor.shape, tensor.dtype, tensor.device)
            plans = self.__dict__.get('_plans')
            p = plans.get(key) if plans else None
            if p is not None and tensor.is_contiguous() and _cur_dev() == p[1]:
                out = torch.empty_like(tensor)
                st = p[2](p[0], tensor.data_ptr(), out.data_ptr(), None, None, None, _stream(p[1]))
                if st:
                    _cabi.check(st, 'lcb_fakequant')
                return out
            if self.calib_algo == 'minmax' and tensor.is_contiguous():
                q, _, _ = self._dynamic(tensor, _cabi.OUT_FAKE, tensor.dtype, False, plan_key=key)
                return q
        org_bit = None
        if 'current_bit' in args:
            # quant.py:754-783 / 833-869: only `self.bit` changes for the duration of the call; qmin / qmax keep their
            # constructor values, so the grid does not move (QUIK's buf_current_bit is bookkeeping in the reference too)
            org_bit, self.bit = self.bit, args['current_bit']
        try:
            dim = 1 if is_weight else 2
            sub = self._split_mixed(tensor, args, dim)
            if sub is not None:
                q, _, _ = self._dynamic(sub, _cabi.OUT_FAKE, sub.dtype, False)
                mix = tensor.clone()
                mix.index_copy_(dim, torch.as_tensor(args['int_indices'], device=tensor.device), q)
                return mix
            if self.calib_algo == 'learnable':   # AWQ clip v2 deploy (base_bq.py:46-65): bounds from the module buffers
                _, sc, zp, qmax, qmin = self.get_tensor_qparams(tensor, args)
                return self._static(tensor, sc, zp, qmax, qmin, _cabi.OUT_FAKE, tensor.dtype)
            if is_weight and 'dim' in args and 'ic' in args['dim']:
                q, _, _ = self._dynamic(tensor.T, _cabi.OUT_FAKE, tensor.dtype, False)
                return q.T
            q, _, _ = self._dynamic(tensor, _cabi.OUT_FAKE, tensor.dtype, False)
            return q
        finally:
            if org_bit is not None:
                self.bit = org_bit

    def _fake_static(self, tensor, args, is_weight):
        # args['current_bit'] (quant.py:785-831) only renames self.bit for the call: no effect on the result
        if 'rounding' in args:
            raise NotImplementedError("args['rounding'] (AdaRound-style floor+offset) is not on the CUDA hot path")
        osf = args.get('output_scale_factor', 1) if is_weight else 1
        dim = 1 if is_weight else 2
        sub = self._split_mixed(tensor, args, dim)
        if sub is not None:
            q = self._static(sub, args['scales'], args['zeros'], args['qmax'], args['qmin'], _cabi.OUT_FAKE,
                             sub.dtype, osf)
            mix = tensor.clone()
            mix.index_copy_(dim, torch.as_tensor(args['int_indices'], device=tensor.device), q)
            return mix
        if is_weight and 'dim' in args and 'ic' in args['dim']:
            return self._static(tensor.T, args['scales'], args['zeros'], args['qmax'], args['qmin'],
                                _cabi.OUT_FAKE, tensor.dtype, osf).T
        return self._static(tensor, args['scales'], args['zeros'], args['qmax'], args['qmin'], _cabi.OUT_FAKE,
                            tensor.dtype, osf)

    def fake_quant_weight_dynamic(self, weight, args={}):
        """quant.py:833-869 / :1142-1159."""
        return self._fake_dynamic(weight, args, True)

    def fake_quant_act_dynamic(self, act, args={}):
        """quant.py:754-783 / :1100-1109."""
        return self._fake_dynamic(act, args, False)

    def fake_quant_weights_dynamic(self, weights):
        """Extension (not in the reference API): fake_quant_weight_dynamic of a list of weights in ONE launch
        (lcb_fakequant_multi) — the deploy-time w_qdq of all linears of a transformer block. Falls back to one launch per
        tensor for specs / layouts outside the multi-tensor envelope; results are identical either way."""
        import ctypes
        ws = list(weights)
        if self._block_fp8_multi_ok(ws):
            res = self._block_fp8_multi(ws, _cabi.OUT_FAKE)
            if res is not None:
                return res
        ok = (len(ws) > 0 and self.granularity == 'per_group' and self.calib_algo == 'minmax'
              and all(w.is_cuda and w.is_contiguous() and w.dtype == ws[0].dtype and w.device == ws[0].device
                      and w.dim() >= 1 and w.shape[-1] >= self.group_size and w.shape[-1] % self.group_size == 0
                      for w in ws))
        if not ok:
            return [self.fake_quant_weight_dynamic(w) for w in ws]
        outs_all = []
        lib = _cabi._lib or _cabi.load()
        dt = ws[0].dtype
        qmin, qmax = self._qrange()
        spec = self._spec(dt, dt, dt, _cabi.OUT_FAKE, qmin, qmax, 1.0, 1.0, _DT[dt] + 1)
        with _on_device(ws[0].device):
            for i0 in range(0, len(ws), 16):
                part = ws[i0:i0 + 16]
                outs = [torch.empty_like(w) for w in part]
                n = len(part)
                ins_a = (ctypes.c_void_p * n)(*[w.data_ptr() for w in part])
                outs_a = (ctypes.c_void_p * n)(*[o.data_ptr() for o in outs])
                rows_a = (ctypes.c_int64 * n)(*[w.numel() // w.shape[-1] for w in part])
                cols_a = (ctypes.c_int64 * n)(*[w.shape[-1] for w in part])
                st = lib.lcb_fakequant_multi(ins_a, outs_a, rows_a, cols_a, n, self.group_size, spec,
                                             _stream(ws[0].device.index))
                if st == -3:   # LCB_E_UNSUPPORTED: spec outside the multi-tensor envelope
                    outs = [self.fake_quant_weight_dynamic(w) for w in part]
                elif st:
                    _cabi.check(st, 'lcb_fakequant_multi')
                outs_all.extend(outs)
        return outs_all

    def _block_fp8_multi_ok(self, ws):
        """the quantizer-side half of the multi-tensor envelope; the per-tensor half is checked by _block_fp8_multi's scan"""
        return (len(ws) > 0 and self.granularity == 'per_block' and self._kind() != _cabi.KIND_INT
                and self.calib_algo != 'hqq' and self.calib_algo != 'mse' and getattr(self, 'block_size', 0) == 128
                and ws[0].is_cuda and ws[0].dtype in (torch.bfloat16, torch.float16))

    def _block_fp8_multi(self, ws, out_mode):
        """FP8 per_block (128 x 128) dynamic quant of a list of 16-bit weights, 64 per launch
        (lcb_fakequant_block_fp8_multi). out_mode OUT_FAKE -> [fake weights]; OUT_FP8 -> [(codes, scales, None)];
        None when a weight is outside the envelope (the caller then loops over the per-tensor entry).

        A MoE block hands over ~200 small matrices (1.4 us of HBM traffic each), so the host side is the cost. One scan
        over the list (shape, layout, address); outputs of equal shape are slices of ONE allocation (rows of a
        `[n, pitch]` pool: two allocator calls per distinct shape instead of two per weight — the slices keep their pool
        alive until the last one is dropped, which is what a deployed model does anyway), their addresses follow from
        the pool's base address, so every launch is issued BEFORE the per-weight tensor objects are created: the GPU
        works while Python builds the 2 x n views."""
        import ctypes
        lib = _cabi._lib or _cabi.load()
        dt, dev = ws[0].dtype, ws[0].device
        real = out_mode == _cabi.OUT_FP8
        n_all = len(ws)
        rows, cols, in_ptr = [0] * n_all, [0] * n_all, [0] * n_all
        by_shape = {}
        for i, w in enumerate(ws):
            sh = w.shape
            if len(sh) != 2 or sh[1] % 8 or w.dtype != dt or w.device != dev or not w.is_contiguous():
                return None
            rows[i], cols[i] = sh
            in_ptr[i] = w.data_ptr()
            by_shape.setdefault(sh, []).append(i)
        qmin, qmax = self._qrange()
        spec = self._spec(dt, torch.float32, torch.float32, out_mode, qmin, qmax, 1.0, 1.0, _DT[torch.float32] + 1)

        def pool_of(k, shape, dtype, view_dtype=None):
            """k tensors of `shape` carved out of one allocation, each starting on a 256-byte boundary (the kernel wants
            16-byte aligned outputs; a [1, 8] matrix of bytes is 8 bytes long). Returns (2-D pool view, base, step)."""
            n = shape[0] * shape[1]
            es = 1 if dtype == torch.uint8 else (4 if dtype == torch.float32 else 2)
            pitch = -(-n * es // 256) * 256 // es
            pool = torch.empty((k, pitch), dtype=dtype, device=dev)
            base = pool.data_ptr()
            if view_dtype is not None:
                pool = pool.view(view_dtype)
            return pool[:, :n].unflatten(1, (shape[0], shape[1])), base, pitch * es

        out_ptr, sc_ptr = [0] * n_all, [0] * n_all
        pools = []
        for sh, idx in by_shape.items():
            k = len(idx)
            pool, base, step = pool_of(k, sh, torch.uint8 if real else dt, self._fp8_dtype() if real else None)
            for j, i in enumerate(idx):
                out_ptr[i] = base + j * step
            spool = None
            if real:
                spool, base, step = pool_of(k, ((sh[0] + 127) // 128, (sh[1] + 127) // 128), torch.float32)
                for j, i in enumerate(idx):
                    sc_ptr[i] = base + j * step
            pools.append((idx, pool, spool))
        stream = _stream(dev.index)
        per_launch = 64
        redo = []
        with _on_device(dev):
            for i0 in range(0, n_all, per_launch):
                i1 = min(n_all, i0 + per_launch)
                n = i1 - i0
                ins_a = (ctypes.c_void_p * n)(*in_ptr[i0:i1])
                outs_a = (ctypes.c_void_p * n)(*out_ptr[i0:i1])
                scs_a = (ctypes.c_void_p * n)(*sc_ptr[i0:i1]) if real else None
                rows_a = (ctypes.c_int64 * n)(*rows[i0:i1])
                cols_a = (ctypes.c_int64 * n)(*cols[i0:i1])
                st = lib.lcb_fakequant_block_fp8_multi(ins_a, outs_a, scs_a, rows_a, cols_a, n, spec, stream)
                if st == -3:   # LCB_E_UNSUPPORTED: outside the multi-tensor envelope -> the per-tensor entry, same results
                    redo.extend(range(i0, i1))
                    continue
                _cabi.check(st, 'lcb_fakequant_block_fp8_multi')
        outs, scs = [None] * n_all, [None] * n_all
        for idx, pool, spool in pools:
            for i, o in zip(idx, pool.unbind(0)):
                outs[i] = o
            if spool is not None:
                for i, sc in zip(idx, spool.unbind(0)):
                    scs[i] = sc
        for i in redo:
            if real:
                outs[i], scs[i], _ = self.real_quant_weight_dynamic(ws[i])
            else:
                outs[i] = self.fake_quant_weight_dynamic(ws[i])
        if real:
            return [(o, sc, None) for o, sc in zip(outs, scs)]
        return outs

    def fake_quant_weight_clipped(self, weight, ratio, clip_sym=False):
        """AutoClipper.fake_quantize_weight, clip_version 'v1' (auto_clip.py:258-260) at one shrink step:
        fake_quant_weight_dynamic(clamp(w, org_min * ratio, org_max * ratio)) per unit (clip_sym: +/- |w|.amax * ratio),
        the clamp fused into the quant kernel."""
        ratio = float(ratio)
        q, _, _ = self._dynamic(weight, _cabi.OUT_FAKE, weight.dtype, False, clip=(-ratio if clip_sym else ratio))
        return q

    def fake_quant_weight_static(self, weight, args):
        """quant.py:785-831 / :1111-1140."""
        return self._fake_static(weight, args, True)

    def fake_quant_act_static(self, act, args={}):
        """quant.py:719-752 / :1083-1098."""
        return self._fake_static(act, args, False)

    def _set_int_range(self):
        pass

    def _qparams_view(self, scales, zeros, weight):
        if self.granularity == 'per_tensor':
            shape = (1,)
        elif self.granularity == 'per_block':
            shape = (scales.shape[0], scales.shape[2])
        else:
            shape = (weight.shape[0], -1)
        return scales.view(shape), (zeros.view(shape) if zeros is not None else None)


class IntegerQuantizer(BaseQuantizer):
    """Mirror of quant.py:661-960."""

    def __init__(self, bit, symmetric, granularity, **kwargs):
        super().__init__(bit, symmetric, granularity, **kwargs)
        self.quant_type = 'int-quant'
        self._set_int_range()
        self.dst_nbins = 2 ** bit

    def _kind(self):
        return _cabi.KIND_INT

    def _set_int_range(self):
        # quant.py:665-677 — note the asymmetric qmin is the *float* 0.0, qmax an int (kept for drop-in parity)
        if 'int_range' in self.kwargs:
            lo, hi = self.kwargs['int_range'][0], self.kwargs['int_range'][1]
        elif self.sym:
            lo, hi = -(2 ** (self.bit - 1)), 2 ** (self.bit - 1) - 1
        else:
            lo, hi = 0.0, 2 ** self.bit - 1
        self.qmin = torch.tensor(lo)
        self.qmax = torch.tensor(hi)

    def dequant(self, tensor, scales, zeros):
        """quant.py:710-712: (codes - zeros) * scales on a reshaped [units, unit_len] tensor of codes."""
        md = _promote(tensor.dtype, *([scales.dtype] if torch.is_tensor(scales) and scales.dim() > 0 else []))
        x = tensor if md == tensor.dtype else tensor.to(md)
        return self._static(x, scales, zeros, None, None, _cabi.OUT_DEQUANT, x.dtype, units=self._bcast_units(x, scales))

    def _code_dtype(self):
        # quant.py:929-936: bit 8 -> int8 (uint8 when qmin == 0); anything else -> unpacked int32
        if self.bit == 8:
            return (torch.int8, _cabi.OUT_I8) if self.qmin != 0 else (torch.uint8, _cabi.OUT_U8)
        return torch.int32, _cabi.OUT_I32

    def real_quant_weight_dynamic(self, weight, args={}):
        """quant.py:916-953 → (codes, scales, zeros|None)."""
        osf = 1
        if 'output_scale_factor' in args:
            osf = args['output_scale_factor']
            del args['output_scale_factor']
        dtype, mode = self._code_dtype()
        codes, scales, zeros = self._dynamic(weight, mode, dtype, True, osf=osf)
        if not self.sym and self.round_zp:
            zeros = zeros.to(dtype)
        elif self.sym:
            zeros = None
        scales, zeros = self._qparams_view(scales, zeros, codes)
        return codes, scales, zeros

    def real_quant_weight_static(self, weight, args):
        """quant.py:871-914."""
        osf = 1
        if 'output_scale_factor' in args:
            osf = args['output_scale_factor']
            del args['output_scale_factor']
        dtype, mode = self._code_dtype()
        scales, zeros = args['scales'], args['zeros']
        codes = self._static(weight, scales, zeros, args['qmax'], args['qmin'], mode, dtype)
        scales = scales * osf
        if not self.sym and self.round_zp:
            zeros = zeros.to(dtype)
        elif self.sym:
            zeros = None
        scales, zeros = self._qparams_view(scales, zeros, codes)
        return codes, scales, zeros

    def __repr__(self):
        return (f'IntegerQuantizer(bit={self.bit}, sym={self.sym},granularity={self.granularity},'
                f'kwargs={self.kwargs}, qmin={self.qmin}, qmax={self.qmax})')


class FloatQuantizer(BaseQuantizer):
    """Mirror of quant.py:963-1229.

    `use_qtorch: True` (scale-based FP8, e4m3 / e5m2): rounding is fp32 -> fp8 round-to-nearest-even (the rounding of
    `Tensor.to(torch.float8_*)`, which ends every reference real-quant call, quant.py:1183/1211, and of the reference's
    own Triton path, kernel.py:8-83). qtorch itself is an un-vendored dependency: parity unpinned for its tie rule, see
    DESIGN.md.

    `use_qtorch` unset (configs/quantization/methods/FP_Quant/rtn_we4m3ae4m3.yml ...): per-ELEMENT scales from
    get_float_qparams (quant.py:1005-1027), any eXmY format, 1-D granularities — csrc/floatq.cu, pinned bit-exactly to the
    unmodified reference on the GPU box (tests/test_floatq_gpu.py). Real quant has no meaning there (one scale per
    element) and raises.
    """

    def __init__(self, bit, symmetric, granularity, **kwargs):
        super().__init__(bit, symmetric, granularity, **kwargs)
        self.sym = True
        self.quant_type = 'float-quant'
        self.e_bits = int(self.bit[1])
        self.m_bits = int(self.bit[-1])
        self.sign_bits = 1
        self.num_bits = self.e_bits + self.m_bits + self.sign_bits
        self.default_bias = 2 ** (self.e_bits - 1)
        self.dst_nbins = 2 ** self.num_bits
        self.use_qtorch = kwargs.get('use_qtorch')
        if not self.use_qtorch:
            self.qmin = self.qmax = None      # the reference never sets them on this path
            return
        if self.bit not in ('e4m3', 'e5m2'):
            raise NotImplementedError('use_qtorch: only FP8 E4M3 / E5M2 are on the CUDA hot path')
        if 'float_range' in kwargs:
            self.qmin, self.qmax = kwargs['float_range']
        else:
            finfo = torch.finfo(torch.float8_e4m3fn if self.e_bits == 4 else torch.float8_e5m2)
            self.qmin, self.qmax = finfo.min, finfo.max
        self.qmax = torch.tensor(self.qmax)
        self.qmin = torch.tensor(self.qmin)

    # ---- use_qtorch unset: per-element scales (quant.py:1005-1027, 1061-1076) -------------------------------------------
    def _elem_units(self, tensor):
        """(n_units, unit_size) of the 1-D granularities; what the reference cannot run raises the same way."""
        if self.calib_algo not in ('minmax',):
            raise NotImplementedError(f'FloatQuantizer without qtorch: calib_algo={self.calib_algo!r} is not on the CUDA hot path')
        if self.granularity == 'per_tensor':
            raise IndexError('tuple index out of range')     # quant.py:1013: maxval.shape[0] of a 0-dim tensor
        if self.granularity == 'per_block':
            raise NotImplementedError('FloatQuantizer without qtorch: per_block is not on the CUDA hot path')
        u = self._units(tensor)
        return u.n_units, u.uc

    def _elem_dynamic(self, tensor, want_out, want_qparams):
        _require_cuda(tensor, 'FloatQuantizer (per-element scales)')
        if tensor.dtype not in _DT:
            raise NotImplementedError(f'dtype {tensor.dtype}')
        x = tensor.contiguous()
        n_units, unit = self._elem_units(x)
        qdt = torch.float32 if (self.e_bits >= 5 or x.dtype == torch.float32) else x.dtype
        out = torch.empty_like(x) if want_out else None
        xc = torch.empty(x.shape, dtype=qdt, device=x.device) if want_qparams else None
        sc = torch.empty(x.shape, dtype=qdt, device=x.device) if want_qparams else None
        lib = _cabi._lib or _cabi.load()
        with _on_device(x.device):
            st = lib.lcb_floatquant_elem(x.data_ptr(), out.data_ptr() if want_out else None,
                                         xc.data_ptr() if want_qparams else None, sc.data_ptr() if want_qparams else None,
                                         n_units, unit, self.e_bits, self.m_bits, _DT[x.dtype], _stream(x.device.index))
        _cabi.check(st, 'lcb_floatquant_elem')
        return out, xc, sc

    def _elem_static(self, tensor, scales):
        """(round(x / s + 0) - 0) * s with per-element scales (the `scales[scales == 0] = 1` of quant.py:1062 is applied
        to the caller's tensor too, as the reference does)."""
        _require_cuda(tensor, 'FloatQuantizer (per-element scales)')
        x = tensor.contiguous()
        if not torch.is_tensor(scales) or scales.numel() != x.numel():
            raise ValueError(f'per-element scales hold {getattr(scales, "numel", lambda: 1)()} values, tensor has {x.numel()}')
        scales[scales == 0] = 1
        sc = scales.contiguous()
        if sc.dtype != x.dtype and sc.dtype != torch.float32:
            sc = sc.float()
        out = torch.empty_like(x)
        lib = _cabi._lib or _cabi.load()
        with _on_device(x.device):
            st = lib.lcb_floatquant_elem_static(x.data_ptr(), sc.data_ptr(), out.data_ptr(), x.numel(), _DT[x.dtype],
                                                _DT[sc.dtype], _stream(x.device.index))
        _cabi.check(st, 'lcb_floatquant_elem_static')
        return out

    def get_float_qparams(self, tensor, tensor_range=None, device=None):
        """quant.py:1005-1027 -> (xc, scales) for a tensor in its reshape_tensor() layout (the range is recomputed: it
        is the min / max of that layout's rows)."""
        _, xc, sc = self._elem_dynamic(tensor, False, True)
        return xc, sc

    def get_tensor_qparams(self, tensor, args={}):
        if self.use_qtorch:
            return super().get_tensor_qparams(tensor, args)
        t = self.reshape_tensor(tensor)
        xc, sc = self.get_float_qparams(t)
        return xc, sc, torch.tensor(0), None, None

    def quant_dequant(self, tensor, scales, zeros, qmax, qmin):
        if self.use_qtorch:
            return super().quant_dequant(tensor, scales, zeros, qmax, qmin)
        out = self._elem_static(tensor.to(scales.dtype) if scales.dtype == torch.float32 and tensor.dtype != torch.float32
                                else tensor, scales)
        return out

    def _fake_dynamic(self, tensor, args, is_weight):
        if self.use_qtorch:
            return super()._fake_dynamic(tensor, args, is_weight)
        t = tensor.T if (is_weight and 'dim' in args and 'ic' in args['dim']) else tensor
        q, _, _ = self._elem_dynamic(t, True, False)
        return q.T if t is not tensor else q

    def _fake_static(self, tensor, args, is_weight):
        if self.use_qtorch:
            return super()._fake_static(tensor, args, is_weight)
        if 'rounding' in args:
            raise NotImplementedError("args['rounding'] (AdaRound-style floor+offset) is not on the CUDA hot path")
        t = tensor.T if (is_weight and 'dim' in args and 'ic' in args['dim']) else tensor
        q = self._elem_static(t, args['scales'])
        return q.T if t is not tensor else q

    def _kind(self):
        return _cabi.KIND_FP8_E4M3 if self.e_bits == 4 else _cabi.KIND_FP8_E5M2

    def _fp8_dtype(self):
        return torch.float8_e4m3fn if self.e_bits == 4 else torch.float8_e5m2

    def _no_real_quant(self):
        if not self.use_qtorch:
            raise NotImplementedError('FloatQuantizer without qtorch has one scale per ELEMENT (quant.py:1005-1027): '
                                      'real quant is not defined on the CUDA hot path')

    def fake_quant_weights_dynamic(self, weights):
        if not self.use_qtorch:
            return [self.fake_quant_weight_dynamic(w) for w in weights]
        return super().fake_quant_weights_dynamic(weights)

    def real_quant_weight_dynamic(self, weight, args={}):
        """quant.py:1195-1221 → (fp8 codes, scales, None)."""
        assert self.bit in ['e4m3', 'e5m2'], 'Only FP8 E4M3 and E5M2 support real quant'
        self._no_real_quant()
        osf = 1
        if 'output_scale_factor' in args:
            osf = args['output_scale_factor']
            del args['output_scale_factor']
        codes, scales, _ = self._dynamic(weight, _cabi.OUT_FP8, torch.uint8, True, osf=osf)
        scales, _ = self._qparams_view(scales, None, codes)
        return codes.view(self._fp8_dtype()), scales, None

    def real_quant_weights_dynamic(self, weights):
        """Extension (not in the reference API): real_quant_weight_dynamic of a list of weights — FP8 per_block weights go
        64 per launch (the expert matrices of a MoE block); anything else loops. Same results as the per-tensor calls."""
        ws = list(weights)
        self._no_real_quant()
        if self.bit in ['e4m3', 'e5m2'] and self._block_fp8_multi_ok(ws):
            res = self._block_fp8_multi(ws, _cabi.OUT_FP8)
            if res is not None:
                return res
        return [self.real_quant_weight_dynamic(w) for w in ws]

    def real_quant_weight_static(self, weight, args):
        """quant.py:1161-1193."""
        assert self.bit in ['e4m3', 'e5m2'], 'Only FP8 E4M3 and E5M2 support real quant'
        self._no_real_quant()
        osf = 1
        if 'output_scale_factor' in args:
            osf = args['output_scale_factor']
            del args['output_scale_factor']
        scales = args['scales']
        codes = self._static(weight, scales, None, args['qmax'], args['qmin'], _cabi.OUT_FP8, torch.uint8)
        scales, _ = self._qparams_view(scales * osf, None, codes)
        return codes.view(self._fp8_dtype()), scales, None

    def __repr__(self):
        return (f'FloatQuantizer(bit={self.bit},e_bits={self.e_bits}, m_bits={self.m_bits},'
                f'granularity={self.granularity},kwargs={self.kwargs}, qmin={self.qmin}, qmax={self.qmax})')


def weight_cast_to_fp8(weight, block_size):
    """quant.py:34-43 / kernel.py:73-83: per-block e4m3 codes + fp32 scales [ceil(R/bs), ceil(C/bs)]."""
    q = FloatQuantizer(bit='e4m3', symmetric=True, granularity='per_block', block_size=block_size, use_qtorch=True)
    codes, scales, _ = q.real_quant_weight_dynamic(weight)
    return codes, scales


def weight_cast_to_bf16(weight, scale, block_size):
    """quant.py:18-31 / kernel.py:115-138: y = fp8 * s[block] -> bf16."""
    _require_cuda(scale, 'weight_cast_to_bf16')
    lib = _cabi.load()
    w = weight.contiguous()
    out = torch.empty(w.shape, dtype=torch.bfloat16, device=w.device)
    kind = _cabi.KIND_FP8_E4M3 if w.dtype == torch.float8_e4m3fn else _cabi.KIND_FP8_E5M2
    sc = scale.float().contiguous()
    with _on_device(w.device):
        st = lib.lcb_fp8_block_dequant(w.data_ptr(), sc.data_ptr(), out.data_ptr(),
                                       w.shape[0], w.shape[1], block_size, kind, _cabi.LCB_BF16,
                                       _stream(w.device.index))
    _cabi.check(st, 'lcb_fp8_block_dequant')
    return out
