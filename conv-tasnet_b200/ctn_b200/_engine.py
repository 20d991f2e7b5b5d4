"""Host-side orchestration of the separation forward pass on one GPU.

PyTorch is used for device memory and streams only; every arithmetic step is a
call into libctn_b200.so.  The sequence mirrors ConvTasNet.forward
(conv_tasnet.py:45-60) but in fused form:

  encoder (+ cLN statistics)                          ctn_encoder_fwd
  cLN folded into the bottleneck 1x1                  ctn_gemm1x1   (in_norm = CLN)
  per TemporalBlock (conv_tasnet.py:212-272):
    conv1x1 (+ statistics of PReLU output)            ctn_gemm1x1   (out_stat)
    PReLU, norm, depthwise, PReLU (+ statistics)      ctn_dwconv_fwd
    norm folded into pointwise 1x1, + residual        ctn_gemm1x1   (in_norm, resid_hi/lo)
  mask 1x1                                            ctn_gemm1x1
  mask non-linearity, decoder, overlap-add, pad       ctn_decoder_fwd

A `_Plan` owns the static workspaces for one (M, T) shape and the prebuilt
argument structs, so a forward is a flat list of C calls on the current stream;
it can be captured into a CUDA graph (`use_graph`) to remove launch gaps.
"""
import ctypes as C

import torch

from . import _native as nat
from ._layout import block_keys, frame_geometry, is_bn

_NORM = {"gLN": nat.NORM_GLN, "cLN": nat.NORM_CLN}
_MASK = {"relu": nat.MASK_RELU, "softmax": nat.MASK_SOFTMAX}


BN_EPS = 1e-5   # nn.BatchNorm1d default (conv_tasnet.py:303 passes no eps)


def _check_cfg(cfg, training=False):
    if training and nat.PRECISION != "bf16x3":
        raise RuntimeError("CTN_PRECISION=%s is an inference-only precision: gradients need the bf16 (hi, lo) "
                           "operands of the default mode" % nat.PRECISION)
    # (any norm_type but gLN / cLN is nn.BatchNorm1d, conv_tasnet.py:300-303: inference folds the running
    # statistics into a per-channel affine (_Plan); training uses batch statistics, csrc/bn_train.cu)
    if cfg["mask_nonlinear"] not in _MASK:
        raise ValueError("Unsupported mask non-linear function")  # conv_tasnet.py:208


def aligned_params(store, params, device):
    """The kernels read parameter vectors with 16-byte loads.  Parameters of an ordinary module are
    separate allocations (256-byte aligned); the replicas torch.nn.DataParallel builds (train.py:121) are
    views into one coalesced broadcast buffer and generally are not.  In that case the values are copied
    (device-to-device, one foreach copy) into engine-owned buffers with stable addresses, so plans and
    CUDA graphs built on them stay valid from call to call."""
    if all(p.data_ptr() % 16 == 0 for p in params.values()):
        return params
    bufs = store.get(device.index)
    if bufs is None or list(bufs) != list(params) or any(bufs[k].shape != p.shape for k, p in params.items()):
        bufs = store[device.index] = {k: torch.empty(p.shape, dtype=p.dtype, device=device) for k, p in params.items()}
    torch._foreach_copy_(list(bufs.values()), [p.detach() for p in params.values()])
    return bufs


def _tensor_bytes(obj, seen):
    if isinstance(obj, torch.Tensor):
        st = obj.untyped_storage()
        if st.data_ptr() in seen:
            return 0
        seen.add(st.data_ptr())
        return st.nbytes()
    if isinstance(obj, (list, tuple)):
        return sum(_tensor_bytes(o, seen) for o in obj)
    if isinstance(obj, dict):
        return sum(_tensor_bytes(o, seen) for o in obj.values())
    return 0


class PlanCache:
    """Shape-keyed plans (workspaces, call lists, CUDA graphs) with a byte budget: least recently used plans are
    dropped once the cached workspaces exceed `limit_bytes`.  The reference's callers produce one tensor
    shape per utterance length -- solver.py:96-97 validates with gradients enabled on the unsegmented loader,
    evaluate.py / separate.py see ~3000 distinct lengths -- and a training plan holds the saved activations of
    every block (about 0.8 GB per utterance of 6 s at the paper configuration), so an unbounded cache runs out
    of memory after a few dozen lengths.  The `keep` most recent plans always stay (the slots of one
    pipelined call are in use at the same time).  An evicted plan that an autograd graph still references
    lives until that graph is freed."""

    def __init__(self, limit_bytes=None, keep=4):
        from collections import OrderedDict
        self.plans = OrderedDict()
        self.bytes = {}
        self.limit_bytes = limit_bytes      # None = a quarter of the device's memory, resolved at first use
        self.keep = keep
        self.evictions = 0

    def __len__(self):
        return len(self.plans)

    def __contains__(self, key):
        return key in self.plans

    def values(self):
        return self.plans.values()

    def clear(self):
        self.plans.clear()
        self.bytes.clear()

    def get(self, key):
        plan = self.plans.get(key)
        if plan is not None:
            self.plans.move_to_end(key)
        return plan

    def total_bytes(self):
        return sum(self.bytes.values())

    def put(self, key, plan, device):
        if self.limit_bytes is None:
            self.limit_bytes = torch.cuda.get_device_properties(device).total_memory // 4
        self.plans[key] = plan
        self.bytes[key] = _tensor_bytes(plan.__dict__, set())
        total = self.total_bytes()
        while total > self.limit_bytes and len(self.plans) > self.keep:
            old, _ = self.plans.popitem(last=False)
            total -= self.bytes.pop(old)
            self.evictions += 1
        return plan


class _Prepared:
    """Weight-derived buffers: bf16 (hi, lo) splits with the preceding norm's affine folded in."""

    def __init__(self, cfg, device, fold_bn=True):
        # fold_bn: BatchNorm1d blocks fold their RUNNING statistics into w2 here (inference); the training plan
        # folds batch statistics per step itself (TrainPlan.bn_w2) and passes False
        self.fold_bn = fold_bn
        N, B, H, C = cfg["N"], cfg["B"], cfg["H"], cfg["C"]
        bf = dict(dtype=torch.bfloat16, device=device)
        f32 = dict(dtype=torch.float32, device=device)
        self.bott = (torch.empty(B, N, **bf), torch.empty(B, N, **bf), torch.empty(B, **f32), torch.empty(B, **f32))
        self.blocks = []
        for _ in range(cfg["R"] * cfg["X"]):
            self.blocks.append(dict(
                w1=(torch.empty(H, B, **bf), torch.empty(H, B, **bf)),
                w2=(torch.empty(B, H, **bf), torch.empty(B, H, **bf), torch.empty(B, **f32), torch.empty(B, **f32))))
        self.mask = (torch.empty(C * N, B, **bf), torch.empty(C * N, B, **bf))
        if is_bn(cfg["norm_type"]) and fold_bn:     # per block: (scale, shift) of both BatchNorm1d layers, folded from the
            for blk in self.blocks:     # running statistics by ctn_bn_fold
                blk["bn1"] = (torch.empty(H, **f32), torch.empty(H, **f32))
                blk["bn2"] = (torch.empty(H, **f32), torch.empty(H, **f32))
        self.fingerprint = None
        self.table = self.table_ptrs = None

    def refresh(self, cfg, params):
        fp = tuple((p.data_ptr(), p._version) for p in params.values())
        if fp == self.fingerprint:
            return
        ptrs = tuple(p.data_ptr() for p in params.values())
        bn = is_bn(cfg["norm_type"]) and self.fold_bn
        if bn:
            lib, st, i = nat.lib(), nat.stream(), 0
            for r in range(cfg["R"]):
                for x in range(cfg["X"]):
                    k = block_keys(r, x, cfg["causal"], cfg["norm_type"])
                    for tag, out in (("1", self.blocks[i]["bn1"]), ("2", self.blocks[i]["bn2"])):
                        nat.check(lib.ctn_bn_fold(params[k["g" + tag]].data_ptr(), params[k["b" + tag]].data_ptr(),
                                                  params[k["rm" + tag]].data_ptr(), params[k["rv" + tag]].data_ptr(),
                                                  BN_EPS, cfg["H"], out[0].data_ptr(), out[1].data_ptr(), st))
                    i += 1
        if self.table is None or self.table_ptrs != ptrs:
            N, B, H, C = cfg["N"], cfg["B"], cfg["H"], cfg["C"]
            t = nat.PrepTable(self.bott[0].device)
            t.add(params["separator.network.1.weight"], B, N, params["separator.network.0.gamma"],
                  params["separator.network.0.beta"], 0, self.bott)
            i = 0
            for r in range(cfg["R"]):
                for x in range(cfg["X"]):
                    k = block_keys(r, x, cfg["causal"], cfg["norm_type"])
                    g2, b2 = self.blocks[i]["bn2"] if bn else (params[k["g2"]], params[k["b2"]])
                    t.add(params[k["w1"]], H, B, None, None, 0, self.blocks[i]["w1"])
                    t.add(params[k["w2"]], B, H, g2, b2, 0, self.blocks[i]["w2"])
                    i += 1
            t.add(params["separator.network.3.weight"], C * N, B, None, None, 0, self.mask)
            self.table, self.table_ptrs = t, ptrs
        self.table.run()
        self.fingerprint = fp


class _Plan:
    """Static workspaces + call list for one (M, T) on one device."""
    FUSE_UPDATE_MIN_ROWS = 8192     # frames (M * K) from which the update rides in the next conv1x1

    def __init__(self, cfg, M, T, device, fuse_max_dilation=0, fuse_update=False):
        self.cfg, self.M, self.T, self.device = cfg, M, T, device
        self.fuse_max_dilation = fuse_max_dilation
        self.fuse_update = fuse_update
        N, L, B, H, C = cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["C"]
        self.hop, self.K, self.T_conv = frame_geometry(T, L)
        K = self.K
        R_ = M * K
        lib = nat.lib()
        f32 = dict(dtype=torch.float32, device=device)
        bf = dict(dtype=torch.bfloat16, device=device)
        self.mix = torch.empty(M, T, **f32)
        self.est = torch.empty(M, C, T, **f32)
        self.w = torch.empty(R_, N, **f32)
        self.w_hl = (torch.empty(R_, N, **bf), torch.empty(R_, N, **bf))
        self.rowstat0 = torch.empty(R_, 2, **f32)
        self.x_hl = (torch.empty(R_, B, **bf), torch.empty(R_, B, **bf))
        self.a = torch.empty(R_, H, **f32)
        self.q_hl = (torch.empty(R_, H, **bf), torch.empty(R_, H, **bf))
        self.score = torch.empty(R_, C * N, **f32)
        self.one = torch.ones(1, **f32)
        # BatchNorm1d blocks (inference) run through the gLN kernels: per-channel (scale, shift) in place of
        # (gamma, beta) and unit statistics (mean 0, rstd 1) in place of the utterance statistics
        self.bn = is_bn(cfg["norm_type"])
        self.norm = nat.NORM_GLN if self.bn else _NORM[cfg["norm_type"]]
        if self.bn:
            self.unit = torch.zeros(M, 1, 2, **f32)
            self.unit[:, 0, 1] = float(K * H)       # (sum, sumsq) = (0, count): mean 0, var 1
        if self.norm == nat.NORM_GLN:
            self.P1 = lib.ctn_gemm_gln_partials(K, H)
            self.P2s = [lib.ctn_dwconv_gln_partials(K, H, cfg["P"], 2 ** x) for x in range(cfg["X"])]
            self.P2 = max(self.P2s)
            self.stat1 = torch.zeros(M, self.P1, 2, **f32)
            self.stat2 = torch.zeros(M, self.P2, 2, **f32)
        else:
            self.P1 = lib.ctn_gemm_cln_groups(H)
            self.P2 = 1
            self.P2s = [1] * cfg["X"]
            self.stat1 = torch.zeros(R_, self.P1, 2, **f32)
            self.stat2 = torch.zeros(R_, 2, **f32)
        # fused depthwise + pointwise path (ctn_dwpw_fwd + ctn_resid_update) for blocks with dilation <=
        # fuse_max_dilation, where the configuration allows it
        self.fused = [False] * cfg["X"]
        if self.norm == nat.NORM_GLN and fuse_max_dilation > 0:
            self.fused = [2 ** x <= fuse_max_dilation and bool(lib.ctn_dwpw_supported(H, B, cfg["P"], 2 ** x, self.norm))
                          for x in range(cfg["X"])]
        if any(self.fused):
            self.y = torch.empty(R_, B, **f32)
            self.P2f = [lib.ctn_dwpw_partials(K, 2 ** x) for x in range(cfg["X"])]
            self.stat2f = torch.zeros(M, max(self.P2f), 2, **f32)
        self.calls = None
        self.param_ptrs = None
        self.graph = None
        self.runs = 0
        self.computed = self.drained = None     # events of the host-to-host pipeline (forward_host)
        self._keep = []

    def _gemm(self, **kw):
        g = nat.GemmArgs()
        g.M, g.K = self.M, self.K
        for k, v in kw.items():
            setattr(g, k, v.data_ptr() if isinstance(v, torch.Tensor) else v)
        self._keep.append(g)
        fn = nat.lib().ctn_gemm1x1
        ref = C.byref(g)
        return lambda s: fn(ref, s)

    def build(self, params, prep):
        cfg = self.cfg
        N, L, B, H, C_ = cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["C"]
        M, K, T = self.M, self.K, self.T
        lib = nat.lib()
        self._keep = []
        calls = []
        info = []   # (kind, algorithmic FLOPs, algorithmic bytes) per call, for profiling / roofline reports
        Rr = M * K
        U = params["encoder.conv1d_U.weight"]
        V = params["decoder.basis_signals.weight"]
        enc = lib.ctn_encoder_fwd
        a = (self.mix.data_ptr(), U.data_ptr(), M, T, N, L, self.w.data_ptr(), self.w_hl[0].data_ptr(),
             self.w_hl[1].data_ptr(), self.rowstat0.data_ptr())
        calls.append(lambda s, a=a: enc(*a, s))
        info.append(("encoder", 2.0 * Rr * N * L, 4.0 * M * T + 8.0 * Rr * N))
        info.append(("gemm_bottleneck", 2.0 * Rr * N * B, 4.0 * Rr * N + 4.0 * Rr * B))
        calls.append(self._gemm(Cin=N, Cout=B, a_hi=self.w_hl[0], a_lo=self.w_hl[1], w_hi=prep.bott[0],
                                w_lo=prep.bott[1], col_g=prep.bott[2], col_bias=prep.bott[3],
                                in_norm=nat.NORM_CLN, in_stat=self.rowstat0,
                                out_hi=self.x_hl[0], out_lo=self.x_hl[1]))
        dw = lib.ctn_dwconv_fwd
        fw, upd, uc1 = lib.ctn_dwpw_fwd, lib.ctn_resid_update, lib.ctn_updc1_fwd

        def emit(kind, flops, nbytes, call):
            info.append((kind, flops, nbytes))
            calls.append(call)

        def emit_update(pend):
            u = (self.x_hl[0].data_ptr(), self.x_hl[1].data_ptr(), self.y.data_ptr(), pend["stat"].data_ptr(),
                 pend["P"], float(K) * H, pend["g"].data_ptr(), pend["b"].data_ptr(), M, K, B)
            emit("resid_update", 2.0 * Rr * B, 12.0 * Rr * B, lambda s, u=u: upd(*u, s))

        # (a row tile per CTA and both column tiles in sequence: below ~3 utterances of 4 s the grid of ctn_updc1_fwd
        # is a fraction of the GPU and the two plain kernels are quicker: 1.14 against 1.26 ms at M = 1, 1.46 against
        # 1.36 ms at M = 3, tools/exp_small_batch.py)
        use_c1 = (self.fuse_update and self.norm == nat.NORM_GLN and bool(lib.ctn_updc1_supported(B, H, self.norm))
                  and M * K >= self.FUSE_UPDATE_MIN_ROWS)
        pending = None    # deferred gLN-2 + residual of the previous (fused) block
        i = 0
        for r in range(cfg["R"]):
            for xb in range(cfg["X"]):
                k = block_keys(r, xb, cfg["causal"], cfg["norm_type"])
                pb = prep.blocks[i]
                i += 1
                g1, b1 = pb["bn1"] if self.bn else (params[k["g1"]], params[k["b1"]])
                st1, st1_P = (self.unit, 1) if self.bn else (self.stat1, self.P1)
                # self.a holds PReLU(conv1x1 output): the consumers skip the first PReLU (identity slope)
                if pending is not None and use_c1:
                    c = nat.Updc1Args()
                    c.M, c.K, c.B, c.H = M, K, B, H
                    c.x_hi, c.x_lo, c.y = self.x_hl[0].data_ptr(), self.x_hl[1].data_ptr(), self.y.data_ptr()
                    c.in_stat, c.in_P, c.in_count = pending["stat"].data_ptr(), pending["P"], float(K) * H
                    c.col_g, c.col_bias = pending["g"].data_ptr(), pending["b"].data_ptr()
                    c.w_hi, c.w_lo = pb["w1"][0].data_ptr(), pb["w1"][1].data_ptr()
                    c.a, c.out_stat, c.slope = self.a.data_ptr(), self.stat1.data_ptr(), params[k["a1"]].data_ptr()
                    self._keep.append(c)
                    emit("gemm_update_conv1x1", 2.0 * Rr * B * H + 2.0 * Rr * B, 12.0 * Rr * B + 4.0 * Rr * H,
                         lambda s, ref=C.byref(c): uc1(ref, s))
                    pending = None
                else:
                    if pending is not None:
                        emit_update(pending)
                        pending = None
                    emit("gemm_conv1x1", 2.0 * Rr * B * H, 4.0 * Rr * B + 4.0 * Rr * H,
                         self._gemm(Cin=B, Cout=H, a_hi=self.x_hl[0], a_lo=self.x_hl[1], w_hi=pb["w1"][0],
                                    w_lo=pb["w1"][1], out_f32=self.a, out_stat=self.norm,
                                    out_stat_buf=self.stat1, stat_prelu=params[k["a1"]], store_prelu=1))
                if self.fused[xb]:
                    f = nat.DwpwArgs()
                    f.M, f.K, f.H, f.B, f.dilation, f.causal = M, K, H, B, 2 ** xb, int(bool(cfg["causal"]))
                    f.a, f.prelu_in = self.a.data_ptr(), None
                    f.gamma_in, f.beta_in = g1.data_ptr(), b1.data_ptr()
                    f.in_stat, f.in_P = st1.data_ptr(), st1_P
                    f.wd, f.prelu_out = params[k["wd"]].data_ptr(), params[k["a2"]].data_ptr()
                    f.w_hi, f.w_lo = pb["w2"][0].data_ptr(), pb["w2"][1].data_ptr()
                    f.y, f.out_stat = self.y.data_ptr(), self.stat2f.data_ptr()
                    self._keep.append(f)
                    emit("gemm_dwpw_fused", 2.0 * Rr * B * H + 2.0 * Rr * H * cfg["P"], 4.0 * Rr * H + 4.0 * Rr * B,
                         lambda s, ref=C.byref(f): fw(ref, s))
                    pending = dict(P=1 if self.bn else self.P2f[xb], stat=self.unit if self.bn else self.stat2f,
                                   g=pb["w2"][2], b=pb["w2"][3])
                    continue
                d = nat.DwconvArgs()
                d.M, d.K, d.H, d.P, d.dilation, d.causal = M, K, H, cfg["P"], 2 ** xb, int(bool(cfg["causal"]))
                d.inp, d.prelu_in = self.a.data_ptr(), self.one.data_ptr()
                d.gamma_in, d.beta_in = g1.data_ptr(), b1.data_ptr()
                d.norm, d.in_stat, d.in_P = self.norm, st1.data_ptr(), st1_P
                d.wd, d.prelu_out = params[k["wd"]].data_ptr(), params[k["a2"]].data_ptr()
                d.out_pre = None
                d.out_hi, d.out_lo = self.q_hl[0].data_ptr(), self.q_hl[1].data_ptr()
                d.out_stat = self.stat2.data_ptr()
                self._keep.append(d)
                emit("dwconv", 2.0 * Rr * H * cfg["P"], 8.0 * Rr * H, lambda s, ref=C.byref(d): dw(ref, s))
                emit("gemm_pointwise", 2.0 * Rr * B * H, 4.0 * Rr * H + 8.0 * Rr * B,
                     self._gemm(Cin=H, Cout=B, a_hi=self.q_hl[0], a_lo=self.q_hl[1], w_hi=pb["w2"][0],
                                w_lo=pb["w2"][1], col_g=pb["w2"][2], col_bias=pb["w2"][3],
                                in_norm=self.norm, in_stat=self.unit if self.bn else self.stat2,
                                in_P=1 if self.bn else self.P2s[xb],
                                in_count=float(K) * H, resid_hi=self.x_hl[0], resid_lo=self.x_hl[1],
                                out_hi=self.x_hl[0], out_lo=self.x_hl[1]))
        if pending is not None and use_c1 and bool(lib.ctn_updc1_supported(B, C_ * N, self.norm)):
            # the mask conv has the shape of a conv1x1 (B -> C N channels): the last block's deferred update rides in
            # it as well (PReLU slope 1 = identity; its gLN statistics go to a scratch buffer nobody reads)
            c = nat.Updc1Args()
            c.M, c.K, c.B, c.H = M, K, B, C_ * N
            c.x_hi, c.x_lo, c.y = self.x_hl[0].data_ptr(), self.x_hl[1].data_ptr(), self.y.data_ptr()
            c.in_stat, c.in_P, c.in_count = pending["stat"].data_ptr(), pending["P"], float(K) * H
            c.col_g, c.col_bias = pending["g"].data_ptr(), pending["b"].data_ptr()
            c.w_hi, c.w_lo = prep.mask[0].data_ptr(), prep.mask[1].data_ptr()
            self.stat_mask = torch.zeros(M, lib.ctn_gemm_gln_partials(K, C_ * N), 2, dtype=torch.float32, device=self.device)
            c.a, c.out_stat, c.slope = self.score.data_ptr(), self.stat_mask.data_ptr(), self.one.data_ptr()
            self._keep.append(c)
            emit("gemm_update_mask", 2.0 * Rr * B * C_ * N + 2.0 * Rr * B, 12.0 * Rr * B + 4.0 * Rr * C_ * N,
                 lambda s, ref=C.byref(c): uc1(ref, s))
        else:
            if pending is not None:
                emit_update(pending)
            calls.append(self._gemm(Cin=B, Cout=C_ * N, a_hi=self.x_hl[0], a_lo=self.x_hl[1], w_hi=prep.mask[0],
                                    w_lo=prep.mask[1], out_f32=self.score))
            info.append(("gemm_mask", 2.0 * Rr * B * C_ * N, 4.0 * Rr * B + 4.0 * Rr * C_ * N))
        info.append(("decoder", 2.0 * Rr * C_ * N * L, 4.0 * Rr * N * (C_ + 1) + 4.0 * M * C_ * T))
        dec = lib.ctn_decoder_fwd
        a = (self.score.data_ptr(), self.w.data_ptr(), V.data_ptr(), M, K, N, L, C_, T,
             _MASK[cfg["mask_nonlinear"]], self.est.data_ptr())
        calls.append(lambda s, a=a: dec(*a, s))
        self.calls = calls
        self.call_info = info
        assert len(info) == len(calls)
        self.param_ptrs = tuple(p.data_ptr() for p in params.values())
        self.graph = None

    def launch(self):
        s = nat.stream()
        check = nat.check
        for c in self.calls:
            check(c(s))

    def launch_timed(self):
        """Runs the call list with a CUDA event pair around every call (profiling aid):
        returns [(kind, flops, bytes, milliseconds)]."""
        s = nat.stream()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(self.calls) + 1)]
        evs[0].record()
        for i, c in enumerate(self.calls):
            nat.check(c(s))
            evs[i + 1].record()
        torch.cuda.synchronize(self.device)
        return [(k, f, b, evs[i].elapsed_time(evs[i + 1])) for i, (k, f, b) in enumerate(self.call_info)]

    def run(self, use_graph, capture_now=False):
        """A shape is captured into a CUDA graph the SECOND time it runs (or at once with capture_now):
        warm-up + capture cost three passes, which a length seen once (evaluate.py / separate.py over a corpus
        of distinct utterance lengths) never earns back."""
        self.runs += 1
        if not use_graph or (self.graph is None and self.runs < 2 and not capture_now):
            self.launch()
            return
        if self.graph is None:
            self.launch()  # warm-up outside capture (first-use attribute setup, lazy module load)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=nat.capture_stream(self.device)):
                self.launch()
            self.graph = g
        self.graph.replay()


class SeparationEngine:
    """Inference forward for one ConvTasNet parameter set on one device."""

    def __init__(self, cfg):
        _check_cfg(cfg)
        self.cfg = dict(cfg)
        self.prepared = {}
        self.plans = PlanCache()
        self.micro_batch = 16
        self.use_graph = False
        # blocks with dilation <= this take the fused depthwise + pointwise kernel where the configuration
        # allows it (0 = never; the default covers every dilation)
        self.fuse_max_dilation = 1 << 30
        # after a fused block, fold its deferred residual update into the next block's conv1x1
        # (ctn_updc1_fwd: 203 us per 64 utterances against 99 + 158 for ctn_resid_update + ctn_gemm1x1)
        self.fuse_update = True
        # micro-batches in flight: with n > 1, consecutive micro-batches (independent utterances, separate
        # workspaces) are issued round-robin on n side streams, so one chain's HBM-bound kernels
        # (residual update, decoder) and kernel tails overlap the other chain's tensor-bound GEMMs
        self.n_streams = 1
        self._streams = {}
        self._host_slot = 0   # workspace slot of the next forward_host micro-batch (alternates across calls)
        self._staged = {}   # per device: aligned copies of unaligned (DataParallel replica) parameters
        self.launches = 0   # kernels of libctn_b200 launched by forward() so far

    def num_launches(self):
        """Kernel launches of one plan execution (encoder + bottleneck + 3 per block + mask + decoder;
        the fused path also has 3 per block: conv1x1, depthwise+pointwise, residual update; with
        fuse_update the update rides in the next block's conv1x1: 2 per block + 1)."""
        n = self.cfg["R"] * self.cfg["X"]
        plans = list(self.plans.values())
        if plans and plans[-1].calls is not None:
            return len(plans[-1].calls)
        return 4 + (2 * n + 1 if self.fuse_update and self.fuse_max_dilation >= 2 ** (self.cfg["X"] - 1) else 3 * n)

    def _plan(self, params, M, T, device, slot=0):
        key = (device.index, M, T, self.fuse_max_dilation, self.fuse_update, slot)
        prep = self.prepared.get(device.index)
        if prep is None:
            prep = self.prepared[device.index] = _Prepared(self.cfg, device)
        old_fp = prep.fingerprint
        prep.refresh(self.cfg, params)
        plan = self.plans.get(key)
        if plan is None:
            plan = self.plans.put(key, _Plan(self.cfg, M, T, device, self.fuse_max_dilation, self.fuse_update), device)
        ptrs = tuple(p.data_ptr() for p in params.values())
        if plan.calls is None or plan.param_ptrs != ptrs:
            plan.build(params, prep)
        elif old_fp != prep.fingerprint:
            pass  # same pointers, refreshed contents: call list and any captured graph stay valid
        return plan

    @torch.no_grad()
    def forward_host(self, mixture, params, out=None, join=True):
        """Bulk separation from HOST memory to HOST memory (BASELINE configs[3]): mixture is a CPU
        float32 [M, T] tensor (pinned for asynchronous copies), the result a pinned CPU [M, C, T] tensor.
        Micro-batch i+1 is copied in and micro-batch i-1 copied out on two copy streams while micro-batch
        i computes (two workspace slots), so PCIe time disappears behind the kernels instead of adding
        to them.  The slots alternate ACROSS calls as well and the per-slot events persist, so with
        join=False consecutive calls pipeline: the copy-out of one call overlaps the compute of the next even
        when a call is a single micro-batch (512 mixtures over 8 GPUs = one 64-utterance micro-batch per rank).
        join=True: stream-ordered on the current stream (synchronise it before reading `out`).
        join=False: returns (out, event); `out` is complete when the event, recorded on the copy-out stream,
        has completed, and the caller must not reuse `mixture` / `out` before that."""
        if mixture.is_cuda or mixture.dim() != 2 or mixture.dtype != torch.float32:
            raise RuntimeError("mixture must be a CPU float32 [M, T] tensor (use forward() for CUDA tensors)")
        M, T = mixture.shape
        cfg = self.cfg
        if T < cfg["L"]:
            raise RuntimeError("Calculated output size is too small: T=%d < L=%d" % (T, cfg["L"]))
        device = next(iter(params.values())).device
        if out is None:
            out = torch.empty(M, cfg["C"], T, dtype=torch.float32).pin_memory()
        elif out.is_cuda or tuple(out.shape) != (M, cfg["C"], T) or out.dtype != torch.float32 or not out.is_contiguous():
            raise RuntimeError("out must be a contiguous CPU float32 [M, C, T] tensor")
        if M == 0:
            return out if join else (out, None)
        with torch.cuda.device(device):
            params = aligned_params(self._staged, params, device)
            mb = max(1, int(self.micro_batch))
            chunks = [(m0, min(M, m0 + mb)) for m0 in range(0, M, mb)]
            first = self._host_slot
            self._host_slot = (first + len(chunks)) % 2
            plans = [self._plan(params, m1 - m0, T, device, slot=(first + i) % 2) for i, (m0, m1) in enumerate(chunks)]
            cur = torch.cuda.current_stream(device)
            if self.use_graph:
                for plan in plans:
                    if plan.graph is None:
                        plan.run(True, capture_now=True)      # capture (and one replay) before the pipeline starts
            side = self._streams.setdefault(("io", device.index), [])
            while len(side) < 2:
                side.append(torch.cuda.Stream(device=device))
            s_in, s_out = side
            if join:
                # stream-ordered call: the copies start behind whatever the caller has queued on its stream
                start = cur.record_event()
                s_in.wait_event(start)
                s_out.wait_event(start)
            # join=False (pipelined calls): the copy-in only waits for its own slot's last compute (plan.computed), so it
            # runs under the PREVIOUS call's kernels - with one micro-batch per call (8 GPUs x 64 mixtures) the wait on the
            # caller's stream above would put every call's 8 MB copy-in in front of its compute
            last = None
            for (m0, m1), plan in zip(chunks, plans):
                # plan.computed: its last compute has finished (mix may be overwritten);
                # plan.drained: its last result has left the device (est may be overwritten)
                with torch.cuda.stream(s_in):
                    if plan.computed is not None:
                        s_in.wait_event(plan.computed)
                    plan.mix.copy_(mixture[m0:m1], non_blocking=True)
                    loaded = s_in.record_event()
                cur.wait_event(loaded)
                if plan.drained is not None:
                    cur.wait_event(plan.drained)
                plan.run(self.use_graph)
                self.launches += len(plan.calls)
                plan.computed = cur.record_event()
                with torch.cuda.stream(s_out):
                    s_out.wait_event(plan.computed)
                    out[m0:m1].copy_(plan.est, non_blocking=True)
                    plan.drained = last = s_out.record_event()
            if not join:
                return out, last
            cur.wait_stream(s_out)
            cur.wait_stream(s_in)
        return out

    @torch.no_grad()
    def forward(self, mixture, params):
        nat.require_cuda(mixture, "mixture")
        if mixture.dim() != 2:
            raise RuntimeError("mixture must be [M, T], got shape %s" % (tuple(mixture.shape),))
        if mixture.dtype != torch.float32:
            raise RuntimeError("mixture must be float32, got %s" % mixture.dtype)
        M, T = mixture.shape
        cfg = self.cfg
        if T < cfg["L"]:
            raise RuntimeError("Calculated output size is too small: T=%d < L=%d" % (T, cfg["L"]))
        device = mixture.device
        out = torch.empty(M, cfg["C"], T, dtype=torch.float32, device=device)
        if M == 0:      # empty batch: the reference's modules return an empty [0, C, T] tensor
            return out
        with torch.cuda.device(device):
            params = aligned_params(self._staged, params, device)
            mb = max(1, int(self.micro_batch))
            n_str = max(1, min(int(self.n_streams), (M + mb - 1) // mb))
            if n_str == 1:
                for m0 in range(0, M, mb):
                    m1 = min(M, m0 + mb)
                    plan = self._plan(params, m1 - m0, T, device)
                    if plan.drained is not None:      # an un-joined forward_host may still be copying est out
                        torch.cuda.current_stream(device).wait_event(plan.drained)
                        plan.drained = None
                    plan.mix.copy_(mixture[m0:m1])
                    plan.run(self.use_graph)
                    self.launches += len(plan.calls)
                    out[m0:m1].copy_(plan.est)
                return out
            # several chains in flight: plans (and their CUDA graphs) are built on the caller's stream first,
            # then each micro-batch runs on its slot's side stream; the caller's stream joins at the end
            chunks = [(m0, min(M, m0 + mb)) for m0 in range(0, M, mb)]
            plans = [self._plan(params, m1 - m0, T, device, slot=i % n_str) for i, (m0, m1) in enumerate(chunks)]
            side = self._streams.setdefault(device.index, [])
            while len(side) < n_str:
                side.append(torch.cuda.Stream(device=device))
            cur = torch.cuda.current_stream(device)
            if self.use_graph:
                for plan in plans:
                    if plan.graph is None:
                        plan.run(True, capture_now=True)      # capture (and one replay) on the caller's stream
            ready = cur.record_event()
            for i, ((m0, m1), plan) in enumerate(zip(chunks, plans)):
                st = side[i % n_str]
                if i < n_str:
                    st.wait_event(ready)
                with torch.cuda.stream(st):
                    if plan.drained is not None:
                        st.wait_event(plan.drained)
                        plan.drained = None
                    plan.mix.copy_(mixture[m0:m1])
                    plan.run(self.use_graph)
                    out[m0:m1].copy_(plan.est)
                self.launches += len(plan.calls)
            for st in side[:n_str]:
                cur.wait_stream(st)
        return out
