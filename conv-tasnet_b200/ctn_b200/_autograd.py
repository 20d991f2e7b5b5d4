"""Training path: forward that keeps what backward needs + hand-written backward.

The reference gets its backward from autograd over ~3000 aten ops per step
(SURVEY.md 3.1); here ONE torch.autograd.Function wraps the whole separation
network.  Its forward is the inference call sequence with per-block saved
tensors (block input x as bf16 hi/lo, conv1x1 output a, depthwise output d and
the gLN statistics); its backward is a fixed list of libctn_b200 kernels
(csrc/backward.cu, wgrad.cu, gemm.cu) writing every parameter gradient into one
flat fp32 buffer -- the buffer the data-parallel allreduce then reduces in one
NCCL call (ctn_b200/_dist.py).
"""
import ctypes as C
import weakref

import torch

from . import _native as nat
from ._engine import _MASK, _NORM, PlanCache, _Plan, _Prepared, _check_cfg, aligned_params
from ._engine import BN_EPS
from ._layout import block_keys, frame_geometry, is_bn, param_spec

BN_MOMENTUM = 0.1   # nn.BatchNorm1d default (conv_tasnet.py:303 passes none)


class _PreparedBwd:
    """Transposed bf16 (hi, lo) weight copies: operands of the data-gradient GEMMs."""

    def __init__(self, cfg, device):
        N, B, H, C_ = cfg["N"], cfg["B"], cfg["H"], cfg["C"]
        bf = dict(dtype=torch.bfloat16, device=device)
        self.bott_t = (torch.empty(N, B, **bf), torch.empty(N, B, **bf))
        self.blocks = [dict(w1t=(torch.empty(B, H, **bf), torch.empty(B, H, **bf)),
                            w2t=(torch.empty(H, B, **bf), torch.empty(H, B, **bf)))
                       for _ in range(cfg["R"] * cfg["X"])]
        self.mask_t = (torch.empty(B, C_ * N, **bf), torch.empty(B, C_ * N, **bf))
        self.table = self.table_ptrs = None

    def refresh(self, cfg, params):
        ptrs = tuple(p.data_ptr() for p in params.values())
        if self.table is None or self.table_ptrs != ptrs:
            N, B, H, C_ = cfg["N"], cfg["B"], cfg["H"], cfg["C"]
            t = nat.PrepTable(self.bott_t[0].device)
            t.add(params["separator.network.1.weight"], B, N, None, None, 1, self.bott_t)
            i = 0
            for r in range(cfg["R"]):
                for x in range(cfg["X"]):
                    k = block_keys(r, x, cfg["causal"], cfg["norm_type"])
                    t.add(params[k["w1"]], H, B, None, None, 1, self.blocks[i]["w1t"])
                    t.add(params[k["w2"]], B, H, None, None, 1, self.blocks[i]["w2t"])
                    i += 1
            t.add(params["separator.network.3.weight"], C_ * N, B, None, None, 1, self.mask_t)
            self.table, self.table_ptrs = t, ptrs
        self.table.run()


class TrainPlan:
    """Workspaces and call lists of one (M, T) training step on one device."""

    fuse_forward = True     # class-wide switch (tests compare both forward routes)

    def __init__(self, cfg, M, T, device, overlap=True):
        _check_cfg(cfg, training=True)
        # BatchNorm1d blocks run through the gLN kernels with a per-channel affine and unit statistics (as in
        # inference, _engine._Plan); csrc/bn_train.cu adds the batch statistics and the backward's mean terms
        self.bn = is_bn(cfg["norm_type"])
        self.norm = nat.NORM_GLN if self.bn else _NORM[cfg["norm_type"]]
        self.cfg, self.M, self.T, self.device = cfg, M, T, device
        self.overlap = bool(overlap)                 # weight gradients on a side stream (see _launch_backward)
        self.side_stream = torch.cuda.Stream(device=device) if overlap else None
        N, L, B, H, C_ = cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["C"]
        self.hop, self.K, self.T_conv = frame_geometry(T, L)
        K = self.K
        Rr = M * K
        self.NB = NB = cfg["R"] * cfg["X"]
        lib = nat.lib()
        f32 = dict(dtype=torch.float32, device=device)
        bf = dict(dtype=torch.bfloat16, device=device)
        z = torch.zeros
        e = torch.empty
        # forward
        self.mix = e(M, T, **f32)
        self.w = e(Rr, N, **f32)
        self.w_hl = (e(Rr, N, **bf), e(Rr, N, **bf))
        self.rowstat0 = e(Rr, 2, **f32)
        self.x = e(Rr, B, **f32)
        self.x_hl = [(e(Rr, B, **bf), e(Rr, B, **bf)) for _ in range(NB + 1)]   # input of block b / of the mask conv
        self.a = [e(Rr, H, **f32) for _ in range(NB)]
        self.d = [e(Rr, H, **f32) for _ in range(NB)]
        self.q_hl = (e(Rr, H, **bf), e(Rr, H, **bf))
        if self.norm == nat.NORM_GLN:       # (sum, sumsq) partials per utterance
            self.P1 = lib.ctn_gemm_gln_partials(K, H)
            self.P2s = [lib.ctn_dwconv_gln_partials(K, H, cfg["P"], 2 ** x) for x in range(cfg["X"])]
            self.stat1 = [z(M, self.P1, 2, **f32) for _ in range(NB)]
            self.stat2 = [z(M, max(self.P2s), 2, **f32) for _ in range(NB)]
        else:                               # cLN: per frame; column-group partials / (mean, rstd)
            self.P1 = lib.ctn_gemm_cln_groups(H)
            self.P2s = [1] * cfg["X"]
            self.stat1 = [z(Rr, self.P1, 2, **f32) for _ in range(NB)]
            self.stat2 = [z(Rr, 2, **f32) for _ in range(NB)]
        self.score = e(Rr, C_ * N, **f32)
        # Forward through the inference kernels (ctn_updc1_fwd + ctn_dwpw_fwd, two launches per block instead of
        # three): gLN models whose shapes those kernels cover, from 8192 frames up (below, the plain
        # kernels are as quick).  They save what autograd needs: every block's input pair (out-of-place update), a
        # before the PReLU, and the depthwise output before the second PReLU.
        self.fused_fwd = (self.fuse_forward and not self.bn and self.norm == nat.NORM_GLN and Rr >= _Plan.FUSE_UPDATE_MIN_ROWS
                          and bool(lib.ctn_updc1_supported(B, H, self.norm))
                          and bool(lib.ctn_updc1_supported(B, C_ * N, self.norm))
                          and all(bool(lib.ctn_dwpw_supported(H, B, cfg["P"], 2 ** x, self.norm)) for x in range(cfg["X"])))
        if self.fused_fwd:
            self.y = e(Rr, B, **f32)
            self.P2s = [lib.ctn_dwpw_partials(K, 2 ** x) for x in range(cfg["X"])]     # the fused kernel's partial layout
            self.stat2 = [z(M, max(self.P2s), 2, **f32) for _ in range(NB)]
            self.stat_mask = z(M, lib.ctn_gemm_gln_partials(K, C_ * N), 2, **f32)
            self.one = torch.ones(1, **f32)
        # backward
        self.ds_hl = (e(Rr, C_ * N, **bf), e(Rr, C_ * N, **bf))
        self.dw_dec = e(Rr, N, **f32)
        self.g = e(Rr, B, **f32)
        self.g_hl = (e(Rr, B, **bf), e(Rr, B, **bf))
        # The weight-gradient GEMMs and the per-block reduction run on a side stream, up to two blocks behind the
        # dgrad chain (_launch_backward): what the chain writes and they read exists twice, by block parity.
        self.g_hl = (self.g_hl, (e(Rr, B, **bf), e(Rr, B, **bf)))
        self.hl_H = ((e(Rr, H, **bf), e(Rr, H, **bf)), (e(Rr, H, **bf), e(Rr, H, **bf)))   # da (hi, lo)
        # u (hi, lo), the operand of dW2: written by the depthwise backward where it can (emit_u), else by
        # ctn_norm_split on the side stream
        self.emit_u = bool(lib.ctn_bwd_dw_emits_u(H, cfg["P"], self.norm))
        self.u_hl = [(e(Rr, H, **bf), e(Rr, H, **bf)) for _ in range(2 if self.emit_u else 1)]
        self.du = e(Rr, H, **f32)
        self.dn1 = e(Rr, H, **f32)
        self.bstat = z(M if self.norm == nat.NORM_GLN else Rr, self.P1, 2, **f32)
        self.dn0 = e(Rr, N, **f32)
        self.dpre = e(Rr, N, **f32)
        self.bd_ctas, self.bd_rows = [], []          # per dilation: scal_part / chan_part rows per utterance
        for x in range(cfg["X"]):
            ctas, rows = C.c_int(), C.c_int()
            lib.ctn_bwd_dw_layout(M, K, H, cfg["P"], 2 ** x, self.norm, C.byref(ctas), C.byref(rows))
            self.bd_ctas.append(ctas.value)
            self.bd_rows.append(rows.value)
        self.chan_part = [e(M * max(self.bd_rows), (4 + cfg["P"]) * H, **f32) for _ in range(2)]
        self.scal_part = [e(M * max(self.bd_ctas), 4, **f32) for _ in range(2)]
        self.tiles_pn = lib.ctn_bwd_prelu_norm_tiles(K)
        self.alpha_part = [e(M * self.tiles_pn, **f32) for _ in range(2)]
        self.tiles_db = lib.ctn_decoder_bwd_tiles(K)
        self.dV_part = e(M * self.tiles_db, L * N, **f32)
        self.tiles_cb = lib.ctn_cln_bwd_tiles(K)
        self.cb_part = e(M * self.tiles_cb, 2 * N, **f32)
        self.tiles_ew = lib.ctn_encoder_wgrad_tiles(K)
        self.dU_part = e(M * self.tiles_ew, N * L, **f32)
        shapes = [(C_ * N, B), (B, H), (H, B), (B, N)]
        self.wg_part = e(max(lib.ctn_wgrad_splits(Rr, i, j) * i * j for i, j in shapes), **f32)
        self.wg_part2 = e(lib.ctn_wgrad_splits(Rr, H, B) * H * B, **f32)
        # flat gradient buffer in state_dict order
        self.spec = param_spec(N, L, B, H, cfg["P"], cfg["X"], cfg["R"], C_, cfg["causal"], cfg["norm_type"])
        self.offsets, off = {}, 0
        for key, (shape, kind) in self.spec.items():
            if kind.startswith("buf_"):       # BatchNorm1d buffers carry no gradient
                continue
            n = 1
            for s_ in shape:
                n *= s_
            self.offsets[key] = (off, n, shape)
            off += n
        self.flat_grad = z(off, **f32)
        if self.bn:
            self.unit = z(M, 1, 2, **f32)
            self.unit[:, 0, 1] = float(K * H)            # (sum, sumsq) = (0, count): mean 0, rstd 1
            self.zero_stat = z(M, 1, 2, **f32)           # backward correction sums: none (bn_bwd_center did it)
            self.zero_scal = z(M * max(self.bd_ctas), 4, **f32)
            self.bn_part = e(lib.ctn_bn_tiles(Rr) * 2 * H, **f32)
            self.bn_sums = z(2 * H, dtype=torch.float64, device=device)
            # per block and norm: scale, shift (the kernels' "gamma, beta"), mean, rstd (the backward's xh)
            self.bn_aff = [[tuple(e(H, **f32) for _ in range(4)) for _ in range(2)] for _ in range(NB)]
            self.bn_w2 = [(e(B, H, **bf), e(B, H, **bf), e(B, **f32), e(B, **f32)) for _ in range(NB)]
        self.bn_mode = None                              # (use running statistics, process group, global rows)
        self.fwd_calls = self.bwd_calls = None
        self.param_ptrs = None
        self._keep = []
        self.generation = 0
        # static endpoints so that each direction is ONE CUDA-graph launch (at the reference's batch
        # size of 3 the ~500 kernels of a step are latency-, not throughput-bound)
        self.est = e(M, C_, T, **f32)
        self.g_est = e(M, C_, T, **f32)
        self.fwd_graph = self.bwd_graph = None
        self.fwd_runs = self.bwd_runs = 0
        self.owner, self.done = None, True      # weakref to the autograd ctx whose activations are saved here

    # ------------------------------------------------------------------ helpers
    def grad_ptr(self, key):
        return self.flat_grad.data_ptr() + 4 * self.offsets[key][0]

    def grad_views(self, flat):
        return [flat[o:o + n].view(shape) for (o, n, shape) in self.offsets.values()]

    def _gemm(self, **kw):
        g = nat.GemmArgs()
        g.M, g.K = self.M, self.K
        for k, v in kw.items():
            setattr(g, k, v.data_ptr() if isinstance(v, torch.Tensor) else v)
        self._keep.append(g)
        fn, ref = nat.lib().ctn_gemm1x1, C.byref(g)
        return lambda s: fn(ref, s)

    def _reduce(self, part, n_part, part_stride, segs, accumulate=0):
        """segs: list of (src_off, len, src_stride, dst_ptr, dst_stride) over ONE partial buffer"""
        return self._reduce_multi([(part, n_part, part_stride) + tuple(sg) for sg in segs], accumulate)

    def _reduce_multi(self, segs, accumulate=0):
        """ONE launch summing the partials of several kernels.
        segs: list of (part tensor, n_part, part_stride, src_off, len, src_stride, dst_ptr, dst_stride)"""
        r = nat.ReduceArgs()
        r.nseg, r.accumulate = len(segs), accumulate
        for i, (part, n_part, ps, so, ln, ss, dp, ds) in enumerate(segs):
            r.seg_part[i], r.seg_n_part[i], r.seg_part_stride[i] = part.data_ptr(), n_part, ps
            r.src_off[i], r.len[i], r.src_stride[i], r.dst[i], r.dst_stride[i] = so, ln, ss, dp, ds
        self._keep.append(r)
        fn, ref = nat.lib().ctn_reduce_partials, C.byref(r)
        nbytes = sum(4.0 * n_part * ln for (_, n_part, _, _, ln, _, _, _) in segs)
        return self._tag(lambda s: fn(ref, s), "bwd_reduce_partials", 0.0, nbytes)

    def _wgrad(self, x_hl, y_hl, Ci, Cj, dst_key, part=None):
        """G = x^T y over all rows -> grad of `dst_key` ([Ci, Cj] row-major).
        Returns (launch, reduce segment); the segment is summed by a later _reduce_multi."""
        lib = nat.lib()
        part = self.wg_part if part is None else part
        Rr = self.M * self.K
        S = lib.ctn_wgrad_splits(Rr, Ci, Cj)
        a = (x_hl[0].data_ptr(), x_hl[1].data_ptr(), y_hl[0].data_ptr(), y_hl[1].data_ptr(), Rr, Ci, Cj,
             part.data_ptr(), 0)
        fn = lib.ctn_wgrad
        call = self._tag(lambda s, a=a: fn(*a, s), "bwd_wgrad", 2.0 * Rr * Ci * Cj, 4.0 * Rr * (Ci + Cj))
        return call, (part, S, Ci * Cj, 0, Ci * Cj, 1, self.grad_ptr(dst_key), 1)

    def _tag(self, call, kind, flops, nbytes):
        """Registers (kind, algorithmic FLOPs, algorithmic HBM bytes) of one launch for launch_timed()."""
        self.meta[call] = (kind, float(flops), float(nbytes))
        return call

    # ------------------------------------------------------------- call lists
    def _build_bn_forward_block(self, fwd, params, bufs, pb, i, xb, k, bn_count, use_running):
        """conv1x1 -> [batch statistics of PReLU(a) -> fold] -> depthwise with the folded affine ->
        [batch statistics of PReLU(d) -> fold -> W2 with the affine folded in] -> pointwise + residual."""
        cfg, lib, tag = self.cfg, nat.lib(), self._tag
        B, H, P, M, K = cfg["B"], cfg["H"], cfg["P"], self.M, self.K
        Rr = M * K

        def fold(j, aff):
            g, b = params[k["g%d" % j]], params[k["b%d" % j]]
            a = (self.bn_sums.data_ptr(), bn_count, g.data_ptr(), b.data_ptr(), BN_EPS, BN_MOMENTUM, int(use_running),
                 bufs[k["rm%d" % j]].data_ptr(), bufs[k["rv%d" % j]].data_ptr(), bufs[k["nb%d" % j]].data_ptr(), H,
                 aff[0].data_ptr(), aff[1].data_ptr(), aff[2].data_ptr(), aff[3].data_ptr())
            return lambda s, a=a: lib.ctn_bn_train_fold(*a, s)

        aff1, aff2 = self.bn_aff[i]
        fwd.append(tag(self._gemm(Cin=B, Cout=H, a_hi=self.x_hl[i][0], a_lo=self.x_hl[i][1], w_hi=pb["w1"][0],
                                  w_lo=pb["w1"][1], out_f32=self.a[i]),
                       "fwd_gemm_conv1x1", 2.0 * Rr * B * H, Rr * (4.0 * B + 4.0 * H)))
        if not use_running:
            fwd.extend(self._bn_stats(None, self.a[i], params[k["a1"]], aff1))
        fwd.append(fold(1, aff1))
        d = nat.DwconvArgs()
        d.M, d.K, d.H, d.P, d.dilation, d.causal = M, K, H, P, 2 ** xb, int(bool(cfg["causal"]))
        d.inp, d.prelu_in = self.a[i].data_ptr(), params[k["a1"]].data_ptr()
        d.gamma_in, d.beta_in = aff1[0].data_ptr(), aff1[1].data_ptr()
        d.norm, d.in_stat, d.in_P = nat.NORM_GLN, self.unit.data_ptr(), 1
        d.wd, d.prelu_out = params[k["wd"]].data_ptr(), params[k["a2"]].data_ptr()
        d.out_pre = self.d[i].data_ptr()
        d.out_hi, d.out_lo = self.q_hl[0].data_ptr(), self.q_hl[1].data_ptr()
        d.out_stat = self.stat2[i].data_ptr()        # (the gLN partials the kernel writes are not used)
        self._keep.append(d)
        dw = lib.ctn_dwconv_fwd
        fwd.append(tag(lambda s, ref=C.byref(d): dw(ref, s), "fwd_dwconv", 2.0 * Rr * H * P, 12.0 * Rr * H))
        if not use_running:
            fwd.extend(self._bn_stats(None, self.d[i], params[k["a2"]], aff2))
        fwd.append(fold(2, aff2))
        w2 = self.bn_w2[i]
        a = (params[k["w2"]].data_ptr(), B, H, aff2[0].data_ptr(), aff2[1].data_ptr(), 0, w2[0].data_ptr(),
             w2[1].data_ptr(), w2[2].data_ptr(), w2[3].data_ptr())
        fwd.append(lambda s, a=a: lib.ctn_weight_prep(*a, s))
        fwd.append(tag(self._gemm(Cin=H, Cout=B, a_hi=self.q_hl[0], a_lo=self.q_hl[1], w_hi=w2[0], w_lo=w2[1],
                                  col_g=w2[2], col_bias=w2[3], in_norm=nat.NORM_GLN, in_stat=self.unit, in_P=1,
                                  in_count=float(K) * H, resid=self.x, out_f32=self.x, out_hi=self.x_hl[i + 1][0],
                                  out_lo=self.x_hl[i + 1][1]),
                       "fwd_gemm_pointwise", 2.0 * Rr * B * H, Rr * (4.0 * H + 12.0 * B)))

    def _bn_backward_norm(self, dy, pre, slope, aff, bn_count, use_running, key_w, key_b):
        """BatchNorm backward around the gLN kernels: d weight = sum dy xh, d bias = sum dy (into the flat
        gradient) and, with batch statistics, dy -= mean(dy) + xh mean(dy xh) in place."""
        lib, Rr, H = nat.lib(), self.M * self.K, self.cfg["H"]
        calls = self._bn_stats(dy, pre, slope, aff)
        a = (dy.data_ptr(), pre.data_ptr(), slope.data_ptr(), aff[2].data_ptr(), aff[3].data_ptr(),
             self.bn_sums.data_ptr(), bn_count * self._bn_weight(), int(not use_running), Rr, H, self.grad_ptr(key_w),
             self.grad_ptr(key_b))
        calls.append(self._tag(lambda s, a=a: lib.ctn_bn_bwd_center(*a, s), "bn_bwd_center", 4.0 * Rr * H, 12.0 * Rr * H))
        return calls

    def _bn_weight(self):
        """n_local / n_global: the weight of this rank's local-mean loss in the global-mean loss."""
        rows_global = self.bn_mode[2]
        return float(self.M * self.K) / float(rows_global) if rows_global else 1.0

    def _bn_stats(self, dy, pre, slope, aff):
        """Column sums over all rows (ctn_bn_colsums) and, in a data-parallel job, their sum over the ranks."""
        lib, Rr, H = nat.lib(), self.M * self.K, self.cfg["H"]
        # backward sums are weighted by this rank's share of the global batch (see include/ctn_b200.h)
        w = self._bn_weight() if dy is not None else 1.0
        a = (ptr_(dy), pre.data_ptr(), slope.data_ptr(), aff[2].data_ptr() if dy is not None else None,
             aff[3].data_ptr() if dy is not None else None, Rr, H, w, self.bn_part.data_ptr(), self.bn_sums.data_ptr())
        calls = [self._tag(lambda s, a=a: lib.ctn_bn_colsums(*a, s), "bn_colsums", 3.0 * Rr * H,
                           (8.0 if dy is not None else 4.0) * Rr * H)]
        group = self.bn_mode[1]
        if group is not None:
            import torch.distributed as dist

            def allreduce(s, sums=self.bn_sums, group=group):
                dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)      # [2, H] fp64 on the current stream
                return 0
            calls.append(allreduce)
        return calls

    def _updc1(self, x_in, x_out, i_prev, w2_prev, w, Cout, out, out_stat, slope, kind):
        """x_out = x_in + deferred gLN-2 + residual of block i_prev (whose raw pointwise product is self.y);
        out = W x_out (before the PReLU: autograd needs it raw) with the gLN statistics of PReLU(out)."""
        cfg, M, K = self.cfg, self.M, self.K
        B, H = cfg["B"], cfg["H"]
        Rr = M * K
        c = nat.Updc1Args()
        c.M, c.K, c.B, c.H = M, K, B, Cout
        c.x_hi, c.x_lo, c.y = x_in[0].data_ptr(), x_in[1].data_ptr(), self.y.data_ptr()
        c.x_out_hi, c.x_out_lo, c.store_raw = x_out[0].data_ptr(), x_out[1].data_ptr(), 1
        c.in_stat, c.in_P, c.in_count = self.stat2[i_prev].data_ptr(), self.P2s[i_prev % cfg["X"]], float(K) * H
        c.col_g, c.col_bias = w2_prev[2].data_ptr(), w2_prev[3].data_ptr()
        c.w_hi, c.w_lo = w[0].data_ptr(), w[1].data_ptr()
        c.a, c.out_stat, c.slope = out.data_ptr(), out_stat.data_ptr(), slope.data_ptr()
        self._keep.append(c)
        fn = nat.lib().ctn_updc1_fwd
        return self._tag(lambda s, ref=C.byref(c): fn(ref, s), kind, 2.0 * Rr * B * Cout + 2.0 * Rr * B,
                         Rr * (12.0 * B + 4.0 * Cout))

    def _build_fused_forward_block(self, fwd, params, prep, i, xb, k):
        """Block i through the inference kernels: [deferred update of block i-1 +] conv1x1 (a before the PReLU,
        statistics of PReLU(a)), then PReLU / gLN-1 / depthwise / PReLU straight into the pointwise GEMM, whose raw
        product y and gLN-2 partials the NEXT block's (or the mask conv's) update consumes.  Saved for the backward:
        x_hl[i], a[i], d[i], stat1[i], stat2[i] - what the three-kernel route saves."""
        cfg, M, K, tag = self.cfg, self.M, self.K, self._tag
        B, H, P = cfg["B"], cfg["H"], cfg["P"]
        Rr = M * K
        pb = prep.blocks[i]
        if i == 0:      # the bottleneck wrote x_hl[0]: plain conv1x1
            fwd.append(tag(self._gemm(Cin=B, Cout=H, a_hi=self.x_hl[0][0], a_lo=self.x_hl[0][1], w_hi=pb["w1"][0],
                                      w_lo=pb["w1"][1], out_f32=self.a[0], out_stat=self.norm,
                                      out_stat_buf=self.stat1[0], stat_prelu=params[k["a1"]]),
                           "fwd_gemm_conv1x1", 2.0 * Rr * B * H, Rr * (4.0 * B + 4.0 * H)))
        else:
            fwd.append(self._updc1(self.x_hl[i - 1], self.x_hl[i], i - 1, prep.blocks[i - 1]["w2"], pb["w1"], H, self.a[i],
                                   self.stat1[i], params[k["a1"]], "fwd_gemm_update_conv1x1"))
        f = nat.DwpwArgs()
        f.M, f.K, f.H, f.B, f.dilation, f.causal = M, K, H, B, 2 ** xb, int(bool(cfg["causal"]))
        f.a, f.prelu_in = self.a[i].data_ptr(), params[k["a1"]].data_ptr()
        f.gamma_in, f.beta_in = params[k["g1"]].data_ptr(), params[k["b1"]].data_ptr()
        f.in_stat, f.in_P = self.stat1[i].data_ptr(), self.P1
        f.wd, f.prelu_out = params[k["wd"]].data_ptr(), params[k["a2"]].data_ptr()
        f.w_hi, f.w_lo = pb["w2"][0].data_ptr(), pb["w2"][1].data_ptr()
        f.y, f.out_stat, f.d_out = self.y.data_ptr(), self.stat2[i].data_ptr(), self.d[i].data_ptr()
        self._keep.append(f)
        fn = nat.lib().ctn_dwpw_fwd
        fwd.append(tag(lambda s, ref=C.byref(f): fn(ref, s), "fwd_gemm_dwpw_fused", 2.0 * Rr * B * H + 2.0 * Rr * H * P,
                       Rr * (8.0 * H + 4.0 * B)))

    def build(self, params, prep, prep_b, bn_mode=(False, None, None), bn_buffers=None):
        cfg = self.cfg
        N, L, B, H, C_, P = cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["C"], cfg["P"]
        M, K, T, NB = self.M, self.K, self.T, self.NB
        lib = nat.lib()
        self._keep = []
        self.meta = {}
        tag, Rr = self._tag, M * K
        GLN = self.norm                      # the block norm (gLN or cLN)
        is_gln = self.norm == nat.NORM_GLN
        bn = self.bn
        self.bn_mode = bn_mode
        if bn:
            use_running, _, rows_global = bn_mode
            bn_count = float(rows_global if rows_global else Rr)
        causal = int(bool(cfg["causal"]))
        U = params["encoder.conv1d_U.weight"]
        V = params["decoder.basis_signals.weight"]
        g0 = params["separator.network.0.gamma"]
        b0 = params["separator.network.0.beta"]
        count = float(K) * H

        # ---------------- forward (keeps x_hl[b], a[b], d[b], stat1[b], stat2[b]) ----------------
        fwd = []
        enc = lib.ctn_encoder_fwd
        ea = (self.mix.data_ptr(), U.data_ptr(), M, T, N, L, self.w.data_ptr(), self.w_hl[0].data_ptr(),
              self.w_hl[1].data_ptr(), self.rowstat0.data_ptr())
        fwd.append(tag(lambda s, a=ea: enc(*a, s), "fwd_encoder", 2.0 * Rr * N * L, 4.0 * M * T + 8.0 * Rr * N))
        fwd.append(tag(self._gemm(Cin=N, Cout=B, a_hi=self.w_hl[0], a_lo=self.w_hl[1], w_hi=prep.bott[0],
                                  w_lo=prep.bott[1], col_g=prep.bott[2], col_bias=prep.bott[3], in_norm=nat.NORM_CLN,
                                  in_stat=self.rowstat0, out_f32=self.x, out_hi=self.x_hl[0][0],
                                  out_lo=self.x_hl[0][1]), "fwd_gemm_bottleneck", 2.0 * Rr * N * B, Rr * (4.0 * N + 8.0 * B)))
        dw = lib.ctn_dwconv_fwd
        blocks = []
        i = 0
        for r in range(cfg["R"]):
            for xb in range(cfg["X"]):
                k = block_keys(r, xb, cfg["causal"], cfg["norm_type"])
                blocks.append((i, xb, k))
                pb = prep.blocks[i]
                if bn:
                    self._build_bn_forward_block(fwd, params, bn_buffers, pb, i, xb, k, bn_count, use_running)
                    i += 1
                    continue
                if self.fused_fwd:
                    self._build_fused_forward_block(fwd, params, prep, i, xb, k)
                    i += 1
                    continue
                fwd.append(tag(self._gemm(Cin=B, Cout=H, a_hi=self.x_hl[i][0], a_lo=self.x_hl[i][1], w_hi=pb["w1"][0],
                                          w_lo=pb["w1"][1], out_f32=self.a[i], out_stat=GLN,
                                          out_stat_buf=self.stat1[i], stat_prelu=params[k["a1"]]),
                               "fwd_gemm_conv1x1", 2.0 * Rr * B * H, Rr * (4.0 * B + 4.0 * H)))
                d = nat.DwconvArgs()
                d.M, d.K, d.H, d.P, d.dilation, d.causal = M, K, H, P, 2 ** xb, causal
                d.inp, d.prelu_in = self.a[i].data_ptr(), params[k["a1"]].data_ptr()
                d.gamma_in, d.beta_in = params[k["g1"]].data_ptr(), params[k["b1"]].data_ptr()
                d.norm, d.in_stat, d.in_P = GLN, self.stat1[i].data_ptr(), self.P1
                d.wd, d.prelu_out = params[k["wd"]].data_ptr(), params[k["a2"]].data_ptr()
                d.out_pre = self.d[i].data_ptr()
                d.out_hi, d.out_lo = self.q_hl[0].data_ptr(), self.q_hl[1].data_ptr()
                d.out_stat = self.stat2[i].data_ptr()
                self._keep.append(d)
                fwd.append(tag(lambda s, ref=C.byref(d): dw(ref, s), "fwd_dwconv", 2.0 * Rr * H * P, 12.0 * Rr * H))
                fwd.append(tag(self._gemm(Cin=H, Cout=B, a_hi=self.q_hl[0], a_lo=self.q_hl[1], w_hi=pb["w2"][0],
                                          w_lo=pb["w2"][1], col_g=pb["w2"][2], col_bias=pb["w2"][3], in_norm=GLN,
                                          in_stat=self.stat2[i], in_P=self.P2s[xb], in_count=count, resid=self.x,
                                          out_f32=self.x, out_hi=self.x_hl[i + 1][0], out_lo=self.x_hl[i + 1][1]),
                               "fwd_gemm_pointwise", 2.0 * Rr * B * H, Rr * (4.0 * H + 12.0 * B)))
                i += 1
        if self.fused_fwd:     # the last block's deferred update rides in the mask conv (PReLU slope 1 = identity)
            fwd.append(self._updc1(self.x_hl[NB - 1], self.x_hl[NB], NB - 1, prep.blocks[NB - 1]["w2"], prep.mask, C_ * N,
                                   self.score, self.stat_mask, self.one, "fwd_gemm_update_mask"))
        else:
            fwd.append(tag(self._gemm(Cin=B, Cout=C_ * N, a_hi=self.x_hl[NB][0], a_lo=self.x_hl[NB][1], w_hi=prep.mask[0],
                                      w_lo=prep.mask[1], out_f32=self.score),
                           "fwd_gemm_mask", 2.0 * Rr * B * C_ * N, Rr * (4.0 * B + 4.0 * C_ * N)))
        self.fwd_calls = fwd
        self.dec_args = (self.score.data_ptr(), self.w.data_ptr(), V.data_ptr(), M, K, N, L, C_, T,
                         _MASK[cfg["mask_nonlinear"]])

        # ---------------- backward ----------------
        bwd = []
        self.dec_bwd_args = (self.score.data_ptr(), self.w.data_ptr(), V.data_ptr(), M, K, N, L, C_, T,
                             _MASK[cfg["mask_nonlinear"]], self.ds_hl[0].data_ptr(), self.ds_hl[1].data_ptr(),
                             self.dw_dec.data_ptr(), self.dV_part.data_ptr())
        call, seg_mask = self._wgrad(self.ds_hl, self.x_hl[NB], C_ * N, B, "separator.network.3.weight")
        bwd.append(call)
        bwd.append(self._reduce_multi([
            (self.dV_part, M * self.tiles_db, L * N, 0, L * N, 1, self.grad_ptr("decoder.basis_signals.weight"), 1),
            seg_mask]))
        bwd.append(tag(self._gemm(Cin=C_ * N, Cout=B, a_hi=self.ds_hl[0], a_lo=self.ds_hl[1], w_hi=prep_b.mask_t[0],
                                  w_lo=prep_b.mask_t[1], out_f32=self.g, out_hi=self.g_hl[0][0], out_lo=self.g_hl[0][1]),
                       "bwd_gemm_mask_dgrad", 2.0 * Rr * B * C_ * N, Rr * (4.0 * C_ * N + 8.0 * B)))
        ns, bd, pn = lib.ctn_norm_split, lib.ctn_bwd_dwconv, lib.ctn_bwd_prelu_norm
        pn_cln = lib.ctn_bwd_prelu_norm_cln
        # Side-stream schedule (see _launch_backward).  Block number j of the backward, parity p = j & 1:
        #   main: dgrad (du = g W2) -> [wait blk:j-2] depthwise backward (also writes u) -> prelu/norm backward (da)
        #         -> [wait w2:j-1] dgrad (g' = da W1 + g)
        #   side, forked behind da: dW2 = g^T u  [mark w2:j] -> dW1 = da^T x -> reduction of the block  [mark blk:j]
        # g (hi, lo) alternates between two pairs (dW2 of block j reads the pair that block j+1's last GEMM rewrites),
        # u, da and the partial buffers of the main-stream kernels likewise; the GEMM partials are touched by the side
        # stream only.  (cLN, P != 3: u comes from ctn_norm_split on the side stream, forked at the top of the block.)
        for j, (i, xb, k) in enumerate(reversed(blocks)):
            pb = prep_b.blocks[i]
            par = j & 1
            g_in, g_out, da_hl = self.g_hl[par], self.g_hl[par ^ 1], self.hl_H[par]
            chan_part, scal_part, alpha_part = self.chan_part[par], self.scal_part[par], self.alpha_part[par]
            P2 = 1 if bn else self.P2s[xb]
            if bn:      # the kernels' (gamma, beta) = the folded BatchNorm affine; statistics = unit
                aff1, aff2 = self.bn_aff[i]
                g1_, b1_, g2_, b2_ = aff1[0], aff1[1], aff2[0], aff2[1]
                st1_, st2_, P1_ = self.unit, self.unit, 1
            else:
                g1_, b1_, g2_, b2_ = params[k["g1"]], params[k["b1"]], params[k["g2"]], params[k["b2"]]
                st1_, st2_, P1_ = self.stat1[i], self.stat2[i], self.P1
            u_hl = self.u_hl[par if self.emit_u else 0]
            call_w2, seg_w2 = self._wgrad(g_in, u_hl, B, H, k["w2"])              # dW2 = g^T u
            if not self.emit_u:
                a_ = (self.d[i].data_ptr(), params[k["a2"]].data_ptr(), g2_.data_ptr(),
                      b2_.data_ptr(), GLN, st2_.data_ptr(), P2, M, K, H, u_hl[0].data_ptr(), u_hl[1].data_ptr())
                bwd.append(_FORK)     # g (hi, lo) of this block is complete
                bwd.append((_SIDE, tag(lambda s, a=a_: ns(*a, s), "bwd_norm_split", 4.0 * Rr * H, 8.0 * Rr * H)))   # u (hi, lo)
                bwd.append((_SIDE, call_w2))
                bwd.append(_Mark("w2:%d" % j))
            if bn:
                bwd.append(tag(self._gemm(Cin=B, Cout=H, a_hi=g_in[0], a_lo=g_in[1], w_hi=pb["w2t"][0],
                                          w_lo=pb["w2t"][1], out_f32=self.du),            # du = g W2
                               "bwd_gemm_dgrad_pointwise", 2.0 * Rr * B * H, Rr * (4.0 * B + 4.0 * H)))
                bwd.extend(self._bn_backward_norm(self.du, self.d[i], params[k["a2"]], aff2, bn_count, use_running,
                                                  k["g2"], k["b2"]))
            else:
                bwd.append(tag(self._gemm(Cin=B, Cout=H, a_hi=g_in[0], a_lo=g_in[1], w_hi=pb["w2t"][0],
                                          w_lo=pb["w2t"][1], out_f32=self.du, out_stat=GLN, out_stat_buf=self.bstat,
                                          bw_pre=self.d[i], bw_gamma=params[k["g2"]], bw_slope=params[k["a2"]],
                                          bw_stat=self.stat2[i], bw_P=P2),                    # du = g W2 (+ S1, S2)
                               "bwd_gemm_dgrad_pointwise", 2.0 * Rr * B * H, Rr * (4.0 * B + 8.0 * H)))
            bwd.append(_Wait("blk:%d" % (j - 2)))   # its dW1 read this parity's da, its reduction this parity's partials
            b = nat.BwdDwArgs()
            b.M, b.K, b.H, b.P, b.dilation, b.causal = M, K, H, P, 2 ** xb, causal
            b.P1, b.P2, b.Pb = P1_, P2, (1 if bn else self.P1)
            b.du, b.d, b.a = self.du.data_ptr(), self.d[i].data_ptr(), self.a[i].data_ptr()
            b.stat1, b.stat2 = st1_.data_ptr(), st2_.data_ptr()
            b.bstat = self.zero_stat.data_ptr() if bn else self.bstat.data_ptr()
            b.g1, b.b1, b.g2 = g1_.data_ptr(), b1_.data_ptr(), g2_.data_ptr()
            b.wd, b.slope1, b.slope2 = params[k["wd"]].data_ptr(), params[k["a1"]].data_ptr(), params[k["a2"]].data_ptr()
            b.dn1, b.chan_part, b.scal_part = self.dn1.data_ptr(), chan_part.data_ptr(), scal_part.data_ptr()
            b.norm = self.norm
            if self.emit_u:     # u of this parity was last read by dW2 two blocks ago: behind the wait above
                b.b2, b.u_hi, b.u_lo = b2_.data_ptr(), u_hl[0].data_ptr(), u_hl[1].data_ptr()
            self._keep.append(b)
            bwd.append(tag(lambda s, ref=C.byref(b): bd(ref, s), "bwd_dwconv", 4.0 * Rr * H * P,
                           (20.0 if self.emit_u else 16.0) * Rr * H))
            if bn:
                bwd.extend(self._bn_backward_norm(self.dn1, self.a[i], params[k["a1"]], aff1, bn_count, use_running,
                                                  k["g1"], k["b1"]))
            if is_gln:
                a_ = (self.dn1.data_ptr(), self.a[i].data_ptr(), st1_.data_ptr(), P1_,
                      (self.zero_scal if bn else scal_part).data_ptr(), self.bd_ctas[xb], g1_.data_ptr(),
                      params[k["a1"]].data_ptr(), M, K, H, da_hl[0].data_ptr(), da_hl[1].data_ptr(),
                      alpha_part.data_ptr())
                bwd.append(tag(lambda s, a=a_: pn(*a, s), "bwd_prelu_norm", 6.0 * Rr * H, 12.0 * Rr * H))   # da (hi, lo)
            else:
                a_ = (self.dn1.data_ptr(), self.a[i].data_ptr(), self.stat1[i].data_ptr(), self.P1,
                      params[k["g1"]].data_ptr(), params[k["a1"]].data_ptr(), M, K, H,
                      da_hl[0].data_ptr(), da_hl[1].data_ptr(), alpha_part.data_ptr())
                bwd.append(tag(lambda s, a=a_: pn_cln(*a, s), "bwd_prelu_norm", 6.0 * Rr * H, 12.0 * Rr * H))
            call, seg_w1 = self._wgrad(da_hl, self.x_hl[i], H, B, k["w1"], self.wg_part2)   # dW1 = da^T x
            bwd.append(_FORK)
            if self.emit_u:
                bwd.append((_SIDE, call_w2))
                bwd.append(_Mark("w2:%d" % j))
            bwd.append((_SIDE, call))
            # every parameter gradient of the block in ONE reduction launch
            cp, cn, cs = chan_part, M * self.bd_rows[xb], (4 + P) * H
            # (BatchNorm: d weight / d bias of both norms come from ctn_bn_bwd_center, not from chan_part)
            segs = [] if bn else [
                (cp, cn, cs, 0, H, 1, self.grad_ptr(k["g2"]), 1), (cp, cn, cs, H, H, 1, self.grad_ptr(k["b2"]), 1),
                (cp, cn, cs, 2 * H, H, 1, self.grad_ptr(k["g1"]), 1),
                (cp, cn, cs, 3 * H, H, 1, self.grad_ptr(k["b1"]), 1)]
            segs += [(cp, cn, cs, (4 + jj) * H, H, 1, self.grad_ptr(k["wd"]) + 4 * jj, P) for jj in range(P)]
            segs.append((scal_part, M * self.bd_ctas[xb], 4, 2, 1, 1, self.grad_ptr(k["a2"]), 1))
            segs.append((alpha_part, M * self.tiles_pn, 1, 0, 1, 1, self.grad_ptr(k["a1"]), 1))
            if len(segs) + 2 <= nat.MAX_SEG:
                bwd.append((_SIDE, self._reduce_multi(segs + [seg_w2, seg_w1])))
            else:                                   # P > 3
                bwd.append((_SIDE, self._reduce_multi(segs[:nat.MAX_SEG])))
                bwd.append((_SIDE, self._reduce_multi(segs[nat.MAX_SEG:] + [seg_w2, seg_w1])))
            bwd.append(_Mark("blk:%d" % j))
            bwd.append(_Wait("w2:%d" % (j - 1)))    # it read the pair that is rewritten now
            bwd.append(tag(self._gemm(Cin=H, Cout=B, a_hi=da_hl[0], a_lo=da_hl[1], w_hi=pb["w1t"][0],
                                      w_lo=pb["w1t"][1], resid=self.g, out_f32=self.g, out_hi=g_out[0],
                                      out_lo=g_out[1]),                               # g = da W1 + g
                           "bwd_gemm_dgrad_conv1x1", 2.0 * Rr * B * H, Rr * (4.0 * H + 12.0 * B)))
        g_fin = self.g_hl[len(blocks) & 1]
        # bottleneck + leading cLN + encoder
        bwd.append(_JOIN)
        a_ = (self.w.data_ptr(), None, g0.data_ptr(), b0.data_ptr(), nat.NORM_CLN, self.rowstat0.data_ptr(), 1,
              M, K, N, self.w_hl[0].data_ptr(), self.w_hl[1].data_ptr())
        bwd.append(tag(lambda s, a=a_: ns(*a, s), "bwd_norm_split", 4.0 * Rr * N, 8.0 * Rr * N))   # n0 (hi, lo) over w_hl
        call, seg_wb = self._wgrad(g_fin, self.w_hl, B, N, "separator.network.1.weight")
        bwd.append(call)
        bwd.append(tag(self._gemm(Cin=B, Cout=N, a_hi=g_fin[0], a_lo=g_fin[1], w_hi=prep_b.bott_t[0],
                                  w_lo=prep_b.bott_t[1], out_f32=self.dn0),
                       "bwd_gemm_dgrad_bottleneck", 2.0 * Rr * B * N, Rr * (4.0 * B + 4.0 * N)))
        cb = lib.ctn_cln_bwd
        a_ = (self.dn0.data_ptr(), self.w.data_ptr(), self.rowstat0.data_ptr(), g0.data_ptr(),
              self.dw_dec.data_ptr(), M, K, N, self.dpre.data_ptr(), self.cb_part.data_ptr())
        bwd.append(tag(lambda s, a=a_: cb(*a, s), "bwd_cln", 8.0 * Rr * N, 16.0 * Rr * N))
        ew = lib.ctn_encoder_wgrad
        a_ = (self.dpre.data_ptr(), self.mix.data_ptr(), M, T, N, L, self.dU_part.data_ptr())
        bwd.append(tag(lambda s, a=a_: ew(*a, s), "bwd_encoder_wgrad", 2.0 * Rr * N * L, 4.0 * Rr * N + 4.0 * M * T))
        ncb = M * self.tiles_cb
        bwd.append(self._reduce_multi([
            seg_wb,
            (self.cb_part, ncb, 2 * N, 0, N, 1, self.grad_ptr("separator.network.0.gamma"), 1),
            (self.cb_part, ncb, 2 * N, N, N, 1, self.grad_ptr("separator.network.0.beta"), 1),
            (self.dU_part, M * self.tiles_ew, N * L, 0, N * L, 1, self.grad_ptr("encoder.conv1d_U.weight"), 1)]))
        self.bwd_calls = bwd
        self.param_ptrs = tuple(p.data_ptr() for p in params.values())
        self.fwd_graph = self.bwd_graph = None     # captured pointers may have changed
        self.fwd_runs = self.bwd_runs = 0          # the next run of each direction is eager again (see _capture)

    def _launch_forward(self):
        s, check, lib = nat.stream(), nat.check, nat.lib()
        for c in self.fwd_calls:
            check(c(s))
        check(lib.ctn_decoder_fwd(*self.dec_args, self.est.data_ptr(), s))

    def _launch_backward(self):
        """The dgrad chain runs on the current stream; weight-gradient GEMMs and the per-block parameter
        reduction (results needed only at the end of the step) run on a side stream, so they fill the SMs the
        chain's kernels leave idle at small batch.  Markers in bwd_calls: _FORK (side waits for everything issued
        to main so far), _Mark(name) (an event on the side stream), _Wait(name) (main waits for that event, if
        any), _JOIN (main waits for all side work).  With overlap off everything runs in list order on one stream
        (the list order is a valid serial order)."""
        s, check, lib = nat.stream(), nat.check, nat.lib()
        check(lib.ctn_decoder_bwd(self.g_est.data_ptr(), *self.dec_bwd_args, s))
        if not self.overlap:
            for c in self.bwd_calls:
                if _is_marker(c):
                    continue
                check((c[1] if isinstance(c, tuple) else c)(s))
            return
        main = torch.cuda.current_stream(self.device)
        side = self.side_stream
        side_ptr = side.cuda_stream
        pending = False     # side work issued since the last join
        marks = {}          # name -> event recorded on the side stream
        for c in self.bwd_calls:
            if c is _FORK:
                ev = torch.cuda.Event()
                ev.record(main)
                side.wait_event(ev)
            elif c is _JOIN:
                if pending:
                    ev = torch.cuda.Event()
                    ev.record(side)
                    main.wait_event(ev)
                    pending = False
                marks.clear()
            elif isinstance(c, _Mark):
                ev = torch.cuda.Event()
                ev.record(side)
                marks[c.name] = ev
            elif isinstance(c, _Wait):
                ev = marks.pop(c.name, None)
                if ev is not None:
                    main.wait_event(ev)
            elif isinstance(c, tuple):
                check(c[1](side_ptr))
                pending = True
            else:
                check(c(s))
        assert not pending

    def launch_timed(self):
        """Profiling aid: one forward + backward on ONE stream (no side-stream overlap) with a CUDA event pair
        around every launch.  Returns [(kind, algorithmic FLOPs, algorithmic bytes, milliseconds)]; est is
        fed back as its own gradient (the timing does not depend on the values)."""
        s, lib = nat.stream(), nat.lib()
        cfg, M, K, T = self.cfg, self.M, self.K, self.T
        N, L, C_ = cfg["N"], cfg["L"], cfg["C"]
        Rr = M * K
        dec_meta = ("fwd_decoder", 2.0 * Rr * C_ * N * L, 4.0 * Rr * N * (C_ + 1) + 4.0 * M * C_ * T)
        decb_meta = ("bwd_decoder", 6.0 * Rr * C_ * N * L, Rr * N * (4.0 * (C_ + 1) + 4.0 * C_ + 4.0) + 4.0 * M * C_ * T)
        seq = [(c, self.meta.get(c, ("other", 0.0, 0.0))) for c in self.fwd_calls]
        seq.append((lambda st: lib.ctn_decoder_fwd(*self.dec_args, self.est.data_ptr(), st), dec_meta))
        seq.append((lambda st: lib.ctn_decoder_bwd(self.est.data_ptr(), *self.dec_bwd_args, st), decb_meta))
        for c in self.bwd_calls:
            if _is_marker(c):
                continue
            c = c[1] if isinstance(c, tuple) else c
            seq.append((c, self.meta.get(c, ("other", 0.0, 0.0))))
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(seq) + 1)]
        evs[0].record()
        for i, (c, _) in enumerate(seq):
            nat.check(c(s))
            evs[i + 1].record()
        torch.cuda.synchronize(self.device)
        return [(m[0], m[1], m[2], evs[i].elapsed_time(evs[i + 1])) for i, (_, m) in enumerate(seq)]

    def _capture(self, launch):
        # No warm-up launch here: a plan is captured on its SECOND run (run_forward / run_backward), so its first,
        # eager run has already done the first-use set-up (kernel attributes, lazy module loading) -- and a
        # BatchNorm training forward must execute exactly once per call, because it moves the running statistics.
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=nat.capture_stream(self.device), capture_error_mode="thread_local"):
            launch()
        return g

    def run_forward(self, use_graph):
        """mix -> est (static buffers).  Returns the number of kernels launched."""
        self.fwd_runs += 1
        use_graph = use_graph and not (self.bn and self.bn_mode[1] is not None)   # all-reduces in the call list
        if not use_graph or (self.fwd_graph is None and self.fwd_runs < 2):
            self._launch_forward()     # a shape is captured the second time it is seen (see _engine._Plan.run)
        else:
            if self.fwd_graph is None:
                self.fwd_graph = self._capture(self._launch_forward)
            self.fwd_graph.replay()
        return len(self.fwd_calls) + 1

    def run_backward(self, use_graph):
        """g_est -> flat_grad (static buffers)."""
        self.bwd_runs += 1
        use_graph = use_graph and not (self.bn and self.bn_mode[1] is not None)
        if not use_graph or (self.bwd_graph is None and self.bwd_runs < 2):
            self._launch_backward()
        else:
            if self.bwd_graph is None:
                self.bwd_graph = self._capture(self._launch_backward)
            self.bwd_graph.replay()
        return sum(1 for c in self.bwd_calls if not _is_marker(c)) + 1


_FORK, _JOIN, _SIDE = "fork", "join", "side"   # markers in TrainPlan.bwd_calls


class _Mark:
    """bwd_calls marker: the side stream has finished everything issued to it so far (named)."""
    def __init__(self, name):
        self.name = name


class _Wait(_Mark):
    """bwd_calls marker: the main stream waits for the _Mark of that name (no-op if there is none)."""


def _is_marker(c):
    return c is _FORK or c is _JOIN or isinstance(c, _Mark)


class TrainEngine:
    """Per-model cache of training plans and prepared weights."""

    def __init__(self, cfg):
        self.cfg = dict(cfg)
        # bounded: solver.py:96-97 validates with gradients enabled, one new (M, T) per minibatch
        self.plans, self.prep, self.prep_b = PlanCache(keep=2), {}, {}
        self.launches = 0
        # data parallelism (ctn_b200/_dist.py): when set, every backward scales the flat gradient by
        # n_local / n_global and all-reduces (sum) it over this torch.distributed group, which equals
        # the gradient of the reference's global-mean loss
        self.use_graph = True
        self.overlap_wgrad = True      # backward: weight-gradient GEMMs + reductions on a side stream
        self.process_group = None
        self.global_batch = None       # utterances per step over all ranks (None = equal shards)
        self.last_flat_grad = None
        self.max_pending = 4           # activation sets per shape that may await their backward at one time
        self._staged = {}              # per device: aligned copies of DataParallel replica parameters

    def plan(self, params, M, T, device, bn_buffers=None, bn_running=False):
        """The plan of the first activation set of this shape that no live autograd graph is waiting on: a
        second forward before the first one's backward (two micro-batches whose losses are summed, or a
        train-mode forward whose graph is kept) gets its own saved activations, as autograd gives the
        reference; sets whose graph was dropped without a backward are reused."""
        for slot in range(self.max_pending):
            plan = self._plan(params, M, T, device, slot, bn_buffers, bn_running)
            owner = plan.owner() if plan.owner is not None else None
            if owner is None or plan.done:
                return plan
        raise RuntimeError("ConvTasNet.forward: %d forwards of shape (%d, %d) are awaiting backward; free their "
                           "graphs or raise train_engine().max_pending" % (self.max_pending, M, T))

    def _plan(self, params, M, T, device, slot=0, bn_buffers=None, bn_running=False):
        di = device.index
        if di not in self.prep:
            self.prep[di] = _Prepared(self.cfg, device, fold_bn=False)
            self.prep_b[di] = _PreparedBwd(self.cfg, device)
        old = self.prep[di].fingerprint
        self.prep[di].refresh(self.cfg, params)
        if old != self.prep[di].fingerprint:
            self.prep_b[di].refresh(self.cfg, params)
        key = (di, M, T, bool(self.overlap_wgrad), slot)
        plan = self.plans.get(key)
        if plan is None:
            plan = self.plans.put(key, TrainPlan(self.cfg, M, T, device, self.overlap_wgrad), device)
        ptrs = tuple(p.data_ptr() for p in params.values())
        bn_mode = (False, None, None)
        if plan.bn:
            # batch statistics span the GLOBAL batch: rows of all ranks (equal shards unless global_batch is set)
            group = self.process_group
            rows = None
            if group is not None and not bn_running:
                import torch.distributed as dist
                n_global = self.global_batch if self.global_batch is not None else M * dist.get_world_size(group)
                rows = n_global * plan.K
            bn_mode = (bool(bn_running), None if bn_running else group, rows)
            ptrs = ptrs + tuple(b.data_ptr() for b in bn_buffers.values())
        if plan.fwd_calls is None or plan.param_ptrs != ptrs or plan.bn_mode != bn_mode:
            plan.build(params, self.prep[di], self.prep_b[di], bn_mode, bn_buffers)
            plan.param_ptrs = ptrs
        return plan


class _Separate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, mixture, *param_list):
        keys = model._keys
        params = {k: p.detach() for k, p in zip(keys, param_list)}
        M, T = mixture.shape
        dev = mixture.device
        eng = model.train_engine()
        with torch.cuda.device(dev):
            params = aligned_params(eng._staged, params, dev)
            bufs = model._bn_buffers() if is_bn(model.norm_type) else None
            plan = eng.plan(params, M, T, dev, bufs, bn_running=not model.training)
            plan.mix.copy_(mixture)
            eng.launches += plan.run_forward(eng.use_graph)
            est = plan.est.clone()
            if bufs and model.training:     # the fold kernels updated the running statistics through raw pointers
                torch.autograd.graph.increment_version(list(bufs.values()))
        plan.generation += 1
        plan.owner, plan.done = weakref.ref(ctx), False
        ctx.plan, ctx.generation, ctx.engine = plan, plan.generation, eng
        ctx.n_params = len(param_list)
        return est

    @staticmethod
    def backward(ctx, g_est):
        plan = ctx.plan
        if plan.generation != ctx.generation:
            raise RuntimeError("backward through a ConvTasNet graph whose saved activations were released "
                               "(a second backward through the same graph)")
        eng = ctx.engine
        with torch.cuda.device(plan.device):
            plan.g_est.copy_(g_est)
            eng.launches += plan.run_backward(eng.use_graph)
            # one private copy per backward: autograd may adopt the returned views as .grad
            flat = plan.flat_grad.clone()
            plan.done = True           # the activation set is free for the next forward of this shape
            if eng.process_group is not None:
                from . import _dist
                _dist.reduce_flat_grad(flat, plan.M, eng.global_batch, eng.process_group)
        eng.last_flat_grad = flat
        return (None, None) + tuple(plan.grad_views(flat))


def ptr_(t):
    return None if t is None else t.data_ptr()


def separate(model, mixture, params):
    if mixture.requires_grad:
        raise NotImplementedError("gradient with respect to the mixture is not implemented")
    if mixture.dim() != 2 or mixture.dtype != torch.float32:
        raise RuntimeError("mixture must be float32 [M, T]")
    if mixture.shape[1] < model.L:
        raise RuntimeError("Calculated output size is too small: T=%d < L=%d" % (mixture.shape[1], model.L))
    plist = [params[k] for k in model._keys]      # the parameters; BatchNorm1d buffers are not autograd inputs
    if not torch.is_grad_enabled() or not any(p.requires_grad for p in plist):
        # BatchNorm1d in train() mode without a graph (torch.no_grad): batch statistics + buffer update only
        return _Separate.forward(_NoCtx(), model, mixture, *plist)
    return _Separate.apply(model, mixture, *plist)


class _NoCtx:
    """Stands in for the autograd context when the training-mode forward runs without a graph."""
