"""Chunked (streaming) separation for causal cLN models.

A causal ConvTasNet with channel-wise layer norm (`causal=True, norm_type="cLN"`,
conv_tasnet.py:175-176, 289, 307-329) has no operation that looks ahead in time and no
statistic that spans the utterance, so the output for samples [0, t) depends on the
mixture up to t + hop only.  The reference has no streaming path (data.py:101-104 drops long
audio; separate.py runs whole files); this module runs the SAME kernels as
`SeparationEngine` on fixed-size chunks of `chunk_frames` encoder frames and carries, per
TemporalBlock, the (P-1)*2^x frames of depthwise-convolution input that the next chunk's
first outputs reach back to (conv_tasnet.py:247-252 with Chomp1d), plus one hop of
mixture samples and one frame of (encoder output, mask score) for the overlap-add
(utils.py:28-46).  Concatenating what `feed()` and `flush()` return reproduces
`ConvTasNet.forward` of the whole stream.

Layout: every TemporalBlock b has its own conv1x1-output buffer `a_b` with
ctx_b = (P-1)*2^x context rows in front of the chunk's rows; the 1x1 GEMMs write / read
at a row offset (they are pointwise in time), the depthwise kernel runs over
ctx_b + chunk rows and its first ctx_b output rows are never read.  The first chunk of a
stream has no context: the reference zero-pads the NORMALISED depthwise input, so the
first chunk runs the depthwise kernel on the chunk rows alone.  Steady-state chunks are
one CUDA graph: state shift (two table-driven copy launches), encoder, 3 kernels per block,
mask GEMM, decoder.
"""
import ctypes as C

import torch

from . import _native as nat
from ._engine import _MASK, _Prepared, _check_cfg
from ._layout import block_keys


class StreamingSeparator:
    """One audio stream on one device.

        sep = model.streaming(chunk_frames=256)
        for piece in pieces:                  # 1-D float32 CUDA tensors of any length
            out = sep.feed(piece)             # [C, n] separated samples (n may be 0)
        tail = sep.flush()                    # the rest, up to the stream length

    Algorithmic latency: one chunk (chunk_frames * L/2 samples) + L/2 samples.
    """

    def __init__(self, model, chunk_frames=256, use_graph=True):
        cfg = model._cfg()
        _check_cfg(cfg)
        if not (cfg["causal"] and cfg["norm_type"] == "cLN"):
            raise ValueError("streaming separation is exact only for causal=True, norm_type='cLN' "
                             "(gLN statistics span the whole utterance; non-causal blocks look ahead)")
        if cfg["L"] % 2:
            raise ValueError("streaming separation needs an even filter length L")
        if chunk_frames < 1:
            raise ValueError("chunk_frames must be positive")
        self.model, self.cfg = model, cfg
        self.Kc = int(chunk_frames)
        self.use_graph = bool(use_graph)
        params = {k: p.detach() for k, p in model._params().items()}
        dev = next(iter(params.values())).device
        if dev.type != "cuda":
            raise RuntimeError("the model must be on a CUDA device: this implementation has no CPU path")
        self.device = dev
        self.params = params
        self.prep = _Prepared(cfg, dev)
        self.launches = 0
        self._alloc()
        self._build()
        self.reset()

    # ------------------------------------------------------------------ buffers
    def _alloc(self):
        cfg, Kc, dev = self.cfg, self.Kc, self.device
        N, L, B, H, C_ = cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["C"]
        hop = self.hop = L // 2
        f32 = dict(dtype=torch.float32, device=dev)
        bf = dict(dtype=torch.bfloat16, device=dev)
        lib = nat.lib()
        self.P1 = lib.ctn_gemm_cln_groups(H)
        self.ctx = [(cfg["P"] - 1) * 2 ** x for x in range(cfg["X"])] * cfg["R"]
        cmax = max(self.ctx)
        self.new = torch.zeros(Kc * hop, **f32)            # staging for the chunk's new samples
        self.inbuf = torch.zeros((Kc + 1) * hop, **f32)    # [previous hop | new samples]
        self.w = torch.zeros(1 + Kc, N, **f32)             # row 0: last frame of the previous chunk
        self.w_hl = (torch.empty(Kc, N, **bf), torch.empty(Kc, N, **bf))
        self.rowstat0 = torch.empty(Kc, 2, **f32)
        self.x_hl = (torch.empty(Kc, B, **bf), torch.empty(Kc, B, **bf))
        self.a = [torch.zeros(c + Kc, H, **f32) for c in self.ctx]
        self.stat1 = [torch.zeros(c + Kc, self.P1, 2, **f32) for c in self.ctx]
        self.q_hl = (torch.empty(cmax + Kc, H, **bf), torch.empty(cmax + Kc, H, **bf))
        self.stat2 = torch.zeros(cmax + Kc, 2, **f32)
        self.score = torch.zeros(1 + Kc, C_ * N, **f32)    # row 0: last frame of the previous chunk
        self.est_first = torch.empty(C_, (Kc + 1) * hop, **f32)
        self.est = torch.empty(C_, (Kc + 2) * hop, **f32)
        self.one = torch.ones(1, **f32)

    def _gemm(self, rows, keep, **kw):
        g = nat.GemmArgs()
        g.M, g.K = 1, rows
        for k, v in kw.items():
            setattr(g, k, v if isinstance(v, (int, float)) or v is None else int(v))
        keep.append(g)
        fn = nat.lib().ctn_gemm1x1
        ref = C.byref(g)
        return lambda s: fn(ref, s)

    @staticmethod
    def _row(t, r):
        """Device address of row r of a 2-D/3-D contiguous tensor."""
        return t.data_ptr() + r * t.stride(0) * t.element_size()

    def _body(self, hist):
        """Call list encoder .. mask GEMM for one chunk that has `hist` frames of the stream before it
        (0: first chunk, no left context; fewer than a block's context: the depthwise kernel starts at
        the first frame of the stream, because the reference zero-pads the NORMALISED input)."""
        cfg, Kc, p, prep = self.cfg, self.Kc, self.params, self.prep
        N, L, B, H, C_ = cfg["N"], cfg["L"], cfg["B"], cfg["H"], cfg["C"]
        lib = nat.lib()
        keep, calls = [], []
        row = self._row
        enc = lib.ctn_encoder_fwd
        a = (self.inbuf.data_ptr(), p["encoder.conv1d_U.weight"].data_ptr(), 1, (Kc + 1) * self.hop, N, L,
             row(self.w, 1), self.w_hl[0].data_ptr(), self.w_hl[1].data_ptr(), self.rowstat0.data_ptr())
        calls.append(lambda s, a=a: enc(*a, s))
        calls.append(self._gemm(Kc, keep, Cin=N, Cout=B, a_hi=self.w_hl[0].data_ptr(), a_lo=self.w_hl[1].data_ptr(),
                                w_hi=prep.bott[0].data_ptr(), w_lo=prep.bott[1].data_ptr(),
                                col_g=prep.bott[2].data_ptr(), col_bias=prep.bott[3].data_ptr(),
                                in_norm=nat.NORM_CLN, in_stat=self.rowstat0.data_ptr(),
                                out_hi=self.x_hl[0].data_ptr(), out_lo=self.x_hl[1].data_ptr()))
        dw = lib.ctn_dwconv_fwd
        i = 0
        for r in range(cfg["R"]):
            for xb in range(cfg["X"]):
                k = block_keys(r, xb, True)
                pb = prep.blocks[i]
                ctx = self.ctx[i]
                ab, sb = self.a[i], self.stat1[i]
                i += 1
                calls.append(self._gemm(Kc, keep, Cin=B, Cout=H, a_hi=self.x_hl[0].data_ptr(),
                                        a_lo=self.x_hl[1].data_ptr(), w_hi=pb["w1"][0].data_ptr(),
                                        w_lo=pb["w1"][1].data_ptr(), out_f32=row(ab, ctx), out_stat=nat.NORM_CLN,
                                        out_stat_buf=row(sb, ctx), stat_prelu=p[k["a1"]].data_ptr(), store_prelu=1))
                off = max(0, ctx - hist)      # rows in front of the stream's first frame are not input
                d = nat.DwconvArgs()
                d.M, d.K, d.H, d.P, d.dilation, d.causal = 1, Kc + ctx - off, H, cfg["P"], 2 ** xb, 1
                d.inp, d.prelu_in = row(ab, off), self.one.data_ptr()
                d.gamma_in, d.beta_in = p[k["g1"]].data_ptr(), p[k["b1"]].data_ptr()
                d.norm, d.in_stat, d.in_P = nat.NORM_CLN, row(sb, off), self.P1
                d.wd, d.prelu_out = p[k["wd"]].data_ptr(), p[k["a2"]].data_ptr()
                d.out_pre = None
                d.out_hi, d.out_lo = row(self.q_hl[0], off), row(self.q_hl[1], off)
                d.out_stat = row(self.stat2, off)
                keep.append(d)
                calls.append(lambda s, ref=C.byref(d): dw(ref, s))
                calls.append(self._gemm(Kc, keep, Cin=H, Cout=B, a_hi=row(self.q_hl[0], ctx), a_lo=row(self.q_hl[1], ctx),
                                        w_hi=pb["w2"][0].data_ptr(), w_lo=pb["w2"][1].data_ptr(),
                                        col_g=pb["w2"][2].data_ptr(), col_bias=pb["w2"][3].data_ptr(),
                                        in_norm=nat.NORM_CLN, in_stat=row(self.stat2, ctx), in_P=1,
                                        in_count=float(H), resid_hi=self.x_hl[0].data_ptr(),
                                        resid_lo=self.x_hl[1].data_ptr(), out_hi=self.x_hl[0].data_ptr(),
                                        out_lo=self.x_hl[1].data_ptr()))
        calls.append(self._gemm(Kc, keep, Cin=B, Cout=C_ * N, a_hi=self.x_hl[0].data_ptr(), a_lo=self.x_hl[1].data_ptr(),
                                w_hi=prep.mask[0].data_ptr(), w_lo=prep.mask[1].data_ptr(), out_f32=row(self.score, 1)))
        return calls, keep

    def _decoder(self, first, frames, T_out, est):
        """Mask non-linearity, basis GEMM and overlap-add over `frames` rows, starting at the carried
        previous frame (row 0) unless this is the first chunk of the stream."""
        cfg = self.cfg
        r0 = 1 if first else 0
        V = self.params["decoder.basis_signals.weight"]
        a = (self._row(self.score, r0), self._row(self.w, r0), V.data_ptr(), 1, frames, cfg["N"], cfg["L"], cfg["C"],
             T_out, _MASK[cfg["mask_nonlinear"]], est.data_ptr())
        dec = nat.lib().ctn_decoder_fwd
        return lambda s, a=a: dec(*a, s)

    def _build(self):
        Kc, hop = self.Kc, self.hop
        self.prep.refresh(self.cfg, self.params)
        self.body_first, self._keep1 = self._body(0)
        self.body_steady, self._keep2 = self._body(max(self.ctx))
        self.dec_first = self._decoder(True, Kc, (Kc + 1) * hop, self.est_first)
        self.dec_steady = self._decoder(False, Kc + 1, (Kc + 2) * hop, self.est)
        self._build_shift()
        self.graph_steady = None
        self._steady_warm = False

    # ------------------------------------------------------------------ state
    def reset(self):
        """Forget the stream: the next feed() starts a new one."""
        self.fifo = []          # fed pieces not yet consumed (device tensors)
        self.fifo_n = 0
        self.first = True       # no chunk of this stream has run yet
        self.hist = 0           # frames of this stream processed so far
        self.consumed = 0       # samples of the stream copied into chunks so far
        self.emitted = 0        # samples of the stream returned so far

    def _build_shift(self):
        """Copy tables (ctn_copy_table: one launch each) that move what the next chunk needs from the end
        of the previous chunk's buffers to their fronts: one hop of mixture, one frame of encoder output
        and score, ctx_b rows of conv1x1 output and statistics per block.  Where the chunk is shorter
        than a block's context the two ranges overlap, so those go through a scratch buffer: table 1 =
        direct moves + moves into scratch, table 2 = moves out of scratch + the chunk's new samples (which
        land where table 1 read the old hop from)."""
        Kc, hop = self.Kc, self.hop
        t1 = [(self.inbuf[:hop], self.inbuf[Kc * hop:]), (self.w[0], self.w[Kc]), (self.score[0], self.score[Kc])]
        t2 = [(self.inbuf[hop:], self.new)]
        self._scratch = []
        for ab, sb, c in zip(self.a, self.stat1, self.ctx):
            for buf in (ab, sb):
                if Kc >= c:
                    t1.append((buf[:c], buf[Kc:]))
                else:
                    tmp = torch.empty_like(buf[:c])
                    self._scratch.append(tmp)
                    t1.append((tmp, buf[Kc:]))
                    t2.append((buf[:c], tmp))
        self.shift1 = nat.CopyTable(self.device, t1)
        self.shift2 = nat.CopyTable(self.device, t2)

    def _shift_state(self):
        self.shift1.run()
        self.shift2.run()
        self.launches += 2

    def _launch(self, calls):
        s = nat.stream()
        for c in calls:
            nat.check(c(s))
        self.launches += len(calls)

    def _chunk_first(self, with_decoder=True):
        self._launch(self.body_first + ([self.dec_first] if with_decoder else []))

    def _chunk_steady(self, with_decoder=True):
        self._shift_state()
        body = self.body_steady
        if self.hist < max(self.ctx):     # ramp-up: some block's context still reaches before the stream start
            body, self._keep3 = self._body(self.hist)
        self._launch(body + ([self.dec_steady] if with_decoder else []))

    def _run(self, which):
        """Runs one chunk.  The first chunk of a stream and the first steady-state chunk run eagerly
        (which also performs the kernels' one-time setup); from the second steady-state chunk on, the
        chunk (state shifts + kernels) is one CUDA graph replay."""
        if which == "first":
            self._chunk_first()
            return
        if self.hist < max(self.ctx):
            self._chunk_steady()
            return
        if not self.use_graph or not self._steady_warm:
            self._chunk_steady()
            self._steady_warm = True
            return
        if self.graph_steady is None:
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            n = self.launches
            with torch.cuda.graph(g, stream=nat.capture_stream(self.device)):   # records only; replayed below
                self._chunk_steady()
            self.launches = n
            self.graph_steady = g
        self.graph_steady.replay()
        self.launches += len(self.body_steady) + 3       # two copy tables, the body, the decoder

    def _take(self, n, dst):
        """Moves the next n fed samples into dst[:n] (device-to-device copies)."""
        got = 0
        while got < n:
            piece = self.fifo[0]
            m = min(n - got, piece.numel())
            dst[got:got + m].copy_(piece[:m])
            got += m
            if m == piece.numel():
                self.fifo.pop(0)
            else:
                self.fifo[0] = piece[m:]
        self.fifo_n -= n
        self.consumed += n

    # ------------------------------------------------------------------ API
    @torch.no_grad()
    def feed(self, samples):
        """samples: 1-D float32 CUDA tensor (any length).  Returns [C, n] separated samples that are
        final (n is a multiple of chunk_frames * L/2, possibly 0)."""
        nat.require_cuda(samples, "samples")
        if samples.dim() != 1 or samples.dtype != torch.float32:
            raise RuntimeError("samples must be a 1-D float32 tensor")
        if samples.device != self.device:
            raise RuntimeError("samples are on %s, the model on %s" % (samples.device, self.device))
        self.prep.refresh(self.cfg, self.params)       # no-op unless the weights were updated in place
        if samples.numel():
            self.fifo.append(samples.contiguous().clone())
            self.fifo_n += samples.numel()
        Kc, hop = self.Kc, self.hop
        outs = []
        with torch.cuda.device(self.device):
            while True:
                if self.first and self.fifo_n >= (Kc + 1) * hop:
                    self._take((Kc + 1) * hop, self.inbuf)
                    self._run("first")
                    outs.append(self.est_first[:, :Kc * hop].clone())
                    self.first = False
                elif not self.first and self.fifo_n >= Kc * hop:
                    self._take(Kc * hop, self.new)
                    self._run("steady")
                    outs.append(self.est[:, hop:(Kc + 1) * hop].clone())
                else:
                    break
                self.emitted += Kc * hop
                self.hist += Kc
        if not outs:
            return torch.empty(self.cfg["C"], 0, dtype=torch.float32, device=self.device)
        return outs[0] if len(outs) == 1 else torch.cat(outs, dim=1)

    @torch.no_grad()
    def flush(self):
        """Ends the stream: returns the remaining [C, n] samples so that everything returned since
        the last reset() has the stream's length T, exactly as ConvTasNet.forward would (frames that
        do not fit are dropped and the tail is zero-filled, conv_tasnet.py:56-59), then resets."""
        cfg, Kc, hop, L = self.cfg, self.Kc, self.hop, self.cfg["L"]
        T = self.consumed + self.fifo_n
        f32 = dict(dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            n_rem = self.fifo_n
            if self.first:
                if T < L:
                    self.reset()
                    raise RuntimeError("Calculated output size is too small: T=%d < L=%d" % (T, L))
                frames = (T - L) // hop + 1
                self.inbuf.zero_()
                self._take(n_rem, self.inbuf)
                self._chunk_first(with_decoder=False)      # causal: the zero-padded frames cannot reach back
                est = torch.empty(cfg["C"], T, **f32)
                self._launch([self._decoder(True, frames, T, est)])
                out = est
            else:
                frames = n_rem // hop                      # new complete frames
                t_loc = n_rem + 2 * hop                    # samples from the carried frame's start to the end
                self.new.zero_()
                self._take(n_rem, self.new)
                if frames > 0:
                    self._chunk_steady(with_decoder=False)
                else:
                    self._shift_state()
                est = torch.empty(cfg["C"], t_loc, **f32)
                self._launch([self._decoder(False, frames + 1, t_loc, est)])
                out = est[:, hop:].clone()
        self.emitted += out.size(1)
        assert self.emitted == T, (self.emitted, T)
        self.reset()
        return out
