"""Drop-in for the reference's src/conv_tasnet.py: same class name, constructor,
attributes, state_dict keys, checkpoint package and forward signature
(conv_tasnet.py:13-94), with the computation done by hand-written sm_100a CUDA
kernels in libctn_b200.so instead of a tree of torch.nn modules.

Put this directory on PYTHONPATH in place of the reference's src/:
    from conv_tasnet import ConvTasNet
"""
import math

import torch
import torch.nn as nn

from ctn_b200 import _engine, _native
from ctn_b200._layout import is_bn, param_spec

EPS = 1e-8


class _Node(nn.Module):
    """Parameter container; exists only so that state_dict keys nest exactly like the
    reference's module tree.  It is never called."""

    def forward(self, *a, **k):
        raise RuntimeError("parameter container, not a layer; call ConvTasNet.forward")


def _attach(root, dotted, param):
    node = root
    parts = dotted.split(".")
    for name in parts[:-1]:
        child = node._modules.get(name)
        if child is None:
            child = _Node()
            node.add_module(name, child)
        node = child
    if isinstance(param, nn.Parameter):
        node.register_parameter(parts[-1], param)
    else:
        node.register_buffer(parts[-1], param)


class ConvTasNet(nn.Module):
    def __init__(self, N, L, B, H, P, X, R, C, norm_type="gLN", causal=False,
                 mask_nonlinear='relu'):
        """
        Args (conv_tasnet.py:14-31):
            N: Number of filters in autoencoder
            L: Length of the filters (in samples)
            B: Number of channels in bottleneck 1 x 1-conv block
            H: Number of channels in convolutional blocks
            P: Kernel size in convolutional blocks
            X: Number of convolutional blocks in each repeat
            R: Number of repeats
            C: Number of speakers
            norm_type: BN, gLN, cLN (BN: inference with running statistics only)
            causal: causal or non-causal
            mask_nonlinear: 'relu' or 'softmax'
        """
        super(ConvTasNet, self).__init__()
        self.N, self.L, self.B, self.H, self.P, self.X, self.R, self.C = N, L, B, H, P, X, R, C
        self.norm_type = norm_type
        self.causal = causal
        self.mask_nonlinear = mask_nonlinear
        _engine._check_cfg(self._cfg())
        spec = param_spec(N, L, B, H, P, X, R, C, causal, norm_type)
        # Same RNG consumption as the reference's construction, so that the same seed gives
        # the same initial weights: first the layers' default initialisers in creation
        # order, then xavier_normal_ on every parameter with more than one dim (:41-43) --
        # which includes the [1, C, 1] gamma / beta of every norm.
        created = []
        for key, (shape, kind) in spec.items():
            if kind.startswith("buf_"):   # BatchNorm1d buffers: running_mean 0, running_var 1, count 0
                _attach(self, key, torch.zeros(shape, dtype=torch.long) if kind == "buf_count" else
                        torch.full(shape, 1.0 if kind == "buf_ones" else 0.0))
                continue
            t = torch.empty(shape)
            if kind == "conv":
                nn.init.kaiming_uniform_(t, a=math.sqrt(5))
            elif kind == "ones":
                t.fill_(1)
            elif kind == "zeros":
                t.zero_()
            else:
                t.fill_(0.25)
            p = nn.Parameter(t)
            _attach(self, key, p)
            created.append(p)
        for p in created:
            if p.dim() > 1:
                nn.init.xavier_normal_(p)
        self._keys = [k for k, (_, kind) in spec.items() if not kind.startswith("buf_")]
        self._buffer_keys = [k for k, (_, kind) in spec.items() if kind in ("buf_zeros", "buf_ones")]
        # created eagerly (no CUDA work) so that DataParallel replicas, whose __dict__ is a shallow copy of
        # this module's, share one orchestrator with per-device plans instead of rebuilding it per call
        self.__dict__["_engine"] = _engine.SeparationEngine(self._cfg())
        from ctn_b200 import _autograd
        self.__dict__["_train_engine"] = _autograd.TrainEngine(self._cfg())

    def _cfg(self):
        return dict(N=self.N, L=self.L, B=self.B, H=self.H, P=self.P, X=self.X, R=self.R, C=self.C,
                    norm_type=self.norm_type, causal=self.causal, mask_nonlinear=self.mask_nonlinear)

    def _params(self):
        """{state_dict key: tensor} of the parameters (and BatchNorm running statistics) in registration
        order.  The owning container and leaf name of every key are resolved once; a forward then costs
        one dict lookup per tensor (the single-utterance latency is host-bound).  The replicas that
        torch.nn.DataParallel builds (train.py:121, separate.py:45) hold their parameter copies as plain
        attributes of replicated containers, not in _parameters, so they walk the tree by attribute."""
        if getattr(self, "_is_replica", False):
            out = {}
            for key in self._keys + self._buffer_keys:
                node = self
                for part in key.split("."):
                    node = getattr(node, part)
                out[key] = node
            return out
        slots = self.__dict__.get("_param_slots")
        if slots is None:
            slots = []
            for key in self._keys + self._buffer_keys:
                node = self
                parts = key.split(".")
                for part in parts[:-1]:
                    node = node._modules[part]
                store = node._parameters if parts[-1] in node._parameters else node._buffers
                slots.append((key, store, parts[-1]))
            self.__dict__["_param_slots"] = slots
        return {key: store[leaf] for key, store, leaf in slots}

    def _bn_buffers(self):
        """{state_dict key: tensor} of the BatchNorm1d buffers (running_mean, running_var, num_batches_tracked)."""
        out = {}
        for key, (_, kind) in param_spec(self.N, self.L, self.B, self.H, self.P, self.X, self.R, self.C, self.causal,
                                         self.norm_type).items():
            if kind.startswith("buf_"):
                node = self
                parts = key.split(".")
                for part in parts[:-1]:
                    node = node._modules[part]
                out[key] = node._buffers[parts[-1]]
        return out

    def engine(self):
        """The per-module kernel orchestrator (workspaces, prepared weights, CUDA graphs)."""
        eng = self.__dict__.get("_engine")
        if eng is None:
            eng = _engine.SeparationEngine(self._cfg())
            self.__dict__["_engine"] = eng
        return eng

    def train_engine(self):
        """Kernel orchestrator of the training path (saved activations, backward call list)."""
        eng = self.__dict__.get("_train_engine")
        if eng is None:
            from ctn_b200 import _autograd
            eng = _autograd.TrainEngine(self._cfg())
            self.__dict__["_train_engine"] = eng
        return eng

    @torch.no_grad()
    def separate(self, mixture, out=None, wait=True):
        """Inference from host memory to host memory (the loop of separate.py:60-71 moves every batch to
        the GPU and every estimate back with blocking copies): mixture is a CPU float32 [M, T] tensor,
        ideally pinned; returns a pinned CPU [M, C, T] tensor (or fills `out`).  Host-device copies of
        neighbouring micro-batches overlap the kernels.  The call returns after the result has arrived.
        wait=False returns (out, event) at once instead: `out` is complete when `event.synchronize()`
        returns (event is None for an empty batch), and consecutive calls pipeline -- one call's copy-out
        runs under the next call's kernels.  Neither `mixture` nor `out` may be touched before that."""
        params = self._params()
        if is_bn(self.norm_type) and self.training:
            raise NotImplementedError("norm_type=%r (BatchNorm1d) is implemented for inference only" % (self.norm_type,))
        dev = next(iter(params.values())).device
        if dev.type != "cuda":
            raise RuntimeError("the model must be on a CUDA device: this implementation has no CPU path")
        if not wait:
            return self.engine().forward_host(mixture, params, out, join=False)
        res = self.engine().forward_host(mixture, params, out)
        torch.cuda.current_stream(dev).synchronize()
        return res

    def streaming(self, chunk_frames=256, use_graph=True):
        """Chunked separation of one audio stream with bounded latency (causal cLN models only; an
        addition to the reference's surface): see ctn_b200/_stream.py."""
        from ctn_b200 import _stream
        return _stream.StreamingSeparator(self, chunk_frames, use_graph)

    def forward(self, mixture):
        """
        Args:
            mixture: [M, T], M is batch size, T is #samples
        Returns:
            est_source: [M, C, T]
        """
        _native.require_cuda(mixture, "mixture")
        params = self._params()
        for k, p in params.items():
            if not p.is_cuda or p.dtype != torch.float32 or p.device != mixture.device:
                raise RuntimeError("parameter %s is %s on %s; float32 on %s is required (call model.cuda())"
                                   % (k, p.dtype, p.device, mixture.device))
        needs_grad = torch.is_grad_enabled() and (mixture.requires_grad or any(p.requires_grad for p in params.values()))
        # nn.BatchNorm1d blocks (conv_tasnet.py:300-303) in train() mode normalise with the statistics of the
        # whole batch and update their running buffers, with or without a graph: the training path does both
        if needs_grad or (is_bn(self.norm_type) and self.training):
            if getattr(self, "_is_replica", False) and is_bn(self.norm_type) and self.training:
                raise NotImplementedError("BatchNorm1d training under torch.nn.DataParallel is not implemented: "
                                          "use one process per GPU (ctn_b200._dist.attach), which also "
                                          "synchronises the batch statistics")
            from ctn_b200 import _autograd
            return _autograd.separate(self, mixture, params)
        # the engine only takes addresses and version counters of the parameters (under no_grad): no detach
        return self.engine().forward(mixture, params)

    def __getstate__(self):
        state = self.__dict__.copy()
        state["_engine"] = None  # workspaces are per process / device
        state.pop("_param_slots", None)
        state.pop("_train_engine", None)
        return state

    @classmethod
    def load_model(cls, path):
        # Load to CPU
        package = torch.load(path, map_location=lambda storage, loc: storage)
        model = cls.load_model_from_package(package)
        return model

    @classmethod
    def load_model_from_package(cls, package):
        model = cls(package['N'], package['L'], package['B'], package['H'],
                    package['P'], package['X'], package['R'], package['C'],
                    norm_type=package['norm_type'], causal=package['causal'],
                    mask_nonlinear=package['mask_nonlinear'])
        model.load_state_dict(package['state_dict'])
        return model

    @staticmethod
    def serialize(model, optimizer, epoch, tr_loss=None, cv_loss=None):
        package = {
            # hyper-parameter
            'N': model.N, 'L': model.L, 'B': model.B, 'H': model.H,
            'P': model.P, 'X': model.X, 'R': model.R, 'C': model.C,
            'norm_type': model.norm_type, 'causal': model.causal,
            'mask_nonlinear': model.mask_nonlinear,
            # state
            'state_dict': model.state_dict(),
            'optim_dict': optimizer.state_dict(),
            'epoch': epoch
        }
        if tr_loss is not None:
            package['tr_loss'] = tr_loss
            package['cv_loss'] = cv_loss
        return package
