"""state_dict key names / shapes of ConvTasNet and frame geometry.

The key names are part of the drop-in boundary: reference checkpoints must load
into this implementation and vice versa (conv_tasnet.py:62-94).  They follow
from the reference's module nesting: encoder.conv1d_U (:106), separator.network
= Sequential(cLN, bottleneck, repeats, mask conv) (:186-189), each TemporalBlock
.net = Sequential(conv1x1, PReLU, norm, dsconv) (:225) and dsconv.net =
Sequential(depthwise, [Chomp1d,] PReLU, norm, pointwise) (:258-263), decoder
.basis_signals (:126).
"""
from collections import OrderedDict


def is_bn(norm_type):
    """chose_norm (conv_tasnet.py:292-303): anything but "gLN" / "cLN" builds nn.BatchNorm1d."""
    return norm_type not in ("gLN", "cLN")


def block_keys(r, x, causal, norm_type="gLN"):
    base = "separator.network.2.%d.%d.net." % (r, x)
    o = 1 if causal else 0  # Chomp1d sits at dsconv index 1 in causal models
    if is_bn(norm_type):
        # nn.BatchNorm1d names its affine weight / bias and carries running statistics as buffers
        n1, n2 = base + "2.", base + "3.net.%d." % (2 + o)
        return {
            "w1": base + "0.weight", "a1": base + "1.weight",
            "g1": n1 + "weight", "b1": n1 + "bias",
            "rm1": n1 + "running_mean", "rv1": n1 + "running_var", "nb1": n1 + "num_batches_tracked",
            "wd": base + "3.net.0.weight",
            "a2": base + "3.net.%d.weight" % (1 + o),
            "g2": n2 + "weight", "b2": n2 + "bias",
            "rm2": n2 + "running_mean", "rv2": n2 + "running_var", "nb2": n2 + "num_batches_tracked",
            "w2": base + "3.net.%d.weight" % (3 + o),
        }
    return {
        "w1": base + "0.weight", "a1": base + "1.weight",
        "g1": base + "2.gamma", "b1": base + "2.beta",
        "wd": base + "3.net.0.weight",
        "a2": base + "3.net.%d.weight" % (1 + o),
        "g2": base + "3.net.%d.gamma" % (2 + o), "b2": base + "3.net.%d.beta" % (2 + o),
        "w2": base + "3.net.%d.weight" % (3 + o),
    }


# how each parameter is initialised before the final xavier pass (conv_tasnet.py:41-43):
# "conv"  -> nn.Conv1d / nn.Linear default (kaiming_uniform_, a = sqrt(5)): consumes RNG
# "ones" / "zeros" -> cLN / gLN reset_parameters (:321-323, :346-348)
# "prelu" -> nn.PReLU() default 0.25
# BatchNorm1d (norm_type other than gLN / cLN): weight "ones" [H], bias "zeros" [H] (1-D: the xavier pass
# skips them) and the BUFFERS running_mean "buf_zeros", running_var "buf_ones", num_batches_tracked "buf_count"
def param_spec(N, L, B, H, P, X, R, C, causal, norm_type="gLN"):
    """Ordered {key: (shape, init kind)} in the reference's state_dict order (kinds starting with
    "buf_" are buffers, everything else is a parameter)."""
    spec = OrderedDict()
    spec["encoder.conv1d_U.weight"] = ((N, 1, L), "conv")
    spec["separator.network.0.gamma"] = ((1, N, 1), "ones")
    spec["separator.network.0.beta"] = ((1, N, 1), "zeros")
    spec["separator.network.1.weight"] = ((B, N, 1), "conv")
    for r in range(R):
        for x in range(X):
            k = block_keys(r, x, causal, norm_type)
            bn = is_bn(norm_type)
            nshape = (H,) if bn else (1, H, 1)
            spec[k["w1"]] = ((H, B, 1), "conv")
            spec[k["a1"]] = ((1,), "prelu")
            spec[k["g1"]] = (nshape, "ones")
            spec[k["b1"]] = (nshape, "zeros")
            if bn:
                spec[k["rm1"]] = ((H,), "buf_zeros")
                spec[k["rv1"]] = ((H,), "buf_ones")
                spec[k["nb1"]] = ((), "buf_count")
            spec[k["wd"]] = ((H, 1, P), "conv")
            spec[k["a2"]] = ((1,), "prelu")
            spec[k["g2"]] = (nshape, "ones")
            spec[k["b2"]] = (nshape, "zeros")
            if bn:
                spec[k["rm2"]] = ((H,), "buf_zeros")
                spec[k["rv2"]] = ((H,), "buf_ones")
                spec[k["nb2"]] = ((), "buf_count")
            spec[k["w2"]] = ((B, H, 1), "conv")
    spec["separator.network.3.weight"] = ((C * N, B, 1), "conv")
    spec["decoder.basis_signals.weight"] = ((L, N), "conv")
    return spec


def frame_geometry(T, L):
    """(hop, K, T_conv): K = (T - L)//hop + 1 frames (conv_tasnet.py:113),
    T_conv = (K-1)*hop + L overlap-added samples (utils.py:35)."""
    hop = L // 2
    K = (T - L) // hop + 1
    return hop, K, (K - 1) * hop + L
