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


def block_keys(r, x, causal):
    base = "separator.network.2.%d.%d.net." % (r, x)
    o = 1 if causal else 0  # Chomp1d sits at dsconv index 1 in causal models
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
def param_spec(N, L, B, H, P, X, R, C, causal):
    """Ordered {key: (shape, init kind)} in the reference's registration order."""
    spec = OrderedDict()
    spec["encoder.conv1d_U.weight"] = ((N, 1, L), "conv")
    spec["separator.network.0.gamma"] = ((1, N, 1), "ones")
    spec["separator.network.0.beta"] = ((1, N, 1), "zeros")
    spec["separator.network.1.weight"] = ((B, N, 1), "conv")
    for r in range(R):
        for x in range(X):
            k = block_keys(r, x, causal)
            spec[k["w1"]] = ((H, B, 1), "conv")
            spec[k["a1"]] = ((1,), "prelu")
            spec[k["g1"]] = ((1, H, 1), "ones")
            spec[k["b1"]] = ((1, H, 1), "zeros")
            spec[k["wd"]] = ((H, 1, P), "conv")
            spec[k["a2"]] = ((1,), "prelu")
            spec[k["g2"]] = ((1, H, 1), "ones")
            spec[k["b2"]] = ((1, H, 1), "zeros")
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
