"""Mirror of the helpers of reference ``bigvgan/utils.py`` that the generator uses (``init_weights`` :45-48,
``get_padding`` :57-58).  The plotting / checkpoint-scanning helpers of that file are training-only and out of
scope."""


def init_weights(m, mean=0.0, std=0.01):
    # reference utils.py:45-48.  With old-style weight_norm this mutates the derived ``.weight`` only (a no-op on
    # weight_g/weight_v) but it DOES advance torch's RNG, which seeded construction must reproduce.
    classname = m.__class__.__name__
    if classname.find("Conv") != -1:
        m.weight.data.normal_(mean, std)


def get_padding(kernel_size, dilation=1):
    # reference utils.py:57-58
    return int((kernel_size * dilation - dilation) / 2)
