"""Fused sm_100a forward/backward of the FreeSurfer MLP."""


def supported(model, x):
    return False


def msannet_forward(model, x):
    raise NotImplementedError
