"""Loss dispatch (fused cross-entropy where available)."""


def maybe_cross_entropy(module, logits, target):
    return None
