from .adam import AdamW, FusedAdam, HFAdamW, clip_grad_norm_  # noqa: F401
