from .adam import AdamW, FusedAdam, HFAdamW  # noqa: F401
