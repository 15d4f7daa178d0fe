"""Utilities: device probes and size formatting (util.py of the reference), synthetic
wikitext-shaped datasets, memory telemetry, checkpointing."""
from .devices import format_size, get_device, get_device_count  # noqa: F401
from .data import (SyntheticTextDataset, TextFileDataset, WordHashTokenizer, load_text_file,  # noqa: F401
                   load_wikitext)
from .memory import memory_status  # noqa: F401
from .trace import DeviceTimer, annotate_model, nvtx_range  # noqa: F401
