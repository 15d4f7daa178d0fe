"""Entry-point helpers kept at the repository root for the scripts (same names as the
reference's util.py; implementations live in dtds_b200.utils)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from dtds_b200.utils.data import (load_wikitext, synthetic_collator,  # noqa: E402,F401
                                  synthetic_tokenizer)
from dtds_b200.utils.devices import format_size, get_device, get_device_count  # noqa: E402,F401
