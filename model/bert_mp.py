"""``from model.bert_mp import BertModelWithMP`` — import path of the reference kept; the
implementation is dtds_b200.models.bert_mp."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from dtds_b200.models.bert_mp import BertModelWithMP, LocalValue  # noqa: E402,F401
