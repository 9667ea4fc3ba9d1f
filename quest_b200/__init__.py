"""quest_b200: sm_100a implementation of Quest's query-aware sparse decode-attention path behind the
reference's ``quest.utils`` / ``quest._kernels`` operator surface."""
from quest_b200 import utils  # noqa: F401

__version__ = "0.1.0"
