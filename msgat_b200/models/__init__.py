"""``from models import msgat48, msgat72, msgat96`` (reference ``src/models/__init__.py:8``)."""
from .attention import ChannelAttention, GraphAttention, TemporalAttention
from .embeddings import TimeEmbedding
from .msgat import CACN, GACN, MEAM, MSGAT, TACN, TPC, Chomp, msgat48, msgat72, msgat96

__all__ = ["msgat48", "msgat72", "msgat96", "MSGAT", "TPC", "MEAM", "GACN", "TACN", "CACN", "Chomp",
           "GraphAttention", "TemporalAttention", "ChannelAttention", "TimeEmbedding"]
