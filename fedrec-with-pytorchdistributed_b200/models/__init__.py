from .attention import AdditiveAttention, ScaledDotProductAttention, MultiHeadAttention  # noqa: F401
from .distilbert import DistilBertTrunk, DistilBertConfigLite  # noqa: F401
from .encoder import TextEncoder, UserEncoder  # noqa: F401
from .model import UserModel, ServerUserModel  # noqa: F401
