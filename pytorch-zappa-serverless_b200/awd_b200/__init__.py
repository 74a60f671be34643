"""Host side of the B200-native AWD-LSTM engine: ctypes over the C ABI in include/awd_b200.h."""
from .engine import Engine, SamplingParams  # noqa: F401
from .text import format_tokens, Detokenizer, NO_SPACE_WORDS, CAPITALIZE_WORDS  # noqa: F401
