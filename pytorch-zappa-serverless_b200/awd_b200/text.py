"""Host detokeniser: the string rules of sample_model (reference main.py:45-46, 52-58, 70-79)."""
from typing import Sequence

NO_SPACE_WORDS = frozenset(["'s", "'ll", ",", "?", ".", "'t", "'m", "n't", "!", "'", "'ve", ";", "http", ":", "/", "\\"])
CAPITALIZE_WORDS = frozenset(['.', '!', '\n'])
EXCLUDE_NAMES = ("xxup", "xxfld", "xxrep")


def format_tokens(prompt_words: Sequence[str], words: Sequence[str]) -> str:
    """' ' + each prompt word; then each generated word, capitalised when the previous emitted
    word ends a phrase, glued without a space when it is punctuation-like."""
    parts = [' ' + s for s in prompt_words]
    last = None
    for w in words:
        if last in CAPITALIZE_WORDS:
            w = w.capitalize()
        parts.append(w if w in NO_SPACE_WORDS else ' ' + w)
        last = w
    return ''.join(parts)


def exclude_token_ids(stoi) -> list:
    """exclude_tokens of main.py:47."""
    return [stoi[t] for t in EXCLUDE_NAMES if t in stoi]
