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


class Detokenizer:
    """format_tokens for a whole batch of generated streams at once.

    With 1024 streams x 256 tokens per device batch the per-word Python loop of format_tokens costs as much
    host time as the GPU needs to generate the batch.  The rules of main.py:70-79 only look at the previous
    word, so every vocabulary entry has exactly two possible renderings -- after a phrase end (capitalised) and
    otherwise -- each with its space decided (main.py:74-77 tests the CAPITALISED word against no_space_words,
    so 'http' after a '.' becomes ' Http').  Both renderings are tabulated once per vocabulary; a batch is then
    two table gathers, one select and one join per stream.  Result equals format_tokens word for word."""

    def __init__(self, itos: Sequence[str]):
        import numpy as np
        self._np = np
        n = len(itos)
        self.plain = np.empty(n, dtype=object)
        self.capped = np.empty(n, dtype=object)
        self.ends_phrase = np.zeros(n, dtype=bool)
        for i, w in enumerate(itos):
            c = w.capitalize()
            self.plain[i] = w if w in NO_SPACE_WORDS else ' ' + w
            self.capped[i] = c if c in NO_SPACE_WORDS else ' ' + c
            # `last` holds the emitted (possibly capitalised) word; capitalising never turns a word into or out
            # of '.', '!' or '\n', so the raw word decides
            self.ends_phrase[i] = w in CAPITALIZE_WORDS or c in CAPITALIZE_WORDS
        self.itos = itos
        self._table = np.concatenate([self.plain, self.capped])     # rendering k of word i at i + k * n

    def format_batch(self, prompts: Sequence[Sequence[str]], ids, lengths: Sequence[int]) -> list:
        """ids: int array (n_new, bs) (numpy or CPU torch); stream b keeps its first lengths[b] tokens."""
        np = self._np
        ids = np.asarray(ids)
        if ids.ndim != 2 or ids.shape[1] != len(prompts):
            raise ValueError('ids must be (n_new, bs) with one column per prompt')
        idx = np.ascontiguousarray(ids.T, dtype=np.int64)                     # (bs, n_new): a stream's words are contiguous
        if idx.shape[1] > 1:
            idx[:, 1:] += len(self.plain) * self.ends_phrase[idx[:, :-1]]     # (reads the raw ids: evaluated before the +=)
        cols = self._table[idx]                                               # object array (bs, n_new)
        out = []
        for b, prompt in enumerate(prompts):
            head = ''.join([' ' + s for s in prompt])
            out.append(head + ''.join(cols[b, :lengths[b]].tolist()))
        return out
