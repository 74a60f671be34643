"""Shared test helpers: golden-file loading and tolerance checks."""
import os

import numpy as np

from oracle.awd_numpy import LMConfig, init_state_dict

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name + '.npz'))
    V, E, H, L, tie, bias = [int(x) for x in d['cfg']]
    cfg = LMConfig(V, E, H, L, bool(tie), bool(bias))
    if 'init_seed' in d.files:
        sd = init_state_dict(cfg, int(d['init_seed']), float(d['decoder_bias_scale']))
    else:
        sd = {k[3:]: d[k] for k in d.files if k.startswith('sd.')}
    return cfg, sd, d


def rel_err(a, b):
    """max |a-b| / max|b|  -- the 'relative' of north_star's 1e-4 / 2e-2 tolerances."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


FORWARD_FIXTURES = ['fwd_tiny', 'fwd_odd', 'fwd_untied_l2', 'fwd_l1_nobias', 'fwd_alias', 'fwd_shipped_dims']
GEN_FIXTURES = ['gen_greedy_small', 'gen_greedy_empty_prompt']
