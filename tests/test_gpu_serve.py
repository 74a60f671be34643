"""The served path end to end on the GPU (reference main.py:84-112): checkpoint + vocabulary files on disk ->
load_lm_and_predict() -> GET /inference through the WSGI callable -> BatchScheduler / ContinuousBatcher -> the real
engine -> text.  With the greedy rule every body must equal what the reference's sample_model returns for that request
(the golden text produced by the reference itself, and the oracle's restatement for the other prompts)."""
import json
import threading
from urllib.parse import quote

import numpy as np
import pytest
import torch

from helpers import load_golden
from oracle import awd_numpy as on

pytestmark = pytest.mark.gpu


def _call(app, query=''):
    """One GET /inference through whichever app object main.py built (Flask test client, or the WSGI stand-in)."""
    if hasattr(app, 'test_client'):
        r = app.test_client().get('/inference' + ('?' + query if query else ''))
        return r.status_code, r.mimetype, json.loads(r.data)
    seen = {}
    out = app({'PATH_INFO': '/inference', 'REQUEST_METHOD': 'GET', 'QUERY_STRING': query},
              lambda status, headers: seen.update(status=status, headers=dict(headers)))
    body = b''.join(out)
    assert seen['status'].startswith('200'), (seen['status'], body[:500])
    return 200, seen['headers']['Content-Type'], json.loads(body)


@pytest.fixture
def served(tmp_path, monkeypatch):
    """A reference-layout checkpoint and itos pickle in the local cache, and a fresh process-lifetime model."""
    import main as product_main
    cfg, sd, d = load_golden('gen_greedy_small')
    root = tmp_path / 'models' / 'rjokes'
    root.mkdir(parents=True)
    torch.save({k: torch.from_numpy(v.copy()) for k, v in sd.items()}, root / 'rjokes.model.pth')
    itos = [str(s) for s in d['itos']]
    product_main.pickle_obj(itos, str(root / 'rjokes.itos.pkl'))
    monkeypatch.setenv('AWD_CACHE_DIR', str(tmp_path))
    monkeypatch.setenv('models_bucket', 'unused-bucket')
    monkeypatch.setenv('AWD_PRECISION', 'fp32')          # the parity mode: identical greedy ids
    monkeypatch.setenv('AWD_MAX_WAIT_MS', '20')
    for name in ('_MODEL', '_SCHED', '_CONT'):
        monkeypatch.setattr(product_main, name, None)
    yield product_main, cfg, sd, d, itos
    for s in (product_main._SCHED, product_main._CONT):
        if s is not None:
            s.close()
    for name in ('_MODEL', '_SCHED', '_CONT'):
        setattr(product_main, name, None)


def test_get_inference_from_32_threads_equals_sample_model(served):
    product_main, cfg, sd, d, itos = served
    p = on.effective_params(sd, cfg)
    rng = np.random.default_rng(5)
    prompts = [[str(w) for w in d['prompt_words']]]                      # the request the reference itself answered (golden text)
    for i in range(31):                                                 # ragged prompts, one unknown word among them
        n = int(rng.integers(1, 6))
        prompts.append([itos[int(t)] for t in rng.integers(6, cfg.vocab_sz, n)] + (['notaword'] if i % 7 == 0 else []))
    l = len(d['ids'])
    bodies = [None] * len(prompts)

    def client(i):
        q = f'greedy=1&l={l}&prompt=' + quote(' '.join(prompts[i]))
        bodies[i] = _call(product_main.app, q)

    threads = [threading.Thread(target=client, args=(i,)) for i in range(len(prompts))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=120)
    assert all(b is not None for b in bodies)
    for i, (status, ctype, body) in enumerate(bodies):
        assert status == 200 and ctype == 'application/json'
        assert set(body) == {'response'} and set(body['response']) == {'text'}
        want = on.sample_model(p, cfg, itos, prompts[i], l)
        assert body['response']['text'] == want, (i, prompts[i])
    assert bodies[0][2]['response']['text'] == str(d['text'])           # what the reference's sample_model wrote
    # the batching really happened: 32 requests, far fewer device batches
    st = product_main._scheduler().stats if hasattr(product_main._scheduler(), 'stats') else None
    if st:
        assert st['requests'] >= 32 and st['batches'] < 32


def test_default_get_inference_and_continuous_batching(served, monkeypatch):
    """No query parameters = the reference's request: prompt [''], 200 words, sampled.  Then the same through the continuous
    batcher (AWD_CONTINUOUS=1)."""
    product_main, cfg, sd, d, itos = served
    banned = {'xxunk', 'xxfld', 'xxup', 'xxrep'}
    for continuous in ('0', '1'):
        monkeypatch.setenv('AWD_CONTINUOUS', continuous)
        status, ctype, body = _call(product_main.app)
        assert status == 200 and ctype == 'application/json'
        text = body['response']['text']
        words = text.replace('\n', ' \n ').split(' ')
        assert len(text) > 200 and not (banned & set(words))
