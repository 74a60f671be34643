"""Drop-in for the reference's ``main.py``: the same ``GET /inference`` contract
(reference main.py:105-112: 200, application/json, ``{"response": {"text": ...}}``, CORS open),
the same helper names, the same ``sample_model(model, input_words, l)`` text rules -- with the
model loaded ONCE per process instead of on every request (reference main.py:84-99 rebuilds it
per request) and the whole token loop running on the GPU (``awd_generate``).

Flask is used when installed; otherwise ``app`` is a minimal WSGI callable with the same route,
so ``zappa_settings.json``'s ``"app_function": "main.app"`` keeps working either way.
"""
from __future__ import annotations

import json
import os
import os.path
import pickle
import threading
from typing import Dict, List, Optional, Sequence

import torch

from pytorch_models import awd_lstm
from awd_b200 import SamplingParams, format_tokens
from awd_b200.text import exclude_token_ids
from awd_b200.scheduler import BatchScheduler, ContinuousBatcher

S3_MODEL_PATH = 'models/rjokes/rjokes.model.pth'
S3_ITOS_PATH = 'models/rjokes/rjokes.itos.pkl'

DEFAULT_LENGTH = 200          # reference main.py:103: sample_model(my_model, [''], l=200)
DEFAULT_PROMPT = ['']


def pickle_obj(data, path):
    with open(path, 'wb+') as f:
        pickle.dump(data, f)


def unpickle_obj(path):
    with open(path, 'rb') as f:
        return pickle.load(f)


def maybe_fetch_s3(bucket, path):
    """Local cache at /tmp/<key>, fetched from the bucket only when absent (reference main.py:32-38).
    AWS I/O itself is out of scope for the engine: boto3 is used when importable."""
    filename = os.path.join(os.environ.get('AWD_CACHE_DIR', '/tmp'), path)
    if not os.path.exists(filename):
        try:
            import boto3
        except ImportError as e:
            raise FileNotFoundError(f'{filename} is not cached and boto3 is not installed to fetch '
                                    f's3://{bucket}/{path}') from e
        os.makedirs(os.path.dirname(filename), exist_ok=True)
        boto3.resource('s3').Object(bucket, path).download_file(filename)
    return filename


def _generation_sampling(model, greedy=False, temperature=1.0, seed=0) -> SamplingParams:
    return SamplingParams(temperature=temperature, greedy=greedy, n_draws=10, ban_zero=True,
                          banned=exclude_token_ids(model.stoi), seed=seed)


def sample_model(model, input_words: Sequence[str], l: int = 50, *, greedy: bool = False, temperature: float = 1.0,
                 seed: Optional[int] = None) -> str:
    """Same contract as the reference sample_model (main.py:40-81): feed `input_words` (unknown
    words -> id 0), then predict `l` words with the 10-candidate / first-allowed rule, and build
    the string with the reference's spacing and capitalisation rules.  The keyword arguments are
    additions; their defaults give the reference behaviour."""
    bs_saved = model[0].bs
    model.eval()
    if seed is None:
        seed = int.from_bytes(os.urandom(8), 'little')
    ids = [model.stoi[s] if s in model.stoi else 0 for s in input_words]
    prompt = torch.tensor(ids, dtype=torch.int64).view(-1, 1)
    out = model.generate_ids(prompt, l, _generation_sampling(model, greedy, temperature, seed))
    words = [model.itos[i] for i in out[:, 0].tolist()]
    model[0].bs = bs_saved
    return format_tokens(input_words, words)


# ------------------------------------------------------------------------- process-lifetime model
_MODEL = None
_SCHED = None
_CONT = None
_LOCK = threading.Lock()


def _infer_hparams(sd: Dict[str, torch.Tensor]):
    """Read vocab/emb/hidden/layers off the checkpoint instead of hard-coding them (the reference
    hard-codes emb_sz=1000, n_hid=1150, n_layers=3 at main.py:96)."""
    V, E = sd['0.encoder.weight'].shape
    n_layers = 1 + max(int(k.split('.')[2]) for k in sd if k.startswith('0.rnns.'))
    n_hid = sd['0.rnns.0.module.weight_hh_l0'].shape[1] if n_layers > 1 else 1150
    return int(V), int(E), int(n_hid), int(n_layers)


def load_lm():
    """The loading half of load_lm_and_predict (reference main.py:88-101), done once."""
    global _MODEL
    with _LOCK:
        if _MODEL is not None:
            return _MODEL
        bucket = os.environ["models_bucket"]
        itos = unpickle_obj(maybe_fetch_s3(bucket, S3_ITOS_PATH))
        stoi = {w: i for i, w in enumerate(itos)}
        sd = torch.load(maybe_fetch_s3(bucket, S3_MODEL_PATH), map_location='cpu')
        V, E, H, L = _infer_hparams(sd)
        if V != len(itos):
            raise RuntimeError(f'vocabulary has {len(itos)} entries but the checkpoint has {V} rows')
        model = awd_lstm.get_language_model(vocab_sz=V, emb_sz=E, n_hid=H, n_layers=L, pad_token=1, tie_weights=True,
                                            bias='1.decoder.bias' in sd, qrnn=False)
        model.load_state_dict(sd)        # strict, as the reference (main.py:99)
        # The service runs the tensor-core mode unless told otherwise (AWD_PRECISION=fp32: the 1e-4 parity mode, ~25x slower
        # at full batches); a bare get_language_model() module keeps fp32, the reference's arithmetic.
        model.precision = os.environ.get('AWD_PRECISION', 'bf16')
        model.itos, model.stoi = itos, stoi
        model.eval()
        _MODEL = model
        return model


def _scheduler() -> BatchScheduler:
    global _SCHED
    model = load_lm()
    with _LOCK:
        if _SCHED is None:
            _SCHED = BatchScheduler(model, max_batch=int(os.environ.get('AWD_MAX_BATCH', '1024')),
                                    max_wait_ms=float(os.environ.get('AWD_MAX_WAIT_MS', '2')))
        return _SCHED


def _continuous() -> ContinuousBatcher:
    """AWD_CONTINUOUS=1: requests with the default sampling rule share stream slots that are refilled every
    AWD_CHUNK tokens (awd_generate_resume) instead of waiting for whole batches."""
    global _CONT
    model = load_lm()
    with _LOCK:
        if _CONT is None:
            _CONT = ContinuousBatcher(model, slots=int(os.environ.get('AWD_SLOTS', '256')),
                                      chunk=int(os.environ.get('AWD_CHUNK', '32')))
        return _CONT


def load_lm_and_predict(prompt: Optional[List[str]] = None, l: int = DEFAULT_LENGTH, **kw):
    """Same return shape as the reference (main.py:84-103): {'text': str}.  Requests from
    concurrent WSGI threads are batched into one device-side generation."""
    words = prompt if prompt is not None else DEFAULT_PROMPT
    if os.environ.get('AWD_CONTINUOUS') == '1' and not kw:
        return {'text': _continuous().submit(words, l)}
    text = _scheduler().submit(words, l, **kw)
    return {'text': text}


def _query_options(args) -> dict:
    """Additive query parameters; with none given the behaviour is the reference's."""
    kw = {}
    if args.get('prompt') is not None:
        kw['prompt'] = args.get('prompt').split(' ')
    if args.get('l') is not None:
        kw['l'] = max(0, min(int(args.get('l')), 4096))
    if args.get('temperature') is not None:
        kw['temperature'] = float(args.get('temperature'))
    if args.get('greedy') is not None:
        kw['greedy'] = args.get('greedy') not in ('0', 'false', '')
    if args.get('seed') is not None:
        kw['seed'] = int(args.get('seed'))
    return kw


try:
    from flask import Flask, Response, request
    try:
        from flask_cors import CORS
    except ImportError:          # CORS-all-origins by hand
        def CORS(app):
            @app.after_request
            def _cors(resp):
                resp.headers['Access-Control-Allow-Origin'] = '*'
                return resp

    app = Flask(__name__)
    CORS(app)

    @app.route('/inference', methods=['GET'])
    def inference():
        '''GET: Performs inference on the language model'''
        response = load_lm_and_predict(**_query_options(request.args))
        return Response(response=json.dumps({"response": response}), status=200, mimetype='application/json')

except ImportError:
    from urllib.parse import parse_qs

    class _WsgiApp:
        """Stand-in for the Flask app when Flask is not installed: same route, same bytes."""

        def __call__(self, environ, start_response):
            cors = [('Access-Control-Allow-Origin', '*')]
            if environ.get('PATH_INFO') != '/inference':
                start_response('404 NOT FOUND', [('Content-Type', 'text/html')] + cors)
                return [b'Not Found']
            if environ.get('REQUEST_METHOD', 'GET') not in ('GET', 'HEAD', 'OPTIONS'):
                start_response('405 METHOD NOT ALLOWED', [('Allow', 'GET, HEAD, OPTIONS')] + cors)
                return [b'Method Not Allowed']
            try:
                args = {k: v[0] for k, v in parse_qs(environ.get('QUERY_STRING', '')).items()}
                body = json.dumps({"response": load_lm_and_predict(**_query_options(args))}).encode()
            except Exception as e:   # Flask turns an exception in a view into a 500
                start_response('500 INTERNAL SERVER ERROR', [('Content-Type', 'text/html')] + cors)
                return [repr(e).encode()]
            start_response('200 OK', [('Content-Type', 'application/json'), ('Content-Length', str(len(body)))] + cors)
            return [body]

        def run(self, host='0.0.0.0', port=8082, **_):
            from wsgiref.simple_server import make_server
            make_server(host, port, self).serve_forever()

    app = _WsgiApp()

    def inference():
        return {"response": load_lm_and_predict()}


if __name__ == '__main__':
    # local runner (reference main.py:115-127): environment from zappa_settings.json, port 8082
    if os.path.exists('zappa_settings.json'):
        with open('zappa_settings.json') as f:
            for key, val in json.load(f)['dev']['aws_environment_variables'].items():
                os.environ[key] = val
    app.run(host='0.0.0.0', port=8082)
