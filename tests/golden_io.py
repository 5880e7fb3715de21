"""Loader for the frozen reference outputs (tests/golden/polls.json.gz)."""
import gzip
import json
import os

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "polls.json.gz")
_cache = None


def load():
    global _cache
    if _cache is None:
        with gzip.open(_PATH) as f:
            _cache = json.loads(f.read())
    return _cache
