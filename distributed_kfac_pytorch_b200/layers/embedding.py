"""Embedding layers are not supported, as in the reference
(kfac/layers/embedding.py:19-20 raises unconditionally)."""
from .base import KFACLayer


class EmbeddingLayer(KFACLayer):
    def __init__(self, *args, **kwargs):
        raise ValueError('Embedding layer does not currently work')
