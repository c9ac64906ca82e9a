"""kvx — B200-native prefill→decode paged-KV handoff (host side over libkvx.so)."""
from .api import (ENGINE, HND, NHD, OPT, Context, KvxError, Pool, Queue, Xform)  # noqa: F401
from .shapes import KV_SHAPES, KVShape, payload_bytes  # noqa: F401
