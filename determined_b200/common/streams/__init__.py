"""Streaming updates (reference ``harness/determined/common/streams/__init__.py``)."""
from determined_b200.common.streams import wire
from determined_b200.common.streams._util import encode_keys, range_encoded_keys
from determined_b200.common.streams._client import (KeyCache, ModelSpec, ModelVersionSpec, PollingStreamTransport, WebSocketStreamTransport, ProjectSpec, Stream,
                                                    StreamWebSocket, Sync)
