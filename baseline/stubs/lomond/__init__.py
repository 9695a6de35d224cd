"""Stand-in for lomond (websocket client, absent from this image); the training path never opens
a websocket."""


class WebSocket:
    def __init__(self, *a, **k):
        raise RuntimeError("lomond is not installed in this image")
