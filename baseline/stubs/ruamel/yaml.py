"""Minimal stand-in for ruamel.yaml (absent from this image), backed by PyYAML.  Only used so the
UNMODIFIED reference harness can be imported for the baseline arm of bench.py; the reference only
uses it for CLI YAML round-tripping, never on the training path."""
import yaml as _yaml


class YAML:
    def __init__(self, typ="safe", pure=True):
        self.typ = typ

    def load(self, stream):
        return _yaml.safe_load(stream)

    def dump(self, data, stream=None):
        return _yaml.safe_dump(data, stream)


def safe_load(stream):
    return _yaml.safe_load(stream)


def safe_dump(data, stream=None, **kw):
    return _yaml.safe_dump(data, stream, **kw)


class YAMLError(Exception):
    pass
