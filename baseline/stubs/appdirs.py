"""Stand-in for appdirs (absent from this image)."""
import os


def user_config_dir(appname=None, *a, **k):
    return os.path.join(os.path.expanduser("~/.config"), appname or "")


def user_data_dir(appname=None, *a, **k):
    return os.path.join(os.path.expanduser("~/.local/share"), appname or "")


def user_cache_dir(appname=None, *a, **k):
    return os.path.join(os.path.expanduser("~/.cache"), appname or "")
