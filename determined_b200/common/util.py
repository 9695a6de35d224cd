"""Small helpers shared by the CLI, the SDK and the harness (public names of reference ``harness/determined/common/util.py``)."""
import datetime
import functools
import os
import pathlib
import random
import re
import time
import warnings
from typing import Any, Callable, IO, Iterator, Optional, Sequence, TypeVar, Union

import yaml

T = TypeVar("T")


def chunks(lst: Sequence[Any], chunk_size: int) -> Iterator[Sequence[Any]]:
    """Consecutive slices of at most ``chunk_size`` elements (strings included)."""
    if chunk_size <= 0:
        raise ValueError("chunk_size must be positive")
    for i in range(0, len(lst), chunk_size):
        yield lst[i:i + chunk_size]


def sizeof_fmt(val: Union[int, float]) -> str:
    """``1024 -> "1.0KB"``, ``36 -> "36.0B"`` (binary multiples, one decimal)."""
    v = float(val)
    for unit in ("", "K", "M", "G", "T", "P", "E", "Z"):
        if abs(v) < 1024.0:
            return f"{v:3.1f}{unit}B"
        v /= 1024.0
    return f"{v:.1f}YB"


def debug_mode() -> bool:
    return os.environ.get("DET_DEBUG", "").lower() in ("true", "1", "yes")


def preserve_random_state(fn: Callable[..., T]) -> Callable[..., T]:
    """Decorator: whatever ``fn`` does to Python's ``random`` state is undone afterwards."""
    @functools.wraps(fn)
    def wrapped(*a: Any, **kw: Any) -> T:
        state = random.getstate()
        try:
            return fn(*a, **kw)
        finally:
            random.setstate(state)
    return wrapped


def yaml_safe_dump(data: Any, stream: Optional[IO[str]] = None, default_flow_style: Optional[bool] = None, **kw: Any) -> Any:
    return yaml.safe_dump(data, stream, default_flow_style=default_flow_style, **kw)


def yaml_safe_load(f: Union[IO[Any], str, bytes]) -> Any:
    return yaml.safe_load(f)


def safe_load_yaml_with_exceptions(yaml_file: Union[IO[Any], str]) -> Any:
    """Load YAML, turning parser errors into one readable message (file name, line, problem)."""
    try:
        return yaml.safe_load(yaml_file)
    except yaml.YAMLError as e:
        name = getattr(yaml_file, "name", "<string>")
        mark = getattr(e, "problem_mark", None)
        where = f" (line {mark.line + 1}, column {mark.column + 1})" if mark is not None else ""
        raise ValueError(f"Error: invalid YAML in {name}{where}: {getattr(e, 'problem', e)}") from e


def get_config_path() -> pathlib.Path:
    """Where the CLI keeps its token cache: ``$DET_CONFIG_DIR``-independent, per-platform user config directory."""
    if os.environ.get("XDG_CONFIG_HOME"):
        return pathlib.Path(os.environ["XDG_CONFIG_HOME"]) / "determined"
    return pathlib.Path(os.path.expanduser("~")) / ".config" / "determined"


_TS = re.compile(r"^(\d{4})-(\d{2})-(\d{2})T(\d{2}):(\d{2}):(\d{2})(?:\.(\d+))?(Z|[+-]\d{2}:\d{2})$")


def is_protobuf_timestamp(ts: str) -> bool:
    return isinstance(ts, str) and _TS.match(ts) is not None


def parse_protobuf_timestamp(ts: str) -> datetime.datetime:
    """RFC 3339 with up to nanosecond fractions (as grpc-gateway prints them) -> aware ``datetime`` (microsecond resolution)."""
    m = _TS.match(ts)
    if m is None:
        raise ValueError(f"not an RFC 3339 timestamp: {ts!r}")
    y, mo, d, h, mi, s, frac, tz = m.groups()
    micro = int((frac or "0")[:6].ljust(6, "0"))
    if tz == "Z":
        tzinfo = datetime.timezone.utc
    else:
        sign = 1 if tz[0] == "+" else -1
        tzinfo = datetime.timezone(sign * datetime.timedelta(hours=int(tz[1:3]), minutes=int(tz[4:6])))
    return datetime.datetime(int(y), int(mo), int(d), int(h), int(mi), int(s), micro, tzinfo=tzinfo)


def wait_for(predicate: Callable[[], T], timeout: Optional[float] = None, interval: float = 0.1) -> T:
    """Poll ``predicate`` until it returns something truthy; ``TimeoutError`` after ``timeout`` seconds."""
    start = time.time()
    while True:
        out = predicate()
        if out:
            return out
        if timeout is not None and time.time() - start > timeout:
            raise TimeoutError("timed out waiting for condition")
        time.sleep(interval)


def deprecated(message: Optional[str] = None) -> Callable[[Callable[..., T]], Callable[..., T]]:
    def deco(fn: Callable[..., T]) -> Callable[..., T]:
        @functools.wraps(fn)
        def wrapped(*a: Any, **kw: Any) -> T:
            warnings.warn(message or f"{fn.__name__} is deprecated and will be removed in a future version.", FutureWarning, stacklevel=2)
            return fn(*a, **kw)
        return wrapped
    return deco


def strtobool(val: str) -> bool:
    v = str(val).strip().lower()
    if v in ("y", "yes", "t", "true", "on", "1"):
        return True
    if v in ("n", "no", "f", "false", "off", "0"):
        return False
    raise ValueError(f"invalid truth value {val!r}")
