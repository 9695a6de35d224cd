"""Runtime correctness checks under the two naming schemes user code of the reference may import (``check.true(...)`` and
``check_true(...)``; reference ``harness/determined/common/check.py``).  New code should simply ``assert``; this module
exists so that model definitions written against the reference keep importing.  Every check raises
:class:`CheckFailedError` with the failed condition and the optional ``reason``.
"""
from typing import Any, Callable, Container, Optional, Sized, Tuple, Type, Union


class CheckFailedError(Exception):
    pass


def _fail(what: str, reason: Optional[str]) -> None:
    raise CheckFailedError(f"CHECK FAILED! {what}" + (f": {reason}" if reason else ""))


def _binary(test: Callable[[Any, Any], bool], words: str) -> Callable[..., None]:
    def check(x: Any, y: Any, reason: Optional[str] = None) -> None:
        if not test(x, y):
            _fail(f"Got {x!r}, expected it to {words} {y!r}", reason)
    return check


def true(val: bool, reason: Optional[str] = None) -> None:
    if not val:
        _fail(f"Got {val!r}, expected a true value", reason)


def false(val: bool, reason: Optional[str] = None) -> None:
    if val:
        _fail(f"Got {val!r}, expected a false value", reason)


def is_none(val: Optional[Any], reason: Optional[str] = None) -> None:
    if val is not None:
        _fail(f"Got {val!r}, expected None", reason)


def is_not_none(val: Optional[Any], reason: Optional[str] = None) -> None:
    if val is None:
        _fail("Got None, expected a value", reason)


eq = _binary(lambda x, y: x == y, "equal")
not_eq = _binary(lambda x, y: x != y, "differ from")
gt = _binary(lambda x, y: x > y, "be greater than")
gt_eq = _binary(lambda x, y: x >= y, "be greater than or equal to")
lt = _binary(lambda x, y: x < y, "be less than")
lt_eq = _binary(lambda x, y: x <= y, "be less than or equal to")
is_in = _binary(lambda x, y: x in y, "be in")
not_in = _binary(lambda x, y: x not in y, "not be in")
is_type = _binary(lambda x, y: type(x) is y, "be exactly of type")
is_instance = _binary(lambda x, y: isinstance(x, y), "be an instance of")
is_not_instance = _binary(lambda x, y: not isinstance(x, y), "not be an instance of")
is_subclass = _binary(lambda x, y: isinstance(x, type) and issubclass(x, y), "be a subclass of")


def equal_lengths(x: Sized, y: Sized, reason: Optional[str] = None) -> None:
    if len(x) != len(y):
        _fail(f"Got lengths {len(x)} and {len(y)}, expected them to be equal", reason)


def len_eq(val: Sized, expected_len: int, reason: Optional[str] = None) -> None:
    if len(val) != expected_len:
        _fail(f"Got length {len(val)}, expected {expected_len}", reason)


# the second naming scheme: the same functions for `from check import check_true, ...`
check_true, check_false, check_none, check_not_none = true, false, is_none, is_not_none
check_eq, check_not_eq, check_gt, check_gt_eq, check_lt, check_lt_eq = eq, not_eq, gt, gt_eq, lt, lt_eq
check_eq_len, check_len, check_in, check_not_in, check_type = equal_lengths, len_eq, is_in, not_in, is_type
check_isinstance, check_not_isinstance, check_issubclass = is_instance, is_not_instance, is_subclass

_Unused = Union[Container[Any], Tuple[Type, ...]]       # (typing names kept importable for annotations in user code)
