"""Typing-only helpers the reference's test utilities import at module level (``mypy_extensions`` is not in this image).
They only ever appear inside ``Callable[[...], ...]`` annotations, so returning the wrapped type is enough."""


def _arg(type=None, name=None):  # noqa: A002
    return type


Arg = DefaultArg = NamedArg = DefaultNamedArg = VarArg = KwArg = _arg
