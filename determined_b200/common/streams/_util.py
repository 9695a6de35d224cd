"""Range-encoded key sets, e.g. ``"1-3,7"`` (reference ``common/streams/_util.py``)."""
from typing import Iterable, List, Set


def range_encoded_keys(known: str) -> Iterable[int]:
    if not known:
        return
    for part in known.split(","):
        if "-" in part:
            a, b = part.split("-", 1)
            yield from range(int(a), int(b) + 1)
        else:
            yield int(part)


def encode_keys(keys: Set[int]) -> str:
    out: List[str] = []
    ks = sorted(keys)
    i = 0
    while i < len(ks):
        j = i
        while j + 1 < len(ks) and ks[j + 1] == ks[j] + 1:
            j += 1
        out.append(str(ks[i]) if i == j else f"{ks[i]}-{ks[j]}")
        i = j + 1
    return ",".join(out)
