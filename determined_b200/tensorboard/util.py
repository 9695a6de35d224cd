"""File-name helpers for TensorBoard syncing (public names of reference ``harness/determined/tensorboard/util.py``).

Workers of one trial write profiler traces with the same file name (the host name); before they are uploaded next to
each other, every rank but 0 gets ``#<rank>`` spliced into the name — in front of the extension TensorBoard's profile
plugin recognises, or right after the worker name for PyTorch-profiler traces.  Event files and unknown extensions
keep their names.
"""
import logging
import pathlib
import re
from typing import List, Optional

logger = logging.getLogger("determined.tensorboard")

# glob patterns of the files TensorBoard (and its profile plugin) reads
TB_FILE_GLOBS = ("*tfevents*", "*.trace.json.gz", "*.trace.json", "*.memory_profile.json.gz", "*.pb")
# multi-part extensions of profile-plugin files, longest match wins
PROFILER_EXTENSIONS = (".input_pipeline.pb", ".memory_profile.json.gz", ".tensorflow_stats.pb", ".kernel_stats.pb", ".overview_page.pb",
                       ".xplane.pb", ".trace.json.gz", ".trace.json")
_PT_TRACE = re.compile(r"^(?P<worker>.*?)(?P<span>\.\d+)?(?P<ext>\.pt\.trace\.json(?:\.gz)?)$")


def find_tb_files(base_dir: pathlib.Path) -> List[pathlib.Path]:
    """Every TensorBoard-relevant file below ``base_dir`` (empty, with a warning, if the directory is missing)."""
    base_dir = pathlib.Path(base_dir)
    if not base_dir.exists():
        logger.warning("%s directory does not exist.", base_dir)
        return []
    return [f for pattern in TB_FILE_GLOBS for f in base_dir.rglob(pattern)]


def get_pytorch_profiler_file_extension(path: pathlib.Path) -> Optional[str]:
    m = _PT_TRACE.match(pathlib.Path(path).name)
    return m.group("ext") if m else None


def get_rank_aware_path(path: pathlib.Path, rank: int) -> pathlib.Path:
    """``host.memory_profile.json.gz`` -> ``host#3.memory_profile.json.gz``; ``host_37.1674.pt.trace.json`` ->
    ``host_37#1.1674.pt.trace.json``; anything else is returned unchanged."""
    path = pathlib.Path(path)
    name = path.name
    m = _PT_TRACE.match(name)
    if m:
        return path.with_name(f"{m.group('worker')}#{rank}{m.group('span') or ''}{m.group('ext')}")
    for ext in sorted(PROFILER_EXTENSIONS, key=len, reverse=True):
        if name.endswith(ext):
            return path.with_name(f"{name[:-len(ext)]}#{rank}{ext}")
    return path
