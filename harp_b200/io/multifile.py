"""File dealing — ``MultiFileInputFormat`` / ``SingleFileInputFormat``.

Reference: ``ml/java/.../fileformat/MultiFileInputFormat.java:49-126`` deals *whole files* to map
tasks: with ``avg = nFiles / nMaps`` and ``rest = nFiles % nMaps`` the (name-sorted) file list is cut
into consecutive runs, the first ``rest`` runs holding ``avg + 1`` files.  Mappers receive file
*names* as both key and value (``MultiFileRecordReader.java:42-76``).
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple


def list_input_files(input_dir: str) -> List[str]:
    """Name-sorted regular files of a directory (hidden and ``_SUCCESS``-style files skipped, as
    Hadoop's ``FileInputFormat.listStatus`` does)."""
    out = []
    for name in sorted(os.listdir(input_dir)):
        if name.startswith(".") or name.startswith("_"):
            continue
        p = os.path.join(input_dir, name)
        if os.path.isfile(p):
            out.append(p)
    return out


def multi_file_splits(files: Sequence[str], num_map_tasks: int) -> List[List[str]]:
    files = list(files)
    n = int(num_map_tasks)
    avg, rest = len(files) // n, len(files) % n
    splits: List[List[str]] = []
    pos = 0
    for i in range(n):
        take = avg + (1 if i < rest else 0)
        if take == 0:
            continue  # the reference creates no split for an empty run
        splits.append(files[pos:pos + take])
        pos += take
    return splits


def shard_files(files: Sequence[str], rank: int, world: int) -> List[str]:
    """The files of worker ``rank`` (empty when there are fewer files than workers)."""
    splits = multi_file_splits(files, world)
    return splits[rank] if rank < len(splits) else []


def records_for(files: Sequence[str]) -> List[Tuple[str, str]]:
    """(key, value) records as the record reader hands them to ``mapCollective``."""
    return [(f, f) for f in files]


def single_file_splits(files: Sequence[str]) -> List[List[str]]:
    """``SingleFileInputFormat``: one file per split."""
    return [[f] for f in files]
