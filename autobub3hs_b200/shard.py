"""Event sharding of the path across GPUs: rank g of G owns the contiguous range
[g*E/G, (g+1)*E/G) -- the same formula as the abub3hs binary (host/abub3hs.cpp) -- and rank-ordered
concatenation of the output rows reproduces the reference's `omp ordered` file order
(AutoBubStart3.cpp:380)."""
from __future__ import annotations

from typing import List, Sequence, Tuple


def event_range(n_events: int, rank: int, world: int) -> Tuple[int, int]:
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return n_events * rank // world, n_events * (rank + 1) // world


def merge_rank_outputs(header: str, parts: Sequence[str]) -> str:
    """parts[r] = rows rank r produced for its event range, in event order."""
    return header + "".join(parts)


def all_ranges(n_events: int, world: int) -> List[Tuple[int, int]]:
    return [event_range(n_events, r, world) for r in range(world)]
