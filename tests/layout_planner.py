"""TEST HELPER (not product code): pure-Python statement of the store's flat-segment layout and
shard ownership (csrc/store.cu ``Store::seal`` / ``Store::plan_for``).  The gloo tests use it to
shard a CPU update the way the store would; tests/test_dense_gpu.py::test_layout_planner_matches_store
pins it against the real store through the addresses of ``b200kv_view``.

Layout rule: keys are laid out in init order, each starting on a 64-element boundary; the segment
is padded to a multiple of ``world * 64`` elements and rank r owns ``[r*slice, (r+1)*slice)``.
"""
ALIGN = 64


def align_up(x, a):
    return (x + a - 1) // a * a


def segment_layout(sizes, world):
    """-> (offsets, total, slice) in elements."""
    offs, off = [], 0
    for n in sizes:
        offs.append(off)
        off = align_up(off + n, ALIGN)
    total = align_up(max(off, ALIGN), world * ALIGN)
    return offs, total, total // world


def pieces(sizes, world):
    """[(key index, rank, begin element inside the key, length)] — the (key ∩ owner-slice) pieces
    the store reduces to / broadcasts from their owners."""
    offs, total, sl = segment_layout(sizes, world)
    out = []
    for k, (o, n) in enumerate(zip(offs, sizes)):
        for r in range(world):
            lo, hi = max(o, r * sl), min(o + n, (r + 1) * sl)
            if lo < hi:
                out.append((k, r, lo - o, hi - lo))
    return out


def shard_ranges(sizes, world, rank):
    """per key: (begin, end) element range inside the key owned by `rank` (may be empty)."""
    offs, total, sl = segment_layout(sizes, world)
    lo, hi = rank * sl, (rank + 1) * sl
    res = []
    for o, n in zip(offs, sizes):
        b = min(max(lo - o, 0), n)
        e = max(min(hi - o, n), b)
        res.append((b, e))
    return res
