"""CPU checks that need neither a GPU nor the library's kernels:

  * the tile dealing of the persistent backward kernel (csrc/tc_kernels.cu: bwd_cta_of_tile,
    bwd_next_unit) restated in Python, against the properties three pieces of device code rely on
    independently of each other -- the kernel's slot numbers, the epilogue kernel's slot count per
    row tile (csrc/epilogue_kernels.cu, 32-bit and 64-bit form) and the planner's bound on it
    (`bwd_splits` of ntxent_b200_plan_info);
  * every file under profiles/ that DESIGN.md, README.md or INTEGRATION.md cites exists."""
import os
import re

import pytest

from conftest import ROOT

import ntxent_b200  # noqa: F401
from ntxent_b200 import capi


def cta_of_tile(t, total, ctas):  # bwd_cta_of_tile
    return ((t + 1) * ctas - 1) // total


def units_of_cta(c, row_tiles, col_tiles, ctas):  # the kernel's walk: bwd_next_unit
    total = row_tiles * col_tiles
    t, t_end = c * total // ctas, (c + 1) * total // ctas
    out = []
    while t < t_end:
        rt = t // col_tiles
        first = rt * col_tiles
        n = min(col_tiles - (t - first), t_end - t)
        out.append(dict(rt=rt, ct0=t - first, n=n, slot=c - cta_of_tile(first, total, ctas)))
        t += n
    return out


def slots_of_row_tile(rt, row_tiles, col_tiles, ctas):  # the epilogue's formula
    total = row_tiles * col_tiles
    first = rt * col_tiles
    return ((first + col_tiles) * ctas - 1) // total - ((first + 1) * ctas - 1) // total + 1


SHAPES = [(256, 256), (32, 256), (128, 256), (64, 256), (1, 1), (2, 2), (3, 7), (1024, 1024), (5, 148), (148, 5),
          (33, 257), (8, 1024), (150, 150), (149, 149), (147, 147)]


@pytest.mark.parametrize("row_tiles,col_tiles", SHAPES)
def test_backward_tile_dealing(row_tiles, col_tiles):
    total = row_tiles * col_tiles
    ctas = min(148, total)
    seen = {}
    per_tile_slots = {rt: [] for rt in range(row_tiles)}
    sizes = []
    for c in range(ctas):
        us = units_of_cta(c, row_tiles, col_tiles, ctas)
        sizes.append(sum(u["n"] for u in us))
        for u in us:
            assert u["n"] >= 1 and u["ct0"] + u["n"] <= col_tiles
            for j in range(u["ct0"], u["ct0"] + u["n"]):
                assert (u["rt"], j) not in seen  # every tile exactly once
                seen[(u["rt"], j)] = c
            per_tile_slots[u["rt"]].append(u["slot"])
    assert len(seen) == total
    assert max(sizes) - min(sizes) <= 1  # equal shares to within one tile
    for rt in range(row_tiles):
        n = slots_of_row_tile(rt, row_tiles, col_tiles, ctas)
        # the kernel's slot numbers of a row tile are exactly 0 .. n-1, n from the epilogue's formula
        assert sorted(per_tile_slots[rt]) == list(range(n)), (rt, per_tile_slots[rt], n)
        # the 32-bit evaluation the epilogue uses when total * ctas < 2^31
        if total * ctas < 0x7fffffff:
            m32 = 0xffffffff
            first = rt * col_tiles
            a = ((((first + col_tiles) & m32) * ctas) & m32) - 1
            b = ((((first + 1) & m32) * ctas) & m32) - 1
            assert (a & m32) // total - (b & m32) // total + 1 == n


@pytest.mark.parametrize("M,rows,D", [(32768, 32768, 128), (32768, 4096, 128), (32768, 16384, 128), (131072, 16384, 256),
                                      (19200, 19200, 128), (8192, 8192, 64), (4096, 512, 128)])
def test_planner_bound_covers_every_row_tile(M, rows, D):
    """`bwd_splits` of the plan (the number of partial slots the workspace holds per row) is the
    largest slot count any row tile reaches under the kernel's dealing."""
    info = capi.plan_info(M, D, rows, capi.BF16)
    if not info["tcgen05"]:
        pytest.skip("not a tcgen05 shape")
    row_tiles, col_tiles = (rows + 127) // 128, (M + 127) // 128
    ctas = min(148, row_tiles * col_tiles)
    worst = max(slots_of_row_tile(rt, row_tiles, col_tiles, ctas) for rt in range(row_tiles))
    assert info["bwd_splits"] == worst


def test_cited_profiles_exist():
    missing = []
    for doc in ("DESIGN.md", "README.md", "INTEGRATION.md"):
        text = open(os.path.join(ROOT, doc)).read()
        for m in re.finditer(r"profiles/[A-Za-z0-9_.\-]+", text):
            name = m.group(0).rstrip(".")
            if "*" in text[m.end():m.end() + 1] or "{" in text[m.end():m.end() + 1]:
                continue  # a glob or a brace list: checked by hand
            if not os.path.exists(os.path.join(ROOT, name)) and not any(
                    f.startswith(os.path.basename(name)) for f in os.listdir(os.path.join(ROOT, "profiles"))):
                missing.append((doc, name))
    assert not missing, missing
