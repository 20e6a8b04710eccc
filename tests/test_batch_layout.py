"""Host-side checks of the batched kernel's shared-memory layouts (ekf-slam_b200/csrc/ekf_batch.cuh): the functions are
`__host__ __device__`, so a tiny nvcc-built host program dumps them (no GPU needed).

DD 0/1: rows on 16-byte boundaries.  DD 2: 8x8 tiles in DMMA accumulator order with swizzled 16-byte chunks - the tile
access must stay conflict-free and a walk down a column / the diagonal must spread over the banks."""
import os
import shutil
import subprocess
from collections import defaultdict

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


@pytest.fixture(scope="module")
def dump(tmp_path_factory):
    if not os.path.exists(NVCC):
        pytest.skip("nvcc not available")
    exe = str(tmp_path_factory.mktemp("layout") / "batch_layout_dump")
    subprocess.check_call([NVCC, "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a",
                           "-I" + os.path.join(ROOT, "ekf-slam_b200", "csrc"),
                           os.path.join(ROOT, "tests", "cpp", "batch_layout_dump.cu"), "-o", exe])

    def run(n, dd):
        idx, chunks, extra = {}, {}, {}
        for line in subprocess.check_output([exe, str(n), str(dd)], text=True).split("\n"):
            p = line.split()
            if not p:
                continue
            if p[0] == "S" or p[0] == "SMEM":
                extra[p[0]] = int(p[1])
            elif p[0] == "CH":
                chunks[(int(p[1]), int(p[2]))] = int(p[3])
            else:
                idx[(int(p[0]), int(p[1]))] = int(p[2])
        return idx, chunks, extra
    return run


def _wavefronts(offsets):
    """8-byte accesses of one warp: 16 double-wide banks; wavefronts = most distinct addresses on one bank"""
    banks = defaultdict(set)
    for o in offsets:
        banks[o % 16].add(o)
    return max(len(v) for v in banks.values())


@pytest.mark.parametrize("n", [43, 103, 131, 8, 9])
@pytest.mark.parametrize("dd", [0, 2])
def test_layout_is_a_bijection_inside_its_storage(dump, n, dd):
    idx, _, extra = dump(n, dd)
    assert len(idx) == n * (n + 1) // 2
    assert len(set(idx.values())) == len(idx)
    assert min(idx.values()) >= 0 and max(idx.values()) < extra["S"]
    assert extra["SMEM"] <= 200 * 1024
    if dd == 0:
        assert all(idx[(i, 0)] % 2 == 0 for i in range(n))          # rows start on 16-byte boundaries
        assert all(idx[(i, c)] == idx[(i, 0)] + c for (i, c) in idx)


def test_tile_access_is_conflict_free_and_matches_the_fragment(dump):
    idx, chunks, _ = dump(43, 2)
    for I in (0, 1):
        for qw in range(4):                                              # a 128-bit access is served per quarter-warp
            ch = [chunks[(I, l)] for l in range(qw * 8, qw * 8 + 8)]
            assert sorted(c % 8 for c in ch) == list(range(8)) and len({c // 8 for c in ch}) == 1
        assert sorted(chunks[(I, l)] for l in range(32)) == list(range(32))
    # lane l of tile (I, J) holds row l / 4, columns 2 (l % 4), +1: element index == tile base + 2 * chunk (+1)
    for I in range(5):
        for J in range(I + 1):
            base = (I * (I + 1) // 2 + J) * 64
            for l in range(32):
                i, c = 8 * I + l // 4, 8 * J + 2 * (l % 4)
                if i < 43 and c + 1 <= i:
                    assert idx[(i, c)] == base + 2 * chunks[(I & 1, l)]
                    assert idx[(i, c + 1)] == idx[(i, c)] + 1


def test_column_walks_and_candidate_blocks_spread_over_the_banks(dump):
    idx, _, _ = dump(43, 2)

    def at(i, c):
        return idx[(i, c)] if i >= c else idx[(c, i)]
    for c in (0, 1, 2):                                                  # G = Sigma H^T gathers columns 0..2 ...
        assert _wavefronts([at(i, c) for i in range(32)]) <= (4 if c == 0 else 5)   # floor: 4 (32 rows, one parity)
        assert _wavefronts([at(i, c) for i in range(32, 43)]) <= 2
    for c in (9, 10, 21, 22, 41, 42):                                    # ... and a, a+1 (part row walk, part column walk)
        assert _wavefronts([at(i, c) for i in range(32)]) <= 6           # unswizzled tiles: 8 to 16
        assert _wavefronts([at(i, c) for i in range(32, 43)]) <= 3
    a = [3 + 2 * k for k in range(20)]                                   # the 20 candidates' rows
    for c in (0, 2):
        assert _wavefronts([at(x, c) for x in a]) <= 3 and _wavefronts([at(x + 1, c) for x in a]) <= 3
    assert _wavefronts([at(x, x) for x in a]) <= 3 and _wavefronts([at(x + 1, x + 1) for x in a]) <= 3
    assert _wavefronts([at(x + 1, x) for x in a]) <= 5
