"""The claim the segment sort (vpic_b200/csrc/vpb_sort.cu, k_seg_pass) rests on, checked on the CPU with numpy:
if position i of the array belongs to the home segment the PREVIOUS sort's partition gives it, then a stable sort
by the CURRENT keys lists, for every voxel k, first the strays whose home is below k, then the stayers of k's own
home segment, then the strays from above and the particles appended since -- each group in array order.  That
holds whatever advance_p (keys change) and boundary_p (holes filled from the tail, particles appended, the array
shortened) did in between, which is what lets the kernel rank the stayers without atomics.  The CUDA code itself
is checked against the oracle's sort in tests/test_gpu_parity.py and tests/multi_rank_check.py --resident."""
import numpy as np
import pytest


def three_region_order(keys, home_partition, home_np, nv):
    n = len(keys)
    pos = np.arange(n)
    # home voxel of every position: the segment of the previous partition it lies in; appended positions: nv
    home = np.searchsorted(home_partition[1:nv + 1], pos, side="right")
    home[pos >= home_np] = nv
    out = []
    for k in range(nv):
        mine = pos[keys == k]
        below, stay, above = mine[home[mine] < k], mine[home[mine] == k], mine[home[mine] > k]
        out += [below, stay, above]          # each already ascending in position
    return np.concatenate(out) if out else np.zeros(0, dtype=np.int64)


@pytest.mark.parametrize("seed", range(8))
def test_three_regions_equal_the_stable_sort(seed):
    rng = np.random.default_rng(seed)
    nv = int(rng.integers(3, 40))
    n0 = int(rng.integers(0, 600))
    old_keys = np.sort(rng.integers(0, nv, n0))                       # the array right after the previous sort
    part = np.concatenate([[0], np.cumsum(np.bincount(old_keys, minlength=nv))])   # partition[0..nv], [nv] = n0
    keys = old_keys.copy()
    # advance_p: some particles change voxel (near or far)
    move = rng.random(n0) < rng.uniform(0, 0.6)
    keys[move] = np.clip(keys[move] + rng.integers(-3, 4, move.sum()), 0, nv - 1)
    far = rng.random(n0) < 0.05
    keys[far] = rng.integers(0, nv, far.sum())
    # boundary_p: holes filled with tail particles (the array shrinks), then arrivals appended
    nrem = int(rng.integers(0, n0 // 4 + 1))
    if nrem and n0:
        holes = rng.choice(n0, size=min(nrem, n0), replace=False)
        keep = np.ones(n0, dtype=bool); keep[holes] = False
        survivors_tail = [i for i in range(n0 - 1, -1, -1) if keep[i]]
        new_n = n0 - len(holes)
        keys2 = keys.copy()
        for hpos in sorted(h for h in holes if h < new_n):
            keys2[hpos] = keys[survivors_tail.pop(0)]
        keys = keys2[:new_n]
    napp = int(rng.integers(0, 50))
    keys = np.concatenate([keys, rng.integers(0, nv, napp)]).astype(np.int64)
    home_np = min(n0, len(keys) - napp)
    got = three_region_order(keys, part, home_np, nv)
    want = np.argsort(keys, kind="stable")
    assert np.array_equal(got, want)
