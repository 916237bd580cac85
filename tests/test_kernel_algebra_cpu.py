"""CPU: the two pieces of algebra the late round-2 kernel changes rest on, restated in numpy next to what they replace.
(The kernels themselves are checked on the GPU against the unmodified reference; these tests pin the reasoning.)

1. blend backward (csrc/blend.cuh): the reference walks a pixel's instances back to front with the state
   (accum_rec, last_alpha, last_color) and forms arc = last_alpha * last_color + (1 - last_alpha) * accum_rec at the next
   contributing instance, skipping the others (DGR/cuda_rasterizer/backward.cu:499-551).  The kernel carries arc itself
   and runs EVERY instance through arc <- alpha * c + (1 - alpha) * arc, T <- T / (1 - alpha), with alpha = 0 for the
   skipped ones.  Both must give the same arc / T in front of every contributing instance, bit for bit in fp32.
2. per-tile sample sort (csrc/tile_sort.cuh: rank_splitters): ranks of 256 (value, index) samples from eight sorted runs
   of 32 -- own position + the number of smaller elements in each other run by a 6-probe lower bound -- are the ranks of a
   full sort, ties broken by index."""
import numpy as np


def test_collapsed_backward_recurrence_equals_the_reference_walk():
    rng = np.random.default_rng(5)
    f32 = np.float32
    one = f32(1)
    for trial in range(200):
        n = int(rng.integers(1, 40))
        alpha = rng.uniform(1 / 255, 0.99, n).astype(f32)
        color = rng.uniform(0, 1, n).astype(f32)
        skip = rng.random(n) < 0.4                        # power > 0, alpha < 1/255, behind the last contributor, padding
        T0 = f32(rng.uniform(1e-4, 1.0))
        # the reference's state (and the kernel's before the change: selects keep skipped instances out of it)
        accum, last_a, last_c, T = f32(0), f32(0), f32(0), T0
        ref = []
        for i in range(n):
            if skip[i]:
                continue
            ra = f32(one / f32(one - alpha[i]))
            T = f32(T * ra)                               # T / (1 - alpha) as the kernel forms it
            arc = f32(f32(last_a * last_c) + f32(f32(one - last_a) * accum))
            ref.append((i, arc, T))
            accum, last_a, last_c = arc, alpha[i], color[i]
        # the kernel now: EVERY instance runs the same update, skipped ones with alpha = 0; arc is carried itself
        arc, T = f32(0), T0
        got = []
        for i in range(n):
            a = f32(0) if skip[i] else alpha[i]
            ra = f32(one / f32(one - a))                  # exactly 1 for a skipped instance
            T = f32(T * ra)
            if not skip[i]:
                got.append((i, arc, T))                   # arc in front of this instance, T behind it
            arc = f32(f32(a * color[i]) + f32(f32(one - a) * arc))
        assert len(ref) == len(got)
        for (i, arc_r, T_r), (j, arc_k, T_k) in zip(ref, got):
            assert i == j
            assert arc_r.tobytes() == arc_k.tobytes(), (trial, i)
            assert T_r.tobytes() == T_k.tobytes(), (trial, i)


def _lower_bound_32(run, k):
    """The kernel's probe sequence over a sorted run of 32: five halving steps and one last probe."""
    lo = 0
    for st in (16, 8, 4, 2, 1):
        if run[lo + st - 1] < k:
            lo += st
    return lo + (1 if run[lo] < k else 0)


def _warp_bitonic(keys):
    """Shuffle bitonic network over 32 lanes as rank_splitters runs it (xor partners, ascending final order)."""
    k = list(keys)
    size = 2
    while size <= 32:
        stride = size >> 1
        while stride > 0:
            new = list(k)
            for lane in range(32):
                other = k[lane ^ stride]
                keep_min = ((lane & size) == 0) == ((lane & stride) == 0)
                other_less = other < k[lane]
                new[lane] = other if keep_min == other_less else k[lane]
            k = new
            stride >>= 1
        size <<= 1
    return k


def test_splitter_ranks_from_sorted_runs_are_the_ranks_of_a_full_sort():
    rng = np.random.default_rng(9)
    for trial in range(50):
        # depth words with many ties (a flat surface), the sentinel of thread 255 as in the kernel
        depth = rng.integers(0, 40 if trial % 2 else 2 ** 31, 256).astype(np.uint64)
        depth[255] = 0xFFFFFFFF
        keys = [(int(depth[t]) << 8) | t for t in range(256)]
        runs = [_warp_bitonic(keys[w * 32:(w + 1) * 32]) for w in range(8)]
        for w in range(8):
            assert runs[w] == sorted(keys[w * 32:(w + 1) * 32])
        samp = [None] * 256
        for w in range(8):
            for lane in range(32):
                k = runs[w][lane]
                r = lane + sum(_lower_bound_32(runs[o], k) for o in range(8) if o != w)
                assert samp[r] is None                    # the ranks are a permutation
                samp[r] = k >> 8
        assert samp == [k >> 8 for k in sorted(keys)]
        assert samp[255] == 0xFFFFFFFF
