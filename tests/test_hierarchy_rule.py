"""The rule behind k_leaves_build / k_build_upper (csrc/bvh.cu), as an executable statement on the CPU: emitting the LBVH bottom-up
— a finished node [l, r] joins the neighbour it shares the longer (key, index) prefix with, and is numbered by its range's right end
if it is a left child, by the left end if it is a right child, 0 if it is the root — produces exactly the arrays of the search-based
construction the reference uses (Karras 2012; `getTheOtherEnd` / `findSplitPosition` / `computeNodeRange`, `src/bvh.cu:64-126,150-218`): children, parents, range ends.  Also the test the kernel uses to
decide whether a parent lies inside a block of 256 leaves: one comparison of prefix lengths.  The GPU counterpart is
tests/test_gpu_round2.py::test_bottom_up_hierarchy_is_the_search_based_tree."""
import random


def _clz(x, bits):
    return bits - x.bit_length()


def delta(keys, n, i, j):
    if j < 0 or j >= n:
        return -1
    a, b = keys[i], keys[j]
    if a != b:
        return _clz(a ^ b, 64)
    return 64 + _clz(i ^ j, 32)            # equal keys: the index decides, as in the kernels


def search_based(keys):
    n = len(keys)
    left, right, other = [0] * (n - 1), [0] * (n - 1), [0] * (n - 1)
    parent_int, parent_leaf = [-2] * (n - 1), [-2] * n
    for i in range(n - 1):
        d = 1 if delta(keys, n, i, i + 1) > delta(keys, n, i, i - 1) else -1
        d_min = delta(keys, n, i, i - d)
        lmax = 2
        while delta(keys, n, i, i + lmax * d) > d_min:
            lmax <<= 1
        length, t = 0, lmax >> 1
        while t > 0:
            if delta(keys, n, i, i + (length + t) * d) > d_min:
                length += t
            t >>= 1
        j = i + length * d
        other[i] = j
        d_node = delta(keys, n, i, j)
        s, t = 0, (length + 1) >> 1
        while True:
            if delta(keys, n, i, i + (s + t) * d) > d_node:
                s += t
            if t == 1:
                break
            t = (t + 1) >> 1
        gamma = i + s * d + min(d, 0)
        lo, hi = min(i, j), max(i, j)
        left[i] = ~gamma if lo == gamma else gamma
        right[i] = ~(gamma + 1) if hi == gamma + 1 else gamma + 1
        for c in (left[i], right[i]):
            if c < 0:
                parent_leaf[~c] = i
            else:
                parent_int[c] = i
    parent_int[0] = -1
    return left, right, other, parent_int, parent_leaf


def bottom_up(keys, rng):
    n = len(keys)
    left, right, other = [None] * (n - 1), [None] * (n - 1), [None] * (n - 1)
    parent_int, parent_leaf = [-2] * (n - 1), [-2] * n
    slot = {}                                    # split position -> what the first child to arrive left there
    leaves = list(range(n))
    rng.shuffle(leaves)                          # any arrival order gives the same tree
    for leaf in leaves:
        l = r = leaf
        cid = ~leaf
        while not (l == 0 and r == n - 1):
            go_right = True if l == 0 else False if r == n - 1 else delta(keys, n, r, r + 1) > delta(keys, n, l - 1, l)
            split = r if go_right else l - 1
            if split not in slot:
                slot[split] = (cid, l if go_right else r)
                break
            sid, send = slot.pop(split)
            lid, rid, l, r = (cid, sid, l, send) if go_right else (sid, cid, send, r)
            if l == 0 and r == n - 1:
                k = 0
            else:
                up = True if l == 0 else False if r == n - 1 else delta(keys, n, r, r + 1) > delta(keys, n, l - 1, l)
                k = r if up else l
            left[k], right[k], other[k] = lid, rid, (l if k == r else r)
            for c in (lid, rid):
                if c < 0:
                    parent_leaf[~c] = k
                else:
                    parent_int[c] = k
            cid = k
    parent_int[0] = -1
    return left, right, other, parent_int, parent_leaf


def test_bottom_up_emission_equals_the_search_based_tree():
    rng = random.Random(1)
    for _ in range(200):
        n = rng.choice([2, 3, 4, 5, 7, 8, 16, 33, 100, 257, 600])
        bits = rng.choice([3, 6, 12, 30, 63])        # few bits: many equal keys
        keys = sorted(rng.getrandbits(bits) for _ in range(n))
        assert bottom_up(keys, rng) == search_based(keys)


def test_parent_inside_block_is_one_prefix_comparison():
    """For a left child ending at `split` inside a block [b0, b_end], its parent ends inside the block iff
    delta(split + 1, b_end + 1) < delta(split, split + 1); for a right child starting at split + 1 the parent starts inside iff
    delta(b0 - 1, split) < delta(split, split + 1).  (k_leaves_build reads both from running minima of the neighbour deltas.)"""
    rng = random.Random(2)
    for _ in range(100):
        n = rng.choice([5, 16, 33, 100, 257, 600])
        keys = sorted(rng.getrandbits(rng.choice([3, 6, 12, 40])) for _ in range(n))
        left, right, other, _, _ = search_based(keys)
        block = rng.choice([4, 8, 32])

        def span(k):
            return min(k, other[k]), max(k, other[k])

        for k in range(n - 1):
            lo, hi = span(k)
            lspan = (~left[k], ~left[k]) if left[k] < 0 else span(left[k])
            rspan = (~right[k], ~right[k]) if right[k] < 0 else span(right[k])
            split = lspan[1]
            assert rspan[0] == split + 1
            q = delta(keys, n, split, split + 1)
            b0 = (lspan[0] // block) * block
            b_end = b0 + block - 1
            if lspan[1] <= b_end:                 # the left child lies in one block
                assert ((b_end + 1 >= n) or delta(keys, n, split + 1, b_end + 1) < q) == (hi <= b_end)
            b0 = (rspan[0] // block) * block
            b_end = b0 + block - 1
            if rspan[1] <= b_end:                 # the right child lies in one block
                assert ((b0 == 0) or delta(keys, n, b0 - 1, split) < q) == (lo >= b0)
