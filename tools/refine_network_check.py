"""numpy emulation of the register/shuffle bitonic network of k_refine_runs (csrc/build.cu): same
thread -> element mapping, same compare-exchange rules, 512 "threads". Checks that every segment of a
1024-key run ends up sorted for all five segment sizes. No GPU needed."""
import numpy as np


def refine_sort(keys, seg_shift):
    seg_size = 1 << seg_shift
    keys = keys.copy()
    tid = np.arange(512)
    lane = tid & 31
    e0 = ((tid >> 5) << 6) + lane
    e1 = e0 + 32

    def in_warp(size_from, size_to):
        a, b = keys[e0].copy(), keys[e1].copy()
        size = size_from
        while size <= size_to:
            asc0 = (size == seg_size) | ((e0 & size) == 0)
            asc1 = (size == seg_size) | ((e1 & size) == 0)
            stride = min(size >> 1, 32)
            if stride == 32:
                lo, hi = np.minimum(a, b), np.maximum(a, b)
                a, b = np.where(asc0, lo, hi), np.where(asc0, hi, lo)
                stride = 16
            while stride > 0:
                src = (tid & ~31) | (lane ^ stride)          # shfl_xor
                pa, pb = a[src], b[src]
                lower = (lane & stride) == 0
                a = np.where((pa < a) == (lower == asc0), pa, a)
                b = np.where((pb < b) == (lower == asc1), pb, b)
                stride >>= 1
            size <<= 1
        keys[e0], keys[e1] = a, b

    in_warp(2, min(64, seg_size))
    size = 128
    while size <= seg_size:
        stride = size >> 1
        while stride >= 64:
            i = 2 * tid - (tid & (stride - 1))
            j = i + stride
            asc = (size == seg_size) | ((i & size) == 0)
            a, b = keys[i].copy(), keys[j].copy()
            swap = (a > b) == asc
            keys[i], keys[j] = np.where(swap, b, a), np.where(swap, a, b)
            stride >>= 1
        in_warp(size, size)
        size <<= 1
    return keys


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for seg_shift in (10, 9, 8, 7, 6):
        for _ in range(50):
            k = rng.permutation(1 << 20)[:1024].astype(np.uint64)
            seg = (np.arange(1024) >> seg_shift).astype(np.uint64)
            keys = (seg << np.uint64(42)) | k
            assert np.array_equal(refine_sort(keys, seg_shift), np.sort(keys)), seg_shift
    print("network ok")
