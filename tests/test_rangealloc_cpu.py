"""CPU test of the first-fit range allocator behind the headless renderer handoff.  The
expected offsets come from a direct restatement of the reference's Freelist (core/utils/freelist.cpp:144-159:
walk the node list from the lowest offset, take the first free node that fits, split it)."""
import numpy as np

from vulkanengine_b200 import capi

UINT64_MAX = 2**64 - 1


class FirstFit:  # restatement of Freelist::AllocateBlock / FreeBlock on a sorted free list
    def __init__(self, total):
        self.free = [[0, total]]

    def alloc(self, size):
        for i, (off, sz) in enumerate(self.free):
            if sz >= size:
                self.free[i] = [off + size, sz - size]
                if sz == size:
                    del self.free[i]
                return off
        return UINT64_MAX

    def release(self, off, size):
        self.free.append([off, size])
        self.free.sort()
        merged = [self.free[0]]
        for o, s in self.free[1:]:
            if merged[-1][0] + merged[-1][1] == o:
                merged[-1][1] += s
            else:
                merged.append([o, s])
        self.free = merged


def test_first_fit_sequence_matches_reference_freelist():
    L = capi.lib()
    total = 64 << 20  # the engine's global vertex buffer size (vulkan.cpp:599-628)
    a = L.vkm_rangealloc_create(total)
    ref = FirstFit(total)
    rng = np.random.default_rng(0)
    live = []
    for step in range(3000):
        if live and rng.random() < 0.4:
            off, size = live.pop(int(rng.integers(0, len(live))))
            assert L.vkm_rangealloc_free(a, off, size) == 0
            ref.release(off, size)
        else:
            size = int(rng.integers(1, 200_000)) * 64
            got, want = L.vkm_rangealloc_alloc(a, size), ref.alloc(size)
            assert got == want
            if got != UINT64_MAX:
                live.append((got, size))
        assert L.vkm_rangealloc_free_space(a) == sum(s for _, s in ref.free)
    # double free / overlap is rejected
    off, size = live[0]
    assert L.vkm_rangealloc_free(a, off, size) == 0
    assert L.vkm_rangealloc_free(a, off, size) != 0
    L.vkm_rangealloc_destroy(a)


def test_allocation_without_frees_is_sequential_like_the_engine():
    # the reference never frees (FreeDataRange is a no-op, vulkan.cpp:29-31): offsets are running sums
    L = capi.lib()
    a = L.vkm_rangealloc_create(1 << 20)
    sizes = [640, 64, 6400, 128]
    offs = [L.vkm_rangealloc_alloc(a, s) for s in sizes]
    assert offs == [0, 640, 704, 7104]
    assert L.vkm_rangealloc_alloc(a, 1 << 21) == UINT64_MAX
    L.vkm_rangealloc_destroy(a)
