"""CPU test of the first-fit range allocator behind the headless renderer handoff.

Pinned to the reference's OWN allocator (Freelist::AllocateBlock / FreeByOffset / FreeSpace, core/utils/freelist.cpp,
compiled unmodified into oracle/_ref) on the sequences it can run: allocation only -- all the engine ever does,
FreeDataRange is a no-op (vulkan.cpp:29-31) -- and last-in-first-out releases.  The reference's FreeBlock
(freelist.cpp:36-72) merges neighbours through dangling pointers: with releases in any other order it crashes inside
FreeByOffset after a few dozen steps (observed here), so random release orders are checked against a plain first-fit
model instead (walk from the lowest offset, take the first free range that fits, split it; merge on release)."""
import numpy as np
import pytest

from oracle import refcpu

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


def test_random_release_order_matches_a_first_fit_model():
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


@pytest.mark.skipif(not refcpu.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("lifo", [False, True], ids=["alloc_only", "lifo_release"])
def test_offsets_and_free_space_equal_the_reference_freelist(lifo):
    L = capi.lib()
    total = 64 << 20
    a = L.vkm_rangealloc_create(total)
    ref = refcpu.Freelist(total)
    rng = np.random.default_rng(7)
    live = []
    for step in range(3000):
        if lifo and live and rng.random() < 0.4:
            off, size = live.pop()
            assert L.vkm_rangealloc_free(a, off, size) == 0
            assert ref.free(off)
        else:
            size = int(rng.integers(1, 200_000 if lifo else 800)) * 64
            if L.vkm_rangealloc_free_space(a) < size:  # (the reference logs through its Logger singleton when it is full)
                continue
            got, want = L.vkm_rangealloc_alloc(a, size), ref.alloc(size)
            assert got == want, (step, got, want)
            live.append((got, size))
        assert L.vkm_rangealloc_free_space(a) == ref.free_space()
    # exact fit of what is left, as the reference does it
    rest = ref.free_space()
    if rest:
        assert L.vkm_rangealloc_alloc(a, rest) == ref.alloc(rest)
        assert L.vkm_rangealloc_free_space(a) == ref.free_space() == 0
    ref.close()
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
