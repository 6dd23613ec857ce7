"""CPU model of the page allocator of rows_ring_kernel (llm-mcp_b200/csrc/etn_kernels.cuh,
`issue_more` and the consumer loop around it): every warp owns a ring of P pages of 512 bytes,
keeps up to R row prefixes in flight, places each contiguously (a row that would straddle the
end of the ring starts over at page 0 and holds the skipped pages too) and frees rows in the
order it issued them.

The replay below follows the kernel's bookkeeping statement by statement and checks, over 10^5
random row-length sequences, what the kernel relies on and what no output comparison can
show directly:
  * no two rows in flight ever share a page, and no row runs past the ring;
  * a row is always issued before the consumer waits for it (no stall / deadlock);
  * a slot (row mod R) is never reused while its row is in flight;
  * every page is given back: the ring is empty when the warp is done.
compute-sanitizer is closed on this pool, so this model plus the bounded mbarrier waits are
what stands behind the ring's memory safety."""
import random

PAGE, R = 512, 8


def replay(lengths, P=16):
    """lengths: staged bytes per row of one warp (0 = row not staged: odd row or longer than the ring)."""
    ring = P * PAGE
    n = len(lengths)
    owner = [-1] * P                       # which row holds each page (-1 = free)
    meta = [None] * R                      # per slot: (pos, bytes, pages held)
    in_flight_slot = [False] * R
    head_row, head_page, used, inflight = 0, 0, 0, 0
    max_inflight = 0

    def issue_more():
        nonlocal head_row, head_page, used, inflight, max_inflight
        while head_row < n and inflight < R:
            nb = lengths[head_row] if 0 < lengths[head_row] <= ring else 0
            slot = head_row % R
            assert not in_flight_slot[slot], "slot reused while its row is in flight"
            if nb == 0:
                meta[slot] = (0, 0, 0)
            else:
                need = -(-nb // PAGE)
                if used == 0:
                    head_page = 0
                pos, waste = head_page, 0
                if pos + need > P:
                    waste, pos = P - pos, 0
                if used + waste + need > P:
                    break
                assert pos + need <= P, "row runs past the ring"
                for pg in range(pos, pos + need):
                    assert owner[pg] == -1, f"page {pg} given to row {head_row} while row {owner[pg]} holds it"
                    owner[pg] = head_row
                meta[slot] = (pos, nb, waste + need)
                used += waste + need
                head_page = 0 if pos + need == P else pos + need
            in_flight_slot[slot] = True
            head_row += 1
            inflight += 1
            max_inflight = max(max_inflight, inflight)

    issue_more()
    for row in range(n):
        assert head_row > row, f"consumer would wait for row {row}, which was never issued (stall)"
        slot = row % R
        pos, nb, held = meta[slot]
        if nb:
            need = -(-nb // PAGE)
            assert all(owner[pg] == row for pg in range(pos, pos + need)), "row's pages were overwritten"
            for pg in range(pos, pos + need):
                owner[pg] = -1
            used -= held
        in_flight_slot[slot] = False
        inflight -= 1
        issue_more()
    assert used == 0 and inflight == 0 and all(o == -1 for o in owner), "pages leaked"
    return max_inflight


def test_ring_allocator_invariants_over_random_sequences():
    rng = random.Random(20261107)
    ks = [64, 128, 256, 512, 1024]                       # config 5's prefixes, float64
    deepest = 0
    for trial in range(100_000):
        n = rng.randint(1, 16) if trial % 10 else rng.randint(17, 64)
        kind = trial % 5
        if kind == 0:                                       # config 5: a few requests with one k each
            seq = []
            while len(seq) < n:
                seq += [rng.choice(ks) * 8] * rng.randint(1, n)
            seq = seq[:n]
        elif kind == 1:                                     # anything 16-byte tileable up to the whole ring
            seq = [16 * rng.randint(1, 512) for _ in range(n)]
        elif kind == 2:                                     # page-boundary cases
            seq = [rng.choice([16, 496, 512, 528, 1024, 4096, 7680, 8176, 8192]) for _ in range(n)]
        elif kind == 3:                                     # mixed with rows that are not staged at all
            seq = [rng.choice([0, 0, 8208, 16 * rng.randint(1, 512)]) for _ in range(n)]
        else:                                               # full-ring rows back to back, then tiny ones
            seq = [8192] * rng.randint(1, 4) + [16 * rng.randint(1, 8) for _ in range(n)]
        deepest = max(deepest, replay(seq))
    assert deepest == R                                     # short rows do fill all eight slots


def test_ring_allocator_other_ring_sizes():
    """The page count is a template parameter (tools/tune.cu tries 16..48 pages)."""
    rng = random.Random(7)
    for P in (16, 24, 32, 48):
        for _ in range(2000):
            replay([16 * rng.randint(1, P * 32) for _ in range(rng.randint(1, 40))], P=P)
