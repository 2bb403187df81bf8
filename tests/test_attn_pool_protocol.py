"""Protocol model of the attention-backward operand pool (`attention_sm100.cu: BwdPool`, CPU-only).

The kernel hands its four 32 KiB shared-memory slots out round robin, skipping the slots pinned by the current item's K/V
pairs; the TMA producer and the MMA thread replay the SAME allocation sequence independently, the producer blocks on a
slot's `slot_empty` barrier (all readers done), the MMA thread on its `slot_full` barrier (data landed).  This model runs
both sequences against each other with the kernel's release rules ({Q_i, dO_i}: after the query tile's last product; K/V:
at the end of the item) and checks what the kernel relies on:

  * the two roles compute the same slot for every load (by construction: same function, same calls) and a slot is never
    filled while a previous content still has readers,
  * no schedule deadlocks, for every mix of item lengths (1..8 query tiles, with and without a second key tile),
  * the next item's first K/V pair is loaded BEFORE the current item ends whenever the item has at least two query tiles
    (the prefetch the pool exists for).
"""
import itertools

POOL = 4


class Pool:
    def __init__(self):
        self.next = 0

    def take(self, skip_mask: int) -> int:                 # mirrors BwdPool::take
        s = self.next
        for _ in range(POOL - 1):
            if (skip_mask >> s) & 1:
                s = (s + 1) % POOL
        self.next = (s + 1) % POOL
        return s


def loads_of(items):
    """The allocation sequence both roles replay: [(kind, item, tile, slot)], kind in {"kva", "kvb", "q"}."""
    pool, pins, out = Pool(), 0, []
    for w, (n_q, has_b) in enumerate(items):
        old = pins
        kva = pool.take(old)
        pins = 1 << kva
        out.append(("kva", w, -1, kva))
        if has_b:
            kvb = pool.take(old | pins)
            pins |= 1 << kvb
            out.append(("kvb", w, -1, kvb))
        for t in range(n_q):
            out.append(("q", w, t, pool.take(pins)))
    return out


def simulate(items):
    """Producer fills in sequence order when the slot is free; the consumer (MMA thread) walks items / tiles in order, needs
    kva + the tile's slot (and kvb from the second tile on) and releases slots by the kernel's rules.  Returns the number of
    loads the producer had completed when each item ended."""
    seq = loads_of(items)
    owner = [None] * POOL                                   # what occupies a slot: index into seq, or None when free
    filled = set()
    prod = 0
    done_at_item_end = []

    def pump():
        nonlocal prod
        while prod < len(seq) and owner[seq[prod][3]] is None:
            owner[seq[prod][3]] = prod
            filled.add(prod)
            prod += 1

    index = {(k, w, t): i for i, (k, w, t, _) in enumerate(seq)}
    for w, (n_q, has_b) in enumerate(items):
        for t in range(n_q):
            need = [index[("kva", w, -1)], index[("q", w, t)]]
            if has_b and t >= 1:
                need.append(index[("kvb", w, -1)])
            pump()
            assert all(i in filled for i in need), f"deadlock: item {w} tile {t} waits for {[seq[i] for i in need if i not in filled]}"
            slot = seq[index[("q", w, t)]][3]
            assert owner[slot] == index[("q", w, t)], "slot overwritten while in use"
            owner[slot] = None                              # dK of the tile's last sub-iteration was the last reader
        pump()
        done_at_item_end.append(prod)
        for k in ("kva", "kvb"):
            if (k, w, -1) in index:
                assert owner[seq[index[(k, w, -1)]][3]] == index[(k, w, -1)], "K/V slot overwritten during its item"
                owner[seq[index[(k, w, -1)]][3]] = None
    pump()
    assert prod == len(seq)
    return seq, done_at_item_end


def test_pinned_slots_are_never_handed_out_and_all_slots_stay_distinct():
    for items in itertools.product([(1, False), (2, True), (3, True), (8, True), (5, False)], repeat=3):
        seq = loads_of(items)
        pins = {}
        for kind, w, t, slot in seq:
            if kind in ("kva", "kvb"):
                pins.setdefault(w, set()).add(slot)
            else:
                assert slot not in pins[w], (items, kind, w, t, slot)
        for w, s in pins.items():
            assert len(s) == (2 if items[w][1] else 1)


def test_no_schedule_deadlocks_and_nothing_is_overwritten_while_in_use():
    shapes = [(n, b) for n in (1, 2, 3, 4, 8) for b in (False, True) if not (b and n < 2)]
    for items in itertools.product(shapes, repeat=3):
        simulate(list(items))


def test_next_items_first_kv_pair_is_prefetched_before_the_item_ends():
    for n_q in (2, 3, 5, 8):
        items = [(n_q, True), (n_q, True), (n_q, True)]
        seq, done = simulate(items)
        for w in range(2):
            first_next = next(i for i, (k, ww, _, _) in enumerate(seq) if ww == w + 1 and k == "kva")
            assert done[w] > first_next, (n_q, w, done, first_next)
