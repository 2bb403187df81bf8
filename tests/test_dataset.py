"""DistriDataset dispatcher properties (reference src/server/dataset.ts:43-67)."""
import pytest
import torch

from distriflow_b200 import DistriDataset


def make(n=10, bs=2, epochs=2, small=False):
    x = torch.arange(n, dtype=torch.float32).reshape(n, 1)
    return DistriDataset(x, x.clone(), {"batchSize": bs, "epochs": epochs, "smallLastBatch": small})


def test_fcfs_order_and_epoch_advance():
    ds = make()
    assert ds.batches == 5
    seen = []
    for _ in range(5):
        b = ds.next()
        seen.append(b.batch)
        assert b.epoch == 0 and torch.equal(b.x[:, 0], torch.tensor([2.0 * b.batch, 2.0 * b.batch + 1]))
        ds.complete_batch(b.batch)
    assert seen == [0, 1, 2, 3, 4]
    b = ds.next()
    assert b.epoch == 1 and b.batch == 0 and len(ds.CompletedEpochs) == 1


def test_incomplete_batches_are_reissued_and_epoch_waits():
    ds = make(epochs=1)
    got = [ds.next().batch for _ in range(5)]
    for b in got:
        if b not in (1, 3):
            ds.complete_batch(b)
    # fresh ones ran out: stragglers 1 and 3 are handed out again, epoch does not advance
    assert ds.next().batch == 1
    assert ds.next().batch == 3
    assert ds.next().batch == 1
    assert ds.epoch == 0
    ds.complete_batch(1)
    assert ds.next().batch == 3
    ds.complete_batch(3)
    assert ds.next().done


def test_complete_is_idempotent_and_epoch_checked():
    ds = make()
    b = ds.next()
    assert ds.complete_batch(b.batch) is True
    assert ds.complete_batch(b.batch) is False and ds.duplicates == 1
    assert ds.complete_batch(1, epoch=7) is False
    assert 1 in ds.incompleteBatches


def test_done_after_last_epoch():
    ds = make(n=4, bs=2, epochs=2)
    n = 0
    for b in ds:
        ds.complete_batch(b.batch)
        n += 1
    assert n == 4 and ds.next().done


def test_ragged_last_batch():
    ds = make(n=5, bs=2, epochs=1, small=True)
    assert ds.batches == 3
    sizes = []
    for b in ds:
        sizes.append(b.x.shape[0])
        ds.complete_batch(b.batch)
    assert sizes == [2, 2, 1]
    ds = make(n=5, bs=2, epochs=1, small=False)
    assert ds.batches == 2


def test_uuid_ids_and_preprocess():
    ds = make()
    assert len(set(ds.epochIDs)) == 2 and len(ds.microBatchIDs) == 5
    ds.add_preprocess_callback(lambda b: type(b)(b.batch, b.epoch, b.x * 2, b.y, b.batchId, b.epochId))
    b = ds.next()
    assert torch.equal(b.x[:, 0], torch.tensor([0.0, 2.0])) and b.batchId == ds.microBatchIDs[0]
    assert ds.queuedMicroBatch == [1, 2, 3, 4]


# ---------------------------------------------------------------------------------------------- randomized schedules
from hypothesis import given, settings, strategies as st    # noqa: E402


@settings(max_examples=60, deadline=None)
@given(n=st.integers(min_value=1, max_value=40), bs=st.integers(min_value=1, max_value=9), epochs=st.integers(min_value=1, max_value=3),
       small=st.booleans(), seed=st.integers(min_value=0, max_value=10 ** 6), loss_rate=st.floats(min_value=0.0, max_value=0.6))
def test_dispatcher_invariants_under_random_workers(n, bs, epochs, small, seed, loss_rate):
    """Workers that take batches, finish them late, twice, or never (reference semantics: `dataset.ts:43-67`):
    every batch of every epoch is completed exactly once, epochs never overlap, `done` comes only after the last one,
    and a batch is re-issued only while it is incomplete."""
    import random
    import torch
    from distriflow_b200 import DistriDataset
    rng = random.Random(seed)
    if n < bs and not small:
        return                                   # constructor rejects it (covered by test_ragged_last_batch)
    ds = DistriDataset(torch.arange(n).float().reshape(n, 1), torch.arange(n).float().reshape(n, 1),
                       {"batchSize": bs, "epochs": epochs, "smallLastBatch": small}, use_uuids=False)
    nb = ds.batches
    completed = [[0] * nb for _ in range(epochs)]
    rows_seen = [set() for _ in range(epochs)]
    in_flight = []                               # (epoch, batch) handed out, not yet reported
    steps = 0
    while True:
        steps += 1
        assert steps < 20000, "dispatcher does not terminate"
        if in_flight and (rng.random() < 0.5 or len(in_flight) > 6):
            e, b = in_flight.pop(rng.randrange(len(in_flight)))
            if rng.random() < loss_rate:
                continue                         # the worker died: never reports
            if ds.complete_batch(b, e):
                completed[e][b] += 1
            if rng.random() < 0.1:               # duplicate report must be ignored
                assert ds.complete_batch(b, e) is False
            continue
        batch = ds.next()
        if batch.done:
            if in_flight:                        # stale reports after the end change nothing
                e, b = in_flight.pop()
                assert ds.complete_batch(b, e) is False
                continue
            break
        if batch.x is None:                      # nothing to hand out right now (everything in flight): report something
            if not in_flight:
                continue
            e, b = in_flight.pop(0)
            if ds.complete_batch(b, e):
                completed[e][b] += 1
            continue
        e, b = batch.epoch, batch.batch
        assert e == ds.epoch and completed[e][b] == 0, "a completed batch was issued again"
        assert all(completed[k][j] == 1 for k in range(e) for j in range(nb)), "epochs overlapped"
        lo = b * bs
        assert batch.x.reshape(-1).tolist() == list(map(float, range(lo, min(lo + bs, n))))[:batch.x.shape[0]]
        rows_seen[e].update(int(v) for v in batch.x.reshape(-1).tolist())
        in_flight.append((e, b))
    assert all(c == 1 for ep in completed for c in ep)
    expect = set(range(n if small else nb * bs))
    assert all(r == expect for r in rows_seen)
