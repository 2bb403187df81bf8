"""Descriptor hand-over between rank processes (the bootstrap of the multicast heap, `parallel/symm.py`).

No reference counterpart: the reference's transport is socket.io; this is the plumbing that replaces the
cudaIpc handle exchange when the heap is created through the driver's VMM calls.  Runs on CPU: pipes stand in for the
exported memory handles.
"""
import multiprocessing as mp
import os


def _worker(rank, world, token, q):
    from distriflow_b200.parallel.symm import exchange_fds
    r, w = os.pipe()
    os.write(w, bytes([rank]) * 4)
    os.close(w)
    extra_r, extra_w = os.pipe()
    os.close(extra_w)
    # rank 0 sends two descriptors (its memory handle and the multicast handle), everybody else one
    send = {d: [r] + ([extra_r] if rank == 0 else []) for d in range(world) if d != rank}
    got = exchange_fds(rank, world, send, token, timeout=30.0)
    counts = {}
    for src, fds in got.items():
        for f in fds:
            os.fstat(f)                      # a live descriptor in this process
        counts[src] = len(fds)
    q.put((rank, counts))


def test_exchange_fds_all_to_all():
    world = 4
    ctx = mp.get_context("fork")
    q = ctx.Queue()
    token = f"test{os.getpid()}"
    procs = [ctx.Process(target=_worker, args=(r, world, token, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=60) for _ in procs)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    for r in range(world):
        assert set(res[r]) == set(range(world)) - {r}
        for src, n in res[r].items():
            assert n == (2 if src == 0 else 1)
