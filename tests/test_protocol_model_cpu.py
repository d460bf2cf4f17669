"""Exhaustive interleaving check of the mailbox handshake used by the momentum / staleness push (csrc/comm.cu K2a/K2b).

One slot = one (worker, chunk) pair.  The kernels' steps, in program order (every cross-GPU flag access is a release store or an
acquire load, so steps of one side are observed in order by the other):

  worker push_mailbox:  w1 read local `busy` (skip if set)   w2 copy accumulator -> owner's mailbox, zero accumulator
                        w3 busy = 1                          w4 release-store owner's `full` = version + 1
  owner  owner_apply :  o1 acquire-load `full` (skip if 0)   o2 read the mailbox, fold it into the master shard
                        o3 full = 0                          o4 release-store the worker's `busy` = 0

The model explores every interleaving of a worker that produces K gradient quanta and an owner that keeps polling, and asserts
 * no data race: the worker never writes the mailbox while the owner is between o1 (saw it full) and o2 (finished reading);
 * nothing is lost or applied twice: applied + still-accumulated + parked-in-mailbox == produced, in every reachable state;
 * liveness: from every reachable state the system can reach "everything applied".
A deliberately wrong owner (o4 before o3: free the worker before clearing `full`) must be caught — that ordering lets the owner wipe
a flag the worker has just raised again, and the chunk stays busy forever.  The GPU stress tests in tests/test_peer_gpu.py check the
same invariants on hardware; compute-sanitizer is closed on the GPU pool used for this project.
"""
from collections import deque


def explore(k_quanta: int, owner_order=("o1", "o2", "o3", "o4")):
    # state: (w_pc, o_pc, busy, full, mailbox, acc, produced, applied, owner_reading)
    start = (0, 0, 0, 0, 0, 0, 0, 0, 0)
    seen = {start}
    todo = deque([start])
    edges = {}
    races = []
    while todo:
        s = todo.popleft()
        w_pc, o_pc, busy, full, mailbox, acc, produced, applied, reading = s
        nxt = []
        # ---- worker: either produce one more gradient quantum into the accumulator (a training step) or advance the push
        if w_pc == 0 and produced < k_quanta:
            nxt.append((0, o_pc, busy, full, mailbox, acc + 1, produced + 1, applied, reading))
        if w_pc == 0 and acc > 0:                                   # w1: look at the local busy flag
            nxt.append(((1 if busy == 0 else 0), o_pc, busy, full, mailbox, acc, produced, applied, reading))
        elif w_pc == 1:                                             # w2: move the accumulator into the mailbox
            if reading:
                races.append(s)
            nxt.append((2, o_pc, busy, full, acc, 0, produced, applied, reading))
        elif w_pc == 2:                                             # w3
            nxt.append((3, o_pc, 1, full, mailbox, acc, produced, applied, reading))
        elif w_pc == 3:                                             # w4
            nxt.append((0, o_pc, busy, 1, mailbox, acc, produced, applied, reading))
        # ---- owner
        step = owner_order[o_pc]
        if step == "o1":                                            # poll: stay at o1 until the flag is seen
            nxt.append((w_pc, (1 if full else 0), busy, full, mailbox, acc, produced, applied, 1 if full else 0))
        elif step == "o2":
            nxt.append((w_pc, o_pc + 1, busy, full, 0, acc, produced, applied + mailbox, 0))
        elif step == "o3":
            nxt.append((w_pc, (o_pc + 1) % 4, busy, 0, mailbox, acc, produced, applied, reading))
        elif step == "o4":
            nxt.append((w_pc, (o_pc + 1) % 4, 0, full, mailbox, acc, produced, applied, reading))
        edges[s] = nxt
        for n in nxt:
            if n not in seen:
                seen.add(n)
                todo.append(n)
    return seen, edges, races


def _can_finish(seen, edges, k):
    """States from which 'all k quanta applied' is reachable (backward closure)."""
    goal = {s for s in seen if s[7] == k}
    rev = {}
    for s, ns in edges.items():
        for n in ns:
            rev.setdefault(n, []).append(s)
    ok, todo = set(goal), deque(goal)
    while todo:
        s = todo.popleft()
        for p in rev.get(s, []):
            if p not in ok:
                ok.add(p)
                todo.append(p)
    return ok


def test_mailbox_handshake_is_race_free_lossless_and_live():
    k = 4
    seen, edges, races = explore(k)
    assert not races
    for s in seen:
        w_pc, o_pc, busy, full, mailbox, acc, produced, applied, reading = s
        assert applied + acc + mailbox == produced, s                  # conservation: nothing lost, nothing applied twice
    assert any(s[7] == k for s in seen)
    assert _can_finish(seen, edges, k) == seen                         # no reachable state is a dead end
    assert len(seen) > 100                                             # the exploration really interleaved the two sides


def test_model_catches_the_wrong_owner_ordering():
    """Freeing the worker (busy = 0) before clearing `full` lets the owner erase a freshly raised flag: updates get stuck."""
    k = 3
    seen, edges, races = explore(k, owner_order=("o1", "o2", "o4", "o3"))
    stuck = seen - _can_finish(seen, edges, k)
    assert stuck or races
