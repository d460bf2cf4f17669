"""Exhaustive interleaving check of the accumulator-free asynchronous scheme (parallel/async_optim.py overlap mode,
csrc/comm.cu push_add_kernel delta form, csrc/train_ops.cu land_pull_kernel).

The worker keeps a SNAPSHOT of its replica instead of an accumulator.  Per element e, the three actors do

  training step (step stream)          params[e] += u                                    one quantum of -lr * grad
  push (communication stream)          p = params[e]; master[e] += p - snap[e]; snap[e] = p
  pull (communication stream)          pulled[e] = master[e]
  landing (step stream)                params[e] = pulled[e] + (params[e] - snap[e]); snap[e] = pulled[e]

Ordering the streams give: a push / pull runs CONCURRENTLY with training steps (element by element, in any interleaving); a landing
runs on the step stream, so never inside a training step, and it waits for the communication stream, so never concurrently with a
push or pull; pushes and pulls of one worker are serial.  Other workers add their own updates to `master` at arbitrary times.

Checked in every reachable state: nothing is lost or sent twice —
    master[e] - theta0[e] - (what the other workers added)[e]  +  (params[e] - snap[e])  ==  sum of this worker's updates to e
and after a final push the unsent part is zero; and the replica after a landing equals the pulled server value plus exactly the
worker's unsent updates.  A deliberately wrong push (advance the snapshot from a SECOND read of params) must be caught: an update
that lands between the two reads is then never sent.
"""
from collections import deque

E = 2            # elements of the vector
U = 3            # training steps (each adds one quantum to every element, element by element)


def explore(broken_push: bool = False, others: int = 2):
    # per element: params, snap, master, pulled, own (sum of own updates), foreign (sum added by other workers)
    # program counters: t = (step index, next element) of training; c = comm program counter over the script below
    # communication stream work in issue order; "gap" = nothing queued until the host, at some later step boundary, issues the next op
    script = ["push", "pull", "gap", "push", "gap"]
    zero = tuple([0] * E)
    # state = (params, snap, master, pulled, own, foreign, t_step, t_elem, c_idx, c_elem, c_tmp, landed, others_left)
    start = (zero, zero, zero, zero, zero, zero, 0, 0, 0, 0, None, 0, others)
    seen, todo, bad = {start}, deque([start]), []

    def check(s):
        params, snap, master, pulled, own, foreign = s[:6]
        for e in range(E):
            if master[e] - foreign[e] + (params[e] - snap[e]) != own[e]:
                bad.append(("lost or duplicated update", s))

    while todo:
        s = todo.popleft()
        params, snap, master, pulled, own, foreign, t_step, t_elem, c_idx, c_elem, c_tmp, landed, others_left = s
        check(s)
        nxt = []

        def put(**kw):
            d = dict(params=params, snap=snap, master=master, pulled=pulled, own=own, foreign=foreign, t_step=t_step, t_elem=t_elem,
                     c_idx=c_idx, c_elem=c_elem, c_tmp=c_tmp, landed=landed, others_left=others_left)
            d.update(kw)
            nxt.append((d["params"], d["snap"], d["master"], d["pulled"], d["own"], d["foreign"], d["t_step"], d["t_elem"], d["c_idx"],
                        d["c_elem"], d["c_tmp"], d["landed"], d["others_left"]))

        def setv(vec, e, v):
            return vec[:e] + (v,) + vec[e + 1:]

        comm_busy = script[c_idx] != "gap"
        # ---- training: one element of the current step
        if t_step < U:
            e = t_elem
            q = t_step + 1                                    # distinct quanta make double counting visible
            ne, ns = (e + 1, t_step) if e + 1 < E else (0, t_step + 1)
            put(params=setv(params, e, params[e] + q), own=setv(own, e, own[e] + q), t_elem=ne, t_step=ns)
        # ---- the host issues the next communication op at a step boundary
        if t_elem == 0 and not comm_busy and c_idx + 1 < len(script):
            put(c_idx=c_idx + 1)
        # ---- landing: on the step stream at a step boundary, only when the communication stream is idle and a pull has been fetched
        if t_elem == 0 and not comm_busy and landed == 0:
            np_, nsn = list(params), list(snap)
            for e in range(E):
                unsent = params[e] - snap[e]
                np_[e] = pulled[e] + unsent
                nsn[e] = pulled[e]
            put(params=tuple(np_), snap=tuple(nsn), landed=1)
        # ---- communication stream: one element of the current op
        if comm_busy:
            op, e = script[c_idx], c_elem
            done = e + 1 == E
            if op == "push":
                if broken_push and c_tmp is None:
                    put(c_tmp=params[e])                      # first read (used for the difference) ...
                else:
                    first = c_tmp if broken_push else params[e]
                    second = params[e]                        # ... the broken variant advances the snapshot from a second read
                    put(master=setv(master, e, master[e] + first - snap[e]), snap=setv(snap, e, second), c_tmp=None,
                        c_elem=0 if done else e + 1, c_idx=c_idx + 1 if done else c_idx)
            else:  # pull
                put(pulled=setv(pulled, e, master[e]), c_elem=0 if done else e + 1, c_idx=c_idx + 1 if done else c_idx)
        # ---- another worker adds to the master at any time
        if others_left > 0:
            for e in range(E):
                put(master=setv(master, e, master[e] + 10), foreign=setv(foreign, e, foreign[e] + 10), others_left=others_left - 1)
        for n in nxt:
            if n not in seen:
                seen.add(n)
                todo.append(n)
    return seen, bad


def test_delta_push_is_lossless_under_every_interleaving():
    seen, bad = explore()
    assert len(seen) > 1000                                   # the model really branches
    assert not bad, bad[:3]
    finals = [s for s in seen if s[6] == U and s[8] == 4 and s[11] == 1]
    assert finals
    for s in finals:
        params, snap, master, pulled, own, foreign = s[:6]
        for e in range(E):
            # replica = fetched server value + exactly what this worker has not sent since the snapshot
            assert params[e] - snap[e] == own[e] - (master[e] - foreign[e])


def test_final_push_leaves_nothing_unsent():
    """After training has stopped, one more push sends the rest: master - theta0 - foreign == every update this worker made."""
    seen, _ = explore()
    for s in seen:
        params, snap, master, pulled, own, foreign, t_step = s[:7]
        if t_step < U:
            continue
        for e in range(E):
            unsent = params[e] - snap[e]
            assert master[e] - foreign[e] + unsent == own[e]


def test_a_push_that_reads_the_replica_twice_is_caught():
    _, bad = explore(broken_push=True)
    assert bad, "advancing the snapshot from a second read must lose an update in some interleaving"
