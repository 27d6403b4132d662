"""Exhaustive interleaving check of the two peer-memory protocols of csrc/p2p.cu / p2p.cuh, as MODELS (the flag logic, not the CUDA
code): every schedule of the ranks' atomic steps is explored and, in every reachable state,

  halo exchange   (k_halo_push / k_halo_wait_copy, and the fused variants that keep the same flag logic: push from the producing
                  kernel's epilogue, wait inside the consuming SpMV) -- a consumer of exchange e reads exactly what its neighbour
                  pushed for exchange e (no landing buffer is overwritten before it was read), and no schedule deadlocks;
  all-reduce      (p2p_allreduce_block: mailboxes double-buffered by the parity of the sequence number) -- every rank adds exactly
                  the partial sums of epoch e, and no schedule deadlocks.

compute-sanitizer's racecheck is closed on the GPU pool; this is the protocol-level half of that evidence, the
bitwise-equal histories across runs (profiles/r2_summary.md) the implementation-level half.  The negative tests remove one wait from
each protocol and show that the checker then finds the overwrite."""
import pytest


# ---------------------------------------------------------------------------------------------------------------------------------
# halo exchange between neighbours.  Per rank r and exchange e = 1..E the stream executes, in order:
#   push(e):  P1  for every neighbour q: consumed[q][r] := e - 1      ("what you pushed last time has been moved out": stream order
#                                                                       guarantees my consumer of e - 1 has finished)
#             P2  wait until consumed[r][q] >= e - 1 for every q       (they moved out what I pushed last time)
#             P3  for every q: landing[q][r] := (r, e)                 (the interface entries of exchange e)
#             P4  for every q: ready[q][r] := e
#   take(e):  T1  wait until ready[r][q] >= e for every q
#             T2  read landing[r][q] for every q: must be (q, e)
# flags[x][y] = the word in rank x's memory that rank y writes.
# ---------------------------------------------------------------------------------------------------------------------------------
def halo_program(E, skip_wait_consumed=False):
    prog = []
    for e in range(1, E + 1):
        prog += [("P1", e)] + ([] if skip_wait_consumed else [("P2", e)]) + [("P3", e), ("P4", e), ("T1", e), ("T2", e)]
    return prog


def explore_halo(neigh, E, skip_wait_consumed=False):
    """neigh: dict rank -> tuple of neighbours (symmetric).  Returns (states explored, violation or None)."""
    ranks = sorted(neigh)
    prog = halo_program(E, skip_wait_consumed)
    pairs = [(x, y) for x in ranks for y in neigh[x]]
    idx = {p: i for i, p in enumerate(pairs)}

    def initial():
        z = tuple(0 for _ in pairs)
        return (tuple(0 for _ in ranks), z, z, tuple((None, 0) for _ in pairs))   # pcs, consumed, ready, landing

    seen, stack = set(), [initial()]
    while stack:
        st = stack.pop()
        if st in seen:
            continue
        seen.add(st)
        pcs, consumed, ready, landing = st
        moved = False
        for ri, r in enumerate(ranks):
            if pcs[ri] == len(prog):
                continue
            op, e = prog[pcs[ri]]
            c, rd, ld = list(consumed), list(ready), list(landing)
            if op == "P1":
                for q in neigh[r]:
                    c[idx[(q, r)]] = e - 1
            elif op == "P2":
                if any(consumed[idx[(r, q)]] < e - 1 for q in neigh[r]):
                    continue                                            # blocked
            elif op == "P3":
                for q in neigh[r]:
                    ld[idx[(q, r)]] = (r, e)
            elif op == "P4":
                for q in neigh[r]:
                    rd[idx[(q, r)]] = e
            elif op == "T1":
                if any(ready[idx[(r, q)]] < e for q in neigh[r]):
                    continue
            elif op == "T2":
                for q in neigh[r]:
                    if landing[idx[(r, q)]] != (q, e):
                        return len(seen), ("rank %d reads %r from %d in exchange %d" % (r, landing[idx[(r, q)]], q, e))
            moved = True
            npcs = list(pcs)
            npcs[ri] += 1
            stack.append((tuple(npcs), tuple(c), tuple(rd), tuple(ld)))
        if not moved and any(pcs[i] < len(prog) for i in range(len(ranks))):
            return len(seen), "deadlock at %r" % (pcs,)
    return len(seen), None


@pytest.mark.parametrize("neigh,E", [({0: (1,), 1: (0,)}, 4), ({0: (1,), 1: (0, 2), 2: (1,)}, 3), ({0: (1, 2), 1: (0, 2), 2: (0, 1)}, 2),
                                     ({0: (1,), 1: (0, 2), 2: (1, 3), 3: (2,)}, 3)],
                         ids=["2 ranks x 4 exchanges", "3 slabs x 3 exchanges", "3 ranks all-to-all x 2 exchanges", "4 slabs x 3 exchanges"])
def test_halo_exchange_protocol_is_race_free_and_deadlock_free(neigh, E):
    n, bad = explore_halo(neigh, E)
    assert bad is None, bad
    assert n > 100                                                     # the schedules were really explored


def test_halo_checker_finds_the_overwrite_when_the_consumed_wait_is_removed():
    n, bad = explore_halo({0: (1,), 1: (0,)}, 3, skip_wait_consumed=True)
    assert bad is not None and "reads" in bad


# ---------------------------------------------------------------------------------------------------------------------------------
# all-reduce through mailboxes (p2p_allreduce_block).  Per rank r and epoch e = 1..E, with par = e & 1:
#   A1  for every rank q (itself included): box[q][par][r] := (r, e)     A2  for every q: flag[q][r] := e
#   A3  wait until flag[r][q] >= e for every q                           A4  read box[r][par][q] for every q: must be (q, e)
# ---------------------------------------------------------------------------------------------------------------------------------
def explore_allreduce(R, E, single_buffer=False):
    prog = [(op, e) for e in range(1, E + 1) for op in ("A1", "A2", "A3", "A4")]
    ranks = range(R)
    nbuf = 1 if single_buffer else 2
    box0 = tuple(None for _ in range(R * nbuf * R))                     # [owner][par][writer]
    flag0 = tuple(0 for _ in range(R * R))                              # [owner][writer]
    bi = lambda owner, par, w: (owner * nbuf + (par % nbuf)) * R + w    # noqa: E731
    seen, stack = set(), [(tuple(0 for _ in ranks), box0, flag0)]
    while stack:
        st = stack.pop()
        if st in seen:
            continue
        seen.add(st)
        pcs, box, flag = st
        moved = False
        for r in ranks:
            if pcs[r] == len(prog):
                continue
            op, e = prog[pcs[r]]
            b, f = list(box), list(flag)
            if op == "A1":
                for q in ranks:
                    b[bi(q, e & 1, r)] = (r, e)
            elif op == "A2":
                for q in ranks:
                    f[q * R + r] = e
            elif op == "A3":
                if any(flag[r * R + q] < e for q in ranks):
                    continue
            elif op == "A4":
                for q in ranks:
                    if box[bi(r, e & 1, q)] != (q, e):
                        return len(seen), "rank %d adds %r for rank %d in epoch %d" % (r, box[bi(r, e & 1, q)], q, e)
            moved = True
            npcs = list(pcs)
            npcs[r] += 1
            stack.append((tuple(npcs), tuple(b), tuple(f)))
        if not moved and any(pcs[r] < len(prog) for r in ranks):
            return len(seen), "deadlock at %r" % (pcs,)
    return len(seen), None


@pytest.mark.parametrize("R,E", [(2, 5), (3, 3)])
def test_mailbox_allreduce_is_race_free_and_deadlock_free(R, E):
    n, bad = explore_allreduce(R, E)
    assert bad is None, bad
    assert n > 100


def test_allreduce_checker_finds_the_overwrite_with_a_single_buffer():
    """Without the parity double buffer a fast rank's epoch e + 1 partial lands on a slot a slow rank has not added yet."""
    n, bad = explore_allreduce(2, 3, single_buffer=True)
    assert bad is not None and "adds" in bad
