"""Model check of the two-slot peer-to-peer halo protocol (mpas-model_b200/csrc/cuda/halo.h).

Every rank runs, per exchange k = 1, 2, ...:   pack(k): store message k into the neighbour's slot k & 1
                                               signal(k): publish sequence number k in the neighbour's window
                                               unpack(k): wait until the neighbour has published k, read own slot k & 1
in stream order.  The claim the implementation relies on: with every pair exchanging in BOTH directions in every
exchange, a pack can never overwrite a slot whose previous message has not been unpacked yet, under ANY interleaving of
the ranks -- and without that symmetry it can, which is why dycore_p2p_setup refuses plans that do not send to exactly the
ranks they receive from.  All interleavings of the per-rank programs are explored exhaustively."""
import sys

K = 6  # exchanges per rank; the slot index has period 2, so this covers every reuse pattern


def explore(sends):
    """sends[r] = True if rank r sends to the other rank (and the other rank waits for it).  Returns the first hazard found."""
    sys.setrecursionlimit(10000)
    prog = []
    for r in (0, 1):
        ops = []
        for k in range(1, K + 1):
            if sends[r]:
                ops += [("pack", k), ("signal", k)]
            if sends[1 - r]:
                ops += [("unpack", k)]
        prog.append(ops)
    seen = set()
    hazard = []

    def step(pc, flag, slots, unpacked):
        # flag[r]: last sequence number published TO rank r; slots[r][p]: message number in rank r's slot p;
        # unpacked[r]: number of messages rank r has consumed
        key = (pc, flag, slots, unpacked)
        if key in seen or hazard:
            return
        seen.add(key)
        for r in (0, 1):
            if pc[r] >= len(prog[r]):
                continue
            op, k = prog[r][pc[r]]
            o = 1 - r
            npc = tuple(pc[i] + (1 if i == r else 0) for i in (0, 1))
            if op == "pack":
                old = slots[o][k & 1]
                if old > unpacked[o]:  # the message in the slot has not been consumed by its owner
                    hazard.append((r, k, old, unpacked[o]))
                    return
                ns = tuple(tuple(k if (i == o and p == (k & 1)) else slots[i][p] for p in (0, 1)) for i in (0, 1))
                step(npc, flag, ns, unpacked)
            elif op == "signal":
                nf = tuple(k if i == o else flag[i] for i in (0, 1))
                step(npc, nf, slots, unpacked)
            else:  # unpack: blocked until the flag arrives
                if flag[r] < k:
                    continue
                assert slots[r][k & 1] == k, "unpack would read the wrong message"
                nu = tuple(k if i == r else unpacked[i] for i in (0, 1))
                step(npc, flag, slots, nu)

    step((0, 0), (0, 0), ((0, 0), (0, 0)), (0, 0))
    return hazard[0] if hazard else None, len(seen)


def test_two_slots_suffice_when_every_pair_exchanges_in_both_directions():
    hazard, nstates = explore((True, True))
    assert nstates > 50
    assert hazard is None, hazard


def test_one_directional_plans_can_overwrite_an_unread_slot():
    hazard, _ = explore((True, False))  # rank 0 only sends, rank 1 only receives: nothing holds rank 0 back
    assert hazard is not None
    r, k, old, consumed = hazard
    assert r == 0 and k == old + 2 and consumed < old
