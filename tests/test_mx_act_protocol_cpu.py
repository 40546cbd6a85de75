"""Exhaustive interleaving check of the shared-memory buffer protocol of csrc/mx_act_quantize.cu (the fused SwiGLU / dSwiGLU + MXFP8
quantise kernels, written without a GPU at hand).

Inside one CTA the threads are phase-locked by __syncthreads, so they are modelled as ONE sequential agent; what runs concurrently
with them is the TMA engine: a load lands in its buffer (and completes its mbarrier phase) at an arbitrary later moment, a store
keeps reading its staging buffer until an arbitrary later moment.  The model walks every interleaving of the agent's steps with
those asynchronous completions and checks, for every reachable state:

  * a load never lands in a buffer the threads are still using, and the threads never consume a buffer whose load has not landed
    or that holds another tile's data (tags);
  * the threads never overwrite a staging buffer a store is still reading;
  * an mbarrier never has two loads outstanding, and a parity wait can only be satisfied by the phase it was meant for;
  * the walk terminates (no deadlock) with every tile stored exactly once.

The step lists below are transcriptions of the kernels' per-tile loops (line references in the comments); if the kernels change,
change them here.  Test infrastructure only — nothing here ships."""
from collections import deque

import pytest

# agent steps of one tile iteration: (op, args...)
#   ("store_wait",)              tid 0: cp.async.bulk.wait_group.read 0   -> blocks until no store is reading
#   ("wait", bar)                mbarrier parity wait for this iteration's phase of `bar`
#   ("use", bufs_read, bufs_write)   generic-proxy accesses by the threads (reads need the tile's data, writes retag)
#   ("load", bar, bufs)          tid 0 issues the NEXT tile's TMA load(s) of `bufs` on `bar` (skipped for the last tile)
#   ("store", bufs)              tid 0 issues the TMA store(s) of this tile's staging buffers
FWD = [  # mx_swiglu_quantize_kernel, the `for (int t = ...)` body
    ("store_wait",),
    ("wait", "full_g"), ("wait", "full_u"),
    ("use", ("g", "u"), ("g",)),            # activation pass, in place into g
    # S1
    ("load", "full_u", ("u",)),
    ("use", ("g",), ("out",)),              # row / col passes
    # fence.proxy.async + S2
    ("store", ("out",)),
    ("load", "full_g", ("g",)),
]
BWD = [  # mx_dswiglu_quantize_kernel
    ("store_wait",),
    ("wait", "full_da"), ("wait", "full_gu"),
    ("use", ("da", "g", "u"), ("g", "u")),  # activation pass: dgate -> g, dup -> u
    # S1
    ("load", "full_da", ("da",)),
    ("use", ("g", "u"), ("out",)),          # both halves' row / col passes
    # S2
    ("store", ("out",)),
    ("load", "full_gu", ("g", "u")),
]
PROLOGUE = {"fwd": [("load", "full_g", ("g",)), ("load", "full_u", ("u",))],
            "bwd": [("load", "full_da", ("da",)), ("load", "full_gu", ("g", "u"))]}


def explore(steps, prologue, bufs, bars, ntiles):
    """BFS over (pc, tile, buffer tags, barrier phases, loads in flight, stores in flight).  Returns the number of states visited."""
    # buffer tag: ("raw", k) = tile k's TMA data landed, ("act", k) = overwritten by the threads for tile k, ("none",) = nothing valid
    init_tags = tuple(("none",) for _ in bufs)
    bi = {b: i for i, b in enumerate(bufs)}
    ri = {b: i for i, b in enumerate(bars)}
    # loads in flight: frozenset of (bar, bufs, tile); stores in flight: frozenset of (bufs, tile); phases: completed-phase count per barrier
    start = ("pro", 0, 0, init_tags, tuple(0 for _ in bars), frozenset(), frozenset(), 0)
    seen = {start}
    todo = deque([start])
    finished = False
    while todo:
        stage, pc, tile, tags, phases, loads, stores, stored = todo.popleft()
        succ = []
        # --- asynchronous completions
        for ld in loads:
            bar, lb, k = ld
            t2 = list(tags)
            for b in lb:
                t2[bi[b]] = ("raw", k)
            p2 = list(phases)
            p2[ri[bar]] += 1
            # hazard: the load may land at ANY time after issue, so the buffers must already be dead for the threads: checked at issue
            succ.append((stage, pc, tile, tuple(t2), tuple(p2), loads - {ld}, stores, stored))
        for st in stores:
            succ.append((stage, pc, tile, tags, phases, loads, stores - {st}, stored))
        # --- the agent
        seq = prologue if stage == "pro" else steps
        if stage == "pro" and pc == len(seq):
            succ.append(("run", 0, 0, tags, phases, loads, stores, stored))
        elif stage == "run" and tile == ntiles:
            if not loads and not stores:
                assert stored == ntiles
                finished = True
        elif stage == "run" and pc == len(seq):
            succ.append(("run", 0, tile + 1, tags, phases, loads, stores, stored))
        else:
            op = seq[pc]
            nxt = None
            if op[0] == "store_wait":
                if not stores:
                    nxt = (stage, pc + 1, tile, tags, phases, loads, stores, stored)
            elif op[0] == "wait":
                done = phases[ri[op[1]]]
                # parity wait of iteration `tile`: passes when the barrier has completed an odd/even number of phases matching tile + 1;
                # aliasing (two phases ahead) would also pass — assert it cannot happen
                assert done <= tile + 1, f"barrier {op[1]} ran ahead: {done} phases completed while waiting for phase {tile}"
                if done == tile + 1:
                    nxt = (stage, pc + 1, tile, tags, phases, loads, stores, stored)
            elif op[0] == "use":
                reads, writes = op[1], op[2]
                for ld in loads:
                    assert not (set(ld[1]) & (set(reads) | set(writes))), f"threads touch {ld[1]} while a load of tile {ld[2]} is in flight"
                for st in stores:
                    assert not (set(st[0]) & set(writes)), f"threads overwrite {st[0]} while the store of tile {st[1]} still reads it"
                for b in reads:
                    assert tags[bi[b]][0] in ("raw", "act") and tags[bi[b]][1] == tile, f"tile {tile} reads {b} holding {tags[bi[b]]}"
                t2 = list(tags)
                for b in writes:
                    t2[bi[b]] = ("act", tile)
                nxt = (stage, pc + 1, tile, tuple(t2), phases, loads, stores, stored)
            elif op[0] == "load":
                k = tile if stage == "pro" else tile + 1
                if stage == "pro" or k < ntiles:
                    assert not any(ld[0] == op[1] for ld in loads), f"two loads outstanding on {op[1]}"
                    assert phases[ri[op[1]]] == k, f"{op[1]}: load of tile {k} issued in phase {phases[ri[op[1]]]}"
                    nxt = (stage, pc + 1, tile, tags, phases, loads | {(op[1], op[2], k)}, stores, stored)
                else:
                    nxt = (stage, pc + 1, tile, tags, phases, loads, stores, stored)
            elif op[0] == "store":
                for b in op[1]:
                    assert tags[bi[b]] == ("act", tile), f"store of tile {tile} reads {b} holding {tags[bi[b]]}"
                nxt = (stage, pc + 1, tile, tags, phases, loads, stores | {(op[1], tile)}, stored + 1)
            if nxt is not None:
                succ.append(nxt)
        if not succ and not (stage == "run" and tile == ntiles):
            raise AssertionError(f"deadlock at {stage} pc={pc} tile={tile} loads={set(loads)} stores={set(stores)} phases={phases}")
        for s in succ:
            if s not in seen:
                seen.add(s)
                todo.append(s)
    assert finished, "the final state was never reached"
    return len(seen)


@pytest.mark.parametrize("ntiles", [1, 2, 3, 5])
def test_forward_kernel_buffer_protocol(ntiles):
    n = explore(FWD, PROLOGUE["fwd"], ("g", "u", "out"), ("full_g", "full_u"), ntiles)
    assert n > ntiles * len(FWD)


@pytest.mark.parametrize("ntiles", [1, 2, 3, 5])
def test_backward_kernel_buffer_protocol(ntiles):
    n = explore(BWD, PROLOGUE["bwd"], ("da", "g", "u", "out"), ("full_da", "full_gu"), ntiles)
    assert n > ntiles * len(BWD)


def test_the_checker_catches_a_broken_protocol():
    """Sanity of the model itself: issuing the next gate load BEFORE the block passes have read the activation must be flagged."""
    broken = [FWD[0], FWD[1], FWD[2], FWD[3], FWD[4], ("load", "full_g", ("g",)), FWD[5], FWD[6]]
    with pytest.raises(AssertionError):
        explore(broken, PROLOGUE["fwd"], ("g", "u", "out"), ("full_g", "full_u"), 3)
    no_wait = [s for s in FWD if s != ("store_wait",)]
    with pytest.raises(AssertionError):
        explore(no_wait, PROLOGUE["fwd"], ("g", "u", "out"), ("full_g", "full_u"), 3)
